/*
 * dpk_oracle.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement, in plain scalar
 * C, of the arithmetic of the reference's post-backward update:
 *
 *   TrainLoop.optimize            /root/reference/utils/trainer.py:237-244
 *     grad_clip -> clip_grad_norm_  :246-255  (torch nn/utils/clip_grad.py:92-108,165-174)
 *     _log_grad_norm                :265-271
 *     opt.step() = torch AdamW      :99,:242  (torch optim/adam.py:396-552 CPU, :683-800 CUDA)
 *     update_ema per rate           :243-244, :360-370
 *
 * The arithmetic itself lives in third-party torch (2.11.0+cu128 on this image;
 * the reference pins no version).  torch's CPU and CUDA kernels associate a few
 * operations differently, so the restatement has two flavours:
 *
 *   ORACLE_CUDA  what torch's CUDA foreach kernels compute (nvcc contracts
 *                a + s*(b op c) into one FMA): the flavour libdpk must match.
 *   ORACLE_CPU   what torch's CPU vectorised kernels compute: the flavour that
 *                is pinned, in this container, against the reference TrainLoop
 *                itself (tests/golden, made by oracle/make_golden.py):
 *                exp_avg and exp_avg_sq bit for bit; params and EMA to a few
 *                ulp, because torch's CPU sqrt is MKL's vsSqrt, which is off
 *                by one ulp on ~20% of inputs (sqrtf here is correctly
 *                rounded, as is CUDA's sqrt.rn).
 *
 * Nothing outside tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may link or call this file.  Compile with
 * -ffp-contract=off: every fmaf() below is deliberate, nothing else may fuse.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define ORACLE_CUDA 0
#define ORACLE_CPU 1

#define ORACLE_F32 0
#define ORACLE_BF16 1

/* ---- bf16 <-> f32 (round to nearest even, NaN kept quiet) ----------------- */

static float bf16_bits_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

static uint16_t f32_to_bf16_bits(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u); /* NaN */
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

/* value after a store to the state dtype */
static float rnd(int dtype, float x) {
  return dtype == ORACLE_BF16 ? bf16_bits_to_f32(f32_to_bf16_bits(x)) : x;
}

static float load(int dtype, const void* a, int64_t i) {
  return dtype == ORACLE_BF16 ? bf16_bits_to_f32(((const uint16_t*)a)[i]) : ((const float*)a)[i];
}

static void store(int dtype, void* a, int64_t i, float x) {
  if (dtype == ORACLE_BF16) ((uint16_t*)a)[i] = f32_to_bf16_bits(x);
  else ((float*)a)[i] = x;
}

uint16_t oracle_f32_to_bf16(float f) { return f32_to_bf16_bits(f); }
float oracle_bf16_to_f32(uint16_t h) { return bf16_bits_to_f32(h); }

/* ---- clip_grad_norm_ ------------------------------------------------------
 * clip_grad.py:96      norms  = _foreach_norm(grads)            one per tensor, grad dtype
 * clip_grad.py:106-108 total  = vector_norm(stack(norms))       grad dtype
 * clip_grad.py:165     coef   = max_norm / (total + 1e-6)       Tensor.__rdiv__ = reciprocal()*max_norm
 * clip_grad.py:169     coef   = clamp(coef, max=1.0)
 * Sums of squares are taken in fp64 here (the exact value); torch accumulates
 * in fp32 with a kernel-specific order, which is what the 1e-5 tolerance on
 * the norm absorbs.  Gradients are first multiplied by inv_scale (upstream
 * loss-scale path: grad.mul_(1/2^lg), stored in the grad dtype).
 * out[0]=total, out[1]=coef, out[2]=nonfinite flag.  seg_norms may be NULL. */
void oracle_grad_norm(int dtype, const void* g, int n_seg, const int64_t* seg_off,
                      const int64_t* seg_len, float inv_scale, float max_norm,
                      double* seg_norms, double* out) {
  double tot2 = 0.0;
  int bad = 0;
  for (int s = 0; s < n_seg; ++s) {
    double acc = 0.0;
    for (int64_t i = 0; i < seg_len[s]; ++i) {
      const float y = rnd(dtype, load(dtype, g, seg_off[s] + i) * inv_scale);
      if (!isfinite(y)) bad = 1;
      acc += (double)y * (double)y;
    }
    const float norm = rnd(dtype, sqrtf((float)acc));
    if (seg_norms) seg_norms[s] = (double)norm;
    tot2 += (double)norm * (double)norm;
  }
  const float total = rnd(dtype, sqrtf((float)tot2));
  float coef = 1.0f;
  if (max_norm > 0.0f) {
    const float denom = rnd(dtype, total + 1e-6f);
    const float rcp = rnd(dtype, 1.0f / denom);
    coef = rnd(dtype, rcp * max_norm);
    if (coef > 1.0f) coef = 1.0f; /* NaN stays NaN, as torch.clamp */
  }
  out[0] = (double)total;
  out[1] = (double)coef;
  out[2] = bad ? 1.0 : 0.0;
}

/* ---- one optimize() step over `n` contiguous elements ----------------------
 * g is taken as rnd(rnd(g*inv_scale)*coef) (clip_grad.py:174 _foreach_mul_).
 * Scalars arrive as the Python doubles the reference computes
 * (adam.py:770-781 / :529-535; trainer.py:257-263,370) and are narrowed to fp32
 * here, like torch narrows a double handed to a float kernel.
 * grad_mode: 0 leave g, 1 write the scaled/clipped g back, 2 zero g.
 * Returns sum over elements of rnd(gc*gc) accumulated in fp64 (what
 * _log_grad_norm's (p.grad**2).sum() adds up, before its own rounding). */
double oracle_adamw_ema_step(int flavour, int dtype, int64_t n, void* p, void* g, void* m, void* v,
                             int n_ema, void* const* ema, const double* ema_rates, double lr,
                             double weight_decay, double beta1, double beta2, double eps,
                             int64_t adam_step, double inv_scale, double coef, int grad_mode) {
  const double step = (double)adam_step;
  const double bc1 = 1.0 - pow(beta1, step);
  const double bc2 = 1.0 - pow(beta2, step);
  const float inv_scale_f = (float)inv_scale, coef_f = (float)coef;
  const float decay = (float)(1.0 - lr * weight_decay);
  const float w1 = (float)(1.0 - beta1);
  const float beta2_f = (float)beta2, omb2 = (float)(1.0 - beta2);
  const float bc2_sqrt = (float)pow(bc2, 0.5);
  const float eps_f = (float)eps;
  const float neg_step = (float)((lr / bc1) * -1.0);
  double sumsq = 0.0;

  for (int64_t i = 0; i < n; ++i) {
    float gi = rnd(dtype, load(dtype, g, i) * inv_scale_f);
    gi = rnd(dtype, gi * coef_f);
    sumsq += (double)rnd(dtype, gi * gi);

    float pi = load(dtype, p, i);
    if (weight_decay != 0.0) pi = rnd(dtype, pi * decay); /* adam.py:695 / :417 */

    /* exp_avg.lerp_(grad, 1-beta1): weight < 0.5 -> m + w*(g-m), fused on both devices */
    float mi = load(dtype, m, i);
    mi = rnd(dtype, fmaf(w1, gi - mi, mi));

    /* exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2) */
    float vi = rnd(dtype, load(dtype, v, i) * beta2_f);
    if (flavour == ORACLE_CUDA) vi = rnd(dtype, fmaf(omb2, gi * gi, vi)); /* v + s*(g*g), contracted */
    else vi = rnd(dtype, fmaf(omb2 * gi, gi, vi));                         /* fmadd(s*g, g, v), vectorised */

    /* denom = (sqrt(v) / bc2_sqrt) + eps : three stores */
    float d = rnd(dtype, sqrtf(vi));
    d = rnd(dtype, d / bc2_sqrt);
    d = rnd(dtype, d + eps_f);

    /* param.addcdiv_(exp_avg, denom, value=-step_size) */
    if (flavour == ORACLE_CUDA) pi = rnd(dtype, fmaf(neg_step, mi / d, pi)); /* p + s*(m/d), contracted */
    else pi = rnd(dtype, pi + (neg_step * mi) / d);                           /* p + s*m/d */

    store(dtype, p, i, pi);
    store(dtype, m, i, mi);
    store(dtype, v, i, vi);
    if (grad_mode == 1) store(dtype, g, i, gi);
    else if (grad_mode == 2) store(dtype, g, i, 0.0f);

    /* update_ema: trg.mul_(rate).add_(src, alpha=1-rate) -- add_ fuses alpha*src+trg */
    for (int k = 0; k < n_ema; ++k) {
      const float rate = (float)ema_rates[k];
      float alpha = (float)(1.0 - ema_rates[k]);
      /* torch's CPU add kernel narrows `alpha` to the tensor dtype before use (bf16: 1-0.9999 becomes
       * 1.00136e-4); its CUDA kernel keeps it in fp32.  Pinned by the bf16 golden fixture recorded from
       * the reference TrainLoop (tests/golden/ref_bf16_clip_ema2_ws1.npz). */
      if (flavour == ORACLE_CPU) alpha = rnd(dtype, alpha);
      float e = rnd(dtype, load(dtype, ema[k], i) * rate);
      e = rnd(dtype, fmaf(alpha, pi, e));
      store(dtype, ema[k], i, e);
    }
  }
  return sumsq;
}

/* _log_grad_norm (trainer.py:265-271) over per-tensor sums of squares: each
 * (p.grad**2).sum() comes back in the grad dtype, .item() widens to double. */
double oracle_logged_norm(int dtype, const double* seg_sumsq, int n_seg) {
  double acc = 0.0;
  for (int s = 0; s < n_seg; ++s) acc += (double)rnd(dtype, (float)seg_sumsq[s]);
  return sqrt(acc);
}

/* stand-alone update_ema on n elements */
void oracle_ema_update(int dtype, int64_t n, void* trg, const void* src, double rate) {
  const float r = (float)rate, a = (float)(1.0 - rate);
  for (int64_t i = 0; i < n; ++i) {
    float e = rnd(dtype, load(dtype, trg, i) * r);
    e = rnd(dtype, fmaf(a, load(dtype, src, i), e));
    store(dtype, trg, i, e);
  }
}
