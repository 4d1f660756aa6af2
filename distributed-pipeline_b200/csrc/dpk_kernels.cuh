// Device kernels of libdpk (sm_100a).  Everything here is HBM-bound streaming
// work: 128-bit coalesced accesses, no data reuse, no tensor cores.
//
// Work decomposition: the flat buffers are cut on the host into *tiles* that
// never straddle a parameter tensor (dpk.cu: build_tiles).  Both passes walk
// the same tile table, so each tile yields one partial sum of squares that
// belongs to exactly one tensor -- the per-tensor norms torch's
// clip_grad_norm_ is defined over (nn/utils/clip_grad.py:92-108) fall out of a
// deterministic second-level reduction, with no float atomics anywhere.
//
// Arithmetic mirrors the op sequence torch's CUDA foreach AdamW runs
// (optim/adam.py:683-800) including where it rounds: every `__f*_rn` below is
// one torch kernel's store.  The file is compiled with -fmad=false; the FMAs
// torch's kernels get from nvcc's contraction are written out explicitly.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/dpk.h"

namespace dpk {

constexpr int kThreads = 256;

struct Tile {
  long long start;  // element offset in the flat buffers (multiple of 16 bytes)
  int len;          // elements in this tile (<= tile_elems)
  int seg;          // parameter tensor it belongs to; kTileFrozen set if that tensor takes no optimizer step
};
constexpr int kTileFrozen = 1 << 30;

struct alignas(8) TilePartial {
  float sumsq;
  unsigned bad;  // 1 if a non-finite gradient element was seen
};

// Host-computed scalars of one AdamW step, each already rounded to fp32 the
// way torch rounds a Python double handed to a CUDA kernel.
struct StepScalars {
  float inv_scale;   // 1/2^lg_loss_scale, 1 when unused
  float decay;       // 1 - lr*wd
  int has_decay;     // torch skips the multiply when wd == 0
  float w1;          // 1 - beta1
  float beta2;
  float omb2;        // 1 - beta2
  float bc2_sqrt;    // sqrt(1 - beta2^t)
  float bc2_rcp;     // fp32(1 / bc2_sqrt), for the exact scalar division below
  float eps;
  float neg_step;    // -(lr / (1 - beta1^t))
  float rate[DPK_MAX_EMA];
  float alpha[DPK_MAX_EMA];  // fp32(1 - rate)
};

// ---------------------------------------------------------------- helpers --

// x / c for a divisor that is the same for every element, correctly rounded
// (identical to div.rn.f32) at a third of its cost: with rc = RN(1/c),
// q = RN(x*rc), r = x - q*c (exact in one FMA), RN(q + r*rc) is the correctly
// rounded quotient (Markstein) as long as neither it nor the remainder is
// subnormal -- true here, where x is a freshly stored square root (0 or at
// least 2^-75) and c lies in [0.03, 1].  inf/NaN fall through unchanged.
// dpk_selftest_arith checks it against div.rn for the square root of every
// fp32 bit pattern.
__device__ __forceinline__ float div_by_scalar(float x, float c, float rc) {
  const float q = __fmul_rn(x, rc);
  const float r = __fmaf_rn(-q, c, x);
  const float q1 = __fmaf_rn(r, rc, q);
  return (fabsf(q) < __int_as_float(0x7f800000)) ? q1 : q;
}

__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

template <typename T> struct ElemTraits;
template <> struct ElemTraits<float> {
  static constexpr int VE = 4;  // elements per 16-byte vector
  static __device__ __forceinline__ void unpack(const uint4& r, float* x) {
    x[0] = __uint_as_float(r.x); x[1] = __uint_as_float(r.y);
    x[2] = __uint_as_float(r.z); x[3] = __uint_as_float(r.w);
  }
  static __device__ __forceinline__ uint4 pack(const float* x) {
    return make_uint4(__float_as_uint(x[0]), __float_as_uint(x[1]),
                      __float_as_uint(x[2]), __float_as_uint(x[3]));
  }
  // value as stored in the state dtype
  static __device__ __forceinline__ float rnd(float x) { return x; }
  // the same for a whole vector, in place
  static __device__ __forceinline__ void rndv(float*) {}
  // sqrt whose result is about to be stored in the state dtype
  static __device__ __forceinline__ float sqrt_for_store(float x) { return __fsqrt_rn(x); }
  // a 16-byte vector is processed in CH chunks of four elements
  static constexpr int CH = 1;
  static __device__ __forceinline__ void unpack4(const uint4& r, int, float* x) { unpack(r, x); }
  static __device__ __forceinline__ void rnd4(float*) {}
  // round four values to the state dtype, leave them (rounded) in x and put them
  // into chunk c of the output vector
  static __device__ __forceinline__ void rnd4_store(float* x, uint4& r, int) { r = pack(x); }
  // the same when the rounded values are not needed afterwards
  static __device__ __forceinline__ void rnd4_store_only(const float* x, uint4& r, int) { r = pack(x); }
  // acc += rnd(t[j]) for the first n of four values, in index order
  static __device__ __forceinline__ void rnd4_accumulate(const float* t, int n, float& acc) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (j < n) acc = __fadd_rn(acc, t[j]);
  }
  // lanes [rem, VE) of a ragged last vector keep their old bytes
  static __device__ __forceinline__ uint4 merge(const uint4& nw, const uint4& old, int rem) {
    return make_uint4(rem > 0 ? nw.x : old.x, rem > 1 ? nw.y : old.y, rem > 2 ? nw.z : old.z, rem > 3 ? nw.w : old.w);
  }
};
template <> struct ElemTraits<__nv_bfloat16> {
  static constexpr int VE = 8;
  // bf16 -> fp32 of one half of a packed pair: x + (-0.0) is exact for every x,
  // and the mixed-precision add (PTX add.f32.bf16, SASS FHADD.BF16 Rd, Rs.H0|H1, c)
  // reads the half straight out of the packed register on the full-rate FMA
  // pipe.  The shift / mask it replaces issue on the half-rate ALU pipe, which
  // is what bounded the bf16 kernels (ncu round 1: ALU 46-55 %, DRAM 70 %).
  static __device__ __forceinline__ float add_lo(unsigned w, float c) {
    float r;
    const unsigned short h = (unsigned short)(w & 0xffffu);
    asm("add.f32.bf16 %0, %1, %2;" : "=f"(r) : "h"(h), "f"(c));
    return r;
  }
  static __device__ __forceinline__ float add_hi(unsigned w, float c) {
    float r;
    const unsigned short h = (unsigned short)(w >> 16);
    asm("add.f32.bf16 %0, %1, %2;" : "=f"(r) : "h"(h), "f"(c));
    return r;
  }
  static __device__ __forceinline__ float cvt_lo(unsigned w) { return add_lo(w, -0.0f); }
  static __device__ __forceinline__ float cvt_hi(unsigned w) { return add_hi(w, -0.0f); }
  static __device__ __forceinline__ void unpack(const uint4& r, float* x) {
    x[0] = cvt_lo(r.x); x[1] = cvt_hi(r.x); x[2] = cvt_lo(r.y); x[3] = cvt_hi(r.y);
    x[4] = cvt_lo(r.z); x[5] = cvt_hi(r.z); x[6] = cvt_lo(r.w); x[7] = cvt_hi(r.w);
  }
  // Round-to-nearest-even of two fp32 values to one packed bf16 pair (a low, b
  // high): SASS F2FP.BF16.F32.PACK_AB.  The scalar conversion would be F2F on the
  // quarter-rate XU pipe.
  static __device__ __forceinline__ unsigned pack2(float a, float b) {
    unsigned w;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(b), "f"(a));
    return w;
  }
  static __device__ __forceinline__ uint4 pack(const float* x) {  // rounds (exact on bf16-representable input)
    return make_uint4(pack2(x[0], x[1]), pack2(x[2], x[3]), pack2(x[4], x[5]), pack2(x[6], x[7]));
  }
  // One value rounded to bf16 and kept as fp32: the same conversion with zero as
  // the low half yields the fp32 bit pattern of the rounded value directly -- one
  // instruction, nothing to unpack.
  static __device__ __forceinline__ float rnd(float x) {
    unsigned w;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(x), "f"(0.0f));
    return __uint_as_float(w);
  }
  static __device__ __forceinline__ void rndv(float* x) {
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = rnd(x[i]);
  }
  // The square root of a bf16 value is never closer than 2^-18 (relative) to the
  // midpoint of two bf16 values (all 32 639 positive inputs enumerated), so the
  // 2^-23-accurate sqrt.approx rounds to the same bf16 as the IEEE one.
  static __device__ __forceinline__ float sqrt_for_store(float x) { return sqrt_approx(x); }
  static constexpr int CH = 2;
  static __device__ __forceinline__ void unpack4(const uint4& r, int c, float* x) {
    const unsigned w0 = c ? r.z : r.x, w1 = c ? r.w : r.y;
    x[0] = cvt_lo(w0); x[1] = cvt_hi(w0); x[2] = cvt_lo(w1); x[3] = cvt_hi(w1);
  }
  static __device__ __forceinline__ void rnd4(float* x) {
#pragma unroll
    for (int j = 0; j < 4; ++j) x[j] = rnd(x[j]);
  }
  static __device__ __forceinline__ void rnd4_store(float* x, uint4& r, int c) {
    const unsigned w0 = pack2(x[0], x[1]), w1 = pack2(x[2], x[3]);
    x[0] = cvt_lo(w0); x[1] = cvt_hi(w0); x[2] = cvt_lo(w1); x[3] = cvt_hi(w1);
    if (c) { r.z = w0; r.w = w1; } else { r.x = w0; r.y = w1; }
  }
  static __device__ __forceinline__ void rnd4_store_only(const float* x, uint4& r, int c) {
    const unsigned w0 = pack2(x[0], x[1]), w1 = pack2(x[2], x[3]);
    if (c) { r.z = w0; r.w = w1; } else { r.x = w0; r.y = w1; }
  }
  static __device__ __forceinline__ void rnd4_accumulate(const float* t, int n, float& acc) {
    const unsigned w0 = pack2(t[0], t[1]), w1 = pack2(t[2], t[3]);
    if (0 < n) acc = add_lo(w0, acc);
    if (1 < n) acc = add_hi(w0, acc);
    if (2 < n) acc = add_lo(w1, acc);
    if (3 < n) acc = add_hi(w1, acc);
  }
  static __device__ __forceinline__ unsigned merge_word(unsigned nw, unsigned old, int rem, int first) {
    const unsigned mask = (rem > first ? 0x0000ffffu : 0u) | (rem > first + 1 ? 0xffff0000u : 0u);
    return (nw & mask) | (old & ~mask);
  }
  static __device__ __forceinline__ uint4 merge(const uint4& nw, const uint4& old, int rem) {
    return make_uint4(merge_word(nw.x, old.x, rem, 0), merge_word(nw.y, old.y, rem, 2),
                      merge_word(nw.z, old.z, rem, 4), merge_word(nw.w, old.w, rem, 6));
  }
};

// streaming 128-bit accesses: every byte is touched once per pass
__device__ __forceinline__ uint4 ld_stream(const uint4* p) { return __ldcs(p); }
__device__ __forceinline__ uint4 ld_readonly(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(uint4* p, const uint4& v) { __stcs(p, v); }

__device__ __forceinline__ bool nonfinite_bits(float x) {
  return (__float_as_uint(x) & 0x7f800000u) == 0x7f800000u;
}

// Block-wide sum in a fixed order (lane tree, then warp 0 over the warp
// totals): same inputs -> same bits, run to run.
__device__ __forceinline__ float block_sum(float v, float* smem /*[kThreads/32]*/) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = __fadd_rn(v, __shfl_down_sync(0xffffffffu, v, o));
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  float r = 0.f;
  if (warp == 0) {
    r = (lane < kThreads / 32) ? smem[lane] : 0.f;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) r = __fadd_rn(r, __shfl_down_sync(0xffffffffu, r, o));
  }
  __syncthreads();
  return r;  // valid in thread 0
}

__device__ __forceinline__ double block_sum_f64(double v, double* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  double r = 0.0;
  if (warp == 0) {
    r = (lane < kThreads / 32) ? smem[lane] : 0.0;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) r += __shfl_down_sync(0xffffffffu, r, o);
  }
  __syncthreads();
  return r;
}

// ---- waiting for gradients that other GPUs are still delivering ---------------
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* addr) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ uint4 ld_sys(const uint4* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}

// The fused W-rank passes (dpk_nvls.cu) read a gradient vector only after the
// arrival counter of its unit (2^shift vectors) shows that every owner rank has
// written this step's average into this rank's buffer.
struct FlagWait {
  const unsigned* flags;   // per-unit arrival counters on this rank
  unsigned target;         // value a counter has once all ranks delivered this step
  int shift;               // log2(vectors per unit)
  unsigned long long timeout_ns;
  double* results;         // result block: DPK_R_COMM_TIMEOUT is raised on give-up
};

// False if the wait was abandoned (a peer is gone).  The acquire orders the
// caller's following gradient load behind the owners' stores.
__device__ __forceinline__ bool flag_wait(const FlagWait& fw, long long vidx) {
  const unsigned* flag = fw.flags + (vidx >> fw.shift);
  if ((int)(ld_acquire_sys_u32(flag) - fw.target) >= 0) return true;
  const unsigned long long t0 = global_timer_ns();
  unsigned spins = 0;
  while ((int)(ld_acquire_sys_u32(flag) - fw.target) < 0) {
    __nanosleep(64);
    if ((++spins & 0x3ffu) == 0) {
      const bool late = global_timer_ns() - t0 > fw.timeout_ns;
      if (late) fw.results[DPK_R_COMM_TIMEOUT] = 1.0;
      if (late || *reinterpret_cast<const volatile double*>(fw.results + DPK_R_COMM_TIMEOUT) != 0.0) return false;
    }
  }
  return true;
}

// One tile's sum of squares and non-finite indicator, per thread (the caller
// reduces across the block).  Shared by the handle path and dpk_grad_norm_static
// so that both add the same numbers in the same order.  Full vectors carry no
// per-lane masking; the ragged last vector of a tensor belongs to the thread
// that would own it anyway and is added last.  `tail_by_element`: read the ragged
// tail element by element (the caller's array may end exactly there) instead of
// as one masked 16-byte vector (flat buffers are padded).
// SRC: where the vectors come from -- kSrcGlobal: read-only global loads;
// kSrcStaged: the tile already sits in shared memory (bulk-copied there);
// kSrcFlagged: global memory that other GPUs are writing during this kernel, each
// vector loaded (coherently) only once its unit has arrived (fw, vbase = index of
// the tile's first vector in the flat buffer).
constexpr int kSrcGlobal = 0, kSrcStaged = 1, kSrcFlagged = 2;

template <typename T, int U, int SRC = kSrcGlobal>
__device__ __forceinline__ void tile_sumsq(const T* __restrict__ g_tile, int len, float inv_scale,
                                           bool tail_by_element, float& acc, float& nonfin,
                                           const FlagWait* fw = nullptr, long long vbase = 0) {
  using Tr = ElemTraits<T>;
  constexpr int VE = Tr::VE;
  const uint4* gv = reinterpret_cast<const uint4*>(g_tile);
  auto ld = [&](const uint4* p) -> uint4 {
    if (SRC == kSrcStaged) return *p;
    if (SRC == kSrcFlagged) {
      flag_wait(*fw, vbase + (p - gv));   // on give-up the flag is raised and the data no longer matter
      return ld_sys(p);
    }
    return ld_readonly(p);
  };
  const int nfull = len / VE;
  for (int v0 = threadIdx.x; v0 < nfull; v0 += kThreads * U) {
    uint4 raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int v = v0 + u * kThreads;
      raw[u] = (v < nfull) ? ld(gv + v) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (v0 + u * kThreads < nfull) {
        if constexpr (sizeof(T) == 2) {
          if (inv_scale == 1.0f) {
            // bf16, no loss scale: both accumulations read the packed halves directly
            // (FHFMA.BF16 acc, w.H0|H1, w.H0|H1, acc: the same a*a+acc in fp32, one
            // rounding, as the FFMA on unpacked values) -- nothing to unpack
            const unsigned w[4] = {raw[u].x, raw[u].y, raw[u].z, raw[u].w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const unsigned short lo = (unsigned short)(w[i] & 0xffffu), hi = (unsigned short)(w[i] >> 16);
              asm("fma.rn.f32.bf16 %0, %1, %1, %0;" : "+f"(acc) : "h"(lo));
              asm("fma.rn.f32.bf16 %0, %1, %2, %0;" : "+f"(nonfin) : "h"(lo), "h"((unsigned short)0));
              asm("fma.rn.f32.bf16 %0, %1, %1, %0;" : "+f"(acc) : "h"(hi));
              asm("fma.rn.f32.bf16 %0, %1, %2, %0;" : "+f"(nonfin) : "h"(hi), "h"((unsigned short)0));
            }
            continue;
          }
        }
        float x[VE];
        Tr::unpack(raw[u], x);
        if (inv_scale != 1.0f) {  // x*1 is exact: the unscale and its store only exist with a loss scale
#pragma unroll
          for (int j = 0; j < VE; ++j) x[j] = __fmul_rn(x[j], inv_scale);
          Tr::rndv(x);
        }
#pragma unroll
        for (int j = 0; j < VE; ++j) {
          nonfin = __fmaf_rn(x[j], 0.0f, nonfin);  // stays 0 unless x is inf/nan, then NaN for good
          acc = __fmaf_rn(x[j], x[j], acc);
        }
      }
    }
  }
  const int rem = len - nfull * VE;
  if (rem > 0 && (int)threadIdx.x == nfull % kThreads) {
    float x[VE];
    if (tail_by_element) {
#pragma unroll
      for (int j = 0; j < VE; ++j) x[j] = (j < rem) ? (float)g_tile[nfull * VE + j] : 0.f;
    } else {
      Tr::unpack(ld(gv + nfull), x);
    }
#pragma unroll
    for (int j = 0; j < VE; ++j) {
      if (j < rem) {
        float y = x[j];
        if (inv_scale != 1.0f) y = Tr::rnd(__fmul_rn(y, inv_scale));
        nonfin = __fmaf_rn(y, 0.0f, nonfin);
        acc = __fmaf_rn(y, y, acc);
      }
    }
  }
}

// ------------------------------------------------- reduction tails (shared) --
// Second level: one tensor's partials summed in fp64 by a whole block, thread-strided
// then the fixed block tree: same inputs -> same bits, whoever runs it.  Partials
// written by other blocks of the same launch are read through L2 (ld.cg).
__device__ __forceinline__ void reduce_segment(const TilePartial* __restrict__ partials, long long b, long long e,
                                               double* red, double& tot, int& any_bad) {
  double acc = 0.0;
  unsigned bad = 0;
  for (long long t = b + threadIdx.x; t < e; t += kThreads) {
    const uint2 raw = __ldcg(reinterpret_cast<const uint2*>(partials + t));
    acc += (double)__uint_as_float(raw.x);
    bad |= raw.y;
  }
  any_bad = __syncthreads_or((int)bad);
  tot = block_sum_f64(acc, red);
}

// Third level (one block): per-tensor norms -> total norm -> clip coefficient,
// rounding where torch rounds (clip_grad.py:96-108 and :165-169; for bf16 each
// intermediate is a bf16 tensor).
template <typename T>
__device__ __forceinline__ void finalize_norm_block(const double* __restrict__ seg_sumsq,
                                                    const unsigned* __restrict__ seg_bad, int n_seg, float max_norm,
                                                    double* __restrict__ seg_norms, double* __restrict__ results,
                                                    double* red) {
  using Tr = ElemTraits<T>;
  double acc = 0.0;
  unsigned bad = 0;
  for (int s = threadIdx.x; s < n_seg; s += kThreads) {
    const float n = Tr::rnd(__fsqrt_rn((float)__ldcg(seg_sumsq + s)));  // _foreach_norm output
    seg_norms[s] = (double)n;
    acc += (double)n * (double)n;
    bad |= __ldcg(seg_bad + s);
  }
  const int any_bad = __syncthreads_or((int)bad);
  const double tot2 = block_sum_f64(acc, red);
  if (threadIdx.x == 0) {
    const float total = Tr::rnd(__fsqrt_rn((float)tot2));  // vector_norm(stack(norms))
    float coef = 1.f;
    if (max_norm > 0.f) {
      // max_norm / (total + 1e-6) is Tensor.__rdiv__: reciprocal() * max_norm
      const float denom = Tr::rnd(__fadd_rn(total, 1e-6f));
      const float rcp = Tr::rnd(__fdiv_rn(1.0f, denom));
      coef = Tr::rnd(__fmul_rn(rcp, max_norm));
      coef = (coef > 1.0f) ? 1.0f : coef;  // clamp(max=1); NaN stays NaN
    }
    results[DPK_R_TOTAL_NORM] = (double)total;
    results[DPK_R_CLIP_COEF] = (double)coef;
    results[DPK_R_NONFINITE] = any_bad ? 1.0 : 0.0;
  }
}

// After the fused pass: what TrainLoop._log_grad_norm computes on the grads the
// step used -- sqrt(sum_i float64((g_i**2).sum())) (utils/trainer.py:265-271).
// A step whose gradients were never reduced (peer timeout) or that predicated
// itself off (non-finite under a loss scale) reports nothing new.
template <typename T>
__device__ __forceinline__ void finalize_logged_block(const double* __restrict__ seg_sumsq, int n_seg,
                                                      int skip_if_nonfinite, double* __restrict__ results,
                                                      double* red) {
  using Tr = ElemTraits<T>;
  const bool dead = results[DPK_R_COMM_TIMEOUT] != 0.0;
  const bool skipped = dead || (skip_if_nonfinite && results[DPK_R_NONFINITE] != 0.0);
  double acc = 0.0;
  if (!skipped)
    for (int s = threadIdx.x; s < n_seg; s += kThreads)
      acc += (double)Tr::rnd((float)__ldcg(seg_sumsq + s));  // .sum() returns the grad dtype
  const double tot = block_sum_f64(acc, red);
  if (threadIdx.x == 0) {
    if (!skipped) results[DPK_R_LOGGED_NORM] = sqrt(tot);
    results[DPK_R_SKIPPED] = skipped ? 1.0 : 0.0;
  }
}

// ------------------------------------------------------- pass 1: sum sq ----
// Reads g once (s bytes/param).  One partial per tile.
// (Finishing the reduction inside this kernel -- the block that completes a
// tensor's last tile sums that tensor, the block that completes the last tensor
// finalises -- was built and measured in round 2: the per-tile fence + ticket +
// extra barrier cost 26 us on DiffuSeq-base (0.057 -> 0.083 ms), more than the
// launch it saves.  The tail is one small launch instead: norm_tail_kernel.)
// The tile loop of one worker block out of n_workers.  One barrier per tile: the
// warp totals and the per-warp non-finite flags travel through the same
// (double-buffered) exchange.  The order of the additions is block_sum's -- lane
// tree, then warp 0 over the eight warp totals -- so the partials are
// bit-identical to the three-barrier form dpk_grad_norm_static uses.
template <typename T, int U, int SRC>
__device__ __forceinline__ void sumsq_tiles(const T* __restrict__ g, const Tile* __restrict__ tiles, int tile_begin,
                                            int tile_end, float inv_scale, TilePartial* __restrict__ partials,
                                            int worker, int n_workers, const FlagWait* fw) {
  __shared__ float red[2][kThreads / 32];
  __shared__ unsigned red_bad[2][kThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int buf = 0;
  for (int t = tile_begin + worker; t < tile_end; t += n_workers) {
    const Tile tile = tiles[t];
    float acc = 0.f, nonfin = 0.f;
    // a tensor without a gradient is not part of the norm (clip_grad_norm_ skips grad=None)
    if (!(tile.seg & kTileFrozen))
      tile_sumsq<T, U, SRC>(g + tile.start, tile.len, inv_scale, false, acc, nonfin, fw,
                            tile.start / ElemTraits<T>::VE);
    const unsigned warp_bad = __any_sync(0xffffffffu, nonfin != nonfin) ? 1u : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc = __fadd_rn(acc, __shfl_down_sync(0xffffffffu, acc, o));
    if (lane == 0) {
      red[buf][warp] = acc;
      red_bad[buf][warp] = warp_bad;
    }
    __syncthreads();
    if (warp == 0) {
      float r = (lane < kThreads / 32) ? red[buf][lane] : 0.f;
      const unsigned b = (lane < kThreads / 32) ? red_bad[buf][lane] : 0u;
      const unsigned any_bad = __any_sync(0xffffffffu, b != 0u) ? 1u : 0u;
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) r = __fadd_rn(r, __shfl_down_sync(0xffffffffu, r, o));
      if (lane == 0) {
        TilePartial p;
        p.sumsq = r;
        p.bad = any_bad;
        partials[t] = p;
      }
    }
    buf ^= 1;
  }
}

template <typename T, int U>
__global__ void __launch_bounds__(kThreads)
sumsq_tiles_kernel(const T* __restrict__ g, const Tile* __restrict__ tiles, int tile_begin,
                   int tile_end, float inv_scale, TilePartial* __restrict__ partials) {
  sumsq_tiles<T, U, kSrcGlobal>(g, tiles, tile_begin, tile_end, inv_scale, partials, blockIdx.x, gridDim.x, nullptr);
}

// The same pass with the tiles staged through shared memory by the TMA engine:
// one thread issues cp.async.bulk (1-D bulk tensor copy, SASS UBLKCP) for the
// tile kStages-1 ahead and arms an mbarrier with the byte count; all threads wait
// on the barrier's phase and sum from shared memory in tile_sumsq's order, so
// the partials are bit-identical to the register path.  Selected with
// DPK_OPT_NORM_STAGING; measured against the register path in DESIGN.md.
constexpr int kTmaStages = 3;

__device__ __forceinline__ unsigned smem_addr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

template <typename T, int U>
__global__ void __launch_bounds__(kThreads)
sumsq_tiles_tma_kernel(const T* __restrict__ g, const Tile* __restrict__ tiles, int tile_begin, int tile_end,
                       int tile_bytes_max, float inv_scale, TilePartial* __restrict__ partials) {
  extern __shared__ __align__(128) unsigned char stage_mem[];
  __shared__ __align__(8) unsigned long long full[kTmaStages];
  __shared__ float red[kThreads / 32];
  __shared__ Tile desc[kTmaStages];
  constexpr int VE = ElemTraits<T>::VE;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kTmaStages; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&full[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const int first = tile_begin + blockIdx.x;
  auto issue = [&](int k) {  // thread 0: start the copy of this block's k-th tile into stage k % kTmaStages
    const int t = first + k * gridDim.x;
    if (t >= tile_end) return;
    const int s = k % kTmaStages;
    const Tile tile = tiles[t];
    desc[s] = tile;
    const unsigned bytes = (unsigned)(((tile.len + VE - 1) / VE) * 16);  // whole vectors; flat buffers are padded
    const unsigned bar = smem_addr(&full[s]);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(stage_mem + (size_t)s * tile_bytes_max)), "l"(g + tile.start), "r"(bytes), "r"(bar)
                 : "memory");
  };
  if (threadIdx.x == 0)
    for (int k = 0; k < kTmaStages - 1; ++k) issue(k);

  int k = 0;
  for (int t = first; t < tile_end; t += gridDim.x, ++k) {
    if (threadIdx.x == 0) issue(k + kTmaStages - 1);  // its stage was released by the barrier ending iteration k-1
    const int s = k % kTmaStages;
    const unsigned parity = (unsigned)(k / kTmaStages) & 1u;
    const unsigned bar = smem_addr(&full[s]);
    unsigned done = 0;
    while (!done) {
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    }
    const int len = (desc[s].seg & kTileFrozen) ? 0 : desc[s].len;
    float acc = 0.f, nonfin = 0.f;
    tile_sumsq<T, U, kSrcStaged>(reinterpret_cast<const T*>(stage_mem + (size_t)s * tile_bytes_max), len, inv_scale, false,
                           acc, nonfin);
    const unsigned bad = (nonfin != nonfin) ? 1u : 0u;
    const int any_bad = __syncthreads_or((int)bad);  // also: every thread is done reading stage s
    const float sum = block_sum(acc, red);
    if (threadIdx.x == 0) {
      TilePartial p;
      p.sumsq = sum;
      p.bad = any_bad ? 1u : 0u;
      partials[t] = p;
    }
  }
}

// Stand-alone forms of the tails (the allocation-free entry point, the TMA-staged
// variant and dpk_grad_norm_finish use them).  `per_tile` partials were written for
// every tile (1 by pass 1, chunks-per-tile by pass 2), laid out tile-major.
static __global__ void __launch_bounds__(kThreads)
segment_reduce_kernel(const TilePartial* __restrict__ partials,
                      const int* __restrict__ seg_tile_begin /*[n_seg+1]*/,
                      double* __restrict__ seg_sumsq, unsigned* __restrict__ seg_bad, int n_seg,
                      int per_tile) {
  __shared__ double red[kThreads / 32];
  for (int s = blockIdx.x; s < n_seg; s += gridDim.x) {
    double tot;
    int any_bad;
    reduce_segment(partials, (long long)seg_tile_begin[s] * per_tile, (long long)seg_tile_begin[s + 1] * per_tile,
                   red, tot, any_bad);
    if (threadIdx.x == 0) {
      seg_sumsq[s] = tot;
      seg_bad[s] = any_bad ? 1u : 0u;
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
finalize_norm_kernel(const double* __restrict__ seg_sumsq, const unsigned* __restrict__ seg_bad,
                     int n_seg, float max_norm, double* __restrict__ seg_norms,
                     double* __restrict__ results) {
  __shared__ double red[kThreads / 32];
  finalize_norm_block<T>(seg_sumsq, seg_bad, n_seg, max_norm, seg_norms, results, red);
}

// The tail of the norm pass as ONE launch: per-tensor sums (a block per tensor),
// then -- in whichever block finishes last ("last block done": store, fence,
// ticket) -- per-tensor norms, total norm and clip coefficient.  Every sum has a
// fixed order, so the result does not depend on which block that is.
template <typename T>
__global__ void __launch_bounds__(kThreads)
norm_tail_kernel(const TilePartial* __restrict__ partials, const int* __restrict__ seg_tile_begin,
                 double* __restrict__ seg_sumsq, unsigned* __restrict__ seg_bad, int n_seg, float max_norm,
                 double* __restrict__ seg_norms, double* __restrict__ results, unsigned* __restrict__ ticket) {
  __shared__ double red[kThreads / 32];
  for (int s = blockIdx.x; s < n_seg; s += gridDim.x) {
    double tot;
    int any_bad;
    reduce_segment(partials, (long long)seg_tile_begin[s], (long long)seg_tile_begin[s + 1], red, tot, any_bad);
    if (threadIdx.x == 0) {
      seg_sumsq[s] = tot;
      seg_bad[s] = any_bad ? 1u : 0u;
    }
  }
  int last = 0;
  if (threadIdx.x == 0) {
    __threadfence();
    last = (atomicAdd(ticket, 1u) + 1u == gridDim.x);
  }
  if (!__syncthreads_or(last)) return;
  __threadfence();
  finalize_norm_block<T>(seg_sumsq, seg_bad, n_seg, max_norm, seg_norms, results, red);
  if (threadIdx.x == 0) *ticket = 0u;
}

// finalize_logged_block as a launch of its own (the sharded update sums the
// per-tensor values over the ranks in between).
template <typename T>
__global__ void __launch_bounds__(kThreads)
finalize_logged_kernel(const double* __restrict__ seg_sumsq, int n_seg, int skip_if_nonfinite,
                       double* __restrict__ results) {
  __shared__ double red[kThreads / 32];
  finalize_logged_block<T>(seg_sumsq, n_seg, skip_if_nonfinite, results, red);
}

// The tail of the fused pass as ONE launch: per-tensor sums of the per-chunk
// partials, then -- in whichever block finishes last -- the logged norm.
template <typename T>
__global__ void __launch_bounds__(kThreads)
logged_tail_kernel(const TilePartial* __restrict__ partials, const int* __restrict__ seg_tile_begin,
                   double* __restrict__ seg_sumsq, int n_seg, int per_tile, int skip_if_nonfinite,
                   double* __restrict__ results, unsigned* __restrict__ ticket) {
  __shared__ double red[kThreads / 32];
  for (int s = blockIdx.x; s < n_seg; s += gridDim.x) {
    double tot;
    int any_bad;
    reduce_segment(partials, (long long)seg_tile_begin[s] * per_tile, (long long)seg_tile_begin[s + 1] * per_tile,
                   red, tot, any_bad);
    if (threadIdx.x == 0) seg_sumsq[s] = tot;
  }
  int last = 0;
  if (threadIdx.x == 0) {
    __threadfence();
    last = (atomicAdd(ticket, 1u) + 1u == gridDim.x);
  }
  if (!__syncthreads_or(last)) return;
  __threadfence();
  finalize_logged_block<T>(seg_sumsq, n_seg, skip_if_nonfinite, results, red);
  if (threadIdx.x == 0) *ticket = 0u;
}

// ------------------------------------------------ pass 2: the fused update --
// Reads g,p,m,v and K EMA copies, writes p,m,v and K EMA copies: (7+2K)*s
// bytes per parameter (+s if the gradient is rewritten or zeroed).

// Internal gradient mode of the two-rank fused pass: the gradient is the average
// of this rank's and the peer's copy, taken in registers; nothing is written back.
constexpr int kGradPair = 3;
// Internal gradient mode of the W-rank fused pass (dpk_nvls.cu: reduce_update_kernel):
// the averaged gradient is being written into this rank's buffer by the ranks
// that own each range WHILE this kernel runs; a vector may be read once the
// per-unit arrival counter says all owners have delivered their part of it.
constexpr int kGradFlagged = 4;
// Internal gradient mode of the sharded update (dpk_shard_update): this rank owns the
// chunk; its gradient is the switch-reduced sum of all W copies (multimem.ld_reduce)
// times 1/W, taken straight into registers, and the new parameters are multicast
// into every rank's replica (multimem.st).  Moments and EMA copies stay local.
constexpr int kGradSwitch = 5;

// ---- NVSwitch multimem accesses: one instruction, W copies ----------------------
template <typename T> struct Multimem;
template <> struct Multimem<float> {
  static __device__ __forceinline__ uint4 ld_reduce(const uint4* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return r;
  }
  static __device__ __forceinline__ uint4 scale(uint4 v, float s) {
    v.x = __float_as_uint(__fmul_rn(__uint_as_float(v.x), s));
    v.y = __float_as_uint(__fmul_rn(__uint_as_float(v.y), s));
    v.z = __float_as_uint(__fmul_rn(__uint_as_float(v.z), s));
    v.w = __float_as_uint(__fmul_rn(__uint_as_float(v.w), s));
    return v;
  }
};
template <> struct Multimem<__nv_bfloat16> {
  static __device__ __forceinline__ uint4 ld_reduce(const uint4* mc) {
    uint4 r;  // W bf16 addends accumulated in fp32 inside the switch, rounded once
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return r;
  }
  static __device__ __forceinline__ uint4 scale(uint4 v, float s) {
    float x[8];
    ElemTraits<__nv_bfloat16>::unpack(v, x);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = __fmul_rn(x[j], s);
    return ElemTraits<__nv_bfloat16>::pack(x);  // rounds to bf16
  }
};

__device__ __forceinline__ void multimem_st(uint4* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// (a+b)/2 per element, one rounding to the state dtype: ncclAvg's / DDP's result for two ranks
template <typename T>
__device__ __forceinline__ uint4 average2(const uint4& a, const uint4& b) {
  using Tr = ElemTraits<T>;
  float x[Tr::VE], y[Tr::VE];
  Tr::unpack(a, x);
  Tr::unpack(b, y);
#pragma unroll
  for (int j = 0; j < Tr::VE; ++j) x[j] = __fmul_rn(__fadd_rn(x[j], y[j]), 0.5f);
  return Tr::pack(x);  // rounds to the state dtype
}

template <typename T, int K>
struct UpdateArgs {
  T* p;
  T* g;
  T* m;
  T* v;
  T* ema[K > 0 ? K : 1];
  const T* peer_g;           // kGradPair only: the other rank's gradient buffer (peer mapping)
  FlagWait fw;               // kGradFlagged only
  const uint4* mc_g;         // kGradSwitch only: multicast addresses of the gradient and parameter buffers
  uint4* mc_p;
  float inv_world;
  const Tile* tiles;
  int tile_begin, tile_end;  // this launch covers tiles [tile_begin, tile_end)
  int chunk_elems;           // elements per work item: one 16-byte vector per thread
  int chunks_per_tile;       // tile_elems / chunk_elems
  TilePartial* partials;     // sum of squares per (tile, chunk), tile-major
  const double* results;   // result block of the norm pass
  int use_clip_coef;
  int skip_if_nonfinite;
  StepScalars s;
};

// Returns true only in kGradFlagged mode, when the wait for a peer was abandoned
// (nothing has been stored for this vector then).
template <typename T, int K, int GM, bool FULL>
__device__ __forceinline__ bool update_vector(const UpdateArgs<T, K>& a, long long vidx, int rem,
                                              float coef, float& acc) {
  using Tr = ElemTraits<T>;
  uint4* pv = reinterpret_cast<uint4*>(a.p) + vidx;
  uint4* gv = reinterpret_cast<uint4*>(a.g) + vidx;
  uint4* mv = reinterpret_cast<uint4*>(a.m) + vidx;
  uint4* vv = reinterpret_cast<uint4*>(a.v) + vidx;

  // all loads of this vector are issued before any arithmetic
  uint4 rg = make_uint4(0, 0, 0, 0);
  if (GM == kGradSwitch) rg = Multimem<T>::ld_reduce(a.mc_g + vidx);  // reduce-scatter: W copies summed in the switch
  else if (GM != kGradFlagged) rg = (GM == DPK_GRAD_KEEP || GM == kGradPair) ? ld_readonly(gv) : ld_stream(gv);
  uint4 rpeer = make_uint4(0, 0, 0, 0);
  if (GM == kGradPair) rpeer = ld_sys(reinterpret_cast<const uint4*>(a.peer_g) + vidx);  // over NVLink
  const uint4 rp = ld_stream(pv);
  const uint4 rm = ld_stream(mv);
  const uint4 rv = ld_stream(vv);
  uint4 re[K > 0 ? K : 1];
#pragma unroll
  for (int k = 0; k < K; ++k) re[k] = ld_stream(reinterpret_cast<uint4*>(a.ema[k]) + vidx);
  if (GM == kGradFlagged) {
    // the state loads above are in flight while this thread waits for the owners of
    // its vector's unit
    if (!flag_wait(a.fw, vidx)) return true;
    rg = ld_sys(gv);  // written by other GPUs during this kernel: not through the non-coherent path
  }

  if (GM == kGradPair) rg = average2<T>(rg, rpeer);  // the all-reduce, without a pass of its own
  if (GM == kGradSwitch) rg = Multimem<T>::scale(rg, a.inv_world);

  uint4 wg = rg, wp = rp, wm = rm, wv = rv;
  uint4 we[K > 0 ? K : 1];
#pragma unroll
  for (int k = 0; k < K; ++k) we[k] = re[k];
  const StepScalars& s = a.s;

#pragma unroll
  for (int c = 0; c < Tr::CH; ++c) {
    float g[4], p[4], m[4], v[4], t[4], d[4];
    Tr::unpack4(rg, c, g);
    Tr::unpack4(rp, c, p);
    Tr::unpack4(rm, c, m);
    Tr::unpack4(rv, c, v);
    // Each block below is one torch kernel: compute in fp32, then the store to the
    // state dtype (rnd4*: nothing for fp32, round-to-nearest-even for bf16).
    // grad.mul_(1/2^lg) (upstream loss-scale) and _foreach_mul_(grads, coef): a
    // factor of exactly 1 changes no bit, so those two kernels are skipped then.
    if (s.inv_scale != 1.0f) {
#pragma unroll
      for (int j = 0; j < 4; ++j) g[j] = __fmul_rn(g[j], s.inv_scale);
      Tr::rnd4(g);
    }
    if (coef != 1.0f) {
#pragma unroll
      for (int j = 0; j < 4; ++j) g[j] = __fmul_rn(g[j], coef);
      Tr::rnd4(g);
    }
    if (GM == DPK_GRAD_CLIPPED) Tr::rnd4_store_only(g, wg, c);  // g is representable: packing is exact
    // (p.grad ** 2).sum(): the square is a tensor of the grad dtype, summed in fp32;
    // addcmul below squares again inside its own kernel, in fp32
#pragma unroll
    for (int j = 0; j < 4; ++j) t[j] = __fmul_rn(g[j], g[j]);
    Tr::rnd4_accumulate(t, FULL ? 4 : rem - c * 4, acc);
    // _foreach_mul_(params, 1 - lr*wd), skipped when wd == 0
    if (s.has_decay) {
#pragma unroll
      for (int j = 0; j < 4; ++j) p[j] = __fmul_rn(p[j], s.decay);
      Tr::rnd4(p);
    }
    // _foreach_lerp_(exp_avg, grad, 1-beta1): m + w*(g-m), w < 0.5
#pragma unroll
    for (int j = 0; j < 4; ++j) m[j] = __fmaf_rn(s.w1, __fsub_rn(g[j], m[j]), m[j]);
    Tr::rnd4_store(m, wm, c);
    // _foreach_mul_(exp_avg_sq, beta2); _foreach_addcmul_(exp_avg_sq, g, g, 1-beta2)
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = __fmul_rn(v[j], s.beta2);
    Tr::rnd4(v);
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = __fmaf_rn(s.omb2, t[j], v[j]);
    Tr::rnd4_store(v, wv, c);
    // _foreach_sqrt; _foreach_div_(bc2_sqrt); _foreach_add_(eps)
#pragma unroll
    for (int j = 0; j < 4; ++j) d[j] = Tr::sqrt_for_store(v[j]);
    Tr::rnd4(d);
#pragma unroll
    for (int j = 0; j < 4; ++j) d[j] = div_by_scalar(d[j], s.bc2_sqrt, s.bc2_rcp);
    Tr::rnd4(d);
#pragma unroll
    for (int j = 0; j < 4; ++j) d[j] = __fadd_rn(d[j], s.eps);
    Tr::rnd4(d);
    // _foreach_addcdiv_(params, exp_avg, denom, -step_size)
#pragma unroll
    for (int j = 0; j < 4; ++j) p[j] = __fmaf_rn(s.neg_step, __fdiv_rn(m[j], d[j]), p[j]);
    Tr::rnd4_store(p, wp, c);

#pragma unroll
    for (int k = 0; k < K; ++k) {
      float e[4];
      Tr::unpack4(re[k], c, e);
      // update_ema: trg.mul_(rate).add_(src, alpha=1-rate)  (utils/trainer.py:370),
      // src being the parameters this step just produced
#pragma unroll
      for (int j = 0; j < 4; ++j) e[j] = __fmul_rn(e[j], s.rate[k]);
      Tr::rnd4(e);
#pragma unroll
      for (int j = 0; j < 4; ++j) e[j] = __fmaf_rn(s.alpha[k], p[j], e[j]);
      Tr::rnd4_store_only(e, we[k], c);
    }
  }
  if (GM == DPK_GRAD_ZERO) wg = make_uint4(0, 0, 0, 0);

  if (!FULL) {  // ragged last vector of a tensor: lanes past its end keep their bytes
    wp = Tr::merge(wp, rp, rem);
    wm = Tr::merge(wm, rm, rem);
    wv = Tr::merge(wv, rv, rem);
    if (GM == DPK_GRAD_CLIPPED || GM == DPK_GRAD_ZERO) wg = Tr::merge(wg, rg, rem);
#pragma unroll
    for (int k = 0; k < K; ++k) we[k] = Tr::merge(we[k], re[k], rem);
  }

  // Store addresses are rebuilt from the (constant-bank) base pointers and an opaque
  // copy of the index instead of being kept alive across the arithmetic: 2 live
  // registers instead of 2 per stream, which is what the 64-register bf16 build spilled.
  long long sidx = vidx;
  asm volatile("" : "+l"(sidx));
  if (GM == kGradSwitch) multimem_st(a.mc_p + sidx, wp);   // all-gather: every replica receives the new parameters
  else st_stream(reinterpret_cast<uint4*>(a.p) + sidx, wp);
  st_stream(reinterpret_cast<uint4*>(a.m) + sidx, wm);
  st_stream(reinterpret_cast<uint4*>(a.v) + sidx, wv);
  if (GM == DPK_GRAD_CLIPPED || GM == DPK_GRAD_ZERO) st_stream(reinterpret_cast<uint4*>(a.g) + sidx, wg);
#pragma unroll
  for (int k = 0; k < K; ++k) st_stream(reinterpret_cast<uint4*>(a.ema[k]) + sidx, we[k]);
  return false;
}

// A vector of a tensor that has no gradient (requires_grad=False): torch.optim.AdamW
// skips such a parameter entirely -- no decay, no moments, no step (optim/adam.py
// collects only parameters with a grad) -- while update_ema (utils/trainer.py:360-370)
// still walks every parameter.  So: p, m, v and the gradient bytes are left alone
// (the latter not even read), only the EMA copies advance.  Kept apart from
// update_vector so that the hot path carries no trace of it.
template <typename T, int K>
__device__ __forceinline__ void frozen_vector(const UpdateArgs<T, K>& a, long long vidx, int rem) {
  using Tr = ElemTraits<T>;
  if (K == 0) return;
  const uint4 rp = ld_stream(reinterpret_cast<const uint4*>(a.p) + vidx);
#pragma unroll
  for (int k = 0; k < K; ++k) {
    uint4* ev = reinterpret_cast<uint4*>(a.ema[k]) + vidx;
    const uint4 re = ld_stream(ev);
    uint4 we = re;
#pragma unroll
    for (int c = 0; c < Tr::CH; ++c) {
      float p[4], e[4];
      Tr::unpack4(rp, c, p);
      Tr::unpack4(re, c, e);
#pragma unroll
      for (int j = 0; j < 4; ++j) e[j] = __fmul_rn(e[j], a.s.rate[k]);
      Tr::rnd4(e);
#pragma unroll
      for (int j = 0; j < 4; ++j) e[j] = __fmaf_rn(a.s.alpha[k], p[j], e[j]);
      Tr::rnd4_store_only(e, we, c);
    }
    st_stream(ev, Tr::merge(we, re, rem));
  }
}

// Work distribution.  A work item is one *chunk*: kThreads consecutive 16-byte
// vectors (4 KiB) of every array, one vector per thread.  Items are dealt to the
// resident blocks round-robin, so at any moment the whole grid streams through
// one compact window of each array -- measured on B200 this is worth 6 % of HBM
// bandwidth over letting each block walk a private 32 KiB tile (DRAM pages stay
// open).  Chunks never straddle a tile, hence never a tensor; the tile
// descriptor of the next item is fetched while the current one is processed.
// resident blocks per SM the register allocation is asked to allow
template <typename T, int K>
constexpr int update_min_blocks() { return sizeof(T) == 4 ? (K <= 2 ? 4 : 3) : (K <= 2 ? 4 : 3); }

// The persistent chunk loop of one worker block out of n_workers.
template <typename T, int K, int GM>
__device__ __forceinline__ void update_chunks(const UpdateArgs<T, K>& a, int worker, int n_workers, float coef,
                                              float (*red)[kThreads / 32]) {
  using Tr = ElemTraits<T>;
  constexpr int VE = Tr::VE;
  // item = tile * chunks_per_tile + chunk, advanced by n_workers per iteration
  // without a division in the loop
  const int cpt = a.chunks_per_tile;
  const int step_tiles = n_workers / cpt, step_chunks = n_workers % cpt;
  int tix = a.tile_begin + worker / cpt;
  int chunk = worker % cpt;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  Tile next;
  next.start = 0; next.len = 0; next.seg = 0;
  if (tix < a.tile_end) next = a.tiles[tix];
  int buf = 0;
  while (tix < a.tile_end) {
    const Tile tile = next;
    const int cur_tile = tix, cur_chunk = chunk;
    chunk += step_chunks;
    tix += step_tiles;
    if (chunk >= cpt) { chunk -= cpt; ++tix; }
    if (tix < a.tile_end) next = a.tiles[tix];

    const int off = cur_chunk * a.chunk_elems;
    int len = tile.len - off;                       // elements of this chunk (<= 0: past a short tile's end)
    len = len > a.chunk_elems ? a.chunk_elems : len;
    const long long vbase = (tile.start + off) / VE;
    const int nfull = len > 0 ? len / VE : 0;       // vectors with every lane valid
    const bool frozen = (tile.seg & kTileFrozen) != 0;
    float acc = 0.f;
    bool aborted = false;
    const int v = threadIdx.x;
    if (frozen) {                                    // block-uniform: the whole tile belongs to one tensor
      if (v < nfull) frozen_vector<T, K>(a, vbase + v, VE);
      else if (v == nfull && len > nfull * VE) frozen_vector<T, K>(a, vbase + v, len - nfull * VE);
    } else if (v < nfull)
      aborted = update_vector<T, K, GM, true>(a, vbase + v, VE, coef, acc);
    else if (v == nfull && len > nfull * VE)        // ragged end of a tensor: one masked vector
      aborted = update_vector<T, K, GM, false>(a, vbase + v, len - nfull * VE, coef, acc);

    // block sum, fixed order, one barrier per item (the exchange buffer alternates)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc = __fadd_rn(acc, __shfl_down_sync(0xffffffffu, acc, o));
    if (lane == 0) red[buf][warp] = acc;
    if (GM == kGradFlagged) {
      if (__syncthreads_or((int)aborted)) return;   // a peer is gone: stop, the host raises
    } else {
      __syncthreads();
    }
    if (warp == 0) {
      float r = (lane < kThreads / 32) ? red[buf][lane] : 0.f;
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) r = __fadd_rn(r, __shfl_down_sync(0xffffffffu, r, o));
      if (lane == 0) {
        TilePartial p;
        p.sumsq = r;
        p.bad = 0;
        a.partials[(long long)cur_tile * cpt + cur_chunk] = p;
      }
    }
    buf ^= 1;
  }
}

template <typename T, int K, int GM, int MB = update_min_blocks<T, K>()>
__global__ void __launch_bounds__(kThreads, MB)
fused_update_kernel(const __grid_constant__ UpdateArgs<T, K> a) {
  __shared__ float red[2][kThreads / 32];
  if (a.skip_if_nonfinite && a.results[DPK_R_NONFINITE] != 0.0) return;
  // the library's own all-reduce gave up on a peer: these gradients were never
  // reduced, so nothing may be updated from them (the host raises with the state intact)
  if (a.results[DPK_R_COMM_TIMEOUT] != 0.0) return;
  const float coef = a.use_clip_coef ? (float)a.results[DPK_R_CLIP_COEF] : 1.0f;
  update_chunks<T, K, GM>(a, blockIdx.x, gridDim.x, coef, red);
}

// ------------------------------------------------------ stand-alone EMA ----
// update_ema(target, source, rate) on arbitrary contiguous arrays.
template <typename T>
__global__ void __launch_bounds__(kThreads)
ema_update_kernel(T* __restrict__ trg, const T* __restrict__ src, long long n, long long head,
                  float rate, float alpha) {
  using Tr = ElemTraits<T>;
  constexpr int VE = Tr::VE;
  auto one = [&](long long i) {
    float e, s;
    if constexpr (sizeof(T) == 4) { e = trg[i]; s = src[i]; }
    else { e = __bfloat162float(trg[i]); s = __bfloat162float(src[i]); }
    const float t = Tr::rnd(__fmul_rn(e, rate));
    const float r = Tr::rnd(__fmaf_rn(alpha, s, t));
    if constexpr (sizeof(T) == 4) trg[i] = r; else trg[i] = __float2bfloat16_rn(r);
  };
  const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
  const long long nthreads = (long long)gridDim.x * kThreads;
  // `head` leading elements bring both pointers to 16-byte alignment
  // (head == n when they cannot be co-aligned: all-scalar path)
  for (long long i = tid; i < head; i += nthreads) one(i);
  const long long nvec = (n - head) / VE;
  uint4* tv = reinterpret_cast<uint4*>(trg + head);
  const uint4* sv = reinterpret_cast<const uint4*>(src + head);
  for (long long v = tid; v < nvec; v += nthreads) {
    float e[VE], s[VE];
    Tr::unpack(ld_stream(tv + v), e);
    Tr::unpack(ld_readonly(sv + v), s);
#pragma unroll
    for (int j = 0; j < VE; ++j) {
      const float t = Tr::rnd(__fmul_rn(e[j], rate));
      e[j] = Tr::rnd(__fmaf_rn(alpha, s[j], t));
    }
    st_stream(tv + v, Tr::pack(e));
  }
  for (long long i = head + nvec * VE + tid; i < n; i += nthreads) one(i);
}

}  // namespace dpk
