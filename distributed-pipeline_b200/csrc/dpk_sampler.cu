// Loss-aware timestep sampler (LossSecondMomentResampler) state on the device.
//
// The reference repository dropped this piece; the semantics are
// those of upstream improved-diffusion `resample.py`, restated in
// oracle/sampler_oracle.py.  Everything is fp64/integer and must match numpy
// bit for bit so that host-side np.random draws pick identical timesteps:
// the summation orders below are numpy's (pairwise_sum in
// numpy/_core/src/umath/loops_utils.h.src), not the fastest ones.
#include <string.h>
#include <new>

#include "dpk_internal.h"

namespace dpk {

struct SamplerRecord {
  long long t;
  float loss;
  int valid;
};
static_assert(sizeof(SamplerRecord) == 16, "packed record is 16 bytes");

// numpy's pairwise sum for a contiguous fp64 run (PW_BLOCKSIZE = 128).
// Leaf: fewer than 8 values are added left to right; otherwise eight running
// lanes, combined as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), then the tail.
__device__ double np_leaf_sum(const double* a, int n) {
  if (n < 8) {
    double res = 0.0;
    for (int i = 0; i < n; ++i) res = __dadd_rn(res, a[i]);
    return res;
  }
  double r[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) r[j] = a[j];
  int i;
  for (i = 8; i < n - (n % 8); i += 8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], a[i + j]);
  }
  double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                         __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
  for (; i < n; ++i) res = __dadd_rn(res, a[i]);
  return res;
}

// Above 128 values numpy splits at n/2 rounded down to a multiple of 8 and adds
// the two halves' sums.  Written with an explicit stack: device recursion would
// depend on the per-thread stack limit.
__device__ double np_pairwise_sum(const double* a, int n) {
  struct Frame { int off, n, stage; double left; };
  Frame st[32];
  int sp = 0;
  st[0].off = 0; st[0].n = n; st[0].stage = 0; st[0].left = 0.0;
  double ret = 0.0;
  while (sp >= 0) {
    Frame& f = st[sp];
    if (f.n <= 128) {
      ret = np_leaf_sum(a + f.off, f.n);
      --sp;
      continue;
    }
    int n2 = f.n / 2;
    n2 -= n2 % 8;
    if (f.stage == 0) {
      f.stage = 1;
      ++sp;
      st[sp].off = f.off; st[sp].n = n2; st[sp].stage = 0; st[sp].left = 0.0;
    } else if (f.stage == 1) {
      f.left = ret;
      f.stage = 2;
      ++sp;
      st[sp].off = f.off + n2; st[sp].n = f.n - n2; st[sp].stage = 0; st[sp].left = 0.0;
    } else {
      ret = __dadd_rn(f.left, ret);
      --sp;
    }
  }
  return ret;
}

// Pack this rank's records into its slot of the gather buffer.
__global__ void sampler_pack_kernel(const long long* __restrict__ ts, const float* __restrict__ losses,
                                    int n_local, int max_local, SamplerRecord* __restrict__ slot) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < max_local; i += gridDim.x * blockDim.x) {
    SamplerRecord r;
    r.t = (i < n_local) ? ts[i] : -1;
    r.loss = (i < n_local) ? losses[i] : 0.f;
    r.valid = (i < n_local) ? 1 : 0;
    slot[i] = r;
  }
}

// update_with_all_losses: records are applied in gathered (rank-major, batch)
// order.  One thread owns one timestep and scans the whole gathered batch, so
// repeats of a timestep land in order without atomics.
__global__ void sampler_update_kernel(const SamplerRecord* __restrict__ recs, int n_recs, int T, int H,
                                      double* __restrict__ hist, int* __restrict__ counts) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  int c = counts[t];
  double* row = hist + (long long)t * H;
  for (int i = 0; i < n_recs; ++i) {
    const SamplerRecord r = recs[i];
    if (!r.valid || r.t != (long long)t) continue;
    const double loss = (double)r.loss;  // .item() of an fp32 loss
    if (c == H) {
      for (int j = 0; j + 1 < H; ++j) row[j] = row[j + 1];
      row[H - 1] = loss;
    } else {
      row[c] = loss;
      ++c;
    }
  }
  counts[t] = c;
}

// weights(), stage 1: w[t] = sqrt(mean(hist[t]**2)); warm[0] = all(counts == H)
__global__ void sampler_rms_kernel(const double* __restrict__ hist, const int* __restrict__ counts, int T,
                                   int H, double* __restrict__ w, int* __restrict__ not_warm) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  if (counts[t] != H) atomicOr(not_warm, 1);
  double sq[128];
  const double* row = hist + (long long)t * H;
  for (int j = 0; j < H; ++j) sq[j] = __dmul_rn(row[j], row[j]);
  const double mean = __ddiv_rn(np_pairwise_sum(sq, H), (double)H);
  w[t] = __dsqrt_rn(mean);
}

// stage 2 (one thread): w /= sum(w); w *= 1-u; w += u/T -- or all ones.
__global__ void sampler_norm_kernel(double* __restrict__ w, int T, double uniform_prob,
                                    int* __restrict__ not_warm, double* __restrict__ sum_out) {
  __shared__ double total;
  if (threadIdx.x == 0) {
    total = (*not_warm) ? 0.0 : np_pairwise_sum(w, T);
    *sum_out = total;
  }
  __syncthreads();
  const bool cold = (*not_warm) != 0;
  const double keep = 1.0 - uniform_prob;
  const double add = uniform_prob / (double)T;
  for (int t = threadIdx.x; t < T; t += blockDim.x) {
    if (cold) {
      w[t] = 1.0;
    } else {
      double x = __ddiv_rn(w[t], total);
      x = __dmul_rn(x, keep);
      w[t] = __dadd_rn(x, add);
    }
  }
}

}  // namespace dpk

using namespace dpk;

struct dpk_sampler {
  int device = 0;
  int T = 0, H = 0, max_local = 0;
  double uniform_prob = 0.001;
  double* d_hist = nullptr;
  int* d_counts = nullptr;
  int* d_flag = nullptr;
  double* d_sum = nullptr;
  SamplerRecord* d_recs = nullptr;  // [capacity_ranks * max_local]
  int capacity_ranks = 0;
};

extern "C" {

int dpk_sampler_create(dpk_sampler** out, int32_t device, int32_t n_timesteps, int32_t history_per_term,
                       double uniform_prob, int32_t max_local_batch) {
  if (!out || n_timesteps < 1 || history_per_term < 1 || history_per_term > 128 || max_local_batch < 1) {
    set_error("dpk_sampler_create: bad argument (history_per_term must be 1..128)");
    return DPK_E_BADARG;
  }
  *out = nullptr;
  dpk_sampler* s = new (std::nothrow) dpk_sampler();
  if (!s) { set_error("dpk_sampler_create: out of host memory"); return DPK_E_NOMEM; }
  s->device = device; s->T = n_timesteps; s->H = history_per_term;
  s->max_local = max_local_batch; s->uniform_prob = uniform_prob;
  {
    int prev = 0;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess) e = cudaSetDevice(device);
    const size_t hb = (size_t)s->T * s->H * sizeof(double);
    if (e == cudaSuccess) e = cudaMalloc(&s->d_hist, hb);
    if (e == cudaSuccess) e = cudaMalloc(&s->d_counts, s->T * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_flag, sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_sum, sizeof(double));
    if (e == cudaSuccess) e = cudaMemset(s->d_hist, 0, hb);
    if (e == cudaSuccess) e = cudaMemset(s->d_counts, 0, s->T * sizeof(int));
    cudaSetDevice(prev);
    if (e != cudaSuccess) {
      int rc = cuda_fail(e, "dpk_sampler_create");
      dpk_sampler_destroy(s);
      return rc;
    }
  }
  *out = s;
  return 0;
}

int dpk_sampler_destroy(dpk_sampler* s) {
  if (!s) return 0;
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(s->device);
  cudaFree(s->d_hist); cudaFree(s->d_counts); cudaFree(s->d_flag); cudaFree(s->d_sum); cudaFree(s->d_recs);
  cudaSetDevice(prev);
  delete s;
  return 0;
}

int dpk_sampler_update(dpk_sampler* s, dpk_handle* h, const int64_t* d_ts, const float* d_losses, int32_t n_local,
                       void* stream) {
  if (!s || n_local < 0 || (n_local > 0 && (!d_ts || !d_losses))) { set_error("dpk_sampler_update: bad argument"); return DPK_E_BADARG; }
  if (n_local > s->max_local) {
    set_error("dpk_sampler_update: n_local %d exceeds max_local_batch %d", n_local, s->max_local);
    return DPK_E_LIMIT;
  }
  const int world = (h && h->comm) ? h->world : 1;
  const int rank = (h && h->comm) ? h->rank : 0;
  int prev = 0;
  DPK_CUDA(cudaGetDevice(&prev));
  DPK_CUDA(cudaSetDevice(s->device));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  cudaStream_t st = (cudaStream_t)stream;
  if (world > s->capacity_ranks) {
    // first call (or a larger communicator): size the gather buffer
    DPK_CUDA(cudaStreamSynchronize(st));
    cudaFree(s->d_recs);
    s->d_recs = nullptr;
    DPK_CUDA(cudaMalloc(&s->d_recs, (size_t)world * s->max_local * sizeof(SamplerRecord)));
    s->capacity_ranks = world;
  }
  SamplerRecord* slot = s->d_recs + (size_t)rank * s->max_local;
  sampler_pack_kernel<<<(s->max_local + 255) / 256, 256, 0, st>>>((const long long*)d_ts, d_losses, n_local,
                                                                   s->max_local, slot);
  DPK_CUDA(cudaGetLastError());
  if (world > 1) {
    const NcclApi* api = nccl_api();
    if (!api) return DPK_E_NCCL;
    std::lock_guard<std::mutex> lock(h->mu);
    DPK_NCCL(api->AllGather(slot, s->d_recs, (size_t)s->max_local * sizeof(SamplerRecord), ncclUint8, h->comm, st));
  }
  sampler_update_kernel<<<(s->T + 127) / 128, 128, 0, st>>>(s->d_recs, world * s->max_local, s->T, s->H, s->d_hist,
                                                            s->d_counts);
  DPK_CUDA(cudaGetLastError());
  return 0;
}

int dpk_sampler_weights(dpk_sampler* s, double* d_weights, void* stream) {
  if (!s || !d_weights) { set_error("dpk_sampler_weights: null argument"); return DPK_E_BADARG; }
  int prev = 0;
  DPK_CUDA(cudaGetDevice(&prev));
  DPK_CUDA(cudaSetDevice(s->device));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
  cudaStream_t st = (cudaStream_t)stream;
  DPK_CUDA(cudaMemsetAsync(s->d_flag, 0, sizeof(int), st));
  sampler_rms_kernel<<<(s->T + 63) / 64, 64, 0, st>>>(s->d_hist, s->d_counts, s->T, s->H, d_weights, s->d_flag);
  DPK_CUDA(cudaGetLastError());
  sampler_norm_kernel<<<1, 256, 0, st>>>(d_weights, s->T, s->uniform_prob, s->d_flag, s->d_sum);
  DPK_CUDA(cudaGetLastError());
  return 0;
}

int dpk_sampler_state_ptrs(dpk_sampler* s, void** d_history, void** d_counts) {
  if (!s || !d_history || !d_counts) { set_error("dpk_sampler_state_ptrs: null argument"); return DPK_E_BADARG; }
  *d_history = s->d_hist;
  *d_counts = s->d_counts;
  return 0;
}

}  // extern "C"
