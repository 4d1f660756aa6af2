// Gradient all-reduce over NVLink SHARP (NVLS): the switch does the additions.
//
// The flat gradient buffer lives in symmetric memory that every rank has mapped
// and that is also bound to one multicast address (set up by the caller with
// torch.distributed._symmetric_memory -- plumbing; the kernels are ours).
// Per bucket, every rank owns a contiguous 1/W of the range and does
//     v = multimem.ld_reduce.add [mc + i]   (switch sums the W copies: reduce-scatter)
//     v *= 1/W                              (DDP's average; exact for W = 2, 4, 8)
//     multimem.st [mc + i], v               (switch writes all W copies: all-gather)
// so each byte crosses this GPU's links once in each direction (+1/W), against
// 2(W-1)/W for a ring and with no intermediate buffers.  This replaces the
// ncclAllReduce of dpk_bucket_ready (torch DDP's reducer,
// /root/reference/utils/trainer.py:118-125) when the caller has bound a
// multicast mapping; results differ from NCCL's only by the order in which the
// W addends are summed (none for W = 2).
//
// Cross-GPU ordering uses one 32-bit counter per block in a symmetric,
// multicast-mapped signal array: block b of every rank adds 1 to counter b on
// all ranks with a single multimem.red and spins until its own copy has
// collected W arrivals.  Only block b of the other ranks is waited for, and the
// grid is small enough to be co-resident, so no block ever waits for a block
// that cannot run.
#include <stdlib.h>

#include "dpk_internal.h"

namespace dpk {

constexpr int kNvlsThreads = 512;

__device__ __forceinline__ void mc_signal_add(unsigned* mc_addr) {
  asm volatile("multimem.red.release.sys.global.add.u32 [%0], 1;" ::"l"(mc_addr) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* addr) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

constexpr unsigned long long kPeerTimeoutNs = 30ull * 1000 * 1000 * 1000;  // a rank that is 30 s late is gone

// All threads of the block call this.  `target` is the value this block's
// counter reaches once all W ranks have arrived at this barrier.  A peer that
// never arrives (crashed rank) must not hang the GPU: after kPeerTimeoutNs the
// wait is abandoned and the sticky DPK_R_COMM_TIMEOUT flag raised for the host.
__device__ __forceinline__ void block_barrier_all_ranks(unsigned* mc_sig, const unsigned* local_sig, unsigned target,
                                                        double* results) {
  __threadfence_system();  // this thread's multimem stores before the signal
  __syncthreads();
  if (threadIdx.x == 0) {
    mc_signal_add(mc_sig + blockIdx.x);
    const unsigned long long t0 = global_timer_ns();
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(local_sig + blockIdx.x) - target) < 0) {
      if ((++spins & 0xfffu) == 0 && global_timer_ns() - t0 > kPeerTimeoutNs) {
        results[DPK_R_COMM_TIMEOUT] = 1.0;
        break;
      }
    }
  }
  __syncthreads();
}

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
    ElemTraits<__nv_bfloat16>::rndv(x);
    return ElemTraits<__nv_bfloat16>::pack(x);
  }
};

__device__ __forceinline__ void multimem_st(uint4* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// In place on vectors [vec_begin, vec_begin + n_vec) of the multicast-mapped
// gradient buffer.  sig_base: value of every block counter before this launch.
template <typename T, int kNvlsUnroll>
__global__ void __launch_bounds__(kNvlsThreads)
nvls_allreduce_kernel(uint4* mc_grads, long long vec_begin, long long n_vec, int rank, int world, float inv_world,
                      unsigned* mc_sig, const unsigned* local_sig, unsigned sig_base, int start_barrier,
                      double* results) {
  // every rank's local gradients for this range are complete once its block b is here
  // (skipped for the later buckets of a chain whose first bucket already established it)
  if (start_barrier) {
    block_barrier_all_ranks(mc_sig, local_sig, sig_base + (unsigned)world, results);
    sig_base += (unsigned)world;
  }

  const long long per_rank = (n_vec + world - 1) / world;
  const long long lo = vec_begin + (long long)rank * per_rank;
  long long hi = lo + per_rank;
  if (hi > vec_begin + n_vec) hi = vec_begin + n_vec;
  const long long stride = (long long)gridDim.x * kNvlsThreads;
  for (long long i = lo + (long long)blockIdx.x * kNvlsThreads + threadIdx.x; i < hi; i += stride * kNvlsUnroll) {
    uint4 v[kNvlsUnroll];
#pragma unroll
    for (int u = 0; u < kNvlsUnroll; ++u)
      if (i + u * stride < hi) v[u] = Multimem<T>::ld_reduce(mc_grads + i + u * stride);
#pragma unroll
    for (int u = 0; u < kNvlsUnroll; ++u)
      if (i + u * stride < hi) multimem_st(mc_grads + i + u * stride, Multimem<T>::scale(v[u], inv_world));
  }

  // every rank's slice has been written into every copy once all block-b's are here
  block_barrier_all_ranks(mc_sig, local_sig, sig_base + (unsigned)world, results);
}

// Two ranks: the switch has nothing to save (a multicast store also echoes the
// sender's own half back), so the pair exchanges halves directly.  Each rank
// reads the peer's copy of its half over NVLink, adds its own, averages, and
// writes the result into both copies: S/2 in each direction per phase, and the
// arithmetic is one correctly rounded (a+b)/2 -- bit-identical to ncclAvg.
__device__ __forceinline__ uint4 ld_peer(const uint4* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_peer(uint4* p, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

template <typename T, int U>
__global__ void __launch_bounds__(kNvlsThreads)
pair_allreduce_kernel(uint4* mine, uint4* peer, long long vec_begin, long long n_vec, int rank, unsigned* mc_sig,
                      const unsigned* local_sig, unsigned sig_base, int start_barrier, double* results) {
  if (start_barrier) {
    block_barrier_all_ranks(mc_sig, local_sig, sig_base + 2u, results);
    sig_base += 2u;
  }
  const long long per_rank = (n_vec + 1) / 2;
  const long long lo = vec_begin + (long long)rank * per_rank;
  long long hi = lo + per_rank;
  if (hi > vec_begin + n_vec) hi = vec_begin + n_vec;
  const long long stride = (long long)gridDim.x * kNvlsThreads;
  for (long long i = lo + (long long)blockIdx.x * kNvlsThreads + threadIdx.x; i < hi; i += stride * U) {
    uint4 a[U], b[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < hi) {
        b[u] = ld_peer(peer + i + u * stride);
        a[u] = ld_peer(mine + i + u * stride);
      }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < hi) {
        const uint4 v = average2<T>(a[u], b[u]);
        st_peer(peer + i + u * stride, v);
        st_peer(mine + i + u * stride, v);
      }
  }
  block_barrier_all_ranks(mc_sig, local_sig, sig_base + 2u, results);
}

// A barrier between the two ranks as a kernel of its own: stream order then keeps
// the big kernel between two of them free of any waiting.  It owns the last
// signal slot, which no all-reduce block ever touches.
__global__ void pair_sync_kernel(unsigned* mc_sig, const unsigned* local_sig, int slot, unsigned target, double* results) {
  __threadfence_system();
  if (threadIdx.x == 0) {
    mc_signal_add(mc_sig + slot);
    const unsigned long long t0 = global_timer_ns();
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(local_sig + slot) - target) < 0) {
      if ((++spins & 0xfffu) == 0 && global_timer_ns() - t0 > kPeerTimeoutNs) {
        results[DPK_R_COMM_TIMEOUT] = 1.0;
        break;
      }
    }
  }
}

int pair_sync(dpk_handle* h, cudaStream_t st) {
  h->pair_sig += 2u;
  pair_sync_kernel<<<1, 32, 0, st>>>(h->mc_signal, h->local_signal, h->nvls_slots - 1, h->pair_sig, h->d_results);
  DPK_CUDA(cudaGetLastError());
  h->launches++;
  return 0;
}

int nvls_allreduce(dpk_handle* h, const Bucket& bk, bool start_barrier, cudaStream_t st) {
  const int ve = h->dtype == DPK_F32 ? 4 : 8;
  const long long vec_begin = bk.start / ve, n_vec = bk.count / ve;
  const float inv_world = 1.0f / (float)h->world;
  // the grid never changes after dpk_bind_nvls: every block owns one counter and
  // all counters must advance in lockstep with h->nvls_sig
  const int grid = h->nvls_blocks, unroll = h->nvls_unroll;
  if (h->world == 2) {
    uint4* peer = (uint4*)h->peer_grads[1 - h->rank];
    if (h->dtype == DPK_F32)
      pair_allreduce_kernel<float, 4><<<grid, kNvlsThreads, 0, st>>>((uint4*)h->g, peer, vec_begin, n_vec, h->rank,
                                                                    h->mc_signal, h->local_signal, h->nvls_sig, (int)start_barrier,
                                                                    h->d_results);
    else
      pair_allreduce_kernel<__nv_bfloat16, 4><<<grid, kNvlsThreads, 0, st>>>((uint4*)h->g, peer, vec_begin, n_vec,
                                                                            h->rank, h->mc_signal, h->local_signal,
                                                                            h->nvls_sig, (int)start_barrier, h->d_results);
    DPK_CUDA(cudaGetLastError());
    h->nvls_sig += start_barrier ? 4u : 2u;
    h->launches++;
    return 0;
  }
#define DPK_NVLS_LAUNCH(T, U)                                                                                        \
  nvls_allreduce_kernel<T, U><<<grid, kNvlsThreads, 0, st>>>((uint4*)h->mc_grads, vec_begin, n_vec, h->rank, h->world, \
                                                            inv_world, h->mc_signal, h->local_signal, h->nvls_sig, (int)start_barrier, h->d_results)
  if (h->dtype == DPK_F32) {
    if (unroll == 8) DPK_NVLS_LAUNCH(float, 8); else if (unroll == 2) DPK_NVLS_LAUNCH(float, 2); else DPK_NVLS_LAUNCH(float, 4);
  } else {
    if (unroll == 8) DPK_NVLS_LAUNCH(__nv_bfloat16, 8); else if (unroll == 2) DPK_NVLS_LAUNCH(__nv_bfloat16, 2); else DPK_NVLS_LAUNCH(__nv_bfloat16, 4);
  }
#undef DPK_NVLS_LAUNCH
  DPK_CUDA(cudaGetLastError());
  h->nvls_sig += (start_barrier ? 2u : 1u) * (unsigned)h->world;  // W arrivals per barrier
  h->launches++;
  return 0;
}

}  // namespace dpk

using namespace dpk;

extern "C" int dpk_bind_nvls(dpk_handle* h, void* mc_grads, void* const* peer_grads, void* mc_signal,
                             void* local_signal, int32_t n_slots) {
  if (!h || !mc_grads || !peer_grads || !mc_signal || !local_signal || n_slots < 1) {
    set_error("dpk_bind_nvls: bad argument");
    return DPK_E_BADARG;
  }
  if (!h->bound || h->world <= 1) {
    set_error("dpk_bind_nvls: bind buffers and initialise the communicator first");
    return DPK_E_STATE;
  }
  if ((h->world & (h->world - 1)) != 0) {
    set_error("dpk_bind_nvls: world size %d is not a power of two (averaging by 1/W after the sum would not be DDP's "
              "arithmetic); keep the NCCL path", h->world);
    return DPK_E_LIMIT;
  }
  if (((uintptr_t)mc_grads & 15u) || (h->n_elems * (h->dtype == DPK_F32 ? 4 : 2)) % 16 != 0) {
    set_error("dpk_bind_nvls: multicast pointer not 16-byte aligned");
    return DPK_E_BADARG;
  }
  if (h->world > DPK_MAX_RANKS) { set_error("dpk_bind_nvls: more than %d ranks", DPK_MAX_RANKS); return DPK_E_LIMIT; }
  if (peer_grads[h->rank] != h->g) {
    set_error("dpk_bind_nvls: peer_grads[rank] is not the bound gradient buffer");
    return DPK_E_BADARG;
  }
  std::lock_guard<std::mutex> lock(h->mu);
  for (int r = 0; r < h->world; ++r) h->peer_grads[r] = peer_grads[r];
  h->mc_grads = mc_grads;
  h->mc_signal = (unsigned*)mc_signal;
  h->local_signal = (const unsigned*)local_signal;
  if (n_slots < 2) { set_error("dpk_bind_nvls: at least 2 signal slots"); return DPK_E_BADARG; }
  h->nvls_slots = n_slots;
  h->pair_sig = 0;
  h->nvls_blocks = std::min<int>(n_slots - 1, 32);  // link-bound: 32 blocks x 512 threads x 4 vectors saturate NVLink (measured)
  h->nvls_unroll = 4;
  if (const char* e = getenv("DPK_NVLS_BLOCKS")) h->nvls_blocks = std::max(1, std::min<int>(n_slots - 1, atoi(e)));  // tuning
  if (const char* e = getenv("DPK_NVLS_UNROLL")) h->nvls_unroll = atoi(e);
  h->nvls_sig = 0;
  return 0;
}
