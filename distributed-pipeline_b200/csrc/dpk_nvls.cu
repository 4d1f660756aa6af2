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
#include <string.h>
#include <algorithm>

#include "dpk_internal.h"

namespace dpk {

constexpr int kNvlsThreads = 512;
constexpr int kBlockSlots = 128;          // signal slots reserved for per-block counters
constexpr int kMinSlotsForFlags = 1024;   // smaller arrays carry no unit counters (no fused W-rank pass)

__device__ __forceinline__ void mc_signal_add(unsigned* mc_addr) {
  asm volatile("multimem.red.release.sys.global.add.u32 [%0], 1;" ::"l"(mc_addr) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* addr) { return ld_acquire_sys_u32(addr); }

// How long a block waits for a peer before giving up: h->peer_timeout_ns, ten
// minutes unless DPK_OPT_PEER_TIMEOUT_MS / $DPK_PEER_TIMEOUT_MS says otherwise
// (the reference's NCCL watchdog default).  A rank may legitimately be late by
// a long time: rank 0 alone runs the eval callbacks (utils/trainer.py:189-191).

// All threads of the block call this.  `target` is the value this block's
// counter reaches once all W ranks have arrived at this barrier.  A peer that
// never arrives (crashed rank) must not hang the GPU: after timeout_ns the
// wait is abandoned and the sticky DPK_R_COMM_TIMEOUT flag raised for the host;
// the update kernels then leave the state alone (dpk_kernels.cuh).
__device__ __forceinline__ void block_barrier_all_ranks(unsigned* mc_sig, const unsigned* local_sig, unsigned target,
                                                        double* results, unsigned long long timeout_ns) {
  __threadfence_system();  // this thread's multimem stores before the signal
  __syncthreads();
  if (threadIdx.x == 0) {
    mc_signal_add(mc_sig + blockIdx.x);
    const unsigned long long t0 = global_timer_ns();
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(local_sig + blockIdx.x) - target) < 0) {
      if ((++spins & 0xfffu) == 0 && global_timer_ns() - t0 > timeout_ns) {
        results[DPK_R_COMM_TIMEOUT] = 1.0;
        break;
      }
    }
  }
  __syncthreads();
}

// In place on vectors [vec_begin, vec_begin + n_vec) of the multicast-mapped
// gradient buffer.  sig_base: value of every block counter before this launch.
template <typename T, int kNvlsUnroll>
__global__ void __launch_bounds__(kNvlsThreads)
nvls_allreduce_kernel(uint4* mc_grads, long long vec_begin, long long n_vec, int rank, int world, float inv_world,
                      unsigned* mc_sig, const unsigned* local_sig, unsigned sig_base, int start_barrier,
                      double* results, unsigned long long timeout_ns) {
  // every rank's local gradients for this range are complete once its block b is here
  // (skipped for the later buckets of a chain whose first bucket already established it)
  if (start_barrier) {
    block_barrier_all_ranks(mc_sig, local_sig, sig_base + (unsigned)world, results, timeout_ns);
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
  block_barrier_all_ranks(mc_sig, local_sig, sig_base + (unsigned)world, results, timeout_ns);
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
                      const unsigned* local_sig, unsigned sig_base, int start_barrier, double* results,
                      unsigned long long timeout_ns) {
  if (start_barrier) {
    block_barrier_all_ranks(mc_sig, local_sig, sig_base + 2u, results, timeout_ns);
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
  block_barrier_all_ranks(mc_sig, local_sig, sig_base + 2u, results, timeout_ns);
}

// A barrier between all ranks as a kernel of its own: stream order then keeps
// the big kernel between two of them free of any waiting.  It owns the last
// signal slot, which no all-reduce block ever touches.
__global__ void pair_sync_kernel(unsigned* mc_sig, const unsigned* local_sig, int slot, unsigned target, double* results,
                                 unsigned long long timeout_ns) {
  __threadfence_system();
  if (threadIdx.x == 0) {
    mc_signal_add(mc_sig + slot);
    const unsigned long long t0 = global_timer_ns();
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(local_sig + slot) - target) < 0) {
      if ((++spins & 0xfffu) == 0 && global_timer_ns() - t0 > timeout_ns) {
        results[DPK_R_COMM_TIMEOUT] = 1.0;
        break;
      }
    }
  }
}

// ---- W ranks: gradient all-reduce AND the update in ONE kernel -----------------
// The grid splits into two roles that run side by side on every rank:
//   * the first n_comm blocks are the collective.  The flat gradient buffer is cut
//     into units of 2^unit_shift vectors; of every unit each rank owns 1/W.  Comm
//     block b walks units b, b+n_comm, ...: it pulls the switch-reduced sum of its
//     slice (multimem.ld_reduce), scales by 1/W and multicasts the average into all
//     W copies (multimem.st) -- the arithmetic of nvls_allreduce_kernel, so the
//     averaged gradient is bit-identical -- then adds 1 to the unit's arrival
//     counter on every rank (multimem.red.release).  These blocks never wait.
//   * the other blocks are the update: the persistent chunk loop of
//     fused_update_kernel in kGradFlagged mode.  Each thread issues the loads of
//     p, m, v and the EMA copies, then waits until its vector's unit has collected
//     W arrivals on THIS rank and only then loads the (averaged) gradient.
// So the update trails the collective by one unit instead of one bucket, there is
// no barrier between buckets and no separately launched last-bucket update.
// Progress: update blocks wait only for comm blocks (of all ranks); comm blocks
// wait for nothing, and being the lowest block indices they are resident first.
// A rank_sync kernel in front establishes "every rank's gradients are complete";
// none is needed behind: a rank leaves the kernel only after every unit arrived,
// i.e. after every owner has both read this rank's contribution and written the
// average back, and nobody touches a peer's buffer again before the next
// rank_sync.  Arrival counters are monotonic: after step s they hold s*W.
struct ReduceArgs {
  uint4* mc_grads;        // multicast address of the flat gradient buffer
  long long n_vec;        // 16-byte vectors in the buffer
  int unit_shift;
  int n_units;
  int rank, world;
  float inv_world;
  unsigned* mc_flags;     // multicast address of the per-unit arrival counters
  int n_comm;
};

// The collective role: this block's units, one after the other.
template <typename T, int U>
__device__ __forceinline__ void reduce_units(const ReduceArgs& r) {
  const long long unit_vecs = 1ll << r.unit_shift;
  for (int u = blockIdx.x; u < r.n_units; u += r.n_comm) {
    const long long ulo = (long long)u << r.unit_shift;
    long long n = r.n_vec - ulo;
    if (n > unit_vecs) n = unit_vecs;
    const long long per = (n + r.world - 1) / r.world;
    const long long lo = ulo + (long long)r.rank * per;
    long long hi = lo + per;
    if (hi > ulo + n) hi = ulo + n;
    for (long long i = lo + threadIdx.x; i < hi; i += (long long)kThreads * U) {
      uint4 v[U];
#pragma unroll
      for (int k = 0; k < U; ++k)
        if (i + k * kThreads < hi) v[k] = Multimem<T>::ld_reduce(r.mc_grads + i + k * kThreads);
#pragma unroll
      for (int k = 0; k < U; ++k)
        if (i + k * kThreads < hi) multimem_st(r.mc_grads + i + k * kThreads, Multimem<T>::scale(v[k], r.inv_world));
    }
    // every thread's stores precede the barrier, the barrier precedes thread 0's
    // release: the arrival is ordered behind the whole slice (cumulativity)
    __syncthreads();
    if (threadIdx.x == 0) mc_signal_add(r.mc_flags + u);
  }
}

template <typename T, int K, int U>
__global__ void __launch_bounds__(kThreads, update_min_blocks<T, K>())
reduce_update_kernel(const __grid_constant__ UpdateArgs<T, K> a, const __grid_constant__ ReduceArgs r) {
  __shared__ float red[2][kThreads / 32];
  if (a.results[DPK_R_COMM_TIMEOUT] != 0.0) return;  // rank_sync gave up: leave gradients and state alone
  if ((int)blockIdx.x >= r.n_comm)
    update_chunks<T, K, kGradFlagged>(a, (int)blockIdx.x - r.n_comm, (int)gridDim.x - r.n_comm, 1.0f, red);
  else
    reduce_units<T, U>(r);
}

// The same split with the norm pass as the consumer: with clipping the update
// cannot start before the total norm is known, but the sums of squares can be
// taken range by range as the averaged gradient arrives (pass 1 of
// dpk_grad_norm, same per-tile partials bit for bit); the tail kernel and the
// fused update follow as ordinary launches.
template <typename T, int U>
__global__ void __launch_bounds__(kThreads)
reduce_norm_kernel(const T* __restrict__ g, const Tile* __restrict__ tiles, int n_tiles, float inv_scale,
                   TilePartial* __restrict__ partials, const FlagWait fw, const __grid_constant__ ReduceArgs r) {
  if (fw.results[DPK_R_COMM_TIMEOUT] != 0.0) return;
  if ((int)blockIdx.x >= r.n_comm)
    sumsq_tiles<T, 4, kSrcFlagged>(g, tiles, 0, n_tiles, inv_scale, partials, (int)blockIdx.x - r.n_comm,
                                   (int)gridDim.x - r.n_comm, &fw);
  else
    reduce_units<T, U>(r);
}

static ReduceArgs reduce_args(dpk_handle* h) {
  ReduceArgs r;
  r.mc_grads = (uint4*)h->mc_grads;
  const int ve = h->dtype == DPK_F32 ? 4 : 8;
  r.n_vec = h->n_elems / ve;
  r.unit_shift = h->unit_shift;
  r.n_units = (int)((r.n_vec + (1ll << h->unit_shift) - 1) >> h->unit_shift);
  r.rank = h->rank;
  r.world = h->world;
  r.inv_world = 1.0f / (float)h->world;
  r.mc_flags = h->mc_signal + h->flag_base;
  r.n_comm = h->fused_comm_blocks;
  return r;
}

// One more step of the arrival counters: what they read once every rank has delivered.
static FlagWait next_flag_wait(dpk_handle* h) {
  FlagWait fw;
  fw.flags = h->local_signal + h->flag_base;
  h->flag_epoch += (unsigned)h->world;
  fw.target = h->flag_epoch;
  fw.shift = h->unit_shift;
  fw.timeout_ns = h->peer_timeout_ns;
  fw.results = h->d_results;
  return fw;
}

template <typename Kern>
static int fused_grid(dpk_handle* h, Kern kern, int n_comm, int* out) {
  int occ = 0;
  DPK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kThreads, 0));
  if (occ < 1) occ = 1;
  // every block must be resident at once (consumer blocks wait for collective blocks)
  int grid = occ * h->sm_count;
  if (h->fused_update_blocks > 0 && n_comm + h->fused_update_blocks < grid) grid = n_comm + h->fused_update_blocks;
  if (grid < n_comm + 1) {
    set_error("fused all-reduce pass: %d resident blocks cannot hold %d collective blocks", grid, n_comm);
    return DPK_E_LIMIT;
  }
  *out = grid;
  return 0;
}

template <typename T, int K>
static int launch_reduce_update_k(dpk_handle* h, const StepScalars& sc, cudaStream_t st) {
  UpdateArgs<T, K> a;
  memset(&a, 0, sizeof(a));
  a.p = (T*)h->p; a.g = (T*)h->g; a.m = (T*)h->m; a.v = (T*)h->v;
  for (int k = 0; k < K; ++k) a.ema[k] = (T*)h->ema[k];
  a.tiles = h->d_tiles;
  a.tile_begin = 0;
  a.tile_end = (int)h->tiles.size();
  a.chunk_elems = h->chunk_elems;
  a.chunks_per_tile = h->chunks_per_tile;
  a.partials = h->d_partials_post;
  a.results = h->d_results;
  a.s = sc;
  const ReduceArgs r = reduce_args(h);
  auto kern = reduce_update_kernel<T, K, 8>;
  int grid = 0;
  if (int rc = fused_grid(h, kern, r.n_comm, &grid)) return rc;
  a.fw = next_flag_wait(h);
  {
    KernelTimer timer(h, 0, st);
    kern<<<grid, kThreads, 0, st>>>(a, r);
  }
  h->launches++;
  DPK_CUDA(cudaGetLastError());
  return 0;
}

template <typename T>
static int launch_reduce_update_t(dpk_handle* h, const StepScalars& sc, cudaStream_t st) {
  switch (h->n_ema) {
    case 0: return launch_reduce_update_k<T, 0>(h, sc, st);
    case 1: return launch_reduce_update_k<T, 1>(h, sc, st);
    case 2: return launch_reduce_update_k<T, 2>(h, sc, st);
    case 3: return launch_reduce_update_k<T, 3>(h, sc, st);
    default: return launch_reduce_update_k<T, 4>(h, sc, st);
  }
}

template <typename T>
static int launch_reduce_norm_t(dpk_handle* h, float inv_scale, cudaStream_t st) {
  const ReduceArgs r = reduce_args(h);
  auto kern = reduce_norm_kernel<T, 8>;
  int grid = 0;
  if (int rc = fused_grid(h, kern, r.n_comm, &grid)) return rc;
  const FlagWait fw = next_flag_wait(h);
  {
    KernelTimer timer(h, 1, st);
    kern<<<grid, kThreads, 0, st>>>((const T*)h->g, h->d_tiles, (int)h->tiles.size(), inv_scale, h->d_partials_pre, fw, r);
  }
  h->launches++;
  DPK_CUDA(cudaGetLastError());
  return 0;
}

int launch_reduce_norm(dpk_handle* h, float inv_scale, cudaStream_t st) {
  if (h->flag_slots <= 0) {
    set_error("dpk_reduce_norm: the signal array bound with dpk_bind_nvls has no room for unit counters "
              "(needs at least %d slots)", kMinSlotsForFlags);
    return DPK_E_STATE;
  }
  return h->dtype == DPK_F32 ? launch_reduce_norm_t<float>(h, inv_scale, st)
                             : launch_reduce_norm_t<__nv_bfloat16>(h, inv_scale, st);
}

int launch_reduce_update(dpk_handle* h, const StepScalars& sc, cudaStream_t st) {
  if (h->flag_slots <= 0) {
    set_error("dpk_reduce_update: the signal array bound with dpk_bind_nvls has no room for unit counters "
              "(needs at least %d slots)", kMinSlotsForFlags);
    return DPK_E_STATE;
  }
  return h->dtype == DPK_F32 ? launch_reduce_update_t<float>(h, sc, st)
                             : launch_reduce_update_t<__nv_bfloat16>(h, sc, st);
}

// ---- sharded (ZeRO-1 style) update: reduce-scatter -> 1/W of the update -> all-gather ----
// Work items (4 KiB chunks of the tile table) are dealt to the ranks round-robin:
// item i belongs to rank i mod W.  A rank runs the fused update on its own items
// only, in kGradSwitch mode: the gradient vector is the switch-reduced sum of all
// ranks' copies (multimem.ld_reduce) times 1/W -- the all-reduce's arithmetic, so
// the parameters come out bit-identical to the replica path -- and the new
// parameter vector is multicast into every replica (multimem.st).  One kernel, HBM
// traffic of the update divided by W, the NVLink volume of one all-reduce.  Moments
// and EMA copies are valid on the owner only until dpk_shard_gather.
template <typename T, int K>
__global__ void __launch_bounds__(kThreads, update_min_blocks<T, K>())
shard_update_kernel(const __grid_constant__ UpdateArgs<T, K> a, int rank, int world) {
  __shared__ float red[2][kThreads / 32];
  if (a.results[DPK_R_COMM_TIMEOUT] != 0.0) return;
  // items rank, rank + W, rank + 2W, ... of which this block takes every gridDim.x-th
  update_chunks<T, K, kGradSwitch>(a, rank + world * (int)blockIdx.x, world * (int)gridDim.x, 1.0f, red);
}

// Owner-only state becomes a full replica again: every chunk this rank does not
// own is zeroed, and a bitwise all-reduce (uint32 sum: x + 0 + ... + 0) follows.
template <int NBUF>
__global__ void __launch_bounds__(kThreads)
shard_zero_foreign_kernel(uint4* b0, uint4* b1, uint4* b2, uint4* b3, uint4* b4, uint4* b5,
                          const Tile* __restrict__ tiles, int n_tiles, int chunk_elems, int chunks_per_tile, int ve,
                          int rank, int world) {
  uint4* bufs[6] = {b0, b1, b2, b3, b4, b5};
  const long long n_items = (long long)n_tiles * chunks_per_tile;
  for (long long item = blockIdx.x; item < n_items; item += gridDim.x) {
    if ((int)(item % world) == rank) continue;
    const Tile tile = tiles[item / chunks_per_tile];
    const int off = (int)(item % chunks_per_tile) * chunk_elems;
    int len = tile.len - off;
    len = len > chunk_elems ? chunk_elems : len;
    if (len <= 0) continue;
    const long long vbase = (tile.start + off) / ve;
    const int nvec = (len + ve - 1) / ve;          // the ragged last vector's pad lanes are zero everywhere already
    if ((int)threadIdx.x < nvec) {
#pragma unroll
      for (int b = 0; b < NBUF; ++b) bufs[b][vbase + threadIdx.x] = make_uint4(0, 0, 0, 0);
    }
  }
}

template <typename T, int K>
static int launch_shard_update_k(dpk_handle* h, const StepScalars& sc, cudaStream_t st) {
  UpdateArgs<T, K> a;
  memset(&a, 0, sizeof(a));
  a.p = (T*)h->p; a.g = (T*)h->g; a.m = (T*)h->m; a.v = (T*)h->v;
  for (int k = 0; k < K; ++k) a.ema[k] = (T*)h->ema[k];
  a.tiles = h->d_tiles;
  a.tile_begin = 0;
  a.tile_end = (int)h->tiles.size();
  a.chunk_elems = h->chunk_elems;
  a.chunks_per_tile = h->chunks_per_tile;
  a.partials = h->d_partials_post;
  a.results = h->d_results;
  a.s = sc;
  a.mc_g = (const uint4*)h->mc_grads;
  a.mc_p = (uint4*)h->mc_params;
  a.inv_world = 1.0f / (float)h->world;
  auto kern = shard_update_kernel<T, K>;
  int occ = 0;
  DPK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kThreads, 0));
  const long long items = ((long long)h->tiles.size() * h->chunks_per_tile + h->world - 1) / h->world;
  const int grid = (int)std::max<long long>(1, std::min<long long>(items, (long long)std::max(occ, 1) * h->sm_count));
  // partial sums of chunks other ranks own must read as zero in the tail
  DPK_CUDA(cudaMemsetAsync(h->d_partials_post, 0, h->tiles.size() * h->chunks_per_tile * sizeof(TilePartial), st));
  {
    KernelTimer timer(h, 0, st);
    kern<<<grid, kThreads, 0, st>>>(a, h->rank, h->world);
  }
  h->launches++;
  DPK_CUDA(cudaGetLastError());
  return 0;
}

int launch_shard_update(dpk_handle* h, const StepScalars& sc, cudaStream_t st) {
#define DPK_SHARD_K(T)                                            \
  switch (h->n_ema) {                                             \
    case 0: return launch_shard_update_k<T, 0>(h, sc, st);        \
    case 1: return launch_shard_update_k<T, 1>(h, sc, st);        \
    case 2: return launch_shard_update_k<T, 2>(h, sc, st);        \
    case 3: return launch_shard_update_k<T, 3>(h, sc, st);        \
    default: return launch_shard_update_k<T, 4>(h, sc, st);       \
  }
  if (h->dtype == DPK_F32) { DPK_SHARD_K(float) } else { DPK_SHARD_K(__nv_bfloat16) }
#undef DPK_SHARD_K
}

int launch_shard_zero_foreign(dpk_handle* h, cudaStream_t st) {
  uint4* b[6] = {(uint4*)h->m, (uint4*)h->v, nullptr, nullptr, nullptr, nullptr};
  for (int k = 0; k < h->n_ema; ++k) b[2 + k] = (uint4*)h->ema[k];
  const int ve = h->dtype == DPK_F32 ? 4 : 8;
  const int n_tiles = (int)h->tiles.size();
  if (n_tiles == 0) return 0;
  const int grid = std::min<long long>((long long)n_tiles * h->chunks_per_tile, 8LL * h->sm_count);
#define DPK_ZF(N) shard_zero_foreign_kernel<N><<<grid, kThreads, 0, st>>>(b[0], b[1], b[2], b[3], b[4], b[5], \
    h->d_tiles, n_tiles, h->chunk_elems, h->chunks_per_tile, ve, h->rank, h->world)
  switch (h->n_ema) {
    case 0: DPK_ZF(2); break;
    case 1: DPK_ZF(3); break;
    case 2: DPK_ZF(4); break;
    case 3: DPK_ZF(5); break;
    default: DPK_ZF(6); break;
  }
#undef DPK_ZF
  h->launches++;
  DPK_CUDA(cudaGetLastError());
  return 0;
}

int pair_sync(dpk_handle* h, cudaStream_t st) {
  h->pair_sig += (unsigned)h->world;  // one arrival per rank
  pair_sync_kernel<<<1, 32, 0, st>>>(h->mc_signal, h->local_signal, h->nvls_slots - 1, h->pair_sig, h->d_results,
                                     h->peer_timeout_ns);
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
  if (h->world == 2 && !h->force_multimem) {
    uint4* peer = (uint4*)h->peer_grads[1 - h->rank];
    if (h->dtype == DPK_F32)
      pair_allreduce_kernel<float, 4><<<grid, kNvlsThreads, 0, st>>>((uint4*)h->g, peer, vec_begin, n_vec, h->rank,
                                                                    h->mc_signal, h->local_signal, h->nvls_sig, (int)start_barrier,
                                                                    h->d_results, h->peer_timeout_ns);
    else
      pair_allreduce_kernel<__nv_bfloat16, 4><<<grid, kNvlsThreads, 0, st>>>((uint4*)h->g, peer, vec_begin, n_vec,
                                                                            h->rank, h->mc_signal, h->local_signal,
                                                                            h->nvls_sig, (int)start_barrier, h->d_results,
                                                                            h->peer_timeout_ns);
    DPK_CUDA(cudaGetLastError());
    h->nvls_sig += start_barrier ? 4u : 2u;
    h->launches++;
    return 0;
  }
#define DPK_NVLS_LAUNCH(T, U)                                                                                        \
  nvls_allreduce_kernel<T, U><<<grid, kNvlsThreads, 0, st>>>((uint4*)h->mc_grads, vec_begin, n_vec, h->rank, h->world, \
                                                            inv_world, h->mc_signal, h->local_signal, h->nvls_sig, (int)start_barrier, h->d_results, \
                                                            h->peer_timeout_ns)
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
  // Layout of the signal array: [0, block counters) one per all-reduce block;
  // [kBlockSlots, n_slots-1) per-unit arrival counters of the fused W-rank pass
  // (only when the array is large enough); the last slot belongs to rank_sync.
  const int block_slots = n_slots >= kMinSlotsForFlags ? kBlockSlots : n_slots - 1;
  h->nvls_blocks = std::min<int>(block_slots, 32);  // link-bound: 32 blocks x 512 threads x 4 vectors saturate NVLink (measured)
  h->nvls_unroll = 4;
  if (const char* e = getenv("DPK_NVLS_BLOCKS")) h->nvls_blocks = std::max(1, std::min<int>(block_slots, atoi(e)));  // tuning
  if (const char* e = getenv("DPK_NVLS_UNROLL")) h->nvls_unroll = atoi(e);
  h->nvls_sig = 0;
  h->flag_base = 0;
  h->flag_slots = 0;
  h->flag_epoch = 0;
  if (n_slots >= kMinSlotsForFlags) {
    h->flag_base = kBlockSlots;
    h->flag_slots = n_slots - 1 - kBlockSlots;
    // units of 2^unit_shift vectors: every rank's slice is a whole number of comm
    // block iterations (kThreads x 8 vectors); grow the unit until the counters suffice
    const int ve = h->dtype == DPK_F32 ? 4 : 8;
    const long long n_vec = h->n_elems / ve;
    int shift = 0;
    long long want = (long long)h->world * kThreads * 8;
    if (const char* e = getenv("DPK_FUSED_UNIT_KB")) want = std::max<long long>(1, atoll(e)) * 1024 / 16;
    while ((1ll << shift) < want) ++shift;
    while (((n_vec + (1ll << shift) - 1) >> shift) > h->flag_slots) ++shift;
    h->unit_shift = shift;
    // Collective blocks: each holds 32 KiB in flight, and a rank's share of the reduction is S/W, so the
    // count that saturates the links scales with 1/W.  Measured (ms/step, fp32 DiffuSeq-base): W=8: 16 -> 1.132,
    // 24 -> 0.972, 32 -> 0.947, 64 -> 0.999, 96 -> 1.094;  W=4: 32 -> 1.094, 64 -> 0.928.
    h->fused_comm_blocks = std::max(16, std::min(128, 256 / std::max(h->world, 1)));
    if (const char* e = getenv("DPK_FUSED_COMM_BLOCKS")) h->fused_comm_blocks = std::max(1, atoi(e));
    h->fused_update_blocks = 0;
    if (const char* e = getenv("DPK_FUSED_UPDATE_BLOCKS")) h->fused_update_blocks = std::max(1, atoi(e));
  }
  return 0;
}
