// libdpk: C ABI (include/dpk.h) over the kernels in dpk_kernels.cuh.
// Host code here only validates, plans (tiles, buckets) and enqueues.
#include <dlfcn.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <new>

#include "dpk_internal.h"

namespace dpk {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return (int)e;
}

int nccl_fail(ncclResult_t r, const char* what) {
  const NcclApi* api = nccl_api();
  set_error("NCCL error %d (%s) in %s", (int)r,
            (api && api->GetErrorString) ? api->GetErrorString(r) : "?", what);
  return DPK_E_NCCL;
}

const NcclApi* nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  static char why[256] = "";
  std::call_once(once, [] {
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) {
      snprintf(why, sizeof(why), "dlopen(libnccl.so.2) failed: %s", dlerror());
      return;
    }
    bool all = true;
#define BIND(field, sym)                                         \
  do {                                                           \
    *(void**)(&api.field) = dlsym(lib, sym);                     \
    if (!api.field) { all = false; snprintf(why, sizeof(why), "missing NCCL symbol %s", sym); } \
  } while (0)
    BIND(GetUniqueId, "ncclGetUniqueId");
    BIND(CommInitRank, "ncclCommInitRank");
    BIND(CommDestroy, "ncclCommDestroy");
    BIND(CommAbort, "ncclCommAbort");
    BIND(AllReduce, "ncclAllReduce");
    BIND(AllGather, "ncclAllGather");
    BIND(Broadcast, "ncclBroadcast");
    BIND(GroupStart, "ncclGroupStart");
    BIND(GroupEnd, "ncclGroupEnd");
    BIND(CommGetAsyncError, "ncclCommGetAsyncError");
    BIND(GetErrorString, "ncclGetErrorString");
#undef BIND
    api.ok = all;
  });
  if (!api.ok) {
    set_error("NCCL unavailable: %s", why);
    return nullptr;
  }
  return &api;
}

static inline int elem_size(int dtype) { return dtype == DPK_F32 ? 4 : 2; }
static inline int vec_elems(int dtype) { return 16 / elem_size(dtype); }

// RAII: make the handle's device current for the duration of a call.
struct DeviceGuard {
  int prev = -1;
  bool switched = false;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int dev) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != dev) {
      err = cudaSetDevice(dev);
      switched = (err == cudaSuccess);
    }
  }
  ~DeviceGuard() {
    if (switched) cudaSetDevice(prev);
  }
};
#define DPK_GUARD(h)                       \
  ::dpk::DeviceGuard guard__((h)->device); \
  if (guard__.err != cudaSuccess) return ::dpk::cuda_fail(guard__.err, "cudaSetDevice")

template <typename Kern>
static int grid_for(Kern kern, int n_work, int sm_count, int* out) {
  int occ = 0;
  DPK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kThreads, 0));
  if (occ < 1) occ = 1;
  long long g = (long long)occ * sm_count;
  if (g > n_work) g = n_work;
  if (g < 1) g = 1;
  *out = (int)g;
  return 0;
}

// ---- pass 1 ---------------------------------------------------------------

template <typename T>
static int launch_sumsq(dpk_handle* h, int tile_begin, int tile_end, float inv_scale, cudaStream_t st) {
  const int n = tile_end - tile_begin;
  if (n <= 0) return 0;
  if (h->norm_staging) {
    // TMA-staged variant: kTmaStages tiles of shared memory per block
    auto kern = sumsq_tiles_tma_kernel<T, 4>;
    const int tile_bytes = h->tile_elems * (int)sizeof(T);
    const int smem = kTmaStages * tile_bytes;
    DPK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int occ = 0;
    DPK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kThreads, smem));
    const int grid = (int)std::max<long long>(1, std::min<long long>(n, (long long)std::max(occ, 1) * h->sm_count));
    {
      KernelTimer timer(h, 1, st);
      kern<<<grid, kThreads, smem, st>>>((const T*)h->g, h->d_tiles, tile_begin, tile_end, tile_bytes, inv_scale,
                                         h->d_partials_pre);
    }
    h->launches++;
    DPK_CUDA(cudaGetLastError());
    return 0;
  }
  auto kern = sumsq_tiles_kernel<T, 4>;
  int grid = 0;
  if (int rc = grid_for(kern, n, h->sm_count, &grid)) return rc;
  {
    KernelTimer timer(h, 1, st);
    kern<<<grid, kThreads, 0, st>>>((const T*)h->g, h->d_tiles, tile_begin, tile_end, inv_scale,
                                    h->d_partials_pre);
  }
  h->launches++;
  DPK_CUDA(cudaGetLastError());
  return 0;
}

// ---- pass 2 ---------------------------------------------------------------

template <typename T, int K, int GM>
static int launch_update_impl(dpk_handle* h, const StepScalars& sc, int use_coef, int skip, int tile_begin,
                              int tile_end, cudaStream_t st) {
  const int n_tiles = tile_end - tile_begin;
  if (n_tiles <= 0) return 0;
  const bool whole_buffer = tile_begin == 0 && tile_end == (int)h->tiles.size();
  UpdateArgs<T, K> a;
  a.p = (T*)h->p; a.g = (T*)h->g; a.m = (T*)h->m; a.v = (T*)h->v;
  for (int k = 0; k < K; ++k) a.ema[k] = (T*)h->ema[k];
  a.peer_g = (GM == kGradPair) ? (const T*)h->peer_grads[1 - h->rank] : nullptr;
  memset(&a.fw, 0, sizeof(a.fw));
  a.tiles = h->d_tiles;
  a.tile_begin = tile_begin;
  a.tile_end = tile_end;
  a.chunk_elems = h->chunk_elems;
  a.chunks_per_tile = h->chunks_per_tile;
  a.partials = h->d_partials_post;
  a.results = h->d_results;
  a.use_clip_coef = use_coef;
  a.skip_if_nonfinite = skip;
  a.s = sc;
  void (*kern)(const UpdateArgs<T, K>) = fused_update_kernel<T, K, GM>;
  if constexpr (sizeof(T) == 2 && K <= 2) {
    // bf16: 4 resident blocks/SM cap the kernel at 64 registers (a few spills), 3 give
    // it 80; which wins is a measurement (tuning knob, same results)
    if (h->update_blocks_per_sm == 3) kern = fused_update_kernel<T, K, GM, 3>;
  }
  int grid = 0;
  const long long items = (long long)n_tiles * h->chunks_per_tile;
  if (int rc = grid_for(kern, (int)std::min<long long>(items, 1 << 30), h->sm_count, &grid)) return rc;
  // A bucket-granular launch runs next to the all-reduce of the following
  // buckets: give it one short-lived block per chunk instead of a persistent
  // grid, so the collective's (higher-priority) blocks get SM slots as soon as
  // they are launched.  Measured at N=2: 1.17 -> 1.10 ms per step.
  if (!whole_buffer)
    grid = h->overlap_blocks_per_sm > 0
               ? (int)std::min<long long>(items, (long long)h->overlap_blocks_per_sm * h->sm_count)
               : (int)std::min<long long>(items, 1 << 30);
  {
    KernelTimer timer(h, 0, st);
  kern<<<grid, kThreads, 0, st>>>(a);
  }
  h->launches++;
  DPK_CUDA(cudaGetLastError());
  return 0;
}

template <typename T, int K>
static int launch_update_gm(dpk_handle* h, const StepScalars& sc, int use_coef, int skip, int gm, int tb, int te,
                            cudaStream_t st) {
  switch (gm) {
    case DPK_GRAD_KEEP: return launch_update_impl<T, K, DPK_GRAD_KEEP>(h, sc, use_coef, skip, tb, te, st);
    case DPK_GRAD_CLIPPED: return launch_update_impl<T, K, DPK_GRAD_CLIPPED>(h, sc, use_coef, skip, tb, te, st);
    case kGradPair: return launch_update_impl<T, K, kGradPair>(h, sc, use_coef, skip, tb, te, st);
    default: return launch_update_impl<T, K, DPK_GRAD_ZERO>(h, sc, use_coef, skip, tb, te, st);
  }
}

template <typename T>
static int launch_update(dpk_handle* h, const StepScalars& sc, int use_coef, int skip, int gm, int tb, int te,
                         cudaStream_t st) {
  switch (h->n_ema) {
    case 0: return launch_update_gm<T, 0>(h, sc, use_coef, skip, gm, tb, te, st);
    case 1: return launch_update_gm<T, 1>(h, sc, use_coef, skip, gm, tb, te, st);
    case 2: return launch_update_gm<T, 2>(h, sc, use_coef, skip, gm, tb, te, st);
    case 3: return launch_update_gm<T, 3>(h, sc, use_coef, skip, gm, tb, te, st);
    default: return launch_update_gm<T, 4>(h, sc, use_coef, skip, gm, tb, te, st);
  }
}

// Everything ensure_device_state creates; leaves the handle as dpk_create made it.
static void release_device_state(dpk_handle* h) {
  if (h->comm_stream) cudaStreamDestroy(h->comm_stream);
  if (h->ev_ready) cudaEventDestroy(h->ev_ready);
  if (h->ev_done) cudaEventDestroy(h->ev_done);
  for (cudaEvent_t e : h->ev_bucket_done) cudaEventDestroy(e);
  h->comm_stream = nullptr;
  h->ev_ready = h->ev_done = nullptr;
  h->ev_bucket_done.clear();
  cudaFree(h->d_tiles);
  cudaFree(h->d_seg_tile_begin);
  cudaFree(h->d_partials_pre);
  cudaFree(h->d_partials_post);
  cudaFree(h->d_seg_sumsq);
  cudaFree(h->d_seg_bad);
  cudaFree(h->d_seg_norms);
  cudaFree(h->d_counters);
  cudaFree(h->d_results);
  h->d_tiles = nullptr; h->d_seg_tile_begin = nullptr;
  h->d_partials_pre = h->d_partials_post = nullptr;
  h->d_seg_sumsq = nullptr; h->d_seg_bad = nullptr; h->d_seg_norms = nullptr; h->d_counters = nullptr;
  h->d_results = nullptr;
  if (h->h_results_pinned) cudaFreeHost(h->h_results_pinned);
  h->h_results_pinned = nullptr;
  for (auto& v : h->tev) {
    for (cudaEvent_t e : v) cudaEventDestroy(e);
    v.clear();
  }
  h->tcount[0] = h->tcount[1] = h->tcount[2] = 0;
}

static void free_handle(dpk_handle* h) {
  if (!h) return;
  if (h->comm) {
    const NcclApi* api = nccl_api();
    if (api) api->CommDestroy(h->comm);
  }
  release_device_state(h);
  delete h;
}

// Device-side scratch, tables, stream and events: created on the first
// dpk_bind (dpk_create itself is pure host planning and needs no GPU).
static int create_device_state(dpk_handle* h);

static int ensure_device_state(dpk_handle* h) {
  if (h->d_results) return 0;
  DPK_GUARD(h);
  const int rc = create_device_state(h);
  if (rc) release_device_state(h);  // a failed attempt leaves nothing half-built behind
  return rc;
}

static int create_device_state(dpk_handle* h) {
  int sm = 0;
  DPK_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, h->device));
  h->sm_count = sm;
  const size_t nt = std::max<size_t>(h->tiles.size(), 1), ns = std::max(h->n_seg, 1);
  DPK_CUDA(cudaMalloc(&h->d_tiles, nt * sizeof(Tile)));
  DPK_CUDA(cudaMalloc(&h->d_seg_tile_begin, (ns + 1) * sizeof(int)));
  DPK_CUDA(cudaMalloc(&h->d_partials_pre, nt * sizeof(TilePartial)));
  DPK_CUDA(cudaMalloc(&h->d_partials_post, nt * h->chunks_per_tile * sizeof(TilePartial)));
  DPK_CUDA(cudaMalloc(&h->d_seg_sumsq, ns * sizeof(double)));
  DPK_CUDA(cudaMalloc(&h->d_seg_bad, ns * sizeof(unsigned)));
  DPK_CUDA(cudaMalloc(&h->d_seg_norms, ns * sizeof(double)));
  DPK_CUDA(cudaMalloc(&h->d_counters, 2 * sizeof(unsigned)));
  DPK_CUDA(cudaMemset(h->d_counters, 0, 2 * sizeof(unsigned)));
  DPK_CUDA(cudaMallocHost(&h->h_results_pinned, DPK_R_COUNT * sizeof(double)));
  if (!h->tiles.empty())
    DPK_CUDA(cudaMemcpy(h->d_tiles, h->tiles.data(), h->tiles.size() * sizeof(Tile), cudaMemcpyHostToDevice));
  DPK_CUDA(cudaMemcpy(h->d_seg_tile_begin, h->seg_tile_begin.data(), (h->n_seg + 1) * sizeof(int),
                      cudaMemcpyHostToDevice));
  int lo = 0, hi = 0;
  DPK_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  DPK_CUDA(cudaStreamCreateWithPriority(&h->comm_stream, cudaStreamNonBlocking, hi));
  DPK_CUDA(cudaEventCreateWithFlags(&h->ev_ready, cudaEventDisableTiming));
  DPK_CUDA(cudaEventCreateWithFlags(&h->ev_done, cudaEventDisableTiming));
  for (size_t b = 0; b < h->buckets.size(); ++b) {
    cudaEvent_t e = nullptr;
    DPK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    h->ev_bucket_done.push_back(e);
  }
  h->bucket_pending.assign(h->buckets.size(), 0);
  // last: d_results doubles as the "initialised" marker
  double* res = nullptr;
  DPK_CUDA(cudaMalloc(&res, DPK_R_COUNT * sizeof(double)));
  const double init[DPK_R_COUNT] = {0, 1.0, 0, 0, 0, 0, 0, 0};
  cudaError_t e = cudaMemcpy(res, init, sizeof(init), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { cudaFree(res); return cuda_fail(e, "cudaMemcpy results"); }
  h->d_results = res;
  return 0;
}

}  // namespace dpk

using namespace dpk;

extern "C" {

int dpk_abi_version(void) { return DPK_ABI_VERSION; }
const char* dpk_last_error(void) { return g_err; }

int dpk_create(dpk_handle** out, const dpk_config* cfg) {
  if (!out || !cfg) { set_error("dpk_create: null argument"); return DPK_E_BADARG; }
  *out = nullptr;
  if (cfg->abi_version != DPK_ABI_VERSION) {
    set_error("dpk_create: abi_version %d, library is %d", cfg->abi_version, DPK_ABI_VERSION);
    return DPK_E_BADARG;
  }
  if (cfg->dtype != DPK_F32 && cfg->dtype != DPK_BF16) {
    set_error("dpk_create: unknown dtype %d", cfg->dtype);
    return DPK_E_BADARG;
  }
  if (cfg->n_ema < 0 || cfg->n_ema > DPK_MAX_EMA) {
    set_error("dpk_create: n_ema %d outside 0..%d", cfg->n_ema, DPK_MAX_EMA);
    return DPK_E_LIMIT;
  }
  if (cfg->n_segments < 0 || cfg->n_elems < 0 ||
      (cfg->n_segments > 0 && (!cfg->seg_offsets || !cfg->seg_numel)) ||
      (cfg->n_ema > 0 && !cfg->ema_rates)) {
    set_error("dpk_create: bad segment table or ema_rates");
    return DPK_E_BADARG;
  }
  const int ve = vec_elems(cfg->dtype);
  int tile_elems = cfg->tile_elems ? cfg->tile_elems : 8192;
  if (tile_elems < 1024 || tile_elems % 1024 != 0 || tile_elems > (1 << 20)) {
    set_error("dpk_create: tile_elems %d must be a multiple of 1024 in [1024, 2^20]", tile_elems);
    return DPK_E_BADARG;
  }
  if (cfg->n_elems % ve != 0) {
    set_error("dpk_create: n_elems %lld is not a whole number of 16-byte vectors", (long long)cfg->n_elems);
    return DPK_E_BADARG;
  }

  dpk_handle* h = new (std::nothrow) dpk_handle();
  if (!h) { set_error("dpk_create: out of host memory"); return DPK_E_NOMEM; }
  h->dtype = cfg->dtype;
  h->device = cfg->device;
  h->n_seg = cfg->n_segments;
  h->n_ema = cfg->n_ema;
  h->tile_elems = tile_elems;
  // pass 2 deals out chunks of one vector per thread; fall back to 1024 elements
  // when the tile size is not a multiple of that
  h->chunk_elems = (tile_elems % (kThreads * ve) == 0) ? kThreads * ve : 1024;
  h->chunks_per_tile = tile_elems / h->chunk_elems;
  h->n_elems = cfg->n_elems;
  h->beta1 = cfg->beta1; h->beta2 = cfg->beta2; h->eps = cfg->eps;
  if (const char* e = getenv("DPK_FORCE_MULTIMEM")) h->force_multimem = atoi(e) != 0;
  if (const char* e = getenv("DPK_UPDATE_BLOCKS_PER_SM")) h->update_blocks_per_sm = atoi(e);
  if (const char* e = getenv("DPK_PEER_TIMEOUT_MS")) {
    const long long ms = atoll(e);
    if (ms > 0) h->peer_timeout_ns = (unsigned long long)ms * 1000000ull;
  }
  for (int k = 0; k < cfg->n_ema; ++k) h->ema_rates[k] = cfg->ema_rates[k];

  // segments: ascending, 16-byte aligned starts, ragged tails stay inside the gap
  long long prev_end = 0;
  long long n_tiles_ll = 0;
  for (int s = 0; s < h->n_seg; ++s) {
    const long long off = cfg->seg_offsets[s], len = cfg->seg_numel[s];
    if (off < 0 || len < 0 || off % ve != 0 || off < prev_end) {
      set_error("dpk_create: segment %d (offset %lld, numel %lld) misaligned or overlapping", s, off, len);
      delete h;
      return DPK_E_BADARG;
    }
    prev_end = off + (len + ve - 1) / ve * ve;
    if (prev_end > h->n_elems) {
      set_error("dpk_create: segment %d ends at %lld past n_elems %lld", s, prev_end, (long long)h->n_elems);
      delete h;
      return DPK_E_BADARG;
    }
    h->seg_off.push_back(off);
    h->seg_len.push_back(len);
    n_tiles_ll += (len + tile_elems - 1) / tile_elems;
  }
  if (n_tiles_ll > (1LL << 28)) {  // tile and chunk indices are 32-bit in the kernels
    set_error("dpk_create: %lld tiles; raise tile_elems", n_tiles_ll);
    delete h;
    return DPK_E_LIMIT;
  }
  h->tiles.reserve((size_t)n_tiles_ll);
  h->seg_tile_begin.resize(h->n_seg + 1);
  for (int s = 0; s < h->n_seg; ++s) {
    h->seg_tile_begin[s] = (int)h->tiles.size();
    for (long long pos = 0; pos < h->seg_len[s]; pos += tile_elems) {
      Tile t;
      t.start = h->seg_off[s] + pos;
      t.len = (int)std::min<long long>(tile_elems, h->seg_len[s] - pos);
      t.seg = s | ((cfg->seg_frozen && cfg->seg_frozen[s]) ? kTileFrozen : 0);
      h->tiles.push_back(t);
    }
  }
  h->seg_tile_begin[h->n_seg] = (int)h->tiles.size();

  // buckets: walk tensors from the last (first gradient ready) to the first,
  // closing a bucket once it reaches the cap -- DDP's rebuilt-bucket policy --
  // except that the first bucket (tail) and the last one (head) are small: the
  // first all-reduce has nothing to run beside yet, and nothing is left to hide
  // the last bucket's consumer behind.
  {
    const long long cap = cfg->bucket_cap_bytes > 0 ? cfg->bucket_cap_bytes : (128LL << 20);
    const long long last_cap = cfg->edge_bucket_bytes > 0 ? cfg->edge_bucket_bytes : cap / 8;
    bool first = last_cap < cap;  // the first bucket closes early
    const int es = elem_size(h->dtype);
    auto push = [&](int seg_begin, int seg_end) {
      Bucket b;
      b.seg_begin = seg_begin;
      b.seg_end = seg_end;
      b.start = h->seg_off[seg_begin];
      const long long end = (seg_end == h->n_seg) ? h->n_elems : h->seg_off[seg_end];
      b.count = end - b.start;
      b.tile_begin = h->seg_tile_begin[seg_begin];
      b.tile_end = h->seg_tile_begin[seg_end];
      h->buckets.push_back(b);
    };
    int seg_end = h->n_seg;
    long long bytes = 0;
    for (int s = h->n_seg - 1; s >= 0; --s) {
      bytes += h->seg_len[s] * es;
      if (bytes >= (first ? last_cap : cap) || s == 0) {
        first = false;
        if (s == 0 && last_cap < cap && bytes > 2 * last_cap) {
          // head bucket [0, seg_end): cut at the first tensor boundary past last_cap
          long long head = 0;
          int cut = 0;
          while (cut < seg_end - 1 && head < last_cap) head += h->seg_len[cut++] * es;
          if (cut > 0 && cut < seg_end && head < bytes) {
            push(cut, seg_end);
            push(0, cut);
          } else {
            push(0, seg_end);
          }
        } else {
          push(s, seg_end);
        }
        seg_end = s;
        bytes = 0;
      }
    }
  }

  *out = h;
  return 0;
}

int dpk_destroy(dpk_handle* h) {
  if (!h) return 0;
  if (h->d_results || h->comm_stream || h->d_tiles) {
    DeviceGuard guard(h->device);
    if (h->comm_stream) cudaStreamSynchronize(h->comm_stream);
    free_handle(h);
  } else {
    free_handle(h);
  }
  return 0;
}

int dpk_bind(dpk_handle* h, void* d_params, void* d_grads, void* d_exp_avg, void* d_exp_avg_sq,
             void* const* d_emas) {
  if (!h) { set_error("dpk_bind: null handle"); return DPK_E_BADARG; }
  void* ptrs[4 + DPK_MAX_EMA] = {d_params, d_grads, d_exp_avg, d_exp_avg_sq};
  if (h->n_ema > 0 && !d_emas) { set_error("dpk_bind: d_emas is null"); return DPK_E_BADARG; }
  for (int k = 0; k < h->n_ema; ++k) ptrs[4 + k] = d_emas[k];
  for (int i = 0; i < 4 + h->n_ema; ++i) {
    if (h->n_elems > 0 && (!ptrs[i] || ((uintptr_t)ptrs[i] & 15u))) {
      set_error("dpk_bind: buffer %d is null or not 16-byte aligned", i);
      return DPK_E_BADARG;
    }
    for (int j = 0; j < i; ++j)
      if (h->n_elems > 0 && ptrs[i] == ptrs[j]) {
        set_error("dpk_bind: buffers %d and %d alias", j, i);
        return DPK_E_BADARG;
      }
  }
  if (int rc = ensure_device_state(h)) return rc;
  h->p = d_params; h->g = d_grads; h->m = d_exp_avg; h->v = d_exp_avg_sq;
  for (int k = 0; k < h->n_ema; ++k) h->ema[k] = d_emas[k];
  h->bound = true;
  return 0;
}

// ---- communicator -----------------------------------------------------------

int dpk_comm_unique_id(void* out128) {
  if (!out128) { set_error("dpk_comm_unique_id: null"); return DPK_E_BADARG; }
  const NcclApi* api = nccl_api();
  if (!api) return DPK_E_NCCL;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  DPK_NCCL(api->GetUniqueId(&id));
  memcpy(out128, &id, 128);
  return 0;
}

int dpk_comm_init(dpk_handle* h, const void* id128, int32_t world_size, int32_t rank) {
  if (!h || !id128 || world_size < 1 || rank < 0 || rank >= world_size) {
    set_error("dpk_comm_init: bad argument");
    return DPK_E_BADARG;
  }
  if (h->comm) { set_error("dpk_comm_init: communicator already initialised"); return DPK_E_STATE; }
  const NcclApi* api = nccl_api();
  if (!api) return DPK_E_NCCL;
  DPK_GUARD(h);
  ncclUniqueId id;
  memcpy(&id, id128, 128);
  DPK_NCCL(api->CommInitRank(&h->comm, world_size, id, rank));
  h->world = world_size;
  h->rank = rank;
  return 0;
}

int dpk_world_size(const dpk_handle* h) { return h ? h->world : 0; }

int dpk_num_buckets(const dpk_handle* h) { return h ? (int)h->buckets.size() : 0; }

int dpk_bucket_range(const dpk_handle* h, int32_t b, int64_t* start, int64_t* count) {
  if (!h || b < 0 || b >= (int)h->buckets.size() || !start || !count) {
    set_error("dpk_bucket_range: bad argument");
    return DPK_E_BADARG;
  }
  *start = h->buckets[b].start;
  *count = h->buckets[b].count;
  return 0;
}

// Caller holds h->mu.  `chained`: a previous bucket of the same chain has already
// made sure every rank's gradients are complete (see dpk_all_buckets_ready).
static int enqueue_bucket(dpk_handle* h, int b, cudaStream_t producer, bool chained) {
  const NcclApi* api = nccl_api();
  if (!api || !h->comm) { set_error("dpk_bucket_ready: communicator not initialised"); return DPK_E_STATE; }
  const Bucket& bk = h->buckets[b];
  if (!chained) {
    DPK_CUDA(cudaEventRecord(h->ev_ready, producer));
    DPK_CUDA(cudaStreamWaitEvent(h->comm_stream, h->ev_ready, 0));
  }
  KernelTimer timer(h, 2, h->comm_stream);  // instrumentation: this bucket's collective alone
  if (h->mc_grads) {
    // our own all-reduce over symmetric memory (dpk_nvls.cu)
    if (int rc = nvls_allreduce(h, bk, !chained, h->comm_stream)) return rc;
  } else {
    char* base = (char*)h->g + bk.start * elem_size(h->dtype);
    // ncclAvg pre-multiplies each rank's contribution by 1/W and sums: the
    // reducer's semantics, in place, no copy-in / copy-out.
    DPK_NCCL(api->AllReduce(base, base, (size_t)bk.count, h->dtype == DPK_F32 ? ncclFloat32 : ncclBfloat16,
                            ncclAvg, h->comm, h->comm_stream));
  }
  timer.stop();
  DPK_CUDA(cudaEventRecord(h->ev_bucket_done[b], h->comm_stream));
  DPK_CUDA(cudaEventRecord(h->ev_done, h->comm_stream));
  h->bucket_pending[b] = 1;
  h->any_enqueued = true;
  return 0;
}

int dpk_bucket_ready(dpk_handle* h, int32_t b, void* producer_stream) {
  if (!h || b < 0 || b >= (int)h->buckets.size()) { set_error("dpk_bucket_ready: bad bucket"); return DPK_E_BADARG; }
  if (!h->bound) { set_error("dpk_bucket_ready: buffers not bound"); return DPK_E_STATE; }
  std::lock_guard<std::mutex> lock(h->mu);
  DPK_GUARD(h);
  if (h->world <= 1) return 0;  // single replica: nothing to exchange
  return enqueue_bucket(h, b, (cudaStream_t)producer_stream, false);
}

int dpk_all_buckets_ready(dpk_handle* h, void* producer_stream) {
  if (!h) { set_error("dpk_all_buckets_ready: null handle"); return DPK_E_BADARG; }
  if (!h->bound) { set_error("dpk_all_buckets_ready: buffers not bound"); return DPK_E_STATE; }
  std::lock_guard<std::mutex> lock(h->mu);
  DPK_GUARD(h);
  if (h->world <= 1) return 0;
  for (int b = 0; b < (int)h->buckets.size(); ++b)
    if (int rc = enqueue_bucket(h, b, (cudaStream_t)producer_stream, b > 0)) return rc;
  return 0;
}

int dpk_wait_bucket(dpk_handle* h, int32_t b, void* consumer_stream) {
  if (!h || b < 0 || b >= (int)h->buckets.size()) { set_error("dpk_wait_bucket: bad bucket"); return DPK_E_BADARG; }
  std::lock_guard<std::mutex> lock(h->mu);
  DPK_GUARD(h);
  if (b < (int)h->bucket_pending.size() && h->bucket_pending[b]) {
    DPK_CUDA(cudaStreamWaitEvent((cudaStream_t)consumer_stream, h->ev_bucket_done[b], 0));
    h->bucket_pending[b] = 0;
  }
  return 0;
}

int dpk_wait_allreduce(dpk_handle* h, void* consumer_stream) {
  if (!h) { set_error("dpk_wait_allreduce: null handle"); return DPK_E_BADARG; }
  std::lock_guard<std::mutex> lock(h->mu);
  DPK_GUARD(h);
  if (h->any_enqueued) {
    DPK_CUDA(cudaStreamWaitEvent((cudaStream_t)consumer_stream, h->ev_done, 0));
    h->any_enqueued = false;
    std::fill(h->bucket_pending.begin(), h->bucket_pending.end(), 0);
    if (h->comm) {
      const NcclApi* api = nccl_api();
      ncclResult_t async = ncclSuccess;
      if (api && api->CommGetAsyncError(h->comm, &async) == ncclSuccess && async != ncclSuccess &&
          async != ncclInProgress)
        return nccl_fail(async, "asynchronous NCCL error");
    }
  }
  return 0;
}

int dpk_broadcast_state(dpk_handle* h, int32_t root, int32_t include_ema, void* stream) {
  if (!h) { set_error("dpk_broadcast_state: null handle"); return DPK_E_BADARG; }
  if (!h->bound) { set_error("dpk_broadcast_state: buffers not bound"); return DPK_E_STATE; }
  if (h->world <= 1) return 0;
  if (root < 0 || root >= h->world) { set_error("dpk_broadcast_state: bad root"); return DPK_E_BADARG; }
  const NcclApi* api = nccl_api();
  if (!api || !h->comm) { set_error("dpk_broadcast_state: communicator not initialised"); return DPK_E_STATE; }
  std::lock_guard<std::mutex> lock(h->mu);
  DPK_GUARD(h);
  const ncclDataType_t dt = h->dtype == DPK_F32 ? ncclFloat32 : ncclBfloat16;
  DPK_NCCL(api->GroupStart());
  ncclResult_t r = api->Broadcast(h->p, h->p, (size_t)h->n_elems, dt, root, h->comm, (cudaStream_t)stream);
  for (int k = 0; r == ncclSuccess && include_ema && k < h->n_ema; ++k)
    r = api->Broadcast(h->ema[k], h->ema[k], (size_t)h->n_elems, dt, root, h->comm, (cudaStream_t)stream);
  ncclResult_t r2 = api->GroupEnd();
  if (r != ncclSuccess) return nccl_fail(r, "ncclBroadcast");
  if (r2 != ncclSuccess) return nccl_fail(r2, "ncclGroupEnd");
  return 0;
}

// ---- the update -------------------------------------------------------------

int dpk_zero_grads(dpk_handle* h, void* stream) {
  if (!h) { set_error("dpk_zero_grads: null handle"); return DPK_E_BADARG; }
  if (!h->bound) { set_error("dpk_zero_grads: buffers not bound"); return DPK_E_STATE; }
  DPK_GUARD(h);
  DPK_CUDA(cudaMemsetAsync(h->g, 0, (size_t)h->n_elems * elem_size(h->dtype), (cudaStream_t)stream));
  return 0;
}

static int bucket_tiles(const dpk_handle* h, int32_t bucket, int* tb, int* te, const char* who) {
  if (bucket < 0) {
    *tb = 0;
    *te = (int)h->tiles.size();
    return 0;
  }
  if (bucket >= (int)h->buckets.size()) { set_error("%s: bad bucket %d", who, bucket); return DPK_E_BADARG; }
  *tb = h->buckets[bucket].tile_begin;
  *te = h->buckets[bucket].tile_end;
  return 0;
}

int dpk_grad_norm_bucket(dpk_handle* h, int32_t bucket, double inv_scale, void* stream) {
  if (!h) { set_error("dpk_grad_norm_bucket: null handle"); return DPK_E_BADARG; }
  if (!h->bound) { set_error("dpk_grad_norm: buffers not bound"); return DPK_E_STATE; }
  int tb = 0, te = 0;
  if (int rc = bucket_tiles(h, bucket, &tb, &te, "dpk_grad_norm_bucket")) return rc;
  DPK_GUARD(h);
  cudaStream_t st = (cudaStream_t)stream;
  return (h->dtype == DPK_F32) ? launch_sumsq<float>(h, tb, te, (float)inv_scale, st)
                               : launch_sumsq<__nv_bfloat16>(h, tb, te, (float)inv_scale, st);
}

int dpk_grad_norm_finish(dpk_handle* h, double max_norm, void* stream) {
  if (!h) { set_error("dpk_grad_norm_finish: null handle"); return DPK_E_BADARG; }
  if (!h->bound) { set_error("dpk_grad_norm: buffers not bound"); return DPK_E_STATE; }
  DPK_GUARD(h);
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = std::max(1, std::min(h->n_seg, 4 * h->sm_count));
  unsigned* ticket = h->d_counters;
  if (h->dtype == DPK_F32)
    norm_tail_kernel<float><<<grid, kThreads, 0, st>>>(h->d_partials_pre, h->d_seg_tile_begin, h->d_seg_sumsq,
                                                       h->d_seg_bad, h->n_seg, (float)max_norm, h->d_seg_norms,
                                                       h->d_results, ticket);
  else
    norm_tail_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>(h->d_partials_pre, h->d_seg_tile_begin, h->d_seg_sumsq,
                                                               h->d_seg_bad, h->n_seg, (float)max_norm,
                                                               h->d_seg_norms, h->d_results, ticket);
  h->launches++;
  DPK_CUDA(cudaGetLastError());
  return 0;
}

int dpk_grad_norm(dpk_handle* h, double inv_scale, double max_norm, void* stream) {
  if (int rc = dpk_grad_norm_bucket(h, -1, inv_scale, stream)) return rc;
  return dpk_grad_norm_finish(h, max_norm, stream);
}

// Python-double arithmetic of torch's _multi_tensor_adam (optim/adam.py:770-781),
// each value then narrowed to fp32 as a kernel argument.
static StepScalars step_scalars(const dpk_handle* h, double lr, double weight_decay, int64_t adam_step,
                                double inv_scale) {
  StepScalars sc;
  memset(&sc, 0, sizeof(sc));
  const double step = (double)adam_step;
  const double bc1 = 1.0 - pow(h->beta1, step);
  const double bc2 = 1.0 - pow(h->beta2, step);
  sc.inv_scale = (float)inv_scale;
  sc.decay = (float)(1.0 - lr * weight_decay);
  sc.has_decay = weight_decay != 0.0;
  sc.w1 = (float)(1.0 - h->beta1);
  sc.beta2 = (float)h->beta2;
  sc.omb2 = (float)(1.0 - h->beta2);
  sc.bc2_sqrt = (float)pow(bc2, 0.5);
  sc.bc2_rcp = (float)(1.0 / (double)sc.bc2_sqrt);  // correctly rounded reciprocal of the fp32 divisor
  sc.eps = (float)h->eps;
  sc.neg_step = (float)((lr / bc1) * -1.0);
  for (int k = 0; k < h->n_ema; ++k) {
    sc.rate[k] = (float)h->ema_rates[k];
    sc.alpha[k] = (float)(1.0 - h->ema_rates[k]);
  }
  return sc;
}

static int update_range(dpk_handle* h, int tb, int te, double lr, double weight_decay, int64_t adam_step,
                        double inv_scale, int32_t use_clip_coef, int32_t skip_if_nonfinite, int32_t grad_mode,
                        cudaStream_t st) {
  const StepScalars sc = step_scalars(h, lr, weight_decay, adam_step, inv_scale);
  return (h->dtype == DPK_F32)
             ? launch_update<float>(h, sc, use_clip_coef, skip_if_nonfinite, grad_mode, tb, te, st)
             : launch_update<__nv_bfloat16>(h, sc, use_clip_coef, skip_if_nonfinite, grad_mode, tb, te, st);
}

int dpk_fused_update_bucket(dpk_handle* h, int32_t bucket, double lr, double weight_decay, int64_t adam_step,
                            double inv_scale, int32_t use_clip_coef, int32_t skip_if_nonfinite, int32_t grad_mode,
                            void* stream) {
  if (!h) { set_error("dpk_fused_update: null handle"); return DPK_E_BADARG; }
  if (!h->bound) { set_error("dpk_fused_update: buffers not bound"); return DPK_E_STATE; }
  if (adam_step < 1) { set_error("dpk_fused_update: adam_step %lld < 1", (long long)adam_step); return DPK_E_BADARG; }
  if (grad_mode < DPK_GRAD_KEEP || grad_mode > DPK_GRAD_ZERO) { set_error("dpk_fused_update: bad grad_mode"); return DPK_E_BADARG; }
  int tb = 0, te = 0;
  if (int rc = bucket_tiles(h, bucket, &tb, &te, "dpk_fused_update_bucket")) return rc;
  DPK_GUARD(h);
  return update_range(h, tb, te, lr, weight_decay, adam_step, inv_scale, use_clip_coef, skip_if_nonfinite, grad_mode,
                      (cudaStream_t)stream);
}

// Two ranks, one pass: every gradient element is read once from this rank's HBM
// and once from the peer's over NVLink, averaged in registers and consumed by
// the update on the spot -- the all-reduce (torch DDP's reducer,
// utils/trainer.py:118-125) and optimize() (:237-244) as ONE kernel, the
// transfer overlapping the update chunk by chunk.  The gradient buffers are
// left as they were (local, not averaged).
int dpk_reduce_update_pair(dpk_handle* h, double lr, double weight_decay, int64_t adam_step, double inv_scale,
                           void* stream) {
  if (!h) { set_error("dpk_reduce_update_pair: null handle"); return DPK_E_BADARG; }
  if (!h->bound || !h->mc_grads || h->world != 2) {
    set_error("dpk_reduce_update_pair: needs two ranks and dpk_bind_nvls");
    return DPK_E_STATE;
  }
  if (adam_step < 1) { set_error("dpk_reduce_update_pair: adam_step %lld < 1", (long long)adam_step); return DPK_E_BADARG; }
  std::lock_guard<std::mutex> lock(h->mu);
  DPK_GUARD(h);
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = pair_sync(h, st)) return rc;  // the peer's gradients are complete
  if (int rc = update_range(h, 0, (int)h->tiles.size(), lr, weight_decay, adam_step, inv_scale, 0, 0, kGradPair, st))
    return rc;
  if (int rc = pair_sync(h, st)) return rc;  // the peer is done reading ours
  return dpk_fused_update_finish(h, 0, stream);
}

// Any number of ranks: the gradient all-reduce (torch DDP's reducer,
// utils/trainer.py:118-125) and optimize() (:237-244) as one kernel.  Two ranks
// take the peer-exchange form above; four or more the switch-reduced form of
// dpk_nvls.cu (reduce_update_kernel), whose gradient buffers end up holding the
// averaged gradient exactly as after dpk_all_buckets_ready.
int dpk_reduce_update(dpk_handle* h, double lr, double weight_decay, int64_t adam_step, double inv_scale,
                      void* stream) {
  if (!h) { set_error("dpk_reduce_update: null handle"); return DPK_E_BADARG; }
  if (!h->bound || !h->mc_grads || h->world < 2) {
    set_error("dpk_reduce_update: needs two or more ranks and dpk_bind_nvls");
    return DPK_E_STATE;
  }
  if (h->world == 2 && !h->force_multimem) return dpk_reduce_update_pair(h, lr, weight_decay, adam_step, inv_scale, stream);
  if (adam_step < 1) { set_error("dpk_reduce_update: adam_step %lld < 1", (long long)adam_step); return DPK_E_BADARG; }
  std::lock_guard<std::mutex> lock(h->mu);
  DPK_GUARD(h);
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = pair_sync(h, st)) return rc;  // every rank's gradients are complete
  const StepScalars sc = step_scalars(h, lr, weight_decay, adam_step, inv_scale);
  if (int rc = launch_reduce_update(h, sc, st)) return rc;
  return dpk_fused_update_finish(h, 0, stream);
}

// Four or more ranks, with clipping: the all-reduce and pass 1 as one kernel (see
// dpk_nvls.cu: reduce_norm_kernel); dpk_fused_update(use_clip_coef=1) follows.
int dpk_reduce_norm(dpk_handle* h, double inv_scale, double max_norm, void* stream) {
  if (!h) { set_error("dpk_reduce_norm: null handle"); return DPK_E_BADARG; }
  if (!h->bound || !h->mc_grads || h->world < 2) {
    set_error("dpk_reduce_norm: needs two or more ranks and dpk_bind_nvls");
    return DPK_E_STATE;
  }
  if (h->world == 2 && !h->force_multimem) {
    // two ranks have no switch-reduced form: pair all-reduce of every bucket, norm bucket by bucket behind it
    if (int rc = dpk_all_buckets_ready(h, stream)) return rc;
    for (int b = 0; b < (int)h->buckets.size(); ++b) {
      if (int rc = dpk_wait_bucket(h, b, stream)) return rc;
      if (int rc = dpk_grad_norm_bucket(h, b, inv_scale, stream)) return rc;
    }
    return dpk_grad_norm_finish(h, max_norm, stream);
  }
  {
    std::lock_guard<std::mutex> lock(h->mu);
    DPK_GUARD(h);
    cudaStream_t st = (cudaStream_t)stream;
    if (int rc = pair_sync(h, st)) return rc;  // every rank's gradients are complete
    if (int rc = launch_reduce_norm(h, (float)inv_scale, st)) return rc;
  }
  return dpk_grad_norm_finish(h, max_norm, stream);
}

// ---- sharded (ZeRO-1 style) update, opt-in ------------------------------------

int dpk_bind_shard(dpk_handle* h, void* mc_params) {
  if (!h || !mc_params || ((uintptr_t)mc_params & 15u)) { set_error("dpk_bind_shard: bad argument"); return DPK_E_BADARG; }
  if (!h->bound || !h->mc_grads || h->world < 2) {
    set_error("dpk_bind_shard: bind buffers, the communicator and dpk_bind_nvls first");
    return DPK_E_STATE;
  }
  std::lock_guard<std::mutex> lock(h->mu);
  h->mc_params = mc_params;
  return 0;
}

int dpk_shard_update(dpk_handle* h, double lr, double weight_decay, int64_t adam_step, double inv_scale,
                     void* stream) {
  if (!h) { set_error("dpk_shard_update: null handle"); return DPK_E_BADARG; }
  if (!h->bound || !h->mc_grads || !h->mc_params || h->world < 2 || !h->comm) {
    set_error("dpk_shard_update: needs two or more ranks, dpk_bind_nvls and dpk_bind_shard");
    return DPK_E_STATE;
  }
  if (adam_step < 1) { set_error("dpk_shard_update: adam_step %lld < 1", (long long)adam_step); return DPK_E_BADARG; }
  const NcclApi* api = nccl_api();
  if (!api) return DPK_E_NCCL;
  std::lock_guard<std::mutex> lock(h->mu);
  DPK_GUARD(h);
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = pair_sync(h, st)) return rc;  // every rank's gradients are complete
  const StepScalars sc = step_scalars(h, lr, weight_decay, adam_step, inv_scale);
  if (int rc = launch_shard_update(h, sc, st)) return rc;
  if (int rc = pair_sync(h, st)) return rc;  // every replica holds every rank's new parameters; gradients are free again
  h->shards_dirty = true;
  // the logged norm: per-tensor sums over this rank's chunks, summed over the ranks, then finalised
  if (h->n_seg > 0) {
    const int grid = std::min(h->n_seg, 4 * h->sm_count);
    segment_reduce_kernel<<<grid, kThreads, 0, st>>>(h->d_partials_post, h->d_seg_tile_begin, h->d_seg_sumsq,
                                                     h->d_seg_bad, h->n_seg, h->chunks_per_tile);
    h->launches++;
    DPK_CUDA(cudaGetLastError());
    DPK_NCCL(api->AllReduce(h->d_seg_sumsq, h->d_seg_sumsq, (size_t)h->n_seg, ncclFloat64, ncclSum, h->comm, st));
  }
  if (h->dtype == DPK_F32)
    finalize_logged_kernel<float><<<1, kThreads, 0, st>>>(h->d_seg_sumsq, h->n_seg, 0, h->d_results);
  else
    finalize_logged_kernel<__nv_bfloat16><<<1, kThreads, 0, st>>>(h->d_seg_sumsq, h->n_seg, 0, h->d_results);
  h->launches++;
  DPK_CUDA(cudaGetLastError());
  return 0;
}

int dpk_shard_gather(dpk_handle* h, void* stream) {
  if (!h) { set_error("dpk_shard_gather: null handle"); return DPK_E_BADARG; }
  if (!h->bound) { set_error("dpk_shard_gather: buffers not bound"); return DPK_E_STATE; }
  if (!h->shards_dirty || h->world < 2) return 0;
  const NcclApi* api = nccl_api();
  if (!api || !h->comm) { set_error("dpk_shard_gather: communicator not initialised"); return DPK_E_STATE; }
  std::lock_guard<std::mutex> lock(h->mu);
  DPK_GUARD(h);
  cudaStream_t st = (cudaStream_t)stream;
  if (int rc = launch_shard_zero_foreign(h, st)) return rc;
  const size_t words = (size_t)h->n_elems * elem_size(h->dtype) / 4;   // n_elems is a whole number of 16-byte vectors
  void* bufs[2 + DPK_MAX_EMA] = {h->m, h->v};
  for (int k = 0; k < h->n_ema; ++k) bufs[2 + k] = h->ema[k];
  DPK_NCCL(api->GroupStart());
  ncclResult_t r = ncclSuccess;
  for (int i = 0; r == ncclSuccess && i < 2 + h->n_ema; ++i)
    r = api->AllReduce(bufs[i], bufs[i], words, ncclUint32, ncclSum, h->comm, st);  // x + 0 + ... + 0: every bit survives
  ncclResult_t r2 = api->GroupEnd();
  if (r != ncclSuccess) return nccl_fail(r, "ncclAllReduce (shard gather)");
  if (r2 != ncclSuccess) return nccl_fail(r2, "ncclGroupEnd");
  h->shards_dirty = false;
  return 0;
}

int dpk_fused_update_finish(dpk_handle* h, int32_t skip_if_nonfinite, void* stream) {
  if (!h) { set_error("dpk_fused_update_finish: null handle"); return DPK_E_BADARG; }
  if (!h->bound) { set_error("dpk_fused_update: buffers not bound"); return DPK_E_STATE; }
  DPK_GUARD(h);
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = std::max(1, std::min(h->n_seg, 4 * h->sm_count));
  unsigned* ticket = h->d_counters + 1;
  if (h->dtype == DPK_F32)
    logged_tail_kernel<float><<<grid, kThreads, 0, st>>>(h->d_partials_post, h->d_seg_tile_begin, h->d_seg_sumsq,
                                                         h->n_seg, h->chunks_per_tile, skip_if_nonfinite, h->d_results,
                                                         ticket);
  else
    logged_tail_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>(h->d_partials_post, h->d_seg_tile_begin,
                                                                 h->d_seg_sumsq, h->n_seg, h->chunks_per_tile,
                                                                 skip_if_nonfinite, h->d_results, ticket);
  h->launches++;
  DPK_CUDA(cudaGetLastError());
  return 0;
}

int dpk_fused_update(dpk_handle* h, double lr, double weight_decay, int64_t adam_step, double inv_scale,
                     int32_t use_clip_coef, int32_t skip_if_nonfinite, int32_t grad_mode, void* stream) {
  if (int rc = dpk_fused_update_bucket(h, -1, lr, weight_decay, adam_step, inv_scale, use_clip_coef,
                                       skip_if_nonfinite, grad_mode, stream))
    return rc;
  return dpk_fused_update_finish(h, skip_if_nonfinite, stream);
}

int dpk_set_option(dpk_handle* h, int32_t option, int64_t value) {
  if (!h) { set_error("dpk_set_option: null handle"); return DPK_E_BADARG; }
  switch (option) {
    case DPK_OPT_OVERLAP_BLOCKS_PER_SM:
      if (value < 0 || value > 32) { set_error("dpk_set_option: blocks per SM must be in 0..32"); return DPK_E_BADARG; }
      h->overlap_blocks_per_sm = (int)value;
      return 0;
    case DPK_OPT_NORM_STAGING:
      if (value != 0 && value != 1) { set_error("dpk_set_option: staging is 0 or 1"); return DPK_E_BADARG; }
      if (value && h->tile_elems * 4 * kTmaStages > 200 * 1024) {
        set_error("dpk_set_option: tile of %d elements too large to stage %d deep in shared memory", h->tile_elems, kTmaStages);
        return DPK_E_LIMIT;
      }
      h->norm_staging = (int)value;
      return 0;
    case DPK_OPT_PEER_TIMEOUT_MS:
      if (value < 1) { set_error("dpk_set_option: timeout must be at least 1 ms"); return DPK_E_BADARG; }
      h->peer_timeout_ns = (unsigned long long)value * 1000000ull;
      return 0;
    default:
      set_error("dpk_set_option: unknown option %d", option);
      return DPK_E_BADARG;
  }
}

int dpk_result_ptr(dpk_handle* h, void** d_out) {
  if (!h || !d_out) { set_error("dpk_result_ptr: null argument"); return DPK_E_BADARG; }
  *d_out = h->d_results;
  return 0;
}

int dpk_read_result_sync(dpk_handle* h, double* out, void* stream) {
  if (!h || !out) { set_error("dpk_read_result_sync: null argument"); return DPK_E_BADARG; }
  DPK_GUARD(h);
  DPK_CUDA(cudaMemcpyAsync(h->h_results_pinned, h->d_results, DPK_R_COUNT * sizeof(double), cudaMemcpyDeviceToHost,
                           (cudaStream_t)stream));
  DPK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  memcpy(out, h->h_results_pinned, DPK_R_COUNT * sizeof(double));
  return 0;
}

int dpk_segment_norms_ptr(dpk_handle* h, void** d_out) {
  if (!h || !d_out) { set_error("dpk_segment_norms_ptr: null argument"); return DPK_E_BADARG; }
  *d_out = h->d_seg_norms;
  return 0;
}

int dpk_timing_enable(dpk_handle* h, int32_t on) {
  if (!h) { set_error("dpk_timing_enable: null handle"); return DPK_E_BADARG; }
  h->timing = on != 0;
  return 0;
}

int dpk_timing_read_sync(dpk_handle* h, int32_t which, float* ms_out, int32_t cap, int32_t* n_out) {
  if (!h || which < 0 || which > 2 || !n_out || (cap > 0 && !ms_out)) {
    set_error("dpk_timing_read_sync: bad argument");
    return DPK_E_BADARG;
  }
  DPK_GUARD(h);
  int n = 0;
  for (int i = 0; i < h->tcount[which]; ++i) {
    cudaEvent_t b = h->tev[which][2 * i], e = h->tev[which][2 * i + 1];
    DPK_CUDA(cudaEventSynchronize(e));
    float ms = 0.f;
    DPK_CUDA(cudaEventElapsedTime(&ms, b, e));
    if (n < cap) ms_out[n++] = ms;
  }
  h->tcount[which] = 0;
  *n_out = n;
  return 0;
}

int dpk_launch_count(const dpk_handle* h, int64_t* out) {
  if (!h || !out) { set_error("dpk_launch_count: null argument"); return DPK_E_BADARG; }
  *out = h->launches;
  return 0;
}

int dpk_ema_update(void* d_target, const void* d_source, int64_t n, int32_t dtype, double rate, void* stream) {
  if (n < 0 || (dtype != DPK_F32 && dtype != DPK_BF16)) { set_error("dpk_ema_update: bad argument"); return DPK_E_BADARG; }
  if (n == 0) return 0;
  if (!d_target || !d_source) { set_error("dpk_ema_update: null pointer"); return DPK_E_BADARG; }
  const int es = elem_size(dtype);
  if (((uintptr_t)d_target % es) || ((uintptr_t)d_source % es)) { set_error("dpk_ema_update: pointer not element-aligned"); return DPK_E_BADARG; }
  // leading scalars until both pointers are 16-byte aligned (possible only if
  // they are congruent modulo 16)
  long long head = n;
  if ((((uintptr_t)d_target ^ (uintptr_t)d_source) & 15u) == 0) {
    head = (long long)(((16 - ((uintptr_t)d_target & 15u)) & 15u) / es);
    if (head > n) head = n;
  }
  const float ratef = (float)rate, alphaf = (float)(1.0 - rate);
  int dev = 0, sm = 148;
  DPK_CUDA(cudaGetDevice(&dev));
  DPK_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
  const long long work = (n + (16 / es) - 1) / (16 / es);
  const int grid = (int)std::max<long long>(1, std::min<long long>((work + kThreads - 1) / kThreads, 8LL * sm));
  if (dtype == DPK_F32)
    ema_update_kernel<float><<<grid, kThreads, 0, (cudaStream_t)stream>>>((float*)d_target, (const float*)d_source, n,
                                                                         head, ratef, alphaf);
  else
    ema_update_kernel<__nv_bfloat16><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        (__nv_bfloat16*)d_target, (const __nv_bfloat16*)d_source, n, head, ratef, alphaf);
  DPK_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
