// dpk_grad_norm_static: the global L2 gradient norm / clip coefficient of one
// flat array, for callers that must not allocate.
//
// Who that is: code that runs in the middle of a training step on a device
// whose free memory is already owned by the framework's caching allocator.
// There a raw cudaMalloc from inside a library can simply fail (the pool has
// taken everything), it fragments what is left, and it stalls the whole device
// because cudaMalloc synchronises implicitly.  dpk_create/dpk_bind do allocate
// (tile table, per-tile partial sums); this entry point does not: its scratch
// is the file-scope __device__ array below, which every GPU that loads the
// library carries until the process exits.
//
// Same arithmetic and reduction order as dpk_grad_norm on a handle whose layout
// is a single tensor of n elements with the default tile size, so the two agree
// bit for bit.
#include <mutex>

#include "dpk_internal.h"

namespace dpk {

constexpr int kStaticTileElems = 8192;                // the default tile of dpk_create
constexpr long long kStaticMaxTiles = 65536;          // -> DPK_STATIC_MAX_ELEMS elements
static_assert(kStaticMaxTiles * kStaticTileElems == DPK_STATIC_MAX_ELEMS, "header and kernel limits differ");

// Layout of the scratch: per-tile partials, then the few words the two small
// reduction kernels exchange.  Every word a call reads is written earlier in
// that same call (tiles [0, n_tiles) by the streaming kernel, the rest by the
// kernels behind it), so nothing depends on what the previous call left here.
struct StaticScratch {
  TilePartial partials[kStaticMaxTiles];
  double seg_sumsq[1];
  double seg_norms[1];
  unsigned seg_bad[1];
  int seg_tile_begin[2];
};

// 256-byte aligned so the partials start on a fresh L2 sector group.  In host
// code this name is only a stand-in for the symbol -- never a device address;
// the address comes from cudaGetSymbolAddress (static_scratch() below).
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

// Streaming pass without a tile table: tile t covers [t*8192, min(n, (t+1)*8192)).
// tile_sumsq is the handle path's own inner loop, so per-tile sums are identical.
template <typename T, int U>
__global__ void __launch_bounds__(kThreads)
sumsq_flat_kernel(const T* __restrict__ g, long long n, int n_tiles, float inv_scale,
                  StaticScratch* __restrict__ sc) {
  __shared__ float red[kThreads / 32];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    sc->seg_tile_begin[0] = 0;
    sc->seg_tile_begin[1] = n_tiles;
  }
  for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const long long start = (long long)t * kStaticTileElems;
    const long long left = n - start;
    const int len = left < kStaticTileElems ? (int)left : kStaticTileElems;
    float acc = 0.f, nonfin = 0.f;
    // the caller's array ends at n: a ragged last vector is read element by element
    tile_sumsq<T, U>(g + start, len, inv_scale, true, acc, nonfin);
    const unsigned bad = (nonfin != nonfin) ? 1u : 0u;
    const int any_bad = __syncthreads_or((int)bad);
    const float s = block_sum(acc, red);
    if (threadIdx.x == 0) {
      TilePartial p;
      p.sumsq = s;
      p.bad = any_bad ? 1u : 0u;
      sc->partials[t] = p;
    }
  }
}

constexpr int kMaxDevices = 64;
static std::mutex g_static_mu;
static StaticScratch* g_static_ptr[kMaxDevices] = {nullptr};
static int g_static_sm[kMaxDevices] = {0};

// Device address of the scratch on the CURRENT device, resolved once per device.
static int static_scratch(StaticScratch** out, int* sm_count) {
  int dev = -1;
  DPK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("dpk_grad_norm_static: device ordinal %d not supported", dev);
    return DPK_E_LIMIT;
  }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[dev]) {
    void* p = nullptr;
    DPK_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    int sm = 0;
    DPK_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    g_static_sm[dev] = sm;
    g_static_ptr[dev] = static_cast<StaticScratch*>(p);
  }
  *out = g_static_ptr[dev];
  *sm_count = g_static_sm[dev];
  return 0;
}

template <typename T>
static int run_static(const void* d_g, long long n, int n_tiles, float inv_scale, float max_norm, double* d_out,
                      StaticScratch* sc, int sm_count, cudaStream_t st) {
  // same grid rule as the handle path: resident blocks per SM x SMs, capped by the
  // work; at least one block, which also records "no tiles" for an empty input
  int occ = 0;
  DPK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sumsq_flat_kernel<T, 4>, kThreads, 0));
  long long grid = (long long)(occ < 1 ? 1 : occ) * sm_count;
  if (grid > n_tiles) grid = n_tiles;
  if (grid < 1) grid = 1;
  sumsq_flat_kernel<T, 4><<<(int)grid, kThreads, 0, st>>>((const T*)d_g, n, n_tiles, inv_scale, sc);
  DPK_CUDA(cudaGetLastError());
  segment_reduce_kernel<<<1, kThreads, 0, st>>>(sc->partials, sc->seg_tile_begin, sc->seg_sumsq, sc->seg_bad, 1, 1);
  DPK_CUDA(cudaGetLastError());
  finalize_norm_kernel<T><<<1, kThreads, 0, st>>>(sc->seg_sumsq, sc->seg_bad, 1, max_norm, sc->seg_norms, d_out);
  DPK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace dpk

using namespace dpk;

extern "C" int dpk_grad_norm_static(const void* d_grads, int64_t n, int32_t dtype, double inv_scale, double max_norm,
                                    double* d_out, void* stream) {
  if (n < 0 || (dtype != DPK_F32 && dtype != DPK_BF16) || !d_out) {
    set_error("dpk_grad_norm_static: bad argument");
    return DPK_E_BADARG;
  }
  if (n > DPK_STATIC_MAX_ELEMS) {
    set_error("dpk_grad_norm_static: %lld elements exceed the static scratch limit of %lld; "
              "use a handle (dpk_create/dpk_bind/dpk_grad_norm) for larger arrays",
              (long long)n, (long long)DPK_STATIC_MAX_ELEMS);
    return DPK_E_LIMIT;
  }
  if (n > 0 && (!d_grads || ((uintptr_t)d_grads & 15u))) {
    set_error("dpk_grad_norm_static: gradient pointer is null or not 16-byte aligned");
    return DPK_E_BADARG;
  }
  StaticScratch* sc = nullptr;
  int sm = 0;
  if (int rc = static_scratch(&sc, &sm)) return rc;
  const int n_tiles = (int)((n + kStaticTileElems - 1) / kStaticTileElems);
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == DPK_F32
             ? run_static<float>(d_grads, n, n_tiles, (float)inv_scale, (float)max_norm, d_out, sc, sm, st)
             : run_static<__nv_bfloat16>(d_grads, n, n_tiles, (float)inv_scale, (float)max_norm, d_out, sc, sm, st);
}
