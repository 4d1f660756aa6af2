// Host-side internals shared by the translation units of libdpk.
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>  // types and enums only; the functions are bound with dlsym
#include <stdint.h>
#include <mutex>
#include <string>
#include <vector>
#include "../../include/dpk.h"
#include "dpk_kernels.cuh"

namespace dpk {

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define DPK_CUDA(call)                                        \
  do {                                                        \
    cudaError_t e__ = (call);                                 \
    if (e__ != cudaSuccess) return ::dpk::cuda_fail(e__, #call); \
  } while (0)

// NCCL is resolved at run time from the copy torch already loaded
// (libnccl.so.2), so the library itself has no link-time NCCL dependency and
// loads on machines without it.
struct NcclApi {
  bool ok = false;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*CommGetAsyncError)(ncclComm_t, ncclResult_t*) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
const NcclApi* nccl_api();  // nullptr (with dpk_last_error set) if unavailable
int nccl_fail(ncclResult_t r, const char* what);

#define DPK_NCCL(call)                                               \
  do {                                                               \
    ncclResult_t r__ = (call);                                       \
    if (r__ != ncclSuccess) return ::dpk::nccl_fail(r__, #call);     \
  } while (0)

struct Bucket;
}  // namespace dpk
struct dpk_handle;
namespace dpk {
int nvls_allreduce(dpk_handle* h, const Bucket& bk, bool start_barrier, cudaStream_t st);
int pair_sync(dpk_handle* h, cudaStream_t st);  // barrier between all ranks, a kernel of its own
int launch_reduce_update(dpk_handle* h, const StepScalars& sc, cudaStream_t st);
int launch_reduce_norm(dpk_handle* h, float inv_scale, cudaStream_t st);
int launch_shard_update(dpk_handle* h, const StepScalars& sc, cudaStream_t st);
int launch_shard_zero_foreign(dpk_handle* h, cudaStream_t st);

struct Bucket {
  int seg_begin, seg_end;     // tensors [seg_begin, seg_end)
  long long start, count;     // flat elements
  int tile_begin, tile_end;   // tiles covering those tensors
};

}  // namespace dpk

struct dpk_handle {
  int dtype = DPK_F32;
  int device = 0;
  int n_seg = 0;
  int n_ema = 0;
  int tile_elems = 0;
  int chunk_elems = 0;      // pass-2 work item; divides tile_elems
  int chunks_per_tile = 1;
  int overlap_blocks_per_sm = 0;  // DPK_OPT_OVERLAP_BLOCKS_PER_SM
  int norm_staging = 0;           // DPK_OPT_NORM_STAGING
  int update_blocks_per_sm = 0;   // $DPK_UPDATE_BLOCKS_PER_SM: 3 = the 80-register build of the bf16 fused kernel
  long long n_elems = 0;
  double beta1 = 0.9, beta2 = 0.999, eps = 1e-8;
  double ema_rates[DPK_MAX_EMA] = {0};
  std::vector<long long> seg_off, seg_len;
  std::vector<dpk::Tile> tiles;          // host copy
  std::vector<int> seg_tile_begin;       // [n_seg + 1]
  std::vector<dpk::Bucket> buckets;
  int sm_count = 148;

  // borrowed
  void* p = nullptr; void* g = nullptr; void* m = nullptr; void* v = nullptr;
  void* ema[DPK_MAX_EMA] = {nullptr};
  bool bound = false;

  // owned device scratch
  dpk::Tile* d_tiles = nullptr;
  int* d_seg_tile_begin = nullptr;
  dpk::TilePartial* d_partials_pre = nullptr;   // pass 1
  dpk::TilePartial* d_partials_post = nullptr;  // pass 2
  double* d_seg_sumsq = nullptr;
  unsigned* d_seg_bad = nullptr;
  double* d_seg_norms = nullptr;
  unsigned* d_counters = nullptr;  // "last block done" tickets of the two tail kernels (zero between launches)
  double* d_results = nullptr;
  double* h_results_pinned = nullptr;

  // communication
  ncclComm_t comm = nullptr;
  int world = 1, rank = 0;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_ready = nullptr, ev_done = nullptr;
  std::vector<cudaEvent_t> ev_bucket_done;  // one per bucket
  std::vector<char> bucket_pending;          // all-reduce enqueued, not yet waited for
  bool any_enqueued = false;
  std::mutex mu;

  // NVLS all-reduce (dpk_nvls.cu): multicast mapping of the gradient buffer and
  // of the per-block signal counters, bound by the caller
  void* mc_grads = nullptr;
  void* mc_params = nullptr;   // dpk_bind_shard: multicast mapping of the parameter buffer
  bool shards_dirty = false;   // moments / EMA copies currently valid on the owner rank only
  void* peer_grads[DPK_MAX_RANKS] = {nullptr};  // every rank's mapping of the gradient buffer
  unsigned* mc_signal = nullptr;
  const unsigned* local_signal = nullptr;
  int nvls_blocks = 0;
  bool force_multimem = false;  // $DPK_FORCE_MULTIMEM=1: two ranks take the switch (W >= 4) kernels too (development)
  int nvls_slots = 0;
  unsigned pair_sig = 0;  // value of the last slot's counter after the pair barriers enqueued so far
  int nvls_unroll = 4;
  unsigned nvls_sig = 0;  // value of every block counter after the launches enqueued so far
  // fused W-rank pass: per-unit arrival counters inside the signal array
  int flag_base = 0, flag_slots = 0;
  unsigned flag_epoch = 0;   // value every unit counter has after the steps enqueued so far
  int unit_shift = 0;        // log2(vectors per unit)
  int fused_comm_blocks = 32, fused_update_blocks = 0;
  unsigned long long peer_timeout_ns = 600ull * 1000 * 1000 * 1000;  // DPK_OPT_PEER_TIMEOUT_MS

  // instrumentation
  bool timing = false;
  std::vector<cudaEvent_t> tev[3];  // [which] -> begin,end,begin,end,...
  int tcount[3] = {0, 0, 0};
  long long launches = 0;
};

// Optional CUDA-event bracket around one streaming kernel (bench instrumentation).
namespace dpk {
struct KernelTimer {
  dpk_handle* h;
  int which;
  cudaStream_t st;
  cudaEvent_t end = nullptr;
  KernelTimer(dpk_handle* h_, int which_, cudaStream_t st_) : h(h_), which(which_), st(st_) {
    if (!h->timing || h->tcount[which] >= DPK_TIMING_SLOTS) return;
    auto& v = h->tev[which];
    const size_t need = 2 * (size_t)(h->tcount[which] + 1);
    while (v.size() < need) {
      cudaEvent_t e = nullptr;
      if (cudaEventCreate(&e) != cudaSuccess) return;
      v.push_back(e);
    }
    if (cudaEventRecord(v[need - 2], st) != cudaSuccess) return;  // no begin mark: this launch stays untimed
    end = v[need - 1];
  }
  void stop() {
    if (end) {
      if (cudaEventRecord(end, st) == cudaSuccess) h->tcount[which]++;  // a pair counts only if both marks exist
      end = nullptr;
    }
  }
  ~KernelTimer() { stop(); }
};
}  // namespace dpk
