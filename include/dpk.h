/*
 * dpk.h -- C ABI of libdpk, the B200 (sm_100a) data-parallel post-backward
 * update path: bucketed gradient all-reduce, loss-scale unscale, global L2
 * grad-norm + clip, AdamW and every Polyak-EMA rate in one fused pass, plus
 * the loss-aware timestep sampler's gather/update.
 *
 * The reference (YAIconvalley/distributed-pipeline) is pure Python and has no
 * FFI of its own; the seam this library sits under is the method surface of
 * `TrainLoop` (utils/trainer.py) and the torch objects it holds.  Each entry
 * point below names the reference lines whose device work it replaces.
 *
 * Conventions
 *   - plain C types only; every pointer named `d_*` or documented "device" is
 *     a CUDA device pointer, everything else is host memory.
 *   - every function returns int: 0 ok, >0 a cudaError_t, <0 a DPK_E_* code.
 *     A human-readable message for the calling thread is kept in
 *     dpk_last_error().  Nothing aborts, nothing throws.
 *   - `stream` arguments are `cudaStream_t` passed as void* (NULL = legacy
 *     default stream).  Calls only enqueue work; the only entry points that
 *     block the host are the ones whose name ends in `_sync` and
 *     dpk_create / dpk_destroy / dpk_comm_init.
 *   - the library never allocates or frees the state buffers: parameters,
 *     gradients, moments and EMA copies stay owned by the caller (torch), so
 *     state_dict / deepcopy / torch.save keep working.
 */
#ifndef DPK_H_
#define DPK_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DPK_ABI_VERSION 2

/* error codes (<0); >0 is a cudaError_t */
#define DPK_E_BADARG  (-1)  /* null / misaligned / out-of-range argument      */
#define DPK_E_NCCL    (-2)  /* an NCCL call failed (text in dpk_last_error)   */
#define DPK_E_STATE   (-3)  /* call made in the wrong order / not bound       */
#define DPK_E_NOMEM   (-4)  /* host allocation failed                         */
#define DPK_E_LIMIT   (-5)  /* size beyond a documented maximum               */

/* element type of params, grads, both moments and all EMA copies.  The
 * reference keeps no fp32 master copy (utils/trainer.py:91-92), so all state
 * shares the model's dtype. */
#define DPK_F32  0
#define DPK_BF16 1

#define DPK_MAX_EMA 4
#define DPK_MAX_RANKS 64   /* ranks a symmetric-memory binding may name */

/* what the fused pass leaves in the gradient buffer */
#define DPK_GRAD_KEEP     0  /* untouched (fastest)                               */
#define DPK_GRAD_CLIPPED  1  /* g*inv_scale*coef written back, as the reference's
                                clip_grad_norm_ leaves them (utils/trainer.py:252) */
#define DPK_GRAD_ZERO     2  /* zeroed: folds TrainLoop._zero_grad (:203-207)     */

/* index into the result block of dpk_grad_norm / dpk_fused_update (doubles) */
#define DPK_R_TOTAL_NORM   0  /* pre-clip total norm (norm of per-tensor norms)     */
#define DPK_R_CLIP_COEF    1  /* min(max_norm/(total+1e-6), 1), or 1 if clip off    */
#define DPK_R_LOGGED_NORM  2  /* sqrt(sum_i sum(g_i^2)) on the grads the step used:
                                 what TrainLoop._log_grad_norm logs (:265-271)      */
#define DPK_R_NONFINITE    3  /* 1.0 if any gradient element was inf/nan            */
#define DPK_R_SKIPPED      4  /* 1.0 if the last fused update predicated itself off */
#define DPK_R_COMM_TIMEOUT 5  /* set to 1.0 (and left there) if one of the library's own
                                 all-reduce kernels gave up waiting for a peer rank:
                                 the gradients of that step are not reduced, and every
                                 later fused update is a no-op (state stays intact) */
#define DPK_R_COUNT        8

typedef struct dpk_handle dpk_handle;
typedef struct dpk_sampler dpk_sampler;

typedef struct dpk_config {
  int32_t abi_version;      /* DPK_ABI_VERSION                                  */
  int32_t dtype;            /* DPK_F32 | DPK_BF16                               */
  int32_t device;           /* CUDA ordinal the buffers live on                 */
  int32_t n_segments;       /* parameter tensors (209 for DiffuSeq-base)        */
  int32_t n_ema;            /* EMA rates, 0..DPK_MAX_EMA                        */
  int32_t tile_elems;       /* 0 = default; elements per work tile, mult. of 1024 */
  int64_t n_elems;          /* flat length in elements, padding included        */
  const int64_t* seg_offsets; /* [n_segments] element offset of each tensor in the
                                 flat buffers; each a multiple of 16 bytes      */
  const int64_t* seg_numel; /* [n_segments] elements per tensor (may be 0)      */
  const double* ema_rates;  /* [n_ema] as parsed by TrainLoop (:72-76)          */
  double beta1, beta2, eps; /* AdamW defaults of the reference: .9, .999, 1e-8  */
  int64_t bucket_cap_bytes; /* all-reduce bucket size; reference: 128 MiB (:123) */
  int64_t edge_bucket_bytes; /* size of the FIRST bucket (the tail of the buffer: its
                                all-reduce starts with nothing to run beside it yet)
                                and of the LAST one (the head: nothing is left to hide
                                its consumer behind).  Both are kept small; torch DDP
                                does the same for its first bucket.  0 = cap/8;
                                >= cap: plain cap-sized buckets.  Results do not
                                depend on the bucket plan. */
  const int32_t* seg_frozen; /* [n_segments] or NULL: non-zero = this tensor has no
                                gradient (requires_grad=False).  As with
                                torch.optim.AdamW it then takes no step at all (no
                                weight decay, no moments) and stays out of the norm;
                                its EMA copies are still advanced, as update_ema
                                (utils/trainer.py:360-370) walks every parameter. */
} dpk_config;

/* ---- library ---------------------------------------------------------- */

int dpk_abi_version(void);
const char* dpk_last_error(void);

/* ---- handle: replaces what TrainLoop.__init__ builds (utils/trainer.py:87-128:
 * AdamW state, K EMA lists, the DDP reducer's bucket plan) ---------------- */

/* Validates cfg and plans tiles and buckets.  Pure host work: no CUDA call is
 * made until dpk_bind, so planning can be inspected on a machine without a GPU. */
int dpk_create(dpk_handle** out, const dpk_config* cfg);
int dpk_destroy(dpk_handle* h);

/* Borrow the caller's flat device buffers (each n_elems long, 16-byte aligned,
 * same dtype).  d_emas has cfg.n_ema entries.  Gaps between tensors are never
 * read as data nor modified.  The first call also allocates the library's own
 * small device scratch (tile table, per-tile partial sums, result block). */
int dpk_bind(dpk_handle* h, void* d_params, void* d_grads, void* d_exp_avg,
             void* d_exp_avg_sq, void* const* d_emas);

/* ---- communicator: the process group itself is still created by
 * basic_utils/dist_util.py:57-85; the caller ships the 128-byte id from rank 0
 * to every rank through that group and then calls dpk_comm_init. ----------- */

int dpk_comm_unique_id(void* out128);
int dpk_comm_init(dpk_handle* h, const void* id128, int32_t world_size, int32_t rank);
int dpk_world_size(const dpk_handle* h);

/* ---- gradient all-reduce: replaces torch DDP's reducer as configured at
 * utils/trainer.py:118-125 (bucket_cap_mb=128; average = pre-multiply by 1/W
 * then SUM), in place on the flat gradient buffer. -------------- */

/* Optional: switch the bucket all-reduce from NCCL to the library's own
 * kernels over symmetric memory: for 4 or more ranks the NVSwitch reduces
 * (multimem.ld_reduce / multimem.st), two ranks exchange halves directly.  The
 * caller maps the gradient buffer as symmetric memory on every rank and passes
 *   mc_grads      multicast address of the gradient buffer,
 *   peer_grads    [world] this process's mapping of every rank's buffer
 *                 (peer_grads[rank] is the address given to dpk_bind),
 *   mc_signal     multicast address of a zero-initialised uint32[n_slots] array
 *                 that is symmetric too, local_signal its local address
 *                 (n_slots >= 2; with >= 1024 slots the array also carries the
 *                 per-range arrival counters dpk_reduce_update needs at W >= 4:
 *                 65 536 slots serve any buffer size).
 * Needs a power-of-two world size (the average is taken as sum * 1/W).  No
 * reference counterpart beyond the DDP reducer both paths replace. */
int dpk_bind_nvls(dpk_handle* h, void* mc_grads, void* const* peer_grads, void* mc_signal,
                  void* local_signal, int32_t n_slots);

int dpk_num_buckets(const dpk_handle* h);
/* bucket b covers flat elements [*start, *start + *count); bucket 0 is the
 * tail of the buffer (gradients become ready in reverse definition order). */
int dpk_bucket_range(const dpk_handle* h, int32_t b, int64_t* start, int64_t* count);
/* All gradients of bucket b have been produced on `producer_stream`: enqueue
 * its all-reduce on the library's communication stream.  Thread-safe (it is
 * called from autograd's engine thread). */
int dpk_bucket_ready(dpk_handle* h, int32_t b, void* producer_stream);
/* Every gradient of every bucket has been produced on `producer_stream` (a
 * post-backward update with nothing left to overlap): enqueue all buckets in
 * index order.  Same as dpk_bucket_ready for b = 0..n-1, except that the
 * library's own all-reduce kernels then need their "all ranks' gradients are
 * complete" barrier only once, for the first bucket. */
int dpk_all_buckets_ready(dpk_handle* h, void* producer_stream);
/* Make `consumer_stream` wait for bucket b's all-reduce only (no-op if none is
 * outstanding): lets the caller consume buckets as they land while later ones
 * are still on the wire. */
int dpk_wait_bucket(dpk_handle* h, int32_t b, void* consumer_stream);
/* Make `consumer_stream` wait for every all-reduce enqueued so far. */
int dpk_wait_allreduce(dpk_handle* h, void* consumer_stream);
/* One coalesced broadcast of params (+ every EMA copy if include_ema) from
 * `root`: replaces DDP's constructor broadcast and the per-tensor loop of
 * basic_utils/dist_util.py:141-152. */
int dpk_broadcast_state(dpk_handle* h, int32_t root, int32_t include_ema, void* stream);

/* ---- the update ------------------------------------------------------- */

/* TrainLoop._zero_grad (utils/trainer.py:203-207) as one fill. */
int dpk_zero_grads(dpk_handle* h, void* stream);

/* Pass 1: TrainLoop.grad_clip -> torch.nn.utils.clip_grad_norm_
 * (utils/trainer.py:246-255; torch nn/utils/clip_grad.py:92-108,165-174) and the
 * non-finite test of the upstream loss-scale path.  Reads g once; gradients are
 * taken as g*inv_scale.  max_norm <= 0 disables clipping (coef = 1).  Results
 * land in the device result block; nothing is written to g. */
int dpk_grad_norm(dpk_handle* h, double inv_scale, double max_norm, void* stream);

/* Pass 2: TrainLoop.optimize after the clip (utils/trainer.py:240-244):
 * AdamW.step() (torch optim/adam.py:683-800) and update_ema for every rate
 * (utils/trainer.py:360-370), one read and one write per state element.
 *   lr            this step's learning rate (after _anneal_lr, :257-263)
 *   adam_step     AdamW's step count for this update (1 on the first call)
 *   use_clip_coef 1: scale g by the coefficient the last dpk_grad_norm left in
 *                 the result block; 0: coefficient 1
 *   skip_if_nonfinite 1: do nothing (and set R_SKIPPED) when the last
 *                 dpk_grad_norm saw inf/nan -- the loss-scale "skip step"
 *   grad_mode     DPK_GRAD_*
 * Also accumulates sum(g^2) of the gradients it used, so R_LOGGED_NORM is
 * valid afterwards even when dpk_grad_norm was not called. */
int dpk_fused_update(dpk_handle* h, double lr, double weight_decay, int64_t adam_step,
                     double inv_scale, int32_t use_clip_coef, int32_t skip_if_nonfinite,
                     int32_t grad_mode, void* stream);

/* The two passes cut along the all-reduce buckets, so that a bucket can be
 * consumed the moment its all-reduce lands while the next one is still in
 * flight (torch DDP overlaps its reducer with backward in the same spirit,
 * utils/trainer.py:118-125; here the update itself overlaps the reducer):
 *   dpk_grad_norm    == dpk_grad_norm_bucket(-1)  + dpk_grad_norm_finish
 *   dpk_fused_update == dpk_fused_update_bucket(-1) + dpk_fused_update_finish
 * bucket = -1 means the whole buffer.  Every bucket must be processed exactly
 * once before the matching _finish; results are identical to the one-call forms. */
int dpk_grad_norm_bucket(dpk_handle* h, int32_t bucket, double inv_scale, void* stream);
int dpk_grad_norm_finish(dpk_handle* h, double max_norm, void* stream);
int dpk_fused_update_bucket(dpk_handle* h, int32_t bucket, double lr, double weight_decay,
                            int64_t adam_step, double inv_scale, int32_t use_clip_coef,
                            int32_t skip_if_nonfinite, int32_t grad_mode, void* stream);
int dpk_fused_update_finish(dpk_handle* h, int32_t skip_if_nonfinite, void* stream);

/* Tuning knobs (never change results).
 *   DPK_OPT_OVERLAP_BLOCKS_PER_SM  grid of a bucket-granular update launch, which
 *       runs beside the all-reduce of later buckets: 0 = one short-lived block
 *       per 4 KiB chunk; n > 0 = a persistent grid of n blocks per SM, i.e. a
 *       deliberately throttled update that leaves HBM bandwidth to the
 *       collective when the step is NVLink-bound.
 *   DPK_OPT_NORM_STAGING  1 = the norm pass stages its tiles through shared
 *       memory with TMA bulk copies (cp.async.bulk + mbarrier, 3 deep) instead of
 *       128-bit register loads.  Same partial sums, bit for bit.
 *   DPK_OPT_PEER_TIMEOUT_MS  how long one of the library's own all-reduce kernels
 *       waits for a peer rank before it gives up and raises DPK_R_COMM_TIMEOUT
 *       (default 600 000 = ten minutes, the NCCL watchdog's default; the
 *       environment variable DPK_PEER_TIMEOUT_MS sets the initial value).  After
 *       a timeout the fused update leaves parameters, moments and EMA copies
 *       untouched, so the caller can raise with the state intact. */
#define DPK_OPT_OVERLAP_BLOCKS_PER_SM 1
#define DPK_OPT_NORM_STAGING 2
#define DPK_OPT_PEER_TIMEOUT_MS 3
int dpk_set_option(dpk_handle* h, int32_t option, int64_t value);

/* Two ranks only (after dpk_bind_nvls): gradient all-reduce AND the fused update
 * as one kernel.  Each gradient element is read from this rank's buffer and, over
 * NVLink, from the peer's, averaged in registers ((a+b)/2, one rounding: what
 * ncclAvg / DDP give) and consumed by AdamW + EMA on the spot, so the exchange
 * overlaps the update chunk by chunk and the averaged gradient never travels
 * through HBM.  Replaces dpk_bucket_ready (all buckets) + dpk_wait_allreduce +
 * dpk_fused_update(grad_mode KEEP, no clip) with identical state afterwards; the
 * gradient buffers themselves stay local (not averaged).  DPK_R_LOGGED_NORM is
 * the norm of the averaged gradient. */
int dpk_reduce_update_pair(dpk_handle* h, double lr, double weight_decay, int64_t adam_step,
                           double inv_scale, void* stream);

/* Any number of ranks >= 2 (after dpk_bind_nvls with a signal array of at least
 * 1024 slots): the same replacement as dpk_reduce_update_pair -- DDP's bucketed
 * all-reduce (utils/trainer.py:118-125) plus optimize() without clipping
 * (:240-244) as ONE kernel.  Two ranks take the pair form.  Four or more: part of
 * the grid reduces through the NVSwitch (multimem.ld_reduce, x 1/W, multimem.st:
 * the arithmetic of the bucket all-reduce, so the averaged gradient is bit-identical
 * and is what the gradient buffers hold afterwards) and publishes a per-range
 * arrival counter; the rest of the grid runs the fused update, each thread waiting
 * only for the range its vector lies in.  State afterwards is bit-identical to
 * dpk_all_buckets_ready + dpk_wait_allreduce + dpk_fused_update(KEEP, no clip).
 * If a peer stops arriving mid-kernel the step is abandoned where it stands and
 * DPK_R_COMM_TIMEOUT is raised (parameters are then partly updated: restart from
 * the last checkpoint); a peer that never shows up at all leaves the state intact. */
int dpk_reduce_update(dpk_handle* h, double lr, double weight_decay, int64_t adam_step,
                      double inv_scale, void* stream);

/* The clipping counterpart: DDP's all-reduce (utils/trainer.py:118-125) and pass 1
 * (grad_clip, :246-255) as one kernel for four or more ranks -- the same grid
 * split, the second part taking the per-tile sums of squares range by range as the
 * averaged gradient arrives; two ranks run the pair all-reduce with the norm pass
 * bucket by bucket behind it.  Result block and gradient buffers afterwards are
 * exactly those of dpk_all_buckets_ready + dpk_wait_allreduce + dpk_grad_norm;
 * dpk_fused_update(use_clip_coef = 1) follows as usual. */
int dpk_reduce_norm(dpk_handle* h, double inv_scale, double max_norm, void* stream);

/* ---- sharded update (opt-in).  The reference itself points at it: its grad_clip
 * keeps a branch for "the sharded optimizer" (utils/trainer.py:248-250).  Instead of
 * all-reducing the gradient and updating W identical replicas, each rank owns every
 * W-th 4 KiB chunk: it pulls the switch-reduced gradient of its chunks
 * (reduce-scatter), runs AdamW + EMA on them from registers and multicasts the new
 * parameters into every replica (all-gather) -- one kernel, the NVLink volume of one
 * all-reduce, 1/W of the update's HBM traffic.  Parameters afterwards are
 * bit-identical to the replica path; exp_avg, exp_avg_sq and the EMA copies are
 * valid on the owner of each chunk only, until dpk_shard_gather makes every rank's
 * copies complete again (call it before state_dict / save / reading EMA weights).
 * No clipping in this mode.  Needs dpk_bind_nvls plus the multicast address of the
 * (symmetric) parameter buffer.  dpk_shard_update and dpk_shard_gather issue small
 * NCCL collectives on the CALLER's stream: do not run them while bucket all-reduces
 * of the same handle are still in flight on the library's communication stream (one
 * communicator serves one stream at a time) -- in this mode there are none, the
 * gradient exchange lives inside dpk_shard_update. */
int dpk_bind_shard(dpk_handle* h, void* mc_params);
int dpk_shard_update(dpk_handle* h, double lr, double weight_decay, int64_t adam_step,
                     double inv_scale, void* stream);
int dpk_shard_gather(dpk_handle* h, void* stream);

/* Device address of the result block (DPK_R_COUNT doubles). */
int dpk_result_ptr(dpk_handle* h, void** d_out);
/* Copy the result block to `out[DPK_R_COUNT]` and wait for it: the single host
 * read that replaces the 209 .item() calls of TrainLoop._log_grad_norm. */
int dpk_read_result_sync(dpk_handle* h, double* out, void* stream);
/* Per-tensor L2 norms of the last dpk_grad_norm (n_segments doubles, device). */
int dpk_segment_norms_ptr(dpk_handle* h, void** d_out);

/* Stand-alone update_ema(target, source, rate) (utils/trainer.py:360-370) on
 * caller-supplied contiguous device arrays of `n` elements, for callers that
 * use the module-level function outside optimize(). */
int dpk_ema_update(void* d_target, const void* d_source, int64_t n, int32_t dtype,
                   double rate, void* stream);

/* ---- allocation-free norm.  No counterpart in the reference: it exists for
 * callers of THIS library that cannot allocate device memory -- code running
 * mid-step on a GPU whose free memory already belongs to the framework's caching
 * allocator, where a raw cudaMalloc (which dpk_create/dpk_bind perform) may fail
 * outright, fragments what is left and synchronises the whole device.
 *
 * Computes what dpk_grad_norm computes for a layout of ONE tensor of n elements
 * (same arithmetic, same reduction order, same bits): the L2 norm of
 * d_grads*inv_scale, the clip coefficient for max_norm (<= 0: coefficient 1) and
 * the non-finite flag, written to d_out[DPK_R_TOTAL_NORM], d_out[DPK_R_CLIP_COEF]
 * and d_out[DPK_R_NONFINITE]; d_out is a caller-owned device array of
 * DPK_R_COUNT doubles, its other entries are left alone.  d_grads must be
 * 16-byte aligned; dtype is DPK_F32 or DPK_BF16.  Runs on the current device.
 *
 * The scratch (per-tile partial sums) is a fixed array inside the library's
 * device image, 512 KiB per GPU for the life of the process; hence the limit:
 * n > DPK_STATIC_MAX_ELEMS returns DPK_E_LIMIT.  The call makes no cudaMalloc /
 * cudaFree and uses no allocating library routine.  The first call on each
 * device resolves the scratch address (cudaGetSymbolAddress): make that call
 * before starting any stream capture.  All calls on a device share the one
 * scratch, which still holds the previous call's values when the next begins
 * (every word a call reads it has first rewritten, so none needs clearing);
 * two calls whose kernels overlap on the GPU -- different streams without an
 * ordering between them -- race on it and return garbage. */
#define DPK_STATIC_MAX_ELEMS 536870912
int dpk_grad_norm_static(const void* d_grads, int64_t n, int32_t dtype, double inv_scale,
                         double max_norm, double* d_out, void* stream);

/* ---- instrumentation (bench.py's roofline line; nothing in the reference
 * corresponds: its logger.profile_kv, basic_utils/logger.py:296-320, times the
 * host only).  With timing enabled every dpk_fused_update / dpk_grad_norm
 * brackets its streaming kernel -- that kernel alone, not the small reductions
 * behind it -- with CUDA events on the caller's stream, up to DPK_TIMING_SLOTS
 * launches between two reads. ---------------------------------------------- */

#define DPK_TIMING_SLOTS 256
int dpk_timing_enable(dpk_handle* h, int32_t on);
/* Waits for the recorded events, writes up to `cap` durations (ms, oldest
 * first) for kernel `which` (0 = fused update, or the fused all-reduce + update
 * kernel; 1 = sum-of-squares; 2 = one bucket's all-reduce on the communication
 * stream) and clears them. */
int dpk_timing_read_sync(dpk_handle* h, int32_t which, float* ms_out, int32_t cap, int32_t* n_out);
/* Kernels this handle has launched since creation. */
int dpk_launch_count(const dpk_handle* h, int64_t* out);
/* Exhaustive device-side check of the fused kernel's two arithmetic shortcuts
 * against the IEEE operations they stand in for: d_out2[0] = number of fp32 bit
 * patterns v (all 2^32) for which the reciprocal-and-correct division of
 * sqrt(v) -- the only dividends the kernel has -- by `divisor` differs from
 * div.rn; d_out2[1] = number of bf16 bit patterns whose
 * approximate square root rounds to a different bf16 than sqrt.rn's, plus the
 * number of fp32 bit patterns (all 2^32) the kernel's packed fp32->bf16 rounding
 * converts differently from cvt.rn.bf16.f32, plus the number of bf16 patterns its
 * bf16->fp32 unpacking does not reproduce exactly.  Both must be 0.
 * d_out2: two device uint64. */
int dpk_selftest_arith(double divisor, unsigned long long* d_out2, void* stream);

/* ---- loss-aware timestep sampler.  Absent from the reference;
 * semantics are upstream improved-diffusion resample.py
 * (LossSecondMomentResampler + LossAwareSampler.update_with_local_losses). -- */

int dpk_sampler_create(dpk_sampler** out, int32_t device, int32_t n_timesteps,
                       int32_t history_per_term, double uniform_prob, int32_t max_local_batch);
int dpk_sampler_destroy(dpk_sampler* s);
/* Gather every rank's (timestep, loss) records (rank-major, batch order) over
 * the handle's communicator (h may be NULL or single-rank: local only) and
 * apply update_with_all_losses.  d_ts int64[n_local], d_losses float[n_local].
 * Asynchronous, except that the first call (and the first one with a larger
 * communicator) synchronises `stream` once to size the gather buffer. */
int dpk_sampler_update(dpk_sampler* s, dpk_handle* h, const int64_t* d_ts,
                       const float* d_losses, int32_t n_local, void* stream);
/* weights() into a device array of n_timesteps doubles. */
int dpk_sampler_weights(dpk_sampler* s, double* d_weights, void* stream);
/* Device views of the state, for checkpointing and tests. */
int dpk_sampler_state_ptrs(dpk_sampler* s, void** d_history, void** d_counts);

#ifdef __cplusplus
}
#endif
#endif /* DPK_H_ */
