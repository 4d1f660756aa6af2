"""Drop-in ``TrainLoop`` whose post-backward update runs on libdpk.

Method surface, constructor keywords, attributes and checkpoint file names are
those of the reference's ``utils/trainer.py`` (cited per method) so that
``run/train.py:103-121`` and user subclasses work unchanged.  What is different
is what happens underneath ``forward_backward()`` and ``optimize()``:

===========================  ==================================================
reference                    here
===========================  ==================================================
209 ``grad.zero_()``         one fill of the flat gradient buffer
torch DDP reducer            in-place ``ncclAllReduce`` per flat bucket, launched
                             from grad hooks on a side stream
``clip_grad_norm_``          pass 1 kernel, coefficient stays on the device
209 ``.item()`` syncs        one asynchronous 64-byte read-back per step
foreach AdamW + 2K EMA ops   one fused kernel (pass 2)
===========================  ==================================================

There is no CPU path: constructing a ``TrainLoop`` around a model that is not
on a CUDA device raises.
"""
import os
import weakref
from typing import Dict, List

import torch

from . import _lib, dist_util
from . import logger as _default_logger
from .engine import ema_update as _ema_update_kernel
from .optim import FusedAdamW
from .reducer import BucketReducer, FlatDDP

INITIAL_LOG_LOSS_SCALE = 20.0  # upstream fp16 path (absent from the reference)

_GRAD_MODES = {"keep": _lib.GRAD_KEEP, "clipped": _lib.GRAD_CLIPPED, "zero": _lib.GRAD_ZERO}

# id(list of EMA tensors) -> (weakref to FusedAdamW, rate index); lets the
# module-level update_ema recognise lists the fused step already advanced.
_FUSED_EMA_LISTS: Dict[int, tuple] = {}


class TrainLoop:

    # ---- user hooks (reference utils/trainer.py:19-43) ------------------------
    def log_loss_dict(self, mode, losses, *args, **kwargs):
        raise NotImplementedError

    def compute_losses(self, micro_batch):
        raise NotImplementedError

    @staticmethod
    def backward_from_losses(losses):
        raise NotImplementedError("override backward_from_losses: reduce `losses` to a scalar and call .backward()")

    @classmethod
    def get_batch_length(cls, batch):
        if isinstance(batch, torch.Tensor):
            return batch.shape[0]
        if isinstance(batch, dict):
            return cls.get_batch_length(next(iter(batch.values())))
        if isinstance(batch, list):
            return cls.get_batch_length(batch[0])
        raise TypeError("Unsupported batch type: {}".format(type(batch).__name__))

    # ---- construction (reference :45-132) -------------------------------------
    def __init__(
            self,
            *,
            model,
            data,
            batch_size,
            microbatch,
            lr,
            ema_rate,
            log_interval,
            save_interval,
            resume_checkpoint,
            weight_decay=0.0,
            learning_steps=0,
            checkpoint_path='',
            gradient_clipping=-1.,
            eval_data=None,
            eval_interval=-1,
            eval_callbacks=(),
            # --- additions (all optional) ---
            grads_after_step="clipped",   # "clipped" = what the reference leaves in .grad; "keep"; "zero"
            use_fp16=False,               # upstream loss-scale state machine
            fp16_scale_growth=1e-3,
            logger=None,                  # module/object with logkv, logkv_mean, dumpkvs, log, get_current
            grad_allreduce="auto",        # "nvls": the library's symmetric-memory kernels; "nccl"; "auto": nvls if possible
            schedule_sampler=None,        # upstream's timestep sampler: None | "uniform" | "lossaware" | a ScheduleSampler
            diffusion_steps=None,         # number of timesteps, needed when the sampler is given by name
            timestep_key="t",             # micro_batch entry the sampled timesteps are placed under
    ):
        self.logger = logger if logger is not None else _default_logger
        self.model = model
        self.data = data
        self.eval_data = eval_data
        self.batch_size = batch_size
        self.microbatch = microbatch if microbatch > 0 else batch_size
        self.lr = float(lr)
        self.ema_rate = [ema_rate] if isinstance(ema_rate, float) else [float(x) for x in ema_rate.split(",")]
        self.log_interval = log_interval
        self.eval_interval = eval_interval
        self.save_interval = save_interval
        self.resume_checkpoint = resume_checkpoint
        self.weight_decay = weight_decay
        self.learning_steps = learning_steps
        self.gradient_clipping = gradient_clipping
        self.use_fp16 = bool(use_fp16)
        self.fp16_scale_growth = fp16_scale_growth
        self.lg_loss_scale = INITIAL_LOG_LOSS_SCALE
        if grads_after_step not in _GRAD_MODES:
            raise ValueError(f"grads_after_step must be one of {sorted(_GRAD_MODES)}")
        if grad_allreduce not in ("auto", "nvls", "nccl"):
            raise ValueError("grad_allreduce must be 'auto', 'nvls' or 'nccl'")
        world = dist_util.get_world_size()
        want_symmetric = (dist_util.is_initialized() and world > 1 and grad_allreduce != "nccl"
                          and (world & (world - 1)) == 0)

        self.step = 0
        self.resume_step = 0
        self.global_batch = self.batch_size * dist_util.get_world_size()
        self.eval_callbacks = list(eval_callbacks)
        self.checkpoint_path = checkpoint_path

        self.model_params = list(self.model.parameters())
        self.master_params = self.model_params  # no fp32 master copy, as the reference (:91-92)
        if not self.model_params or self.model_params[0].device.type != "cuda":
            raise RuntimeError("TrainLoop needs the model on a CUDA device: the update path is CUDA-only "
                               "(move it with model.to(dist_util.dev()) as run/train.py:75 does)")

        self._load_and_sync_parameters()

        # From here on parameters, gradients, moments and EMA copies are views
        # of flat buffers; the lists above keep pointing at the same Parameters.
        self.opt = FusedAdamW(self.master_params, lr=self.lr, weight_decay=self.weight_decay,
                              ema_rates=self.ema_rate, bucket_cap_mb=128,
                              grad_mode=_GRAD_MODES[grads_after_step], symmetric_grads=want_symmetric)
        self.ema_params = self.opt.flat.ema_views  # K lists of per-parameter tensors (:111-113)
        for k, lst in enumerate(self.ema_params):
            _FUSED_EMA_LISTS[id(lst)] = (weakref.ref(self.opt), k)
        if self.resume_step:
            self._load_optimizer_state()
            for k, rate in enumerate(self.ema_rate):
                self._load_ema_parameters(rate, k)

        if dist_util.is_initialized():
            self.use_ddp = True
            # DDP's constructor broadcast (rank 0's parameters win), one coalesced
            # collective; EMA copies follow the broadcast parameters on every rank
            # (the reference leaves rank != 0 EMA copies at their local init, F4).
            self.opt.engine.init_comm()
            self.opt.engine.broadcast_state(root=0, include_ema=True)
            if want_symmetric:
                # the gradient all-reduce runs on the library's own kernels over
                # symmetric memory; NCCL stays the path when the NVSwitch multicast
                # mapping cannot be set up (and "nvls" was not demanded)
                # (decided by all ranks together: see Engine.enable_nvls_agreed)
                enabled, why = self.opt.engine.enable_nvls_agreed(self.opt.flat)
                if not enabled:
                    if grad_allreduce == "nvls":
                        raise RuntimeError(f"grad_allreduce='nvls' was demanded but is unavailable: {why}")
                    self.logger.log(f"symmetric-memory all-reduce unavailable ({why}); using NCCL")
            self._reducer = BucketReducer(self.opt.flat, self.opt.engine)
            self.ddp_model = FlatDDP(self.model, self._reducer)
            # Two ranks, nobody needs the averaged gradient afterwards, no clipping: the
            # exchange moves out of backward and into the update kernel itself
            # (dpk_reduce_update_pair); .grad then keeps this rank's local gradient.
            self._pair_fused = (self.opt.engine.nvls and world == 2 and grads_after_step == "keep"
                                and not self.use_fp16 and not self.gradient_clipping > 0)
            if self._pair_fused:
                self._reducer.enabled = False
        else:
            self.use_ddp = False
            self._reducer = None
            self._pair_fused = False
            self.ddp_model = self.model

        # Upstream DiffuSeq's schedule sampler (train_util.py; stripped from the reference at
        # utils/trainer.py:230-235): timesteps are drawn per microbatch, the loss-aware
        # kind is fed every rank's (t, loss) records, and losses are weighted by 1/(T p(t)).
        self.timestep_key = timestep_key
        self.schedule_sampler = self._make_sampler(schedule_sampler, diffusion_steps)
        self._loss_log = _DeferredLossLog(self.logger)

        self._in_optimize = False
        self._norm_requested = False
        self._norm_inflight: List[tuple] = []
        # pinned landing slots for the per-step result read-back (reused round-robin:
        # allocating pinned memory every step would cost more than the update)
        self._norm_ring = [(torch.empty(_lib.R_COUNT, dtype=torch.float64, pin_memory=True), torch.cuda.Event())
                           for _ in range(16)]
        self._norm_ring_pos = 0
        self._grads_are_zero = True  # fresh buffer
        # any forward that records a graph may be followed by a backward, whoever calls it (a subclass
        # with its own forward_backward included): from then on the buffer is no longer known to be zero
        self.model.register_forward_pre_hook(self._mark_grads_dirty)
        torch.cuda.empty_cache()

    # ---- resume (reference :136-173) -----------------------------------------
    def _load_and_sync_parameters(self):
        resume_checkpoint = self.find_resume_checkpoint(self.logger) or self.resume_checkpoint
        if not resume_checkpoint:
            return
        self.resume_step = self.parse_resume_step_from_filename(resume_checkpoint)
        if dist_util.get_rank() == 0:
            self.logger.log(f"loading model from checkpoint: {resume_checkpoint}...")
            self.model.load_state_dict(dist_util.load_state_dict(resume_checkpoint, map_location=dist_util.dev()))
        dist_util.sync_params(self.model.parameters())

    def _load_ema_parameters(self, rate, k=None):
        if k is None:
            k = self.ema_rate.index(rate)
        views = self.ema_params[k]
        main_checkpoint = self.find_resume_checkpoint(self.logger) or self.resume_checkpoint
        if not main_checkpoint:
            return views
        ema_checkpoint = self.find_ema_checkpoint(main_checkpoint, self.resume_step, rate)
        if ema_checkpoint and dist_util.get_rank() == 0:
            self.logger.log(f"loading EMA from checkpoint: {ema_checkpoint}...")
            state_dict = dist_util.load_state_dict(ema_checkpoint, map_location=dist_util.dev())
            with torch.no_grad():
                for dst, src in zip(views, self._state_dict_to_master_params(state_dict)):
                    dst.copy_(src)
        dist_util.sync_params(views)
        return views

    def _load_optimizer_state(self):
        main_checkpoint = self.find_resume_checkpoint(self.logger) or self.resume_checkpoint
        if not main_checkpoint:
            return
        opt_checkpoint = self.find_opt_checkpoint(main_checkpoint, self.resume_step)
        if opt_checkpoint and os.path.exists(opt_checkpoint):
            self.logger.log(f"loading optimizer state from checkpoint: {opt_checkpoint}")
            self.opt.load_state_dict(dist_util.load_state_dict(opt_checkpoint, map_location=dist_util.dev()))

    # ---- the loop (reference :175-201) ---------------------------------------
    def _steps_remain(self):
        return not self.learning_steps or self.step + self.resume_step < self.learning_steps

    def _periodic_work(self):
        """What the reference does between steps (utils/trainer.py:182-193), in its order:
        dump the logger, evaluate, checkpoint."""
        if self.step % self.log_interval == 0:
            self.flush_logs()
            self.logger.dumpkvs()
        if self.eval_data is not None and self.step % self.eval_interval == 0:
            self.forward_only(next(self.eval_data))
            print('eval on validation set')
            self.logger.dumpkvs()
            if dist_util.get_rank() == 0:
                for callback in self.eval_callbacks:
                    callback(self)
        if self.step > 0 and self.step % self.save_interval == 0:
            self.save()

    def run_loop(self):
        while self._steps_remain():
            self.run_step(next(self.data))
            self._periodic_work()
            self.step += 1
        last_done = self.step - 1
        if last_done % self.save_interval != 0:   # not already saved by the loop
            self.save()

    __call__ = run_loop

    def run_step(self, batch):
        self.forward_backward(batch)
        self.optimize()
        self.log_step()

    def _mark_grads_dirty(self, *_):
        if torch.is_grad_enabled():
            self._grads_are_zero = False

    def _make_sampler(self, spec, diffusion_steps):
        if spec is None:
            return None
        if isinstance(spec, str):
            from .sampler import create_named_schedule_sampler
            if diffusion_steps is None:
                raise ValueError("schedule_sampler given by name needs diffusion_steps (the number of timesteps)")
            kw = {}
            if spec == "lossaware":
                kw = dict(device=dist_util.dev(), engine=self.opt.engine if self.use_ddp else None,
                          max_local_batch=max(int(self.microbatch), 1))
            return create_named_schedule_sampler(spec, diffusion_steps, **kw)
        return spec

    def log_loss_dict_deferred(self, losses, ts=None, num_timesteps=None, prefix=""):
        """What upstream's ``log_loss_dict`` logs -- the mean of every entry and, given the timesteps, the
        per-quartile means -- without its ``.item()`` per value: the few sums are reduced on the device,
        copied to pinned memory asynchronously and handed to ``logger.logkv_mean`` when the logs are
        flushed (``log_interval``), so a subclass can call this from its ``log_loss_dict`` hook and keep
        the step free of host syncs.  The logged numbers are the same running means."""
        self._loss_log.add(losses, ts, num_timesteps, prefix)

    def _zero_grad(self):
        # reference :203-207 zeroes 209 tensors one by one; one fill here, and
        # none at all when the fused pass already left the buffer zeroed.
        if not self._grads_are_zero:
            self.opt.engine.zero_grads()
        self._grads_are_zero = True

    def _common_forward(self, total_batch, start_index):
        micro_batch = {
            k: v[start_index: start_index + self.microbatch].to(dist_util.dev(), non_blocking=True)
            for k, v in total_batch.items()
        }
        weights = None
        if self.schedule_sampler is not None:
            t, weights = self.schedule_sampler.sample(self.get_batch_length(micro_batch), dist_util.dev())
            micro_batch[self.timestep_key] = t
        last_batch = (start_index + self.microbatch) >= self.get_batch_length(total_batch)
        if last_batch or not self.use_ddp:
            losses = self.compute_losses(micro_batch)
        else:
            with self.ddp_model.no_sync():
                losses = self.compute_losses(micro_batch)
        if weights is not None:
            from .sampler import LossAwareSampler
            if torch.is_grad_enabled() and isinstance(self.schedule_sampler, LossAwareSampler):
                # one packed all-gather + one kernel, no host sync (upstream: 3 all_gathers + 2*W*B .item())
                self.schedule_sampler.update_with_local_losses(micro_batch[self.timestep_key], losses["loss"].detach())
            losses = {k: v * weights for k, v in losses.items()}   # upstream: (losses["loss"] * weights).mean()
        return losses

    def _microbatch_losses(self, batch):
        """_zero_grad, then the losses of every microbatch in turn, each logged the way the
        reference logs it -- with mode="eval" even while training (sic, :234)."""
        self._zero_grad()
        for start in range(0, self.get_batch_length(batch), self.microbatch):
            losses = self._common_forward(batch, start)
            self.log_loss_dict(mode="eval", losses=losses)
            yield losses

    @torch.no_grad()
    def forward_only(self, batch):
        for _ in self._microbatch_losses(batch):
            pass

    def forward_backward(self, batch):
        for losses in self._microbatch_losses(batch):
            if self.use_fp16:
                scale = 2.0 ** self.lg_loss_scale
                losses = {k: v * scale for k, v in losses.items()}
            self._grads_are_zero = False
            self.backward_from_losses(losses)

    # ---- the update (reference :237-275) -------------------------------------
    def optimize(self):
        self._drain_norms(block=False)
        # a flag left over from a direct opt.step() nobody consumed must not swallow this step's EMA
        self.opt.ema_done = [False] * len(self.opt.ema_done)
        if self.use_fp16:
            return self._optimize_fp16()
        if self._pair_fused and self.gradient_clipping > 0:
            raise RuntimeError("gradient_clipping was switched on after construction: this loop exchanges gradients "
                               "inside the update kernel (grads_after_step='keep', 2 ranks); build it with clipping set")
        self._in_optimize = True
        try:
            if self.gradient_clipping > 0:
                self.grad_clip()
            self._log_grad_norm()
            self._anneal_lr()
            if self._pair_fused and not self.gradient_clipping > 0:
                self.opt.reduce_and_step()   # all-reduce + AdamW + EMA, one kernel
            else:
                self.opt.step()
            for rate, params in zip(self.ema_rate, self.ema_params):
                update_ema(params, self.master_params, rate=rate)
        finally:
            self._in_optimize = False
        self._after_fused_step()

    def _optimize_fp16(self):
        """Upstream ``optimize_fp16``: skip the step and halve the scale when any
        gradient is non-finite, else unscale, step, and grow the scale.  The test
        and the skip both happen on the device; the host reads one flag."""
        opt = self.opt
        opt.inv_scale = 1.0 / (2.0 ** self.lg_loss_scale)
        opt.skip_if_nonfinite = True
        try:
            if self.gradient_clipping > 0:
                self.grad_clip()
            else:
                opt.check_finite()
            self._anneal_lr()
            opt.step()
            res = opt.engine.read_results_sync()
        finally:
            opt.inv_scale = 1.0
            opt.skip_if_nonfinite = False
        if res[_lib.R_COMM_TIMEOUT]:
            raise RuntimeError("gradient all-reduce timed out waiting for a peer rank (is a rank down?)")
        if res[_lib.R_SKIPPED]:
            opt.undo_step_count()
            self.lg_loss_scale -= 1
            self.logger.log(f"Found NaN, decreased lg_loss_scale to {self.lg_loss_scale}")
            return
        self.logger.logkv_mean("grad_norm", res[_lib.R_LOGGED_NORM])
        for rate, params in zip(self.ema_rate, self.ema_params):
            update_ema(params, self.master_params, rate=rate)
        self.lg_loss_scale += self.fp16_scale_growth
        if opt.grad_mode == _lib.GRAD_ZERO:
            self._grads_are_zero = True

    def grad_clip(self):
        # the reference's own seam for optimizers that clip themselves (:248-250);
        # FusedAdamW does: pass 1 runs, the coefficient stays on the device
        clip_own = getattr(self.opt, "clip_grad_norm", None)
        if clip_own is not None:
            clip_own(self.gradient_clipping)
        else:
            torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.gradient_clipping)

    def _anneal_lr(self):
        """Linear decay to zero over learning_steps, from TrainLoop's own step counter (:257-263)."""
        if self.learning_steps:
            frac_done = (self.step + self.resume_step) / self.learning_steps
            for group in self.opt.param_groups:
                group["lr"] = self.lr * (1 - frac_done)

    def _log_grad_norm(self):
        """Reference :265-271 reads 209 sums back one by one.  Inside optimize()
        the value comes for free from the fused pass and is read back
        asynchronously; called on its own it costs one norm pass and one read."""
        if self._in_optimize:
            self._norm_requested = True
            return
        self.opt.check_finite()
        res = self.opt.engine.read_results_sync()
        self.logger.logkv_mean("grad_norm", res[_lib.R_TOTAL_NORM])

    def _after_fused_step(self):
        # known here and only here: the fused pass either zeroed the buffer or it did not
        self._grads_are_zero = self.opt.grad_mode == _lib.GRAD_ZERO
        if not self._norm_requested:
            return
        self._norm_requested = False
        if len(self._norm_inflight) == len(self._norm_ring):
            self._drain_norms(block=False, at_least=1)  # every slot is in flight: resolve the oldest first
        host, ev = self._norm_ring[self._norm_ring_pos]
        self._norm_ring_pos = (self._norm_ring_pos + 1) % len(self._norm_ring)
        host.copy_(self.opt.engine.results, non_blocking=True)
        ev.record()
        self._norm_inflight.append((ev, host))

    def _drain_norms(self, block, at_least=0):
        """Feed finished read-backs to logger.logkv_mean("grad_norm", ...) -- the one
        float per step the reference logs (utils/trainer.py:271) -- oldest first.
        Waits for a read-back only if `block`, or until `at_least` have been resolved."""
        done = 0
        while self._norm_inflight:
            ev, host = self._norm_inflight[0]
            if not ev.query():
                if not block and done >= at_least:
                    break
                ev.synchronize()
            if host[_lib.R_COMM_TIMEOUT] != 0:
                raise RuntimeError("gradient all-reduce timed out waiting for a peer rank (is a rank down?)")
            self.logger.logkv_mean("grad_norm", float(host[_lib.R_LOGGED_NORM]))
            self._norm_inflight.pop(0)
            done += 1

    def flush_logs(self):
        """Resolve every pending read-back (grad norms, deferred loss statistics) into the logger."""
        self._drain_norms(block=True)
        self._loss_log.drain(block=True)

    def log_step(self):
        self.logger.logkv("step", self.step + self.resume_step)
        self.logger.logkv("samples", (self.step + self.resume_step + 1) * self.global_batch)

    # ---- checkpoints (reference :277-355): same file names and layouts ---------
    def _checkpoint_file(self, stem):
        return os.path.join(self.checkpoint_path, f"{stem}_{(self.step + self.resume_step):06d}.pt")

    def save(self):
        self.flush_logs()
        for rate, params in [(0, self.master_params), *zip(self.ema_rate, self.ema_params)]:
            self._save_checkpoint(rate, params)
        self._save_opt()
        dist_util.barrier()

    def _save_checkpoint(self, rate, params):
        state_dict = self._master_params_to_state_dict(params)   # on every rank, like the reference
        if dist_util.get_rank() != 0:
            return
        self.logger.log(f"saving model {rate}...")
        path = self._checkpoint_file(f"ema_{rate}" if rate else "model")
        print('writing to', path)
        with open(path, "wb") as f:
            torch.save(state_dict, f)

    def _save_opt(self):
        if dist_util.get_rank() != 0:
            return
        self.logger.log("saving optimizer...")
        path = self._checkpoint_file("opt")
        print('writing to', path)
        with open(path, "wb") as f:
            torch.save(self.opt.state_dict(), f)

    def _master_params_to_state_dict(self, master_params, key=None):
        state_dict = self.model.state_dict()
        for i, (name, _value) in enumerate(self.model.named_parameters()):
            assert name in state_dict
            if key is not None and key == name:
                return master_params[i]
            # clone: the checkpoint must not alias (and so serialise) a whole flat buffer
            state_dict[name] = master_params[i].detach().clone()
        if key is not None:
            raise KeyError(key)
        for name in state_dict:
            if state_dict[name]._base is not None:
                state_dict[name] = state_dict[name].detach().clone()
        return state_dict

    def _state_dict_to_master_params(self, state_dict):
        return [state_dict[name] for name, _ in self.model.named_parameters()]

    @staticmethod
    def parse_resume_step_from_filename(filename):
        """path/to/model_NNNNNN.pt -> NNNNNN"""
        filename = os.path.basename(filename)
        assert filename.startswith('model') and filename[-3:] == '.pt', "Invalid model name"
        return int(filename[-9:-3])

    @staticmethod
    def find_resume_checkpoint(logger=None):
        log_dir = (logger or _default_logger).get_current().dir
        if not log_dir or not os.path.isdir(log_dir):
            return None
        names = sorted(s for s in os.listdir(log_dir) if s.endswith(".pt") and s.startswith("model"))
        return os.path.join(log_dir, names[-1]) if names else None

    @staticmethod
    def find_ema_checkpoint(main_checkpoint, step, rate):
        if not main_checkpoint:
            return None
        path = os.path.join(os.path.dirname(main_checkpoint), f"ema_{rate}_{step:06d}.pt")
        return path if os.path.exists(path) else None

    @staticmethod
    def find_opt_checkpoint(main_checkpoint, step):
        if not main_checkpoint:
            return None
        path = os.path.join(os.path.dirname(main_checkpoint), f"opt_{step:06d}.pt")
        return path if os.path.exists(path) else None


class _DeferredLossLog:
    """Pinned ring of per-microbatch loss statistics (f-3: the logging hand-off).  Each ``add`` leaves
    2 + 8 numbers per key on the device (sum, count, 4 quartile sums, 4 quartile counts), starts an
    asynchronous copy and returns; ``drain`` feeds finished copies to ``logkv_mean`` with the same
    running-mean arithmetic as one call per sample (basic_utils/logger.py:350-356)."""

    SLOTS = 64

    def __init__(self, logger):
        self.logger = logger
        self._free = []
        self._inflight = []

    def add(self, losses, ts, num_timesteps, prefix):
        keys = list(losses.keys())
        if not keys:
            return
        rows = []
        for k in keys:
            v = losses[k].detach().reshape(-1).double()
            row = torch.zeros(10, dtype=torch.float64, device=v.device)
            row[0] = v.sum()
            row[1] = v.numel()
            if ts is not None and num_timesteps:
                q = (4 * ts.reshape(-1).to(v.device) // int(num_timesteps)).clamp_(0, 3).long()
                row[2:6].index_add_(0, q, v)
                row[6:10].index_add_(0, q, torch.ones_like(v))
            rows.append(row)
        dev_block = torch.stack(rows)
        if len(self._inflight) >= self.SLOTS:
            self.drain(block=False, at_least=1)
        host = None
        for i, h in enumerate(self._free):
            if h.shape == dev_block.shape:
                host = self._free.pop(i)
                break
        if host is None:
            host = torch.empty(dev_block.shape, dtype=torch.float64, pin_memory=True)
        host.copy_(dev_block, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._inflight.append((ev, host, [prefix + k for k in keys]))

    def _mean_n(self, key, mean, n):
        weighted = getattr(self.logger, "logkv_mean_n", None)
        if weighted is not None:
            weighted(key, mean, n)
        else:                      # a logger without the weighted form: n equal samples give the same mean
            for _ in range(n):
                self.logger.logkv_mean(key, mean)

    def drain(self, block, at_least=0):
        done = 0
        while self._inflight:
            ev, host, keys = self._inflight[0]
            if not ev.query():
                if not block and done >= at_least:
                    break
                ev.synchronize()
            for key, row in zip(keys, host.tolist()):
                if row[1] > 0:
                    self.logger.logkv_mean(key, row[0] / row[1])      # upstream: values.mean().item()
                for q in range(4):
                    if row[6 + q] > 0:                                 # upstream: one logkv_mean per sample
                        self._mean_n(f"{key}_q{q}", row[2 + q] / row[6 + q], int(row[6 + q]))
            self._inflight.pop(0)
            self._free.append(host)
            done += 1


def update_ema(target_params, source_params, rate=0.99):
    """Move target parameters towards source parameters:
    ``trg = trg*rate + src*(1-rate)`` (reference utils/trainer.py:360-370).

    When ``target_params`` is one of a TrainLoop's ``ema_params`` lists and the
    fused step has already advanced it this step, this is a no-op -- so the
    reference's ``for rate, params in zip(...): update_ema(...)`` loop can stay
    as written.  Otherwise one kernel per tensor (CUDA tensors only).
    """
    entry = _FUSED_EMA_LISTS.get(id(target_params))
    if entry is not None:
        opt = entry[0]()
        k = entry[1]
        if opt is not None and opt.ema_done[k] and opt.ema_rates[k] == float(rate) \
                and len(source_params) == len(opt.flat.param_list) \
                and all(a is b for a, b in zip(source_params, opt.flat.param_list)):
            opt.ema_done[k] = False
            return
    for trg, src in zip(target_params, source_params):
        _ema_update_kernel(trg.detach(), src.detach(), rate)
