"""``FusedAdamW``: the object ``TrainLoop.opt`` holds.

It stands where ``torch.optim.AdamW(master_params, lr, weight_decay)`` stands in
the reference (``utils/trainer.py:99``) and offers the parts of the optimizer
surface the reference touches: ``param_groups[i]["lr"]`` (``:262-263``),
``step()`` (``:242``), ``state_dict()/load_state_dict()`` in torch's own layout
(``:173``, ``:302``) and the optional ``clip_grad_norm`` hook ``grad_clip``
looks for (``:248-250``).  ``step()`` is one kernel launch: unscale, clip,
AdamW and every EMA rate.
"""
from typing import Dict, List, Sequence

import torch

from . import _lib
from .engine import Engine
from .flat import FlatState


class FusedAdamW:
    def __init__(self, params: Sequence[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999),
                 eps: float = 1e-8, weight_decay: float = 1e-2, ema_rates: Sequence[float] = (),
                 bucket_cap_mb: float = 128, grad_mode: int = _lib.GRAD_CLIPPED, tile_elems: int = 0,
                 symmetric_grads: bool = False, sharded: bool = False):
        """``sharded``: ZeRO-1 style update (SURVEY f-4; the reference's own nod to it is the
        "sharded optimizer" branch of grad_clip, utils/trainer.py:248-250).  Each rank updates every
        W-th chunk and multicasts the new parameters; moments and EMA copies are owner-only until
        ``gather_shards()`` (called by ``state_dict``).  Enable after ``engine.init_comm()`` and
        ``engine.enable_nvls(flat.grads)`` with ``engine.enable_shard(flat.params)``.  No clipping."""
        params = list(params)
        self.sharded = bool(sharded)
        self.flat = FlatState(params, n_ema=len(ema_rates), symmetric_grads=symmetric_grads or sharded,
                              symmetric_params=sharded)
        self.engine = Engine(self.flat.layout, self.flat.dtype, self.flat.device, ema_rates, betas, eps,
                             bucket_cap_bytes=int(bucket_cap_mb * 1024 * 1024), tile_elems=tile_elems,
                             frozen=self.flat.frozen)
        f = self.flat
        self.engine.bind(f.params, f.grads, f.exp_avg, f.exp_avg_sq, f.emas)
        self.defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay, amsgrad=False,
                             maximize=False, foreach=None, capturable=False, differentiable=False, fused=None,
                             decoupled_weight_decay=True)
        self.param_groups: List[Dict] = [dict(self.defaults, params=params)]
        self.ema_rates = [float(r) for r in ema_rates]
        self.grad_mode = grad_mode
        self.step_count = 0          # AdamW's per-parameter `step`, identical for all
        self.inv_scale = 1.0         # 1/2^lg_loss_scale for the next step
        self.skip_if_nonfinite = False
        self.fused_reduce = True     # reduce_and_step may use the one-kernel all-reduce + update
        self._clip_pending = False
        # one flag per EMA rate: set by step(), consumed by trainer.update_ema
        self.ema_done = [False] * len(self.ema_rates)

    # ---- what TrainLoop calls ----------------------------------------------
    def clip_grad_norm(self, max_norm: float):
        """Pass 1.  The coefficient stays on the device; step() applies it."""
        self.engine.grad_norm(self.inv_scale, float(max_norm))
        self._clip_pending = True

    def check_finite(self):
        """Pass 1 without clipping: fills the non-finite flag for a loss-scaled step."""
        self.engine.grad_norm(self.inv_scale, 0.0)
        self._clip_pending = False

    @torch.no_grad()
    def step(self, closure=None):
        if closure is not None:
            raise NotImplementedError("FusedAdamW.step does not take a closure")
        g = self.param_groups[0]
        b1, b2 = g["betas"]
        if (b1, b2) != self.engine.betas or g["eps"] != self.engine.eps:
            raise ValueError("betas/eps are fixed when the optimizer is created")
        self.step_count += 1
        self.engine.fused_update(float(g["lr"]), float(g["weight_decay"]), self.step_count,
                                 inv_scale=self.inv_scale, use_clip_coef=self._clip_pending,
                                 skip_if_nonfinite=self.skip_if_nonfinite, grad_mode=self.grad_mode)
        self._clip_pending = False
        self.ema_done = [True] * len(self.ema_rates)

    @torch.no_grad()
    def reduce_and_step(self, max_norm: float = -1.0):
        """The whole post-backward update when every local gradient is already in
        the flat buffer: enqueue the all-reduce of every bucket, then consume
        buckets as they land -- bucket b is updated (or, when clipping, summed
        into the norm) while bucket b+1 is still on NVLink.  Same results as
        all-reduce -> clip_grad_norm -> step."""
        eng = self.engine
        g = self.param_groups[0]
        n = len(eng.buckets)
        if self.sharded and eng.sharded:
            if max_norm > 0 or self.skip_if_nonfinite or self.grad_mode != _lib.GRAD_KEEP:
                raise ValueError("the sharded update supports neither clipping nor a loss scale nor rewritten gradients")
            self.step_count += 1
            eng.shard_update(float(g["lr"]), float(g["weight_decay"]), self.step_count, self.inv_scale)
            self._clip_pending = False
            self.ema_done = [True] * len(self.ema_rates)
            return
        if (eng.world_size >= 2 and eng.nvls and self.fused_reduce and max_norm <= 0
                and self.grad_mode == _lib.GRAD_KEEP and not self.skip_if_nonfinite):
            # no clipping, gradients not rewritten: exchange and update in ONE kernel -- two ranks
            # average each vector with the peer's in registers, four or more reduce through
            # the switch in part of the grid while the rest updates range by range
            self.step_count += 1
            eng.reduce_update(float(g["lr"]), float(g["weight_decay"]), self.step_count, self.inv_scale)
            self._clip_pending = False
            self.ema_done = [True] * len(self.ema_rates)
            return
        if max_norm > 0 and eng.world_size >= 2 and eng.nvls and self.fused_reduce:
            # clipping: the all-reduce and the norm pass together, then the update
            eng.reduce_norm(self.inv_scale, float(max_norm))
            self._clip_pending = True
            return self.step()
        if eng.world_size >= 4:
            # every gradient is already in place: one "peers complete" barrier serves all buckets
            eng.all_buckets_ready()
        elif eng.world_size > 1:
            for b in range(n):
                eng.bucket_ready(b)
        if max_norm > 0:
            for b in range(n):
                eng.wait_bucket(b)
                eng.grad_norm_bucket(b, self.inv_scale)
            eng.grad_norm_finish(float(max_norm))
            self._clip_pending = True
            return self.step()
        self.step_count += 1
        for b in range(n):
            eng.wait_bucket(b)
            eng.fused_update_bucket(b, float(g["lr"]), float(g["weight_decay"]), self.step_count,
                                    inv_scale=self.inv_scale, use_clip_coef=False,
                                    skip_if_nonfinite=False, grad_mode=self.grad_mode)
        eng.fused_update_finish(False)
        self._clip_pending = False
        self.ema_done = [True] * len(self.ema_rates)

    @torch.no_grad()
    def step_from_host(self, host_grads: torch.Tensor, max_norm: float = -1.0):
        """One update whose gradients arrive from (pinned) host memory, laid out like
        the flat gradient buffer.  The buffer is uploaded bucket by bucket on a copy
        stream and every bucket is all-reduced (multi-rank) and consumed as soon as
        it has landed, so the update hides behind the PCIe transfer of the next
        bucket instead of following the whole copy."""
        eng, flat = self.engine, self.flat
        if host_grads.numel() != flat.layout.n_elems or host_grads.dtype != flat.dtype:
            raise ValueError("host_grads must be a flat tensor with the layout of the gradient buffer")
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=flat.device)
        cur = torch.cuda.current_stream(flat.device)
        copy = self._copy_stream
        copy.wait_stream(cur)                    # the previous step is done with the gradient buffer
        buckets = eng.buckets
        landed = []
        with torch.cuda.stream(copy):
            for b, (s, c) in enumerate(buckets):
                flat.grads[s:s + c].copy_(host_grads[s:s + c], non_blocking=True)
                if eng.world_size > 1:
                    eng.bucket_ready(b, stream=copy)   # all-reduce follows the upload of this bucket
                else:
                    ev = torch.cuda.Event()
                    ev.record(copy)
                    landed.append(ev)
        g = self.param_groups[0]

        def arrived(b):
            if eng.world_size > 1:
                eng.wait_bucket(b)
            else:
                cur.wait_event(landed[b])

        if max_norm > 0:
            for b in range(len(buckets)):
                arrived(b)
                eng.grad_norm_bucket(b, self.inv_scale)
            eng.grad_norm_finish(float(max_norm))
            self._clip_pending = True
            return self.step()
        self.step_count += 1
        for b in range(len(buckets)):
            arrived(b)
            eng.fused_update_bucket(b, float(g["lr"]), float(g["weight_decay"]), self.step_count,
                                    inv_scale=self.inv_scale, grad_mode=self.grad_mode)
        eng.fused_update_finish(False)
        self._clip_pending = False
        self.ema_done = [True] * len(self.ema_rates)

    def undo_step_count(self):
        """A predicated-off (non-finite) step must not advance AdamW's step."""
        self.step_count -= 1
        self.ema_done = [False] * len(self.ema_rates)

    def zero_grad(self, set_to_none: bool = False):
        if set_to_none:
            raise ValueError("gradients are views of the flat buffer and cannot be set to None")
        self.engine.zero_grads()

    # ---- torch.optim.AdamW-compatible checkpoint layout -----------------------
    @property
    def state(self) -> Dict[torch.nn.Parameter, Dict[str, torch.Tensor]]:
        if self.step_count == 0:
            return {}
        self.gather_shards()
        f = self.flat
        return {p: {"step": torch.tensor(float(self.step_count)), "exp_avg": m, "exp_avg_sq": v}
                for p, m, v, fr in zip(f.param_list, f.exp_avg_views, f.exp_avg_sq_views, f.frozen) if not fr}

    def gather_shards(self):
        """After sharded steps: complete every rank's moments and EMA copies (collective)."""
        if self.sharded:
            self.engine.shard_gather()

    def state_dict(self) -> Dict:
        self.gather_shards()
        f = self.flat
        state = {}
        if self.step_count:
            for i, (m, v) in enumerate(zip(f.exp_avg_views, f.exp_avg_sq_views)):
                if f.frozen[i]:
                    continue  # torch.optim.AdamW keeps no state for a parameter that never had a gradient
                state[i] = {"step": torch.tensor(float(self.step_count)), "exp_avg": m, "exp_avg_sq": v}
        groups = []
        for g in self.param_groups:
            d = {k: v for k, v in g.items() if k != "params"}
            d["params"] = list(range(len(g["params"])))
            groups.append(d)
        return {"state": state, "param_groups": groups}

    @torch.no_grad()
    def load_state_dict(self, sd: Dict):
        f = self.flat
        groups = sd["param_groups"]
        if len(groups) != 1 or len(groups[0]["params"]) != len(f.param_list):
            raise ValueError("loaded state dict has a different parameter-group layout")
        for k, v in groups[0].items():
            if k != "params":
                self.param_groups[0][k] = v
        state = sd["state"]
        if not state:
            self.step_count = 0
            f.exp_avg.zero_()
            f.exp_avg_sq.zero_()
            return
        steps = set()
        for i, (m, v) in enumerate(zip(f.exp_avg_views, f.exp_avg_sq_views)):
            st = state.get(i, state.get(str(i)))
            if st is None:
                # no entry: torch never stepped this parameter (frozen / never received a gradient)
                if not f.frozen[i]:
                    m.zero_()
                    v.zero_()
                continue
            m.copy_(st["exp_avg"])
            v.copy_(st["exp_avg_sq"])
            steps.add(int(float(st["step"])))
        if len(steps) != 1:
            raise ValueError("parameters with different AdamW step counts cannot share one fused step")
        self.step_count = steps.pop()
