"""Bucketed gradient all-reduce on the flat gradient buffer.

Takes the place of ``torch.nn.parallel.DistributedDataParallel`` as the
reference configures it (``utils/trainer.py:118-125``: ``bucket_cap_mb=128``,
``broadcast_buffers=False``, ``find_unused_parameters=False``).  Differences
that matter for speed, none for results:

* buckets are contiguous ranges of the flat gradient buffer (tail first, the
  order gradients become ready), so there is no copy-in / copy-out and no
  per-parameter scaling kernel: each bucket is one in-place ``ncclAllReduce``
  with ``ncclAvg`` (pre-multiply by 1/W then sum, the reducer's arithmetic);
* buckets are launched from post-accumulate-grad hooks on a high-priority side
  stream, overlapping the rest of backward;
* nothing here synchronises the host.
"""
import contextlib
import threading
from typing import List

import torch

from .engine import Engine
from .flat import FlatState


class BucketReducer:
    """Counts gradient readiness per bucket and fires the library's all-reduce."""

    def __init__(self, flat: FlatState, engine: Engine):
        self.flat = flat
        self.engine = engine
        self.enabled = True           # False inside no_sync()
        # DDP decides at *forward* time whether the coming backward syncs: the
        # reference leaves no_sync() before calling backward (trainer.py:219-235)
        self.armed = False
        self._callback_queued = False
        self._lock = threading.Lock()
        ranges = engine.buckets
        self.bucket_params: List[List[int]] = [flat.layout.segment_of(s, c) for s, c in ranges]
        self.param_bucket = {}
        for b, idxs in enumerate(self.bucket_params):
            for i in idxs:
                self.param_bucket[i] = b
        missing = set(range(len(flat.param_list))) - set(self.param_bucket)
        if missing:
            raise RuntimeError(f"bucket plan does not cover parameters {sorted(missing)[:5]}...")
        # only parameters that will receive a gradient (and so fire a hook) are waited for;
        # a bucket of frozen tensors only is ready from the start
        self._hooked = [sum(1 for i in idxs if flat.param_list[i].requires_grad) for idxs in self.bucket_params]
        self._pending = list(self._hooked)
        self._ready = [n == 0 for n in self._hooked]
        self._next = 0
        self._handles = []
        for i, p in enumerate(flat.param_list):
            if p.requires_grad:
                self._handles.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))

    def _make_hook(self, index: int):
        b = self.param_bucket[index]

        def hook(param):
            if not self.armed or self.engine.world_size <= 1:
                return
            launch = []
            with self._lock:
                if not self._callback_queued:
                    # like DDP's reducer: finish once this backward pass is over
                    self._callback_queued = True
                    torch.autograd.Variable._execution_engine.queue_callback(self.finish)
                self._pending[b] -= 1
                if self._pending[b] == 0:
                    self._ready[b] = True
                # Collectives must be issued in the same order on every rank, whatever
                # order the hooks happen to run in: buckets launch strictly by index,
                # a ready bucket waiting for its predecessors (DDP's next_bucket rule).
                while self._next < len(self._ready) and self._ready[self._next]:
                    launch.append(self._next)
                    self._next += 1
            for nb in launch:
                # runs on autograd's engine thread; the current stream there is
                # the one the gradients were produced on
                self.engine.bucket_ready(nb)
        return hook

    def finish(self):
        """After backward(): launch buckets whose parameters did not all receive
        a gradient, make the current stream wait for every all-reduce, re-arm."""
        if self.armed and self.engine.world_size > 1:
            for b in range(self._next, len(self._ready)):
                self.engine.bucket_ready(b)
            self.engine.wait_allreduce()
        self._pending = list(self._hooked)
        self._ready = [n == 0 for n in self._hooked]
        self._next = 0
        self._callback_queued = False
        self.armed = False

    def remove(self):
        for h in self._handles:
            h.remove()
        self._handles = []


class FlatDDP(torch.nn.Module):
    """What ``TrainLoop.ddp_model`` is: callable like the model, has ``.module``
    and a ``no_sync()`` context manager (``utils/trainer.py:216-220``)."""

    def __init__(self, module: torch.nn.Module, reducer: BucketReducer):
        super().__init__()
        self.module = module
        self._reducer = reducer

    def forward(self, *args, **kwargs):
        if torch.is_grad_enabled():
            self._reducer.armed = self._reducer.enabled
        return self.module(*args, **kwargs)

    @contextlib.contextmanager
    def no_sync(self):
        old = self._reducer.enabled
        self._reducer.enabled = False
        try:
            yield
        finally:
            self._reducer.enabled = old
