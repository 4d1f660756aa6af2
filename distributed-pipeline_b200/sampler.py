"""Timestep schedule samplers with the upstream DiffuSeq / improved-diffusion
interface (``create_named_schedule_sampler``, ``.sample(batch, device)``,
``.update_with_local_losses(ts, losses)``).

The reference repository removed them; ``north_star`` asks for the
loss-aware one's per-step gather.  Upstream does three ``all_gather`` calls and
``2*W*B`` ``.item()`` syncs per microbatch; here one packed ``ncclAllGather`` and
one kernel update the fp64 loss history on the device, in gathered (rank-major)
order, without any host sync.  ``sample()`` still draws from numpy's global
``RandomState`` on the host so timesteps are bit-identical to upstream for a
given RNG state; it reads the 8*T-byte weight vector back to do so.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .engine import _DeviceArray, _stream_ptr


class ScheduleSampler:
    def weights(self) -> np.ndarray:
        raise NotImplementedError

    def sample(self, batch_size, device):
        w = self.weights()
        p = w / np.sum(w)
        indices_np = np.random.choice(len(p), size=(batch_size,), p=p)
        indices = torch.from_numpy(indices_np).long().to(device)
        weights_np = 1 / (len(p) * p[indices_np])
        weights = torch.from_numpy(weights_np).float().to(device)
        return indices, weights


class UniformSampler(ScheduleSampler):
    def __init__(self, n_timesteps):
        self._weights = np.ones([int(n_timesteps)])

    def weights(self):
        return self._weights


class LossAwareSampler(ScheduleSampler):
    def update_with_local_losses(self, local_ts, local_losses):
        raise NotImplementedError


class LossSecondMomentResampler(LossAwareSampler):
    def __init__(self, n_timesteps, history_per_term=10, uniform_prob=0.001, device=None, engine=None,
                 max_local_batch=1024):
        """``engine``: the TrainLoop's ``opt.engine`` (its NCCL communicator is
        used for the gather); None or a single-rank engine means local only."""
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("LossSecondMomentResampler keeps its history on a CUDA device (no CPU implementation)")
        self.T = int(n_timesteps)
        self.history_per_term = int(history_per_term)
        self.uniform_prob = float(uniform_prob)
        self.engine = engine
        self.lib = _lib.lib()
        self._s = C.c_void_p()
        index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        _lib.check(self.lib.dpk_sampler_create(C.byref(self._s), index, self.T, self.history_per_term,
                                               self.uniform_prob, int(max_local_batch)))
        self._w = torch.empty(self.T, dtype=torch.float64, device=self.device)
        self._w_host = torch.empty(self.T, dtype=torch.float64, pin_memory=True)

    def close(self):
        if getattr(self, "_s", None):
            self.lib.dpk_sampler_destroy(self._s)
            self._s = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def update_with_local_losses(self, local_ts, local_losses):
        ts = local_ts.detach().to(device=self.device, dtype=torch.int64).contiguous()
        losses = local_losses.detach().to(device=self.device, dtype=torch.float32).contiguous()
        if ts.numel() != losses.numel():
            raise ValueError("timesteps and losses differ in length")
        h = self.engine._h if self.engine is not None else None
        _lib.check(self.lib.dpk_sampler_update(self._s, h, ts.data_ptr(), losses.data_ptr(), ts.numel(),
                                               _stream_ptr(None, self.device)))

    def weights_device(self) -> torch.Tensor:
        _lib.check(self.lib.dpk_sampler_weights(self._s, self._w.data_ptr(), _stream_ptr(None, self.device)))
        return self._w

    def weights(self) -> np.ndarray:
        self._w_host.copy_(self.weights_device(), non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return self._w_host.numpy().copy()

    def state(self):
        """(history float64[T,H], counts int32[T]) device tensors aliasing the library's state."""
        hp, cp = C.c_void_p(), C.c_void_p()
        _lib.check(self.lib.dpk_sampler_state_ptrs(self._s, C.byref(hp), C.byref(cp)))
        with torch.cuda.device(self.device):
            hist = torch.as_tensor(_DeviceArray(hp.value, (self.T, self.history_per_term), "<f8", self),
                                   device=self.device)
            counts = torch.as_tensor(_DeviceArray(cp.value, (self.T,), "<i4", self), device=self.device)
        return hist, counts


def create_named_schedule_sampler(name, diffusion_or_steps, **kwargs):
    """``diffusion_or_steps``: an object with ``num_timesteps`` (upstream passes the
    diffusion) or the number of timesteps."""
    n = getattr(diffusion_or_steps, "num_timesteps", diffusion_or_steps)
    if name == "uniform":
        return UniformSampler(n)
    if name == "lossaware":
        return LossSecondMomentResampler(n, **kwargs)
    raise NotImplementedError(f"unknown schedule sampler: {name}")
