"""Shared test plumbing: random flat states, oracle drivers, comparisons."""
import numpy as np
import torch

import oracle
from distributed_pipeline_b200.engine import Engine
from distributed_pipeline_b200.flat import FlatLayout

NP_DT = {torch.float32: oracle.F32, torch.bfloat16: oracle.BF16}


def to_np(t: torch.Tensor) -> np.ndarray:
    """torch tensor -> numpy in the oracle's convention (bf16 as uint16 bits)."""
    t = t.detach().cpu().contiguous()
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    return t.numpy().copy()


def from_np(a: np.ndarray, dtype: torch.dtype, device) -> torch.Tensor:
    if dtype == torch.bfloat16:
        return torch.from_numpy(a.view(np.int16).copy()).view(torch.bfloat16).to(device)
    return torch.from_numpy(a.copy()).to(device)


class FlatProblem:
    """One engine plus flat torch buffers plus a numpy mirror driven by the oracle."""

    def __init__(self, numels, dtype=torch.float32, ema_rates=(0.9999,), seed=0, device="cuda", tile_elems=0,
                 grad_std=1e-3, bucket_cap_bytes=128 << 20, mirror=True):
        self.dtype, self.device = dtype, torch.device(device)
        self.ema_rates = list(ema_rates)
        es = 4 if dtype == torch.float32 else 2
        self.layout = FlatLayout.from_numels(numels, es)
        n = self.layout.n_elems
        gen = torch.Generator(device="cpu").manual_seed(seed)
        self.gen = gen
        self.grad_std = grad_std
        mask = torch.zeros(n, dtype=torch.bool)
        for o, c in zip(self.layout.offsets, self.layout.numels):
            mask[o:o + c] = True
        self.mask = mask
        p = (torch.randn(n, generator=gen) * 0.02).to(dtype)
        p[~mask] = 0
        self.p = p.to(self.device)
        self.g = torch.zeros(n, dtype=dtype, device=self.device)
        self.m = torch.zeros(n, dtype=dtype, device=self.device)
        self.v = torch.zeros(n, dtype=dtype, device=self.device)
        self.emas = [self.p.clone() for _ in self.ema_rates]
        self.engine = Engine(self.layout, dtype, self.device, self.ema_rates, tile_elems=tile_elems,
                             bucket_cap_bytes=bucket_cap_bytes)
        self.engine.bind(self.p, self.g, self.m, self.v, self.emas)
        self.mirror = mirror
        if mirror:
            self.np = {"p": to_np(self.p), "m": to_np(self.m), "v": to_np(self.v),
                       "emas": [to_np(e) for e in self.emas]}

    def new_grads(self, std=None, poison_padding=True):
        std = self.grad_std if std is None else std
        g = (torch.randn(self.layout.n_elems, generator=self.gen) * std).to(self.dtype)
        if poison_padding:
            g[~self.mask] = float("nan")  # the kernels must never read padding as data
        else:
            g[~self.mask] = 0
        self.g.copy_(g)
        return to_np(g)

    def oracle_norm(self, g_np, inv_scale=1.0, max_norm=0.0):
        return oracle.grad_norm(NP_DT[self.dtype], g_np, self.layout.offsets, self.layout.numels, inv_scale, max_norm)

    def oracle_step(self, g_np, lr, wd, step, inv_scale=1.0, coef=1.0, grad_mode=0, flavour=oracle.CUDA):
        """Applies the oracle per tensor (so padding stays untouched); returns per-tensor sumsq list."""
        dt = NP_DT[self.dtype]
        out = []
        for o, c in zip(self.layout.offsets, self.layout.numels):
            sl = slice(o, o + c)
            views = [self.np["p"][sl], g_np[sl], self.np["m"][sl], self.np["v"][sl]]
            em = [e[sl] for e in self.np["emas"]]
            out.append(oracle.adamw_ema_step(dt, views[0], views[1], views[2], views[3], em, self.ema_rates, lr, wd,
                                             step, inv_scale=inv_scale, coef=coef, grad_mode=grad_mode,
                                             flavour=flavour))
        return out

    def assert_state_equal(self, what=("p", "m", "v", "emas")):
        got = {"p": to_np(self.p), "m": to_np(self.m), "v": to_np(self.v), "emas": [to_np(e) for e in self.emas]}
        for k in what:
            if k == "emas":
                for i, (a, b) in enumerate(zip(got["emas"], self.np["emas"])):
                    assert_bits_equal(a, b, f"ema[{i}]")
            else:
                assert_bits_equal(got[k], self.np[k], k)


def assert_bits_equal(a: np.ndarray, b: np.ndarray, name=""):
    ai = a.view(np.uint32) if a.dtype == np.float32 else a
    bi = b.view(np.uint32) if b.dtype == np.float32 else b
    bad = np.nonzero(ai != bi)[0]
    if bad.size:
        i = int(bad[0])
        raise AssertionError(f"{name}: {bad.size}/{a.size} elements differ; first at {i}: {a[i]!r} vs {b[i]!r}")


def rel_err(a, b, floor=1e-30):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), floor)
