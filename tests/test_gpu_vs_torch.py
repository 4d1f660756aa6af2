"""libdpk against torch-CUDA eager running the reference's op sequence
(oracle/torch_port.py = clip_grad_norm_ + torch.optim.AdamW + mul_/add_ EMA on
per-tensor lists): the tight-tolerance oracle of SURVEY §8c (T1).

Tolerances are north_star's: relative 1e-6 on fp32 parameters and EMA after N
steps, 1e-5 on the grad norm, ulp-level on bf16.
"""
import numpy as np
import pytest
import torch

from distributed_pipeline_b200 import _lib
from helpers import FlatProblem, rel_err
from oracle.torch_port import ReferenceUpdatePort

pytestmark = pytest.mark.gpu

SHAPES = [(300, 128), (300,), (64, 128), (64,), (128, 64), (128,), (777,), (5,), (1000, 33)]


def _make(dtype, rates, wd, clip, seed, learning_steps=0, grad_std=1e-3, lr=1e-3):
    numels = [int(np.prod(s)) for s in SHAPES]
    pb = FlatProblem(numels, dtype, rates, seed=seed, grad_std=grad_std, mirror=False)
    tensors = [torch.nn.Parameter(pb.p[o:o + n].clone().view(s))
               for o, n, s in zip(pb.layout.offsets, pb.layout.numels, SHAPES)]
    port = ReferenceUpdatePort(tensors, lr=lr, ema_rates=rates, weight_decay=wd, learning_steps=learning_steps,
                               gradient_clipping=clip)
    return pb, port


def _step(pb, port, lr0, wd, clip, step, learning_steps):
    g = pb.new_grads(poison_padding=False)
    for t, o, n in zip(port.params, pb.layout.offsets, pb.layout.numels):
        t.grad = pb.g[o:o + n].clone().view(t.shape)
    port.optimize()
    lr = lr0 * (1 - (step - 1) / learning_steps) if learning_steps else lr0
    if clip > 0:
        pb.engine.grad_norm(1.0, clip)
    pb.engine.fused_update(lr, wd, step, use_clip_coef=clip > 0, grad_mode=_lib.GRAD_CLIPPED)
    return pb.engine.read_results_sync()


def _gather(pb, port):
    cat = lambda ts: torch.cat([t.detach().reshape(-1) for t in ts])
    lay = pb.layout
    pick = lambda flat: torch.cat([flat[o:o + n] for o, n in zip(lay.offsets, lay.numels)])
    pairs = {"p": (pick(pb.p), cat(port.params)),
             "m": (pick(pb.m), cat([port.opt.state[p]["exp_avg"] for p in port.params])),
             "v": (pick(pb.v), cat([port.opt.state[p]["exp_avg_sq"] for p in port.params])),
             "g": (pick(pb.g), cat([p.grad for p in port.params]))}
    for k, e in enumerate(pb.emas):
        pairs[f"ema{k}"] = (pick(e), cat(port.ema_params[k]))
    return pairs


@pytest.mark.parametrize("clip,wd,learning_steps", [(-1.0, 0.0, 0), (1.0, 0.01, 200)])
@pytest.mark.parametrize("n_steps", [1, 10, 100])
def test_fp32_matches_torch_cuda(clip, wd, learning_steps, n_steps):
    rates = [0.9999, 0.999]
    pb, port = _make(torch.float32, rates, wd, clip, seed=n_steps, learning_steps=learning_steps,
                     grad_std=1e-2 if clip > 0 else 1e-3)
    for step in range(1, n_steps + 1):
        res = _step(pb, port, 1e-3, wd, clip, step, learning_steps)
        assert rel_err(res[_lib.R_LOGGED_NORM], port.last_logged_norm) < 1e-5
        if clip > 0:
            assert rel_err(res[_lib.R_TOTAL_NORM], float(port.last_total_norm)) < 1e-5
    stats = {}
    for name, (a, b) in _gather(pb, port).items():
        a64, b64 = a.double().cpu().numpy(), b.double().cpu().numpy()
        # north_star: relative 1e-6 after N steps.  Entries that pass through zero
        # have no meaningful relative error, hence the usual allclose form with the
        # absolute floor tied to the tensor's own scale: |a-b| <= 1e-6*(|b| + rms(b)).
        rms = float(np.sqrt(np.mean(b64 ** 2)))
        err = np.abs(a64 - b64) / (np.abs(b64) + rms)
        stats[name] = (float(err.max()), float((a64 != b64).mean()))
        assert err.max() < 1e-6, (name, stats[name])
    print(f"\nfp32 clip={clip} steps={n_steps}: max rel err / fraction of elements not bit-equal: {stats}")


@pytest.mark.parametrize("clip,wd", [(-1.0, 0.0), (1.0, 0.01)])
@pytest.mark.parametrize("n_steps", [1, 10])
def test_bf16_matches_torch_cuda_ulp_level(clip, wd, n_steps):
    rates = [0.9999, 0.999]
    pb, port = _make(torch.bfloat16, rates, wd, clip, seed=40 + n_steps, grad_std=1e-2 if clip > 0 else 1e-3)
    for step in range(1, n_steps + 1):
        res = _step(pb, port, 1e-3, wd, clip, step, 0)
        if clip > 0:
            # total norm and coefficient are bf16 tensors in torch
            assert rel_err(res[_lib.R_TOTAL_NORM], float(port.last_total_norm)) <= 2 ** -7
    stats = {}
    for name, (a, b) in _gather(pb, port).items():
        ai = a.view(torch.int16).int().cpu().numpy()
        bi = b.view(torch.int16).int().cpu().numpy()
        ulp = np.abs(ai - bi)
        stats[name] = (int(ulp.max()), float((ulp != 0).mean()))
    print(f"\nbf16 clip={clip} steps={n_steps}: max ulp / fraction differing: {stats}")
    for name, (mx, frac) in stats.items():
        if clip > 0:
            # a clip coefficient that lands on the other side of a bf16 rounding
            # boundary rescales every gradient by one ulp; still ulp-level
            assert mx <= 2 * n_steps + 2, (name, stats[name])
        else:
            assert mx <= 1 and frac < 0.01 * n_steps, (name, stats[name])
