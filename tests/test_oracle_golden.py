"""Pins the oracle: the C restatement (CPU flavour) and the torch port are
replayed on the gradients recorded from the REAL reference TrainLoop
(tests/golden/*.npz, written by oracle/make_golden.py from /root/reference) and
must reproduce what the reference produced."""
import glob
import os

import numpy as np
import pytest
import torch

import oracle
from oracle.torch_port import ReferenceUpdatePort

ALL_GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "ref_*.npz")))
GOLDEN = [p for p in ALL_GOLDEN if "bf16" not in os.path.basename(p)]       # fp32 state
GOLDEN_BF16 = [p for p in ALL_GOLDEN if "bf16" in os.path.basename(p)]      # every state tensor bf16 (F2/F7)


def test_fixtures_present():
    assert len(GOLDEN) == 3 and len(GOLDEN_BF16) == 1


def _close(a, b, rtol=1e-6):
    a, b = a.astype(np.float64), b.astype(np.float64)
    rms = np.sqrt(np.mean(b ** 2))
    return np.abs(a - b) <= rtol * (np.abs(b) + rms)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_c_oracle_reproduces_reference(path):
    z = np.load(path)
    numels = z["numels"]
    off = np.concatenate([[0], np.cumsum(numels)[:-1]])
    p = z["init"].copy()
    m, v = np.zeros_like(p), np.zeros_like(p)
    e = [p.copy(), p.copy()]
    rates, wd, clip = z["ema_rates"], float(z["wd"]), float(z["clip"])
    for s in range(len(z["lr"])):
        g = z["grads"][s].copy()
        coef = 1.0
        if clip > 0:
            r = oracle.grad_norm(oracle.F32, g, off, numels, 1.0, clip)
            # torch-CPU's fp32 norm is itself only good to ~1e-7 here (1e-4 on big tensors)
            assert abs(r["total"] - z["total_norm"][s]) <= 1e-6 * z["total_norm"][s]
            # continue from the coefficient the reference used, so the rest can be compared bit for bit
            t = np.float32(z["total_norm"][s])
            coef = float(min(np.float32(np.float32(1.0) / np.float32(t + np.float32(1e-6))) * np.float32(clip),
                             np.float32(1.0)))
        oracle.adamw_ema_step(oracle.F32, p, g, m, v, e, rates, float(z["lr"][s]), wd, s + 1, coef=coef,
                              grad_mode=1, flavour=oracle.CPU)
        seg_ss = [float((g[o:o + n].astype(np.float64) ** 2).sum()) for o, n in zip(off, numels)]
        assert abs(oracle.logged_norm(oracle.F32, seg_ss) - z["logged"][s]) <= 1e-6 * z["logged"][s]
        # both moments: bit for bit, every step (they chain through the clipped gradient)
        assert np.array_equal(m.view(np.uint32), z["m"][s].view(np.uint32))
        assert np.array_equal(v.view(np.uint32), z["v"][s].view(np.uint32))
        # parameters and EMA: torch-CPU's sqrt is MKL vsSqrt (1 ulp off on ~20% of inputs)
        assert _close(p, z["p"][s]).all()
        assert _close(e[0], z["ema0"][s]).all() and _close(e[1], z["ema1"][s]).all()
        # carry the reference's parameters forward so one-ulp sqrt differences do not compound
        p[:] = z["p"][s]
        e[0][:] = z["ema0"][s]
        e[1][:] = z["ema1"][s]


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_torch_port_reproduces_reference(path):
    """Same torch, same ops, same inputs -> the port must land on the reference's bits."""
    z = np.load(path)
    torch.set_num_threads(1)
    numels = [int(n) for n in z["numels"]]
    parts = np.split(z["init"], np.cumsum(numels)[:-1])
    params = [torch.nn.Parameter(torch.from_numpy(x.copy())) for x in parts]
    port = ReferenceUpdatePort(params, lr=float(z["lr0"]), ema_rates=list(z["ema_rates"]), weight_decay=float(z["wd"]),
                               learning_steps=int(z["learning_steps"]), gradient_clipping=float(z["clip"]))
    for s in range(len(z["lr"])):
        for p, g in zip(params, np.split(z["grads"][s], np.cumsum(numels)[:-1])):
            p.grad = torch.from_numpy(g.copy())
        port.optimize()
        flat = lambda ts: np.concatenate([t.detach().reshape(-1).numpy() for t in ts])
        assert np.array_equal(flat(params).view(np.uint32), z["p"][s].view(np.uint32))
        assert np.array_equal(flat(port.ema_params[0]).view(np.uint32), z["ema0"][s].view(np.uint32))
        assert np.array_equal(flat(port.ema_params[1]).view(np.uint32), z["ema1"][s].view(np.uint32))
        assert abs(port.last_logged_norm - z["logged"][s]) <= 1e-7 * z["logged"][s]
        assert port.opt.param_groups[0]["lr"] == pytest.approx(float(z["lr"][s]), rel=0, abs=0)


def _bf16_ulp_distance(a_bits, b_bits):
    def ordered(b):
        i = b.astype(np.int64)
        return np.where(i >= 0x8000, -(i & 0x7fff), i)
    return np.abs(ordered(a_bits) - ordered(b_bits))


@pytest.mark.parametrize("path", GOLDEN_BF16, ids=[os.path.basename(p)[:-4] for p in GOLDEN_BF16])
def test_c_oracle_bf16_reproduces_reference_trainloop(path):
    """The reference TrainLoop run with a bf16 model (no fp32 master copy, utils/trainer.py:91-92,111-113):
    bf16 per-tensor norms, bf16 total norm and clip coefficient (:252-255 -> clip_grad_norm_), bf16 moments,
    bf16 EMA.  The C oracle's BF16 flavour replayed on the recorded gradients must land on the reference's
    total norm, both moments, the parameters and both EMA copies bit for bit at every step (an fp32 ulp of
    MKL's square root does not survive the rounding to bf16 anywhere in this fixture)."""
    z = np.load(path)
    assert str(z["dtype"]) == "bf16"
    numels = z["numels"]
    off = np.concatenate([[0], np.cumsum(numels)[:-1]])
    bits = oracle.f32_to_bf16_bits
    assert np.array_equal(oracle.bf16_bits_to_f32(bits(z["init"])), z["init"])   # fixture values are bf16 values
    p = bits(z["init"])
    m, v = np.zeros_like(p), np.zeros_like(p)
    e = [p.copy(), p.copy()]
    rates, wd, clip = z["ema_rates"], float(z["wd"]), float(z["clip"])
    worst = 0
    for s in range(len(z["lr"])):
        g = bits(z["grads"][s])
        r = oracle.grad_norm(oracle.BF16, g, off, numels, 1.0, clip)
        assert r["total"] == z["total_norm"][s], (s, r["total"], z["total_norm"][s])   # a bf16 value, exactly
        seg_ss = oracle.adamw_ema_step(oracle.BF16, p, g, m, v, e, rates, float(z["lr"][s]), wd, s + 1, coef=r["coef"],
                                       grad_mode=1, flavour=oracle.CPU)
        assert np.array_equal(m, bits(z["m"][s])), f"exp_avg differs at step {s}"
        assert np.array_equal(v, bits(z["v"][s])), f"exp_avg_sq differs at step {s}"
        for got, want in ((p, z["p"][s]), (e[0], z["ema0"][s]), (e[1], z["ema1"][s])):
            d = _bf16_ulp_distance(got, bits(want))
            worst = max(worst, int(d.max()))
            assert d.max() == 0
        # carry the reference's values forward so a one-ulp difference does not compound
        p[:] = bits(z["p"][s])
        e[0][:] = bits(z["ema0"][s])
        e[1][:] = bits(z["ema1"][s])
    assert worst == 0


def test_c_oracle_flavours_agree_to_rounding_and_bf16_roundtrip():
    rng = np.random.RandomState(0)
    n = 50000
    base = [rng.randn(n).astype(np.float32) * s for s in (0.02, 1e-3)]
    out = {}
    for fl in (oracle.CUDA, oracle.CPU):
        p, g = base[0].copy(), base[1].copy()
        m, v, e = np.zeros(n, np.float32), np.zeros(n, np.float32), base[0].copy()
        for step in (1, 2, 3):
            oracle.adamw_ema_step(oracle.F32, p, g, m, v, [e], [0.999], 1e-3, 0.01, step, flavour=fl)
        out[fl] = (p, m, v, e)
    assert np.array_equal(out[0][1], out[1][1])                 # exp_avg: same formula
    for a, b in zip(out[0], out[1]):
        assert _close(a, b).all()
    x = (rng.randn(10000) * 3).astype(np.float32)
    bits = oracle.f32_to_bf16_bits(x)
    want = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(bits, want)
    assert np.array_equal(oracle.bf16_bits_to_f32(bits), torch.from_numpy(x).to(torch.bfloat16).float().numpy())


def test_c_oracle_bf16_matches_torch_cpu_bf16_rounding_points():
    """with bf16 state every torch op rounds to bf16.  torch-CPU bf16 ops
    compute in fp32 and round once, like the oracle; sqrt is the one op allowed to
    differ (MKL), so compare everything that does not pass through it."""
    rng = np.random.RandomState(1)
    n = 20000
    g32 = (rng.randn(n) * 1e-2).astype(np.float32)
    g_bits = oracle.f32_to_bf16_bits(g32)
    g_t = torch.from_numpy(g_bits.view(np.int16).copy()).view(torch.bfloat16)
    # clip: norms / coefficient are bf16 tensors
    total = torch.nn.utils.get_total_norm([g_t[:7000], g_t[7000:]])
    r = oracle.grad_norm(oracle.BF16, g_bits, [0, 7000], [7000, n - 7000], 1.0, 0.05)
    assert total.dtype == torch.bfloat16 and float(total) == r["total"]
    coef = torch.clamp(0.05 / (total + 1e-6), max=1.0)
    assert float(coef) == r["coef"]
    # moments after one step, bf16 state
    m = torch.zeros(n, dtype=torch.bfloat16)
    v = torch.zeros(n, dtype=torch.bfloat16)
    gc = g_t * coef
    m.lerp_(gc, 1 - 0.9)
    v.mul_(0.999).addcmul_(gc, gc, value=1 - 0.999)
    p_bits = oracle.f32_to_bf16_bits((rng.randn(n) * 0.02).astype(np.float32))
    mo, vo = np.zeros(n, np.uint16), np.zeros(n, np.uint16)
    oracle.adamw_ema_step(oracle.BF16, p_bits, g_bits.copy(), mo, vo, [], [], 1e-3, 0.0, 1, coef=float(coef),
                          flavour=oracle.CPU)
    assert np.array_equal(mo, m.view(torch.int16).numpy().view(np.uint16))
    assert np.array_equal(vo, v.view(torch.int16).numpy().view(np.uint16))


def test_c_oracle_tracks_the_float64_truth():
    """T0: the fp32 restatement stays within fp32 rounding of the same recurrence carried in float64."""
    from oracle import truth_fp64
    rng = np.random.RandomState(3)
    n, steps = 30000, 20
    p32 = (rng.randn(n) * 0.02).astype(np.float32)
    m32, v32, e32 = np.zeros(n, np.float32), np.zeros(n, np.float32), p32.copy()
    p64, m64, v64, e64 = (a.astype(np.float64) for a in (p32, m32, v32, e32))
    for t in range(1, steps + 1):
        g = (rng.randn(n) * 0.05).astype(np.float32)
        r = oracle.grad_norm(oracle.F32, g, [0], [n], 1.0, 1.0)
        nrm, coef = truth_fp64.step(p64, g.astype(np.float64), m64, v64, [e64], [0.999], 1e-3, 0.01, t, max_norm=1.0)
        assert abs(r["total"] - nrm) <= 1e-6 * nrm and abs(r["coef"] - coef) <= 1e-6 * coef
        oracle.adamw_ema_step(oracle.F32, p32, g.copy(), m32, v32, [e32], [0.999], 1e-3, 0.01, t, coef=r["coef"])
    for a32, a64 in ((p32, p64), (m32, m64), (v32, v64), (e32, e64)):
        assert _close(a32, a64, rtol=2e-6).all()


def test_uniform_sampler_draws_like_upstream():
    from distributed_pipeline_b200.sampler import UniformSampler, create_named_schedule_sampler
    s = create_named_schedule_sampler("uniform", 2000)
    assert isinstance(s, UniformSampler)
    np.random.seed(11)
    idx, w = s.sample(64, "cpu")
    np.random.seed(11)
    want = np.random.choice(2000, size=(64,), p=np.ones(2000) / 2000)
    assert np.array_equal(idx.numpy(), want) and idx.dtype == torch.int64
    assert torch.equal(w, torch.ones(64)) and w.dtype == torch.float32
