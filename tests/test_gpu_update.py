"""Parity of the fused update (pass 2) and the norm pass (pass 1) with the C
oracle, through the C ABI.  fp32 and bf16 state, bit-exact where the oracle is
given the same clip coefficient; the norm itself to the 1e-5 `north_star` states.
"""
import math

import pytest
import torch

import oracle
from distributed_pipeline_b200 import _lib
from helpers import FlatProblem, assert_bits_equal, from_np, rel_err, to_np

pytestmark = pytest.mark.gpu

RAGGED = [5, 0, 1, 130, 7, 4096, 8192, 8193, 33, 20000, 3, 64, 12289]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n_ema", [0, 1, 2, 3, 4])
def test_update_bit_exact_multi_step(dtype, n_ema):
    rates = [0.9999, 0.999, 0.5, 0.99][:n_ema]
    pb = FlatProblem(RAGGED, dtype, rates, seed=n_ema, tile_elems=1024)
    for step in range(1, 6):
        g = pb.new_grads()
        lr = 1e-3 * (1 - step / 50)
        pb.engine.fused_update(lr, 0.01, step, grad_mode=_lib.GRAD_KEEP)
        pb.oracle_step(g, lr, 0.01, step)
        pb.assert_state_equal()
        # gradients untouched in KEEP mode, padding included
        assert_bits_equal(to_np(pb.g), g, "g")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("grad_mode", [_lib.GRAD_CLIPPED, _lib.GRAD_ZERO])
@pytest.mark.parametrize("wd", [0.0, 0.05])
def test_grad_modes_clip_and_unscale(dtype, grad_mode, wd):
    pb = FlatProblem(RAGGED, dtype, [0.9999, 0.999], seed=7, grad_std=3.0)
    inv_scale = 1.0 / 2 ** 7
    for step in range(1, 4):
        g = pb.new_grads()
        pb.engine.grad_norm(inv_scale, 1.0)
        res = pb.engine.read_results_sync()
        ref = pb.oracle_norm(g, inv_scale, 1.0)
        assert res[_lib.R_NONFINITE] == 0.0
        assert rel_err(res[_lib.R_TOTAL_NORM], ref["total"]) < (1e-5 if dtype == torch.float32 else 2 ** -8)
        coef = res[_lib.R_CLIP_COEF]
        assert coef < 1.0  # clip active
        pb.engine.fused_update(1e-4, wd, step, inv_scale=inv_scale, use_clip_coef=True, grad_mode=grad_mode)
        seg_ss = pb.oracle_step(g, 1e-4, wd, step, inv_scale=inv_scale, coef=coef, grad_mode=grad_mode)
        pb.assert_state_equal()
        got_g = to_np(pb.g)
        m = pb.mask.numpy()
        assert_bits_equal(got_g[m], g[m], "g (valid lanes)")
        # padding untouched (still NaN) in CLIPPED mode
        if grad_mode == _lib.GRAD_CLIPPED and (~m).any():
            pad = from_np(got_g, dtype, "cpu")[torch.from_numpy(~m)]
            assert torch.isnan(pad.float()).all()
        res = pb.engine.read_results_sync()
        logged = oracle.logged_norm(oracle.F32 if dtype == torch.float32 else oracle.BF16, seg_ss)
        # bf16: each (g**2).sum() is a bf16 value; a sum on a rounding boundary moves by one bf16 ulp
        tol = 1e-5 if dtype == torch.float32 else 4e-3
        assert rel_err(res[_lib.R_LOGGED_NORM], logged) < tol
        assert res[_lib.R_SKIPPED] == 0.0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_norm_matches_oracle_and_segments(dtype):
    pb = FlatProblem([3, 100000, 777, 0, 8192 * 3 + 5, 1], dtype, [], seed=3, tile_elems=2048)
    g = pb.new_grads(std=0.5)
    pb.engine.grad_norm(1.0, 2.5)
    res = pb.engine.read_results_sync()
    ref = pb.oracle_norm(g, 1.0, 2.5)
    seg = pb.engine.segment_norms.cpu().numpy()
    if dtype == torch.float32:
        assert rel_err(seg, ref["seg_norms"], 1e-30).max() < 1e-6
        assert rel_err(res[_lib.R_TOTAL_NORM], ref["total"]) < 1e-6
        assert rel_err(res[_lib.R_CLIP_COEF], ref["coef"]) < 1e-6
    else:
        # bf16: per-tensor norms, total and coefficient are bf16 values;
        # the fp32 sums underneath agree to ~1e-7, so the rounded values coincide
        # except on a rounding boundary (one bf16 ulp)
        assert (rel_err(seg, ref["seg_norms"], 1e-30) <= 2 ** -7).all()
        assert (seg == ref["seg_norms"]).mean() >= 0.8
        assert rel_err(res[_lib.R_TOTAL_NORM], ref["total"]) <= 2 ** -7
        c = torch.tensor(res[_lib.R_CLIP_COEF], dtype=torch.float64)
        assert c.to(torch.bfloat16).double() == c  # coefficient is bf16-representable


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_bucket_granular_passes_equal_whole_buffer_passes(dtype):
    """dpk_*_bucket over every bucket + _finish == the one-call forms, bit for bit."""
    numels = [5000, 3, 70000, 128, 9000, 33, 40000]
    a = FlatProblem(numels, dtype, [0.999, 0.9], seed=9, bucket_cap_bytes=64 * 1024)
    b = FlatProblem(numels, dtype, [0.999, 0.9], seed=9, bucket_cap_bytes=64 * 1024, mirror=False)
    n_buckets = len(a.engine.buckets)
    assert n_buckets >= 3
    for step in (1, 2):
        g = a.new_grads(std=0.3)
        b.g.copy_(a.g)
        a.engine.grad_norm(1.0, 1.0)
        a.engine.fused_update(1e-3, 0.01, step, use_clip_coef=True, grad_mode=_lib.GRAD_CLIPPED)
        for k in reversed(range(n_buckets)):           # any order
            b.engine.wait_bucket(k)                    # no all-reduce outstanding: no-op
            b.engine.grad_norm_bucket(k, 1.0)
        b.engine.grad_norm_finish(1.0)
        for k in range(n_buckets):
            b.engine.fused_update_bucket(k, 1e-3, 0.01, step, use_clip_coef=True, grad_mode=_lib.GRAD_CLIPPED)
        b.engine.fused_update_finish()
        ra, rb = a.engine.read_results_sync(), b.engine.read_results_sync()
        assert ra == rb
        for x, y, name in zip((a.p, a.m, a.v, a.g, *a.emas), (b.p, b.m, b.v, b.g, *b.emas), "pmvgee"):
            m = a.mask.to(x.device)
            assert torch.equal(x[m], y[m]), name
    from distributed_pipeline_b200 import DpkError
    with pytest.raises(DpkError):
        b.engine.fused_update_bucket(n_buckets, 1e-3, 0.0, 1)


def test_clip_inactive_gives_unit_coef():
    pb = FlatProblem([1000, 50], torch.float32, [0.99], seed=5)
    g = pb.new_grads(std=1e-4)
    pb.engine.grad_norm(1.0, 1.0)
    res = pb.engine.read_results_sync()
    assert res[_lib.R_CLIP_COEF] == 1.0
    pb.engine.fused_update(1e-3, 0.0, 1, use_clip_coef=True)
    pb.oracle_step(g, 1e-3, 0.0, 1, coef=1.0)
    pb.assert_state_equal()


@pytest.mark.parametrize("bad", [float("inf"), float("-inf"), float("nan")])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_nonfinite_flag_and_predicated_skip(dtype, bad):
    pb = FlatProblem([4096, 17, 9000], dtype, [0.999], seed=11)
    g = pb.new_grads(poison_padding=False)
    before = {k: to_np(t) for k, t in (("p", pb.p), ("m", pb.m), ("v", pb.v), ("e", pb.emas[0]))}
    pb.g[pb.layout.offsets[2] + 8999] = bad  # last valid element of the last tensor
    pb.engine.grad_norm(1.0 / 1024, 0.0)
    res = pb.engine.read_results_sync()
    assert res[_lib.R_NONFINITE] == 1.0
    pb.engine.fused_update(1e-3, 0.01, 1, inv_scale=1.0 / 1024, skip_if_nonfinite=True, grad_mode=_lib.GRAD_ZERO)
    res = pb.engine.read_results_sync()
    assert res[_lib.R_SKIPPED] == 1.0
    for k, t in (("p", pb.p), ("m", pb.m), ("v", pb.v), ("e", pb.emas[0])):
        assert_bits_equal(to_np(t), before[k], k)  # the step did not happen
    # a clean gradient afterwards steps normally
    g = pb.new_grads(poison_padding=False)
    pb.engine.grad_norm(1.0 / 1024, 0.0)
    pb.engine.fused_update(1e-3, 0.01, 1, inv_scale=1.0 / 1024, skip_if_nonfinite=True)
    res = pb.engine.read_results_sync()
    assert res[_lib.R_NONFINITE] == 0.0 and res[_lib.R_SKIPPED] == 0.0
    pb.oracle_step(g, 1e-3, 0.01, 1, inv_scale=1.0 / 1024)
    pb.assert_state_equal()


def test_padding_nonfinite_is_ignored():
    pb = FlatProblem([5, 3, 1000], torch.float32, [], seed=2)
    pb.new_grads(poison_padding=True)
    pb.engine.grad_norm(1.0, 1.0)
    res = pb.engine.read_results_sync()
    assert res[_lib.R_NONFINITE] == 0.0 and math.isfinite(res[_lib.R_TOTAL_NORM])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_standalone_ema_update_alignment_cases(dtype):
    from distributed_pipeline_b200 import ema_update
    dt = oracle.F32 if dtype == torch.float32 else oracle.BF16
    base_t = (torch.randn(5000) * 0.1).to(dtype).cuda()
    base_s = (torch.randn(5000) * 0.1).to(dtype).cuda()
    for off_t, off_s, n in [(0, 0, 5000), (1, 1, 4000), (3, 3, 1), (1, 2, 3000), (0, 0, 0), (5, 5, 7)]:
        t = base_t[off_t:off_t + n].clone() if off_t == 0 else base_t[off_t:off_t + n]
        s = base_s[off_s:off_s + n]
        want_t, src = to_np(t), to_np(s)
        oracle.ema_update(dt, want_t, src, 0.9999)
        ema_update(t, s, 0.9999)
        assert_bits_equal(to_np(t), want_t, f"ema off=({off_t},{off_s}) n={n}")


def test_empty_and_degenerate_layouts():
    for numels in ([0], [0, 0, 0], [1]):
        pb = FlatProblem(numels, torch.float32, [0.9], seed=1)
        g = pb.new_grads(poison_padding=False)
        pb.engine.grad_norm(1.0, 1.0)
        pb.engine.fused_update(1e-3, 0.0, 1, use_clip_coef=True)
        res = pb.engine.read_results_sync()
        ref = pb.oracle_norm(g, 1.0, 1.0)
        assert abs(res[_lib.R_TOTAL_NORM] - ref["total"]) <= 1e-6 * max(ref["total"], 1e-30)
        pb.oracle_step(g, 1e-3, 0.0, 1, coef=res[_lib.R_CLIP_COEF])
        pb.assert_state_equal()


def test_error_paths():
    from distributed_pipeline_b200 import DpkError
    pb = FlatProblem([100], torch.float32, [0.9], seed=1)
    with pytest.raises(DpkError) as e:
        pb.engine.fused_update(1e-3, 0.0, 0)
    assert e.value.code == _lib.E_BADARG
    with pytest.raises(DpkError):
        pb.engine.fused_update(1e-3, 0.0, 1, grad_mode=9)
    from distributed_pipeline_b200.engine import Engine
    eng = Engine(pb.layout, torch.float32, "cuda:0", [0.9])
    with pytest.raises(DpkError) as e:
        eng.grad_norm()
    assert e.value.code == _lib.E_STATE
    with pytest.raises(ValueError):
        Engine(pb.layout, torch.float32, "cuda:0", [0.1] * 5)
    with pytest.raises(RuntimeError):
        Engine(pb.layout, torch.float32, "cpu", [])


def test_arithmetic_shortcuts_are_exact_for_every_input():
    """The fused kernel divides by the per-step scalar sqrt(1-beta2^t) with a
    reciprocal + two FMAs and, for bf16 state, takes sqrt.approx before rounding
    to bf16.  Both must be indistinguishable from div.rn / sqrt.rn: checked on
    the device for the square root of every fp32 bit pattern (the only dividends
    the kernel has), per divisor, and for all 2^16 bf16 inputs."""
    lib = _lib.lib()
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    steps = [1, 2, 3, 10, 100, 1000, 5000, 100000]
    for t in steps:
        c = (1 - 0.999 ** t) ** 0.5
        _lib.check(lib.dpk_selftest_arith(c, out.data_ptr(), stream))
        bad_div, bad_sqrt = out.cpu().tolist()
        assert bad_div == 0, f"scalar division differs from div.rn for {bad_div} inputs at step {t} (c={c})"
        assert bad_sqrt == 0
    for c in (0.3, 1.0, 0.7071067811865476, 0.0316, 0.999):
        _lib.check(lib.dpk_selftest_arith(c, out.data_ptr(), stream))
        assert out.cpu().tolist() == [0, 0], c


@pytest.mark.parametrize("clip", [-1.0, 0.5])
def test_step_from_host_equals_resident_step(clip):
    """FusedAdamW.step_from_host (bucket-pipelined upload) == copy everything, then step."""
    from distributed_pipeline_b200 import FusedAdamW
    torch.manual_seed(3)
    shapes = [(5000,), (3,), (300, 257), (128,), (9000,), (33,), (40000,)]

    def make():
        torch.manual_seed(3)
        ps = [torch.nn.Parameter(torch.randn(s, device="cuda") * 0.02) for s in shapes]
        return FusedAdamW(ps, lr=1e-3, weight_decay=0.01, ema_rates=[0.999], bucket_cap_mb=0.05, grad_mode=_lib.GRAD_KEEP)
    a, b = make(), make()
    assert len(a.engine.buckets) >= 3
    host = torch.empty(a.flat.layout.n_elems, pin_memory=True)
    for step in range(3):
        host.normal_(0, 0.1)
        a.step_from_host(host, clip)
        b.flat.grads.copy_(host, non_blocking=True)
        if clip > 0:
            b.clip_grad_norm(clip)
        b.step()
        torch.cuda.synchronize()
        assert a.engine.read_results_sync() == b.engine.read_results_sync()
        for x, y in ((a.flat.params, b.flat.params), (a.flat.exp_avg, b.flat.exp_avg),
                     (a.flat.exp_avg_sq, b.flat.exp_avg_sq), (a.flat.emas[0], b.flat.emas[0])):
            assert torch.equal(x, y)
    assert a.step_count == b.step_count == 3


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_tma_staged_norm_pass_is_bit_identical(dtype):
    """DPK_OPT_NORM_STAGING: tiles bulk-copied to shared memory by the TMA engine; same sums, same bits."""
    pb = FlatProblem(RAGGED + [100_000, 8192 * 5, 17], dtype, [], seed=21)
    for inv_scale, max_norm in ((1.0, 1.0), (1.0 / 32, 0.5)):
        g = pb.new_grads(std=0.7)
        pb.engine.set_option(_lib.OPT_NORM_STAGING, 0)
        pb.engine.grad_norm(inv_scale, max_norm)
        a = pb.engine.read_results_sync()
        seg_a = pb.engine.segment_norms.clone()
        pb.engine.set_option(_lib.OPT_NORM_STAGING, 1)
        pb.engine.grad_norm(inv_scale, max_norm)
        b = pb.engine.read_results_sync()
        assert a == b and torch.equal(seg_a, pb.engine.segment_norms)
        ref = pb.oracle_norm(g, inv_scale, max_norm)
        assert rel_err(b[_lib.R_TOTAL_NORM], ref["total"]) < (1e-5 if dtype == torch.float32 else 2 ** -7)
    pb.g[pb.layout.offsets[-2] + 5] = float("nan")
    pb.engine.grad_norm(1.0, 1.0)
    assert pb.engine.read_results_sync()[_lib.R_NONFINITE] == 1.0
