"""Property test: random layouts (empty, 1-element, ragged, tile-straddling tensors), dtypes, EMA counts, gradient
modes, decay, loss scale and clipping -- the fused path must land on the C oracle's bits every time."""
import pytest
import torch
from hypothesis import HealthCheck, given, settings, strategies as st

import oracle
from distributed_pipeline_b200 import _lib
from helpers import FlatProblem, assert_bits_equal, to_np

pytestmark = pytest.mark.gpu

sizes = st.lists(st.one_of(st.integers(0, 40), st.integers(1000, 1100), st.integers(2040, 2060), st.integers(8180, 8200),
                           st.integers(16380, 16390), st.integers(30000, 70000)), min_size=1, max_size=9)


@settings(max_examples=25, deadline=None, derandomize=True, database=None,
          suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])
@given(numels=sizes, bf16=st.booleans(), n_ema=st.integers(0, 4), grad_mode=st.sampled_from([0, 1, 2]),
       wd=st.sampled_from([0.0, 0.01]), lg_scale=st.sampled_from([0, 3, 14]), clip=st.sampled_from([0.0, 0.3, 1e6]),
       tile=st.sampled_from([0, 1024, 2048, 4096]), seed=st.integers(0, 10 ** 6))
def test_random_layouts_match_oracle(numels, bf16, n_ema, grad_mode, wd, lg_scale, clip, tile, seed):
    dtype = torch.bfloat16 if bf16 else torch.float32
    rates = [0.9999, 0.999, 0.5, 0.99][:n_ema]
    pb = FlatProblem(numels, dtype, rates, seed=seed, tile_elems=tile, grad_std=0.05 * 2.0 ** lg_scale,
                     bucket_cap_bytes=64 * 1024)
    inv_scale = 2.0 ** -lg_scale
    n_buckets = len(pb.engine.buckets)
    for step in (1, 2):
        g = pb.new_grads()
        # alternate between the one-call and the bucket-granular forms
        if step == 1:
            pb.engine.grad_norm(inv_scale, clip)
        else:
            for b in range(n_buckets):
                pb.engine.grad_norm_bucket(b, inv_scale)
            pb.engine.grad_norm_finish(clip)
        res = pb.engine.read_results_sync()
        ref = pb.oracle_norm(g, inv_scale, clip)
        assert res[_lib.R_NONFINITE] == 0.0
        tol = 1e-5 if dtype == torch.float32 else 2 ** -7
        assert abs(res[_lib.R_TOTAL_NORM] - ref["total"]) <= tol * max(ref["total"], 1e-30)
        if step == 1:
            pb.engine.fused_update(3e-4, wd, step, inv_scale=inv_scale, use_clip_coef=clip > 0, grad_mode=grad_mode)
        else:
            for b in reversed(range(n_buckets)):
                pb.engine.fused_update_bucket(b, 3e-4, wd, step, inv_scale=inv_scale, use_clip_coef=clip > 0,
                                              grad_mode=grad_mode)
            pb.engine.fused_update_finish()
        coef = res[_lib.R_CLIP_COEF] if clip > 0 else 1.0
        seg_ss = pb.oracle_step(g, 3e-4, wd, step, inv_scale=inv_scale, coef=coef, grad_mode=grad_mode)
        pb.assert_state_equal()
        m = pb.mask.numpy()
        assert_bits_equal(to_np(pb.g)[m], g[m], "g")
        logged = oracle.logged_norm(oracle.F32 if dtype == torch.float32 else oracle.BF16, seg_ss)
        got = pb.engine.read_results_sync()[_lib.R_LOGGED_NORM]
        assert abs(got - logged) <= (1e-5 if dtype == torch.float32 else 4e-3) * max(logged, 1e-30)
    pb.engine.close()
