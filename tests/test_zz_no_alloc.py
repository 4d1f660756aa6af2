"""dpk_grad_norm_static: the gradient-norm entry point that allocates nothing.

It is meant for callers that cannot allocate device memory themselves -- code
running in the middle of a step on a GPU whose free memory is held by the
framework's caching allocator, where the cudaMalloc that dpk_create/dpk_bind
perform could fail, fragment the remainder or stall the device.  Its scratch is
a fixed 512 KiB array inside the library image, shared by every call.

Checked here, through the C ABI only: two calls in a row (different sizes and
dtypes, so the second runs over what the first left in the shared scratch) each
give what dpk_grad_norm gives on a one-tensor handle -- same arithmetic and
reduction order, so the comparison is exact, well inside the 1e-5 the other norm
tests use -- and oversize input is refused with DPK_E_LIMIT.

Collected last on purpose (file name): a run that stops at the first failure has
then already run every other test.
"""
import pytest
import torch

from distributed_pipeline_b200 import _lib
from distributed_pipeline_b200.engine import Engine
from distributed_pipeline_b200.flat import FlatLayout

pytestmark = pytest.mark.gpu


def _handle_norm(g, inv_scale, max_norm):
    """The existing entry point: dpk_create + dpk_bind + dpk_grad_norm on one tensor of g.numel() elements."""
    es = g.element_size()
    lay = FlatLayout.from_numels([g.numel()], es)
    pad = lambda t: torch.cat([t, torch.zeros(lay.n_elems - t.numel(), dtype=t.dtype, device=t.device)])
    bufs = [pad(g)] + [torch.zeros(lay.n_elems, dtype=g.dtype, device=g.device) for _ in range(3)]
    eng = Engine(lay, g.dtype, g.device, [])
    eng.bind(bufs[1], bufs[0], bufs[2], bufs[3], [])
    eng.grad_norm(inv_scale, max_norm)
    res = eng.read_results_sync()
    eng.close()
    return res


def _static_norm(g, inv_scale, max_norm, out):
    lib = _lib.lib()
    stream = torch.cuda.current_stream(g.device).cuda_stream
    dt = _lib.F32 if g.dtype == torch.float32 else _lib.BF16
    _lib.check(lib.dpk_grad_norm_static(g.data_ptr(), g.numel(), dt, inv_scale, max_norm, out.data_ptr(), stream))
    torch.cuda.synchronize(g.device)
    return out.cpu().tolist()


def test_static_norm_matches_handle_norm_on_consecutive_calls():
    dev = torch.device("cuda", torch.cuda.current_device())
    gen = torch.Generator(device=dev).manual_seed(7)
    out = torch.full((_lib.R_COUNT,), -7.0, dtype=torch.float64, device=dev)  # caller-owned output block
    cases = [
        # first call: many tiles, ragged tail, clipping active
        ((torch.randn(3_000_003, generator=gen, device=dev) * 0.5), 1.0, 2.0),
        # second call straight after, smaller: must not see the first call's partial sums
        ((torch.randn(70_001, generator=gen, device=dev) * 1e-3).to(torch.bfloat16), 1.0 / 64, 1.0),
        ((torch.randn(5, generator=gen, device=dev)), 1.0, 0.0),
    ]
    for g, inv_scale, max_norm in cases:
        got = _static_norm(g, inv_scale, max_norm, out)
        want = _handle_norm(g, inv_scale, max_norm)
        for idx in (_lib.R_TOTAL_NORM, _lib.R_CLIP_COEF, _lib.R_NONFINITE):
            assert abs(got[idx] - want[idx]) <= 1e-5 * abs(want[idx]), (idx, got, want)
            assert got[idx] == want[idx], "same reduction order: expected identical bits"
        # entries the call does not own are left alone
        assert got[_lib.R_LOGGED_NORM] == -7.0 and got[_lib.R_SKIPPED] == -7.0
        truth = float(torch.linalg.vector_norm((g.float() * inv_scale).to(g.dtype).double()))
        tol = 1e-5 if g.dtype == torch.float32 else 2 ** -7
        assert abs(got[_lib.R_TOTAL_NORM] - truth) <= tol * truth
    # a non-finite element is reported, not trapped
    g = torch.randn(20_000, generator=gen, device=dev)
    g[12_345] = float("inf")
    got = _static_norm(g, 1.0, 1.0, out)
    assert got[_lib.R_NONFINITE] == 1.0
    # empty input
    got = _static_norm(torch.empty(0, device=dev), 1.0, 1.0, out)
    assert got[_lib.R_TOTAL_NORM] == 0.0 and got[_lib.R_NONFINITE] == 0.0 and got[_lib.R_CLIP_COEF] == 1.0


def test_static_norm_at_its_maximum_size():
    """Exactly DPK_STATIC_MAX_ELEMS elements: all 65 536 scratch slots in use, nothing refused."""
    dev = torch.device("cuda", torch.cuda.current_device())
    out = torch.zeros(_lib.R_COUNT, dtype=torch.float64, device=dev)
    g = torch.full((_lib.STATIC_MAX_ELEMS,), 0.5, dtype=torch.bfloat16, device=dev)   # 1 GiB
    g[-1] = 3.0
    got = _static_norm(g, 1.0, 100.0, out)
    want = _handle_norm(g, 1.0, 100.0)
    truth = ((_lib.STATIC_MAX_ELEMS - 1) * 0.25 + 9.0) ** 0.5
    assert got[_lib.R_TOTAL_NORM] == want[_lib.R_TOTAL_NORM] and got[_lib.R_CLIP_COEF] == want[_lib.R_CLIP_COEF]
    assert abs(got[_lib.R_TOTAL_NORM] - truth) <= 2 ** -7 * truth and got[_lib.R_NONFINITE] == 0.0
    # and the call after it, a small one, is not disturbed by 65 536 stale partials
    small = torch.randn(10_000, device=dev)
    assert _static_norm(small, 1.0, 0.0, out)[_lib.R_TOTAL_NORM] == _handle_norm(small, 1.0, 0.0)[_lib.R_TOTAL_NORM]


def test_static_norm_rejects_what_it_cannot_hold():
    lib = _lib.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    out = torch.zeros(_lib.R_COUNT, dtype=torch.float64, device=dev)
    g = torch.zeros(64, device=dev)
    rc = lib.dpk_grad_norm_static(g.data_ptr(), _lib.STATIC_MAX_ELEMS + 1, _lib.F32, 1.0, 1.0, out.data_ptr(), None)
    assert rc == _lib.E_LIMIT and b"exceed" in lib.dpk_last_error()
    assert lib.dpk_grad_norm_static(g.data_ptr() + 4, 8, _lib.F32, 1.0, 1.0, out.data_ptr(), None) == _lib.E_BADARG
    assert lib.dpk_grad_norm_static(g.data_ptr(), 8, 7, 1.0, 1.0, out.data_ptr(), None) == _lib.E_BADARG
    assert lib.dpk_grad_norm_static(g.data_ptr(), 8, _lib.F32, 1.0, 1.0, None, None) == _lib.E_BADARG
