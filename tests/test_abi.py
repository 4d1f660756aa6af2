"""The C-ABI library loads without a GPU and exports every symbol include/dpk.h
declares; the ctypes binding lists exactly those; argument validation that needs
no device behaves as documented."""
import ctypes as C
import os
import re

import pytest

from distributed_pipeline_b200 import _lib
from distributed_pipeline_b200.engine import plan_buckets
from distributed_pipeline_b200.flat import FlatLayout

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "dpk.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dpk_[a-z0-9_]+)\s*\(", text)))


def test_header_binding_and_library_agree():
    decl = declared_symbols()
    assert len(decl) >= 25
    assert sorted(_lib.SIGNATURES) == decl
    lib = C.CDLL(_lib.LIB_PATH)
    for name in decl:
        assert hasattr(lib, name), f"{name} declared in dpk.h but not exported"


def test_constants_match_header():
    text = open(os.path.join(ROOT, "include", "dpk.h")).read()
    defs = dict(re.findall(r"#define\s+(DPK_[A-Z0-9_]+)\s+\(?(-?\d+)\)?", text))
    assert int(defs["DPK_ABI_VERSION"]) == _lib.ABI_VERSION == _lib.lib().dpk_abi_version()
    assert int(defs["DPK_MAX_EMA"]) == _lib.MAX_EMA
    assert (int(defs["DPK_F32"]), int(defs["DPK_BF16"])) == (_lib.F32, _lib.BF16)
    assert [int(defs[k]) for k in ("DPK_GRAD_KEEP", "DPK_GRAD_CLIPPED", "DPK_GRAD_ZERO")] == [0, 1, 2]
    assert int(defs["DPK_R_COUNT"]) == _lib.R_COUNT
    assert [int(defs[k]) for k in ("DPK_E_BADARG", "DPK_E_NCCL", "DPK_E_STATE", "DPK_E_NOMEM", "DPK_E_LIMIT")] == \
        [_lib.E_BADARG, _lib.E_NCCL, _lib.E_STATE, _lib.E_NOMEM, _lib.E_LIMIT]


def _cfg(offsets, numels, n_elems, dtype=0, n_ema=0, tile=0, abi=_lib.ABI_VERSION):
    n = len(offsets)
    keep = ((C.c_int64 * max(n, 1))(*offsets), (C.c_int64 * max(n, 1))(*numels), (C.c_double * 4)(0.9, 0.9, 0.9, 0.9))
    cfg = _lib.Config(abi_version=abi, dtype=dtype, device=0, n_segments=n, n_ema=n_ema, tile_elems=tile,
                      n_elems=n_elems, seg_offsets=keep[0], seg_numel=keep[1], ema_rates=keep[2], beta1=0.9,
                      beta2=0.999, eps=1e-8, bucket_cap_bytes=0, edge_bucket_bytes=0)
    return cfg, keep


@pytest.mark.parametrize("kw,code", [
    (dict(abi=99), _lib.E_BADARG),
    (dict(dtype=5), _lib.E_BADARG),
    (dict(n_ema=5), _lib.E_LIMIT),
    (dict(tile=1000), _lib.E_BADARG),
])
def test_create_rejects_bad_config(kw, code):
    lib = _lib.lib()
    cfg, keep = _cfg([0], [100], 128, **kw)
    h = C.c_void_p()
    assert lib.dpk_create(C.byref(h), C.byref(cfg)) == code
    assert lib.dpk_last_error()


@pytest.mark.parametrize("offsets,numels,n_elems", [
    ([0, 2], [1, 1], 8),       # second tensor not on a 16-byte boundary
    ([0, 4], [8, 1], 16),      # overlap
    ([0], [100], 64),          # runs past the buffer
    ([0], [5], 6),             # n_elems not a whole number of vectors
])
def test_create_rejects_bad_layout(offsets, numels, n_elems):
    lib = _lib.lib()
    cfg, keep = _cfg(offsets, numels, n_elems)
    h = C.c_void_p()
    assert lib.dpk_create(C.byref(h), C.byref(cfg)) == _lib.E_BADARG


def test_calls_before_bind_report_state_error():
    lib = _lib.lib()
    cfg, keep = _cfg([0], [100], 128)
    h = C.c_void_p()
    assert lib.dpk_create(C.byref(h), C.byref(cfg)) == 0
    assert lib.dpk_grad_norm(h, 1.0, 1.0, None) == _lib.E_STATE
    assert lib.dpk_fused_update(h, 1e-3, 0.0, 1, 1.0, 0, 0, 0, None) == _lib.E_STATE
    assert lib.dpk_zero_grads(h, None) == _lib.E_STATE
    assert b"not bound" in lib.dpk_last_error()
    assert lib.dpk_world_size(h) == 1
    # options and the symmetric-memory binding validate their arguments without a device
    assert lib.dpk_set_option(h, _lib.OPT_OVERLAP_BLOCKS_PER_SM, 2) == 0
    assert lib.dpk_set_option(h, _lib.OPT_OVERLAP_BLOCKS_PER_SM, -1) == _lib.E_BADARG
    assert lib.dpk_set_option(h, 99, 0) == _lib.E_BADARG
    assert lib.dpk_bind_nvls(h, None, None, None, None, 0) == _lib.E_BADARG
    ptrs = (C.c_void_p * 2)(16, 32)
    assert lib.dpk_bind_nvls(h, 16, ptrs, 16, 16, 8) == _lib.E_STATE      # not bound, single rank
    assert lib.dpk_wait_bucket(h, 7, None) == _lib.E_BADARG
    assert lib.dpk_grad_norm_bucket(h, 0, 1.0, None) == _lib.E_STATE
    assert lib.dpk_fused_update_finish(h, 0, None) == _lib.E_STATE
    assert lib.dpk_destroy(h) == 0
    with pytest.raises(_lib.DpkError):
        _lib.check(lib.dpk_grad_norm(None, 1.0, 1.0, None))


def test_bucket_plan_tail_first_and_capped():
    # 4-byte elements; cap 1000 bytes -> a bucket closes once it holds >= 250 elements
    lay = FlatLayout.from_numels([100, 300, 10, 10, 500, 20], 4)
    ranges = plan_buckets(lay, __import__("torch").float32, bucket_cap_bytes=1000, edge_bucket_bytes=1000)
    # walk from the last tensor: [20,500] closes (520>=250); [10,10,300] closes; [100] is what is left
    assert ranges[0][0] == lay.offsets[4] and ranges[0][0] + ranges[0][1] == lay.n_elems
    assert ranges[1] == (lay.offsets[1], lay.offsets[4] - lay.offsets[1])
    assert ranges[2] == (0, lay.offsets[1])
    # buckets tile the whole buffer without overlap
    covered = sorted(ranges)
    assert covered[0][0] == 0
    for (s0, c0), (s1, _) in zip(covered, covered[1:]):
        assert s0 + c0 == s1
    assert covered[-1][0] + covered[-1][1] == lay.n_elems
    # DiffuSeq-base at the reference's 128 MiB cap: 3 buckets fp32, 2 bf16, like DDP's rebuilt
    # buckets ...
    import torch
    from distributed_pipeline_b200.synthetic import diffuseq_shapes, numel
    numels = [numel(s) for s in diffuseq_shapes()]
    cap = 128 << 20
    lay4, lay2 = FlatLayout.from_numels(numels, 4), FlatLayout.from_numels(numels, 2)
    assert len(plan_buckets(lay4, torch.float32, cap, edge_bucket_bytes=cap)) == 3
    assert len(plan_buckets(lay2, torch.bfloat16, cap, edge_bucket_bytes=cap)) == 2
    # ... plus, by default, a small first bucket (tail) and a small last one (head)
    r4 = plan_buckets(lay4, torch.float32)
    assert len(r4) == 5 and r4[-1][0] == 0 and 16 << 20 <= r4[-1][1] * 4 <= 32 << 20
    assert r4[0][0] + r4[0][1] == lay4.n_elems and 16 << 20 <= r4[0][1] * 4 <= 32 << 20
    assert sorted(r4)[0][0] == 0 and sum(c for _, c in r4) == lay4.n_elems
    # the split never produces an empty or out-of-order bucket
    tiny = plan_buckets(FlatLayout.from_numels([10, 10, 10], 4), torch.float32, bucket_cap_bytes=1 << 20)
    assert tiny == [(0, FlatLayout.from_numels([10, 10, 10], 4).n_elems)]
