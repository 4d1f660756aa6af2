"""ctypes binding of ``libdpk.so`` (C ABI declared in ``include/dpk.h``).

This is the whole "extension": plain pointers and sizes go down, ints come
back, and a non-zero return is turned into ``DpkError`` so torch-elastic's
``@record`` (reference ``run/train.py:9``) sees an ordinary Python exception.
ctypes drops the GIL around every foreign call, which matters because
``dpk_bucket_ready`` is invoked from autograd's engine thread.

There is no CPU fallback: if the library cannot be loaded the import of this
module's ``lib()`` raises.
"""
import ctypes as C
import os
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdpk.so")

ABI_VERSION = 2
F32, BF16 = 0, 1
MAX_EMA = 4
GRAD_KEEP, GRAD_CLIPPED, GRAD_ZERO = 0, 1, 2
R_TOTAL_NORM, R_CLIP_COEF, R_LOGGED_NORM, R_NONFINITE, R_SKIPPED, R_COMM_TIMEOUT, R_COUNT = 0, 1, 2, 3, 4, 5, 8
E_BADARG, E_NCCL, E_STATE, E_NOMEM, E_LIMIT = -1, -2, -3, -4, -5
STATIC_MAX_ELEMS = 536870912
OPT_OVERLAP_BLOCKS_PER_SM = 1
OPT_NORM_STAGING = 2
OPT_PEER_TIMEOUT_MS = 3


class DpkError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"libdpk error {code}: {message}")
        self.code = code


class Config(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32),
        ("dtype", C.c_int32),
        ("device", C.c_int32),
        ("n_segments", C.c_int32),
        ("n_ema", C.c_int32),
        ("tile_elems", C.c_int32),
        ("n_elems", C.c_int64),
        ("seg_offsets", C.POINTER(C.c_int64)),
        ("seg_numel", C.POINTER(C.c_int64)),
        ("ema_rates", C.POINTER(C.c_double)),
        ("beta1", C.c_double),
        ("beta2", C.c_double),
        ("eps", C.c_double),
        ("bucket_cap_bytes", C.c_int64),
        ("edge_bucket_bytes", C.c_int64),
        ("seg_frozen", C.POINTER(C.c_int32)),
    ]


# name -> (restype, argtypes); must list every symbol include/dpk.h declares
# (tests/test_abi.py checks the two against each other).
_VP = C.c_void_p
SIGNATURES = {
    "dpk_abi_version": (C.c_int, []),
    "dpk_last_error": (C.c_char_p, []),
    "dpk_create": (C.c_int, [C.POINTER(_VP), C.POINTER(Config)]),
    "dpk_destroy": (C.c_int, [_VP]),
    "dpk_bind": (C.c_int, [_VP, _VP, _VP, _VP, _VP, C.POINTER(_VP)]),
    "dpk_comm_unique_id": (C.c_int, [_VP]),
    "dpk_comm_init": (C.c_int, [_VP, _VP, C.c_int32, C.c_int32]),
    "dpk_world_size": (C.c_int, [_VP]),
    "dpk_bind_nvls": (C.c_int, [_VP, _VP, C.POINTER(_VP), _VP, _VP, C.c_int32]),
    "dpk_num_buckets": (C.c_int, [_VP]),
    "dpk_bucket_range": (C.c_int, [_VP, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "dpk_bucket_ready": (C.c_int, [_VP, C.c_int32, _VP]),
    "dpk_all_buckets_ready": (C.c_int, [_VP, _VP]),
    "dpk_wait_bucket": (C.c_int, [_VP, C.c_int32, _VP]),
    "dpk_wait_allreduce": (C.c_int, [_VP, _VP]),
    "dpk_broadcast_state": (C.c_int, [_VP, C.c_int32, C.c_int32, _VP]),
    "dpk_zero_grads": (C.c_int, [_VP, _VP]),
    "dpk_grad_norm": (C.c_int, [_VP, C.c_double, C.c_double, _VP]),
    "dpk_fused_update": (C.c_int, [_VP, C.c_double, C.c_double, C.c_int64, C.c_double, C.c_int32,
                                    C.c_int32, C.c_int32, _VP]),
    "dpk_grad_norm_bucket": (C.c_int, [_VP, C.c_int32, C.c_double, _VP]),
    "dpk_grad_norm_finish": (C.c_int, [_VP, C.c_double, _VP]),
    "dpk_fused_update_bucket": (C.c_int, [_VP, C.c_int32, C.c_double, C.c_double, C.c_int64, C.c_double, C.c_int32,
                                           C.c_int32, C.c_int32, _VP]),
    "dpk_fused_update_finish": (C.c_int, [_VP, C.c_int32, _VP]),
    "dpk_reduce_update_pair": (C.c_int, [_VP, C.c_double, C.c_double, C.c_int64, C.c_double, _VP]),
    "dpk_reduce_update": (C.c_int, [_VP, C.c_double, C.c_double, C.c_int64, C.c_double, _VP]),
    "dpk_reduce_norm": (C.c_int, [_VP, C.c_double, C.c_double, _VP]),
    "dpk_bind_shard": (C.c_int, [_VP, _VP]),
    "dpk_shard_update": (C.c_int, [_VP, C.c_double, C.c_double, C.c_int64, C.c_double, _VP]),
    "dpk_shard_gather": (C.c_int, [_VP, _VP]),
    "dpk_set_option": (C.c_int, [_VP, C.c_int32, C.c_int64]),
    "dpk_result_ptr": (C.c_int, [_VP, C.POINTER(_VP)]),
    "dpk_read_result_sync": (C.c_int, [_VP, C.POINTER(C.c_double), _VP]),
    "dpk_segment_norms_ptr": (C.c_int, [_VP, C.POINTER(_VP)]),
    "dpk_grad_norm_static": (C.c_int, [_VP, C.c_int64, C.c_int32, C.c_double, C.c_double, _VP, _VP]),
    "dpk_timing_enable": (C.c_int, [_VP, C.c_int32]),
    "dpk_timing_read_sync": (C.c_int, [_VP, C.c_int32, C.POINTER(C.c_float), C.c_int32, C.POINTER(C.c_int32)]),
    "dpk_launch_count": (C.c_int, [_VP, C.POINTER(C.c_int64)]),
    "dpk_selftest_arith": (C.c_int, [C.c_double, _VP, _VP]),
    "dpk_ema_update": (C.c_int, [_VP, _VP, C.c_int64, C.c_int32, C.c_double, _VP]),
    "dpk_sampler_create": (C.c_int, [C.POINTER(_VP), C.c_int32, C.c_int32, C.c_int32, C.c_double, C.c_int32]),
    "dpk_sampler_destroy": (C.c_int, [_VP]),
    "dpk_sampler_update": (C.c_int, [_VP, _VP, _VP, _VP, C.c_int32, _VP]),
    "dpk_sampler_weights": (C.c_int, [_VP, _VP, _VP]),
    "dpk_sampler_state_ptrs": (C.c_int, [_VP, C.POINTER(_VP), C.POINTER(_VP)]),
}

_lock = threading.Lock()
_lib = None


def lib():
    """The loaded library; builds it from source once if the .so is absent."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            from . import _build
            _build.build()
        try:
            handle = C.CDLL(LIB_PATH)
        except OSError as exc:  # no silent degradation
            raise RuntimeError(
                f"libdpk.so could not be loaded from {LIB_PATH}: {exc}. "
                "Build it with `python -c 'import __graft_entry__ as g; g.build()'`."
            ) from exc
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        got = handle.dpk_abi_version()
        if got != ABI_VERSION:
            raise RuntimeError(f"libdpk ABI {got}, binding expects {ABI_VERSION}; rebuild")
        _lib = handle
    return _lib


def check(rc):
    if rc != 0:
        msg = lib().dpk_last_error()
        raise DpkError(rc, msg.decode("utf-8", "replace") if msg else "")
    return rc
