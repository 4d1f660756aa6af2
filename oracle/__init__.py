"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the libdpk update path.

``oracle/dpk_oracle.c``   scalar C restatement of the reference's arithmetic
``oracle/torch_port.py``  the same path spelled with the torch ops the reference
                          calls (CPU baseline for bench.py; tight oracle on GPU)
``oracle/truth_fp64.py``  the same recurrence in float64 numpy (no rounding points): tier T0
``oracle/sampler_oracle.py``  numpy restatement of the upstream loss-aware sampler
``oracle/make_golden.py`` runs the *real* reference TrainLoop (from
                          /root/reference, only in the build container) and
                          writes tests/golden/*.npz

Only tests/, ``__graft_entry__.smoke()`` and bench.py's cpu_baseline /
``--impl reference`` legs may import this package; the product
(distributed-pipeline_b200/) never does.

The reference is pure Python, so there is no ``oracle/_ref`` build: nothing to
compile from /root/reference.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "dpk_oracle.c")
LIB = os.path.join(HERE, "_build", "liboracle.so")

CUDA, CPU = 0, 1
F32, BF16 = 0, 1


def build(force=False):
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    # -ffp-contract=off: only the fmaf() calls in the source may fuse
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math", "-std=c11",
           SRC, "-o", LIB + ".tmp", "-lm"]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("gcc failed:\n" + proc.stderr)
    os.replace(LIB + ".tmp", LIB)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        h = C.CDLL(build())
        h.oracle_grad_norm.restype = None
        h.oracle_grad_norm.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_float,
                                       C.c_float, C.c_void_p, C.c_void_p]
        h.oracle_adamw_ema_step.restype = C.c_double
        h.oracle_adamw_ema_step.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_double),
                                            C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                                            C.c_int64, C.c_double, C.c_double, C.c_int]
        h.oracle_logged_norm.restype = C.c_double
        h.oracle_logged_norm.argtypes = [C.c_int, C.c_void_p, C.c_int]
        h.oracle_ema_update.restype = None
        h.oracle_ema_update.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_double]
        h.oracle_f32_to_bf16.restype = C.c_uint16
        h.oracle_f32_to_bf16.argtypes = [C.c_float]
        h.oracle_bf16_to_f32.restype = C.c_float
        h.oracle_bf16_to_f32.argtypes = [C.c_uint16]
        _lib = h
    return _lib


def _np_dtype(dtype):
    return np.float32 if dtype == F32 else np.uint16  # bf16 travels as raw bits


def _check(a, dtype):
    if a.dtype != _np_dtype(dtype) or not a.flags.c_contiguous:
        raise TypeError(f"expected contiguous {_np_dtype(dtype).__name__} array")
    return a.ctypes.data


def grad_norm(dtype, g, seg_off, seg_len, inv_scale=1.0, max_norm=0.0):
    """-> dict(total, coef, nonfinite, seg_norms).  `g` flat array (uint16 bits for bf16)."""
    off = np.ascontiguousarray(seg_off, dtype=np.int64)
    ln = np.ascontiguousarray(seg_len, dtype=np.int64)
    seg_norms = np.zeros(max(len(off), 1), dtype=np.float64)
    out = np.zeros(3, dtype=np.float64)
    lib().oracle_grad_norm(dtype, _check(g, dtype), len(off), off.ctypes.data, ln.ctypes.data,
                           float(inv_scale), float(max_norm), seg_norms.ctypes.data, out.ctypes.data)
    return {"total": out[0], "coef": out[1], "nonfinite": out[2], "seg_norms": seg_norms[:len(off)]}


def adamw_ema_step(dtype, p, g, m, v, emas, ema_rates, lr, weight_decay, adam_step, beta1=0.9, beta2=0.999,
                   eps=1e-8, inv_scale=1.0, coef=1.0, grad_mode=0, flavour=CUDA):
    """In-place on the numpy arrays.  Returns fp64 sum of rnd(g_used**2)."""
    n = p.size
    for a in (g, m, v, *emas):
        if a.size != n:
            raise ValueError("size mismatch")
    ptrs = (C.c_void_p * max(len(emas), 1))(*[_check(e, dtype) for e in emas])
    rates = (C.c_double * max(len(emas), 1))(*[float(r) for r in ema_rates])
    return lib().oracle_adamw_ema_step(flavour, dtype, n, _check(p, dtype), _check(g, dtype), _check(m, dtype),
                                       _check(v, dtype), len(emas), ptrs, rates, float(lr), float(weight_decay),
                                       float(beta1), float(beta2), float(eps), int(adam_step), float(inv_scale),
                                       float(coef), int(grad_mode))


def logged_norm(dtype, seg_sumsq):
    s = np.ascontiguousarray(seg_sumsq, dtype=np.float64)
    return lib().oracle_logged_norm(dtype, s.ctypes.data, len(s))


def ema_update(dtype, trg, src, rate):
    lib().oracle_ema_update(dtype, trg.size, _check(trg, dtype), _check(src, dtype), float(rate))


def f32_to_bf16_bits(x):
    """numpy float32 array -> uint16 bf16 bits, round to nearest even."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    nan = (u & 0x7fffffff) > 0x7f800000
    r = ((u + 0x7fff + ((u >> 16) & 1)) >> 16).astype(np.uint16)
    r[nan] = ((u[nan] >> 16) | 0x40).astype(np.uint16)
    return r


def bf16_bits_to_f32(b):
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)
