"""TEST INFRASTRUCTURE ONLY -- the update in float64 numpy, no rounding points at all: the mathematical object the
reference's fp32/bf16 arithmetic approximates (tier T0).  Follows /root/reference/utils/trainer.py:237-271
(clip -> log-norm -> anneal -> AdamW -> EMA) and torch optim/adam.py:396-552 for the AdamW formulas."""
import numpy as np


def total_norm(g, seg_off, seg_len):
    return float(np.sqrt(sum(float(np.sum(g[o:o + n].astype(np.float64) ** 2)) for o, n in zip(seg_off, seg_len))))


def step(p, g, m, v, emas, ema_rates, lr, weight_decay, adam_step, beta1=0.9, beta2=0.999, eps=1e-8, inv_scale=1.0,
         max_norm=0.0, norm=None):
    """All arrays float64, updated in place.  Returns (pre-clip norm, coefficient)."""
    g = g * inv_scale
    nrm = float(np.sqrt(np.sum(g ** 2))) if norm is None else norm
    coef = min(max_norm / (nrm + 1e-6), 1.0) if max_norm > 0 else 1.0
    g = g * coef
    if weight_decay != 0:
        p *= 1 - lr * weight_decay
    m += (1 - beta1) * (g - m)
    v *= beta2
    v += (1 - beta2) * g * g
    bc1, bc2 = 1 - beta1 ** adam_step, 1 - beta2 ** adam_step
    p -= (lr / bc1) * m / (np.sqrt(v) / np.sqrt(bc2) + eps)
    for e, r in zip(emas, ema_rates):
        e *= r
        e += (1 - r) * p
    return nrm, coef
