"""Like-for-like GPU baselines on the same B200: the reference's optimize() as torch-CUDA eager
(oracle/torch_port.py: clip_grad_norm_ + 209 .item() syncs + foreach AdamW + per-tensor EMA), and the strongest library
variant (AdamW(fused=True) + torch._foreach EMA, no per-tensor norm logging), next to libdpk.  DiffuSeq-base shapes."""
import os, sys, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_pipeline_b200 import FusedAdamW, _lib
from distributed_pipeline_b200.synthetic import diffuseq_shapes, numel
from oracle.torch_port import ReferenceUpdatePort
dev = torch.device("cuda", 0)
out = {}
for name, dtype, rates, clip in (("fp32_ema1_noclip", torch.float32, [0.9999], -1.0), ("fp32_ema1_clip", torch.float32, [0.9999], 1.0),
                                 ("bf16_ema2_clip", torch.bfloat16, [0.9999, 0.999], 1.0)):
    shapes = diffuseq_shapes(); n = sum(numel(s) for s in shapes)
    gen = torch.Generator(device=dev).manual_seed(0)
    mk = lambda: [torch.nn.Parameter((torch.randn(s, generator=gen, device=dev) * 0.02).to(dtype)) for s in shapes]
    grads = [(torch.randn(s, generator=gen, device=dev) * 1e-3).to(dtype) for s in shapes]
    def wall(fn, iters=20, warm=5):
        for _ in range(warm): fn()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(iters): fn()
        torch.cuda.synchronize(); return (time.perf_counter() - t0) / iters * 1e3
    # (a) the reference's optimize(), torch-CUDA eager
    params = mk(); port = ReferenceUpdatePort(params, lr=1e-4, ema_rates=rates, gradient_clipping=clip)
    def ref_step():
        for p, g in zip(params, grads): p.grad = g
        port.optimize()
    a = wall(ref_step)
    del port, params
    # (b) library stretch: fused AdamW + foreach EMA + one clip_grad_norm_ (no 209-sync logging)
    params = mk(); opt = torch.optim.AdamW(params, lr=1e-4, weight_decay=0.0, fused=True)
    emas = [[p.detach().clone() for p in params] for _ in rates]
    def fused_step():
        for p, g in zip(params, grads): p.grad = g
        if clip > 0: torch.nn.utils.clip_grad_norm_(params, clip)
        else: torch.nn.utils.get_total_norm(grads)
        opt.step()
        with torch.no_grad():
            for r, e in zip(rates, emas):
                torch._foreach_mul_(e, r); torch._foreach_add_(e, [p.detach() for p in params], alpha=1 - r)
    b = wall(fused_step)
    del opt, params, emas
    # (c) libdpk through the optimizer API
    params = mk(); ours = FusedAdamW(params, lr=1e-4, weight_decay=0.0, ema_rates=rates, grad_mode=_lib.GRAD_CLIPPED if clip > 0 else _lib.GRAD_KEEP)
    flat_g = torch.cat([torch.nn.functional.pad(g.reshape(-1), (0, 0)) for g in grads])
    for gv, g in zip(ours.flat.grad_views, grads): gv.copy_(g)
    def our_step():
        if clip > 0: ours.clip_grad_norm(clip)
        ours.step()
    c = wall(our_step, iters=100)
    out[name] = {"params": n, "reference_optimize_torch_cuda_eager_ms": round(a, 3), "torch_fused_adamw_foreach_ema_ms": round(b, 3),
                 "libdpk_ms": round(c, 4), "speedup_vs_reference_eager": round(a / c, 1), "speedup_vs_torch_fused": round(b / c, 2)}
    print(name, json.dumps(out[name]), flush=True)
    del ours, params
