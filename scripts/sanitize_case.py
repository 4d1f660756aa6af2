"""Small, fast exercise of every libdpk kernel family on one GPU, for compute-sanitizer."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from distributed_pipeline_b200 import _lib, ema_update
from distributed_pipeline_b200.sampler import LossSecondMomentResampler
from helpers import FlatProblem
for dtype in (torch.float32, torch.bfloat16):
    pb = FlatProblem([5, 0, 1, 130, 7, 4096, 8193, 33, 20000, 3], dtype, [0.999, 0.9], seed=1, mirror=False,
                     bucket_cap_bytes=16 * 1024)
    for gm in (0, 1, 2):
        pb.new_grads(poison_padding=False)
        pb.engine.grad_norm(0.5, 1.0)
        pb.engine.fused_update(1e-3, 0.01, 1, inv_scale=0.5, use_clip_coef=True, grad_mode=gm)
    pb.engine.set_option(_lib.OPT_NORM_STAGING, 1)
    pb.engine.grad_norm(1.0, 1.0)
    for b in range(len(pb.engine.buckets)):
        pb.engine.grad_norm_bucket(b, 1.0)
        pb.engine.fused_update_bucket(b, 1e-3, 0.0, 2)
    pb.engine.grad_norm_finish(1.0); pb.engine.fused_update_finish()
    out = torch.zeros(_lib.R_COUNT, dtype=torch.float64, device="cuda")
    g = torch.randn(70001, device="cuda").to(dtype)
    _lib.check(_lib.lib().dpk_grad_norm_static(g.data_ptr(), g.numel(), 0 if dtype == torch.float32 else 1, 1.0, 1.0,
                                              out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    t = torch.randn(5003, device="cuda").to(dtype); s = torch.randn(5003, device="cuda").to(dtype)
    ema_update(t[1:], s[1:], 0.99); ema_update(t[1:-1], s[2:], 0.99)
    print(dtype, pb.engine.read_results_sync()[:3])
smp = LossSecondMomentResampler(200, 4, max_local_batch=32)
for _ in range(30):
    smp.update_with_local_losses(torch.randint(0, 200, (32,), device="cuda"), torch.rand(32, device="cuda"))
print("weights sum", float(smp.weights().sum()))
torch.cuda.synchronize()
print("SANITIZE_CASE_DONE")
