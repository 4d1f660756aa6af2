"""Whole TrainLoop.run_step on DiffuSeq-base (synthetic tokens, batch 64 x seq 128, microbatch 64, fp32, clip off,
EMA 0.9999) on one GPU: the drop-in loop vs the reference's loop restated on torch-CUDA eager (oracle/torch_port.py)."""
import os, sys, time, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_pipeline_b200 import TrainLoop, logger
from distributed_pipeline_b200.synthetic import SyntheticDiffuSeq
from oracle.torch_port import ReferenceUpdatePort

class Loop(TrainLoop):
    def compute_losses(self, mb):
        out = self.ddp_model(mb["ids"], mb["t"]); return {"loss": ((out - mb["target"]) ** 2).mean(dim=(1, 2))}
    @staticmethod
    def backward_from_losses(losses): losses["loss"].mean().backward()
    def log_loss_dict(self, mode, losses, *a, **k): pass

torch.manual_seed(102)
model = SyntheticDiffuSeq().cuda(); ref_model = copy.deepcopy(model)
gen = torch.Generator().manual_seed(0)
B, L = 64, 128
batch = {"ids": torch.randint(0, 30522, (B, L), generator=gen), "t": torch.randint(0, 2000, (B,), generator=gen),
         "target": torch.randn(B, L, 128, generator=gen)}
logger.configure(dir="/tmp/dpk-tl")
loop = Loop(model=model, data=iter([]), batch_size=B, microbatch=B, lr=1e-4, ema_rate="0.9999", log_interval=10**9,
            save_interval=10**9, resume_checkpoint="", grads_after_step="zero")
def ours(): loop.run_step(batch); loop.step += 1
port = ReferenceUpdatePort(list(ref_model.parameters()), lr=1e-4, ema_rates=[0.9999])
def ref():
    port.zero_grad()
    mb = {k: v.cuda() for k, v in batch.items()}
    out = ref_model(mb["ids"], mb["t"]); ((out - mb["target"]) ** 2).mean(dim=(1, 2)).mean().backward()
    port.optimize()
def fwd_bwd_only():
    mb = {k: v.cuda() for k, v in batch.items()}
    out = ref_model(mb["ids"], mb["t"]); ((out - mb["target"]) ** 2).mean(dim=(1, 2)).mean().backward()
def wall(fn, n=15, w=4):
    for _ in range(w): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
a, b, c = wall(ours), wall(ref), wall(fwd_bwd_only)
loop.flush_logs()
print(f"DiffuSeq-base fp32 batch {B}x{L}: forward+backward alone {c:.1f} ms; reference loop step {b:.1f} ms; "
      f"drop-in TrainLoop step {a:.1f} ms (post-backward part: reference {b-c:.1f} ms -> ours {a-c:.1f} ms)")
