"""NVLink traffic of the one-kernel all-reduce + update, from the driver's own link counters.

torchrun, N ranks.  Rank 0 reads `nvidia-smi nvlink -gt d` (cumulative data KiB per link, tx and rx) for its GPU before
and after K steps of FusedAdamW.reduce_and_step on the DiffuSeq-base fp32 buffers and prints bytes per step next to the
message size S and the algorithmic wire volume (pair form: S received; switch forms: S(1+1/W) each way)."""
import os, re, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from distributed_pipeline_b200 import FusedAdamW, _lib
from distributed_pipeline_b200.synthetic import diffuseq_shapes
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)

def counters():
    out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(lr)], capture_output=True, text=True).stdout
    tx = sum(int(x) for x in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out))
    rx = sum(int(x) for x in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out))
    return tx * 1024, rx * 1024, out

params = [torch.nn.Parameter(torch.zeros(s, device=dev).normal_(0, 0.02)) for s in diffuseq_shapes()]
opt = FusedAdamW(params, lr=1e-4, weight_decay=0.0, ema_rates=[0.9999], grad_mode=_lib.GRAD_KEEP, bucket_cap_mb=64,
                 symmetric_grads=True)
opt.engine.init_comm(); opt.engine.broadcast_state(0, True); opt.engine.enable_nvls(opt.flat.grads)
opt.flat.grads.normal_(0, 1e-3)
for _ in range(5): opt.reduce_and_step()
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
K = 200
t0, r0, raw = counters()
dist.barrier()
for _ in range(K): opt.reduce_and_step()
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
t1, r1, _ = counters()
if rank == 0:
    S = opt.flat.layout.n_elems * 4
    forced = os.environ.get("DPK_FORCE_MULTIMEM") == "1"
    want = S if (world == 2 and not forced) else S * (1 + 1 / world)
    if t1 == t0 and r1 == r0:
        print("nvidia-smi nvlink counters not available here:\n" + raw[:600])
    else:
        print(f"W={world} {'switch' if (world > 2 or forced) else 'pair'} form, S = {S/1e6:.1f} MB per rank, {K} steps: "
              f"NVLink tx {(t1-t0)/K/1e6:.1f} MB/step, rx {(r1-r0)/K/1e6:.1f} MB/step "
              f"(algorithmic {want/1e6:.1f} MB per direction; tx/S = {(t1-t0)/K/S:.3f}, rx/S = {(r1-r0)/K/S:.3f})", flush=True)
dist.barrier(); dist.destroy_process_group()
