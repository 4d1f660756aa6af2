"""Per-bucket timeline of reduce_and_step (main-stream CUDA events), averaged over steps."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from distributed_pipeline_b200 import FusedAdamW, _lib
from distributed_pipeline_b200.synthetic import diffuseq_shapes
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
cap = float(sys.argv[1]) if len(sys.argv) > 1 else 128.0
mode = sys.argv[2] if len(sys.argv) > 2 else "nvls"
params = [torch.nn.Parameter(torch.zeros(s, device=dev).normal_(0, 0.02)) for s in diffuseq_shapes()]
opt = FusedAdamW(params, lr=1e-4, weight_decay=0.0, ema_rates=[0.9999], grad_mode=_lib.GRAD_KEEP, bucket_cap_mb=cap,
                 symmetric_grads=(mode == "nvls"))
eng, flat = opt.engine, opt.flat
eng.init_comm(); eng.broadcast_state(0, True)
if mode == "nvls": eng.enable_nvls(flat.grads)
if len(sys.argv) > 3: eng.set_option(_lib.OPT_OVERLAP_BLOCKS_PER_SM, int(sys.argv[3]))
flat.grads.normal_(0, 1e-3)
nb = len(eng.buckets)
sizes = [c * 4 / 2**20 for _, c in eng.buckets]
def step(record=None):
    ev = lambda: torch.cuda.Event(enable_timing=True)
    t0 = ev(); t0.record()
    if len(sys.argv) > 4 and sys.argv[4] == "chained": eng.all_buckets_ready()
    else:
        for b in range(nb): eng.bucket_ready(b)
    opt.step_count += 1
    marks = []
    for b in range(nb):
        eng.wait_bucket(b); w = ev(); w.record()
        eng.fused_update_bucket(b, 1e-4, 0.0, opt.step_count); u = ev(); u.record()
        marks.append((w, u))
    eng.fused_update_finish(False); f = ev(); f.record()
    if record is not None: record.append((t0, marks, f))
for _ in range(5): step()
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
rec = []
for _ in range(20): step(rec)
torch.cuda.synchronize()
if rank == 0:
    import statistics as st
    print(f"W={world} mode={mode} cap={cap} overlap_blocks_per_sm={sys.argv[3] if len(sys.argv) > 3 else 0} {sys.argv[4] if len(sys.argv) > 4 else "per-bucket"} buckets MB={[round(s,1) for s in sizes]}")
    for b in range(nb):
        tw = st.mean(t0.elapsed_time(m[b][0]) for t0, m, f in rec); tu = st.mean(t0.elapsed_time(m[b][1]) for t0, m, f in rec)
        print(f"  bucket {b} ({sizes[b]:6.1f} MB): all-reduce visible at {tw:.3f} ms, update done at {tu:.3f} ms (update span {tu-tw:.3f})")
    print(f"  step end {st.mean(t0.elapsed_time(f) for t0, m, f in rec):.3f} ms")
dist.barrier(); dist.destroy_process_group()
