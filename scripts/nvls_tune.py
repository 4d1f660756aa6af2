"""Times the library's NVLS all-reduce alone (whole DiffuSeq-base gradient buffer) for a few launch shapes,
next to ncclAllReduce on the same buffer."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from distributed_pipeline_b200 import FusedAdamW
from distributed_pipeline_b200.synthetic import diffuseq_shapes
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()
configs = [(b, u) for b in (32, 64, 128, 256) for u in (4, 8)]
if len(sys.argv) > 1:
    configs = [tuple(int(x) for x in c.split("x")) for c in sys.argv[1].split(",")]
for dtype in (torch.float32,):
    params = [torch.nn.Parameter(torch.zeros(s, device=dev, dtype=dtype)) for s in diffuseq_shapes()]
    for cap in (512.0, 128.0):
        for blocks, unroll in configs:
            os.environ["DPK_NVLS_BLOCKS"] = str(blocks); os.environ["DPK_NVLS_UNROLL"] = str(unroll)
            opt = FusedAdamW(params, lr=1e-3, ema_rates=[], bucket_cap_mb=cap, symmetric_grads=True)
            eng, flat = opt.engine, opt.flat
            eng.init_comm(); eng.enable_nvls(flat.grads)
            flat.grads.normal_(0, 1e-3)
            nb = len(eng.buckets); nbytes = flat.grads.numel() * flat.grads.element_size()
            def ar():
                for b in range(nb): eng.bucket_ready(b)
                eng.wait_allreduce()
            t = timeit(ar)
            if rank == 0: print(f"{str(dtype):15s} W={world} buckets={nb} blocks={blocks:3d} unroll={unroll}: {t:.3f} ms  algbw {nbytes/t/1e6:6.1f} GB/s busbw {nbytes/t/1e6*2*(world-1)/world:6.1f}", flush=True)
            for p in params: p.grad = None
            torch.cuda.synchronize(); dist.barrier()
            eng.close(); del opt, eng, flat
        ref = torch.empty(91231936, device=dev, dtype=dtype).normal_(0, 1e-3)
        t = timeit(lambda: dist.all_reduce(ref, op=dist.ReduceOp.AVG))
        if rank == 0: print(f"{str(dtype):15s} W={world} ncclAllReduce whole buffer: {t:.3f} ms  algbw {ref.numel()*ref.element_size()/t/1e6:6.1f} GB/s", flush=True)
        del ref
dist.barrier(); dist.destroy_process_group()
