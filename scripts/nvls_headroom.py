"""Is there link bandwidth left beside the switch-reduced all-reduce?  Times, per rank and as max over ranks:
  (a) the library's NVLS all-reduce of the whole DiffuSeq-base fp32 gradient buffer (365 MB), alone;
  (b) a unicast peer write of the same size to the next rank (copy engine, torch copy_ into the symmetric buffer
      of rank+1), alone;
  (c) both at once on two streams.
If (c) ~ max(a, b) the links still have room while multimem traffic is at its limit (the ~480 GB/s per direction
the all-reduce reaches is then a property of the switch's reduction path, not of the links); if (c) ~ a + b they share
one bottleneck."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
from distributed_pipeline_b200 import FusedAdamW
from distributed_pipeline_b200.synthetic import diffuseq_shapes
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
params = [torch.nn.Parameter(torch.zeros(s, device=dev)) for s in diffuseq_shapes()]
opt = FusedAdamW(params, lr=1e-3, ema_rates=[], bucket_cap_mb=512.0, symmetric_grads=True)
eng, flat = opt.engine, opt.flat
eng.init_comm(); eng.enable_nvls(flat.grads)
flat.grads.normal_(0, 1e-3)
n = flat.grads.numel(); nbytes = n * 4
land = symm_mem.empty(n, dtype=torch.float32, device=dev); land.zero_()
hdl = symm_mem.rendezvous(land, dist.group.WORLD)
peer = hdl.get_buffer((rank + 1) % world, (n,), torch.float32)
src = torch.randn(n, device=dev)
side = torch.cuda.Stream(device=dev)
nb = len(eng.buckets)
def ar():
    for b in range(nb): eng.bucket_ready(b)
    eng.wait_allreduce()
def cp():
    with torch.cuda.stream(side): peer.copy_(src)
def both():
    side.wait_stream(torch.cuda.current_stream())
    cp(); ar()
    torch.cuda.current_stream().wait_stream(side)
def cp_main():
    side.wait_stream(torch.cuda.current_stream()); cp(); torch.cuda.current_stream().wait_stream(side)
def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()
ta, tb, tc = timeit(ar), timeit(cp_main), timeit(both)
if rank == 0:
    print(f"W={world} S={nbytes/1e6:.0f} MB: all-reduce alone {ta:.3f} ms (wire {(1+1/world)*nbytes/ta/1e6:.0f} GB/s/dir); "
          f"peer write alone {tb:.3f} ms ({nbytes/tb/1e6:.0f} GB/s); both at once {tc:.3f} ms "
          f"(sum {ta+tb:.3f}, max {max(ta,tb):.3f})", flush=True)
dist.barrier(); dist.destroy_process_group()
