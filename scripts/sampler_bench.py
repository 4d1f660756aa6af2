"""BASELINE.json configs[3]: loss-aware second-moment sampler, T=2000, 64 records per rank per microbatch.
Times update_with_local_losses (pack + gather + in-order history update) and weights() on the device, and the
upstream-style host implementation (3 all_gathers + 2*W*B .item() syncs + Python loop) next to it."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from distributed_pipeline_b200 import FusedAdamW
from distributed_pipeline_b200.sampler import LossSecondMomentResampler
from oracle.sampler_oracle import LossSecondMomentResamplerOracle
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
eng = None
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
    opt = FusedAdamW([torch.nn.Parameter(torch.zeros(1024, device=dev))], ema_rates=[])
    opt.engine.init_comm(); eng = opt.engine
T, B = 2000, 64
s = LossSecondMomentResampler(T, 10, 0.001, device=dev, engine=eng, max_local_batch=B)
gen = torch.Generator(device=dev).manual_seed(rank)
ts = torch.randint(0, T, (B,), device=dev, generator=gen); ls = torch.rand(B, device=dev, generator=gen)
def t_dev(fn, iters=200):
    for _ in range(10): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / iters * 1e3
upd = t_dev(lambda: s.update_with_local_losses(ts, ls))
wts = t_dev(lambda: s.weights_device())
# upstream-style host path
ref = LossSecondMomentResamplerOracle(T, 10, 0.001)
def upstream():
    if world > 1:
        sizes = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(world)]
        dist.all_gather(sizes, torch.tensor([B], dtype=torch.int32, device=dev))
        mx = max(int(x.item()) for x in sizes)
        tb = [torch.zeros(mx, dtype=ts.dtype, device=dev) for _ in range(world)]; lb = [torch.zeros(mx, device=dev) for _ in range(world)]
        dist.all_gather(tb, ts); dist.all_gather(lb, ls)
        t_all = [x.item() for y, n in zip(tb, sizes) for x in y[: int(n.item())]]
        l_all = [x.item() for y, n in zip(lb, sizes) for x in y[: int(n.item())]]
    else:
        t_all = [x.item() for x in ts]; l_all = [x.item() for x in ls]
    ref.update_with_all_losses(t_all, l_all)
for _ in range(3): upstream()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(20): upstream()
torch.cuda.synchronize(); host = (time.perf_counter() - t0) / 20 * 1e6
if rank == 0:
    print(f"W={world} T={T} B={B}: device update {upd:.1f} us, weights {wts:.1f} us; upstream-style host update {host:.0f} us ({host/upd:.0f}x)", flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
