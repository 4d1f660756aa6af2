import os, sys, time
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
n = 91231936
t = symm_mem.empty(n, dtype=torch.float32, device=dev)
hdl = symm_mem.rendezvous(t, dist.group.WORLD)
if rank == 0:
    print("multicast_ptr", hex(hdl.multicast_ptr), "signal_pad_size", hdl.signal_pad_size, "world", hdl.world_size, "buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs][:3], flush=True)
t.normal_(0, 1)
ref = t.clone()
def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters
gn = dist.group.WORLD.group_name
res = {}
res["nccl_allreduce_365MB"] = timeit(lambda: dist.all_reduce(ref))
try:
    res["multimem_all_reduce_"] = timeit(lambda: torch.ops.symm_mem.multimem_all_reduce_(t, "sum", gn))
except Exception as ex:
    res["multimem_all_reduce_"] = repr(ex)[:200]
try:
    res["two_shot_all_reduce_"] = timeit(lambda: torch.ops.symm_mem.two_shot_all_reduce_(t, "sum", gn))
except Exception as ex:
    res["two_shot_all_reduce_"] = repr(ex)[:200]
chunk = ref[: 128 * 1024 * 1024 // 4]
res["nccl_allreduce_128MB"] = timeit(lambda: dist.all_reduce(chunk))
if rank == 0:
    for k, v in res.items():
        if isinstance(v, float):
            nbytes = (n if "128" not in k else chunk.numel()) * 4
            print(f"{k:28s} {v:8.3f} ms  algbw {nbytes/v/1e6:7.1f} GB/s  busbw {nbytes/v/1e6*2*(world-1)/world:7.1f}", flush=True)
        else:
            print(k, v, flush=True)
dist.barrier(); dist.destroy_process_group()
