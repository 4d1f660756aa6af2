"""N ranks: the library's NVLS all-reduce vs ncclAllReduce(avg) on the same data."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from distributed_pipeline_b200 import FusedAdamW
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
ok = True
for dtype in (torch.float32, torch.bfloat16):
    torch.manual_seed(0)
    shapes = [(1000, 33), (7,), (300000,), (128, 128), (5,), (2_000_003,)]
    params = [torch.nn.Parameter(torch.randn(s, device=dev).to(dtype)) for s in shapes]
    opt = FusedAdamW(params, lr=1e-3, ema_rates=[0.999], bucket_cap_mb=1.0, symmetric_grads=True)
    eng, flat = opt.engine, opt.flat
    eng.init_comm(); eng.broadcast_state(0, True)
    nb = len(eng.buckets)
    gen = torch.Generator(device=dev).manual_seed(100 + rank)
    for it in range(3):
        g = (torch.randn(flat.layout.n_elems, device=dev, generator=gen) * 0.1).to(dtype)
        # reference: ncclAvg through torch
        ref = g.clone(); dist.all_reduce(ref, op=dist.ReduceOp.AVG)
        flat.grads.copy_(g)
        if it == 0 and not eng.nvls:
            eng.enable_nvls(flat.grads)
        if it == 2:
            eng.all_buckets_ready()        # chained form: one "gradients complete" barrier for the whole buffer
        else:
            for b in range(nb): eng.bucket_ready(b)
        eng.wait_allreduce()
        torch.cuda.synchronize()
        gl = [torch.empty_like(g) for _ in range(world)]
        dist.all_gather(gl, g)
        exact = (sum(x.double() for x in gl) / world).to(dtype)       # one rounding of the exact mean
        n_ours = (flat.grads != exact).sum().item(); n_nccl = (ref != exact).sum().item()
        if rank == 0: print(f"   elements differing from the correctly rounded mean: ours {n_ours}, ncclAvg {n_nccl}", flush=True)
        if rank == 0 and n_ours and it == 0:
            idx = (flat.grads != exact).nonzero().flatten()[:6]
            for i in idx.tolist():
                h = lambda t: hex(t[i].view(torch.int16).item() & 0xffff)
                print(f"     i={i} a={gl[0][i].item()!r} {h(gl[0])} b={gl[1][i].item()!r} {h(gl[1])} ours={flat.grads[i].item()!r} {h(flat.grads)} exact={exact[i].item()!r} {h(exact)}", flush=True)
            print("     first mismatch positions", idx.tolist(), "count in first/second half", (flat.grads != exact)[:flat.layout.n_elems//2].sum().item(), (flat.grads != exact)[flat.layout.n_elems//2:].sum().item(), flush=True)
        same = torch.equal(flat.grads, ref)
        diff = (flat.grads.float() - ref.float()).abs().max().item()
        if rank == 0: print(f"{dtype} iter {it} buckets {nb}: bit-equal to ncclAvg: {same}, max abs diff {diff:.3e}", flush=True)
        if world == 2:
            ok = ok and same                      # one correctly rounded (a+b)/2 on both paths
        elif dtype == torch.float32:
            ok = ok and n_ours <= n_nccl and diff <= 1e-6
        else:
            # bf16, W >= 4: the switch accumulates in fp32; NCCL's ring rounds to bf16 at every hop
            ok = ok and n_ours <= n_nccl and diff <= 4e-3
    # replicas identical
    chk = flat.grads.clone(); dist.broadcast(chk, 0); ok = ok and torch.equal(chk, flat.grads)
dist.barrier()
if rank == 0: print("NVLS_CHECK", "OK" if ok else "FAIL", flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
