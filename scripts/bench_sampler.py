"""bench.py --workload sampler-t2000  (BASELINE.json configs[3]).

Loss-aware second-moment timestep sampler, T = 2000 timesteps, history 10, 64 (t, loss) records per rank per
microbatch, W ranks.  One "step" = what one microbatch costs: ``update_with_local_losses`` (pack, ONE packed
all-gather over the library's communicator, in-order history update -- all on the device, no host sync) plus
``weights()`` on the device.  Next to it the upstream host path the reference's ancestor runs per microbatch
(improved-diffusion resample.py: three ``all_gather`` calls, 2*W*B ``.item()`` syncs, a Python loop over the
records), restated here with torch.distributed + numpy on the same box.  The sampler is absent from the
reference itself: parity for it is unpinned and says so in DESIGN.md.
"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

T, B, HIST = 2000, 64, 10


def run(args, ClockSampler):
    from distributed_pipeline_b200 import FusedAdamW
    from distributed_pipeline_b200.sampler import LossSecondMomentResampler

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    eng = None
    if world > 1:
        dist.init_process_group("nccl", init_method="env://", device_id=dev)
        holder = FusedAdamW([torch.nn.Parameter(torch.zeros(1024, device=dev))], ema_rates=[])
        holder.engine.init_comm()
        eng = holder.engine
    s = LossSecondMomentResampler(T, HIST, 0.001, device=dev, engine=eng, max_local_batch=B)
    gen = torch.Generator(device=dev).manual_seed(rank)
    ts = torch.randint(0, T, (B,), device=dev, generator=gen)
    ls = torch.rand(B, device=dev, generator=gen)

    def step():
        s.update_with_local_losses(ts, ls)
        s.weights_device()

    steps, warmup = max(args.steps, 10), max(args.warmup, 3)
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        a.record()
        for _ in range(steps):
            step()
        b.record()
        torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps

    # end to end: the records come from host memory and the weights go back to the host (what sample() needs)
    h_ts, h_ls = ts.cpu().pin_memory(), ls.cpu().pin_memory()
    h_w = torch.empty(T, dtype=torch.float64, pin_memory=True)

    def e2e_step():
        s.update_with_local_losses(h_ts.to(dev, non_blocking=True), h_ls.to(dev, non_blocking=True))
        h_w.copy_(s.weights_device(), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(3):
        e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    e_steps = min(steps, 100)
    for _ in range(e_steps):
        e2e_step()
    e2e_ms = (time.perf_counter() - t0) / e_steps * 1e3

    # ---- upstream's host path, restated (3 all_gathers + .item() per record + a Python loop) -----------
    hist = np.zeros([T, HIST], dtype=np.float64)
    counts = np.zeros([T], dtype=np.int64)

    def upstream():
        if world > 1:
            sizes = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(world)]
            dist.all_gather(sizes, torch.tensor([B], dtype=torch.int32, device=dev))
            sizes = [int(x.item()) for x in sizes]
            mx = max(sizes)
            tb = [torch.zeros(mx, dtype=ts.dtype, device=dev) for _ in range(world)]
            lb = [torch.zeros(mx, device=dev) for _ in range(world)]
            dist.all_gather(tb, ts)
            dist.all_gather(lb, ls)
            t_all = [x.item() for y, n in zip(tb, sizes) for x in y[:n]]
            l_all = [x.item() for y, n in zip(lb, sizes) for x in y[:n]]
        else:
            t_all = [x.item() for x in ts]
            l_all = [x.item() for x in ls]
        for t, loss in zip(t_all, l_all):
            if counts[t] == HIST:
                hist[t, :-1] = hist[t, 1:]
                hist[t, -1] = loss
            else:
                hist[t, counts[t]] = loss
                counts[t] += 1
        if (counts == HIST).all():
            w = np.sqrt(np.mean(hist ** 2, axis=-1))
            w /= np.sum(w)
            w *= 1 - 0.001
            w += 0.001 / len(w)

    for _ in range(3):
        upstream()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    n_up = 20
    for _ in range(n_up):
        upstream()
    torch.cuda.synchronize()
    host_ms = (time.perf_counter() - t0) / n_up * 1e3

    def mx(x):
        if world == 1:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ms, e2e_ms, host_ms = mx(ms), mx(e2e_ms), mx(host_ms)
    if rank == 0:
        line = {
            "metric": "sampler_update_us", "value": ms * 1e3, "unit": "us", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": ms, "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "sampler-t2000", "timesteps": T, "history_per_term": HIST,
                       "records_per_rank_per_microbatch": B, "gathered_records": B * world,
                       "l2_policy": "state is 160 KB: L2-resident by nature (latency-bound path, no roofline claimed)"},
            "upstream_host_path_us": host_ms * 1e3,
            "upstream_host_path": "3 all_gather + 2*W*B .item() + Python loop + numpy weights (improved-diffusion resample.py), same box",
            "e2e": {"value": e2e_ms * 1e3, "unit": "us", "h2d_bytes_per_step": B * 12, "d2h_bytes_per_step": T * 8,
                    "api": "LossSecondMomentResampler.update_with_local_losses + weights(): records from pinned host memory, weights read back"},
            "gpu_launches": 4 * steps, "clocks": clocks.summary(),
            "parity": "unpinned (the reference has no sampler; bit-exact against the numpy restatement in tests/)",
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
