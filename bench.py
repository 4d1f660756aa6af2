#!/usr/bin/env python
"""Benchmark of the data-parallel post-backward update path (BASELINE.json metric:
step-update ms & HBM GB/s for grad-allreduce + AdamW + EMA).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one post-backward update over one set of synthetic gradients:
(N>1: bucketed in-place all-reduce of the flat gradient buffer) -> [norm+clip
pass if the workload clips] -> fused unscale/clip/AdamW/EMA pass.

Default workload = BASELINE.json configs[1]: DiffuSeq-base (12-layer d=768,
209 tensors, 91 225 274 params), fp32 state, AdamW + one EMA rate 0.9999, clip
off (the norm is still produced, inside the fused pass, for logging).

Prints ONE JSON line (rank 0).  `value` = algorithmic HBM bytes of the update
processed by all ranks per second with gradients resident in HBM; `e2e` = the
same through the public optimizer API with the step's gradients arriving from
pinned host memory and the result block read back, every step.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (shape kwargs, dtype, ema rates, clip, weight decay)
    "diffuseq-base-fp32-ema1": (dict(), "f32", [0.9999], -1.0, 0.0),
    "diffuseq-base-fp32-ema1-clip": (dict(), "f32", [0.9999], 1.0, 0.0),
    "diffuseq-base-bf16-ema2-clip": (dict(), "bf16", [0.9999, 0.999], 1.0, 0.0),
    "toy-fp32-ema2-clip": (dict(n_layers=2, d=128, ffn=512, vocab=4096), "f32", [0.9999, 0.999], 1.0, 0.01),
}
LR = 1e-4


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", default="diffuseq-base-fp32-ema1", choices=sorted(WORKLOADS))
    ap.add_argument("--flat-params", type=int, default=0,
                    help="sweep mode: one flat tensor of this many parameters instead of the model shapes")
    ap.add_argument("--grad-mode", default="keep", choices=["keep", "clipped", "zero"])
    ap.add_argument("--tile-elems", type=int, default=0)
    ap.add_argument("--bucket-mb", type=float, default=64.0,
                    help="all-reduce bucket size for N>1 (reference DDP: 128; results do not depend on it)")
    ap.add_argument("--allreduce", choices=["nvls", "nccl"], default="nvls",
                    help="N>1: the library's own NVLS (multimem) all-reduce kernel, or ncclAllReduce")
    ap.add_argument("--no-overlap", action="store_true",
                    help="N>1: finish the whole all-reduce before the update instead of consuming buckets as they land")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def workload_shapes(args):
    from distributed_pipeline_b200.synthetic import diffuseq_shapes
    kw, dtype, rates, clip, wd = WORKLOADS[args.workload]
    shapes = [(args.flat_params,)] if args.flat_params else diffuseq_shapes(**kw)
    return shapes, dtype, rates, clip, wd


def bytes_per_param(dtype, n_ema, clip, grad_mode):
    """Algorithmic HBM bytes per parameter of one step."""
    s = 4 if dtype == "f32" else 2
    b = (7 + 2 * n_ema) * s          # fused pass: read g,p,m,v,K ema; write p,m,v,K ema
    if grad_mode != "keep":
        b += s                       # gradient rewritten (clipped) or zeroed
    if clip > 0:
        b += s                       # pass 1 reads g once more
    return b


class ClockSampler:
    """Polls NVML for SM clock and throttle reasons while the timed region runs."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop = [], set(), None, threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def run_reference(args):
    """The reference's own CPU implementation of the path, timed on this box's
    host cores.  /root/reference (pure Python) cannot travel to the GPU box, so
    this is its port onto the very torch CPU ops it calls (oracle/torch_port.py,
    checked against the real reference via tests/golden): kind = "port"."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    os.environ["CUDA_VISIBLE_DEVICES"] = ""
    import torch
    from oracle.torch_port import ReferenceUpdatePort
    from distributed_pipeline_b200.synthetic import numel
    shapes, dtype, rates, clip, wd = workload_shapes(args)
    tdt = torch.float32 if dtype == "f32" else torch.bfloat16
    gen = torch.Generator().manual_seed(102)
    params = [torch.nn.Parameter((torch.randn(s, generator=gen) * 0.02).to(tdt)) for s in shapes]
    n_params = sum(numel(s) for s in shapes)
    port = ReferenceUpdatePort(params, lr=LR, ema_rates=rates, weight_decay=wd, gradient_clipping=clip)
    grads = [(torch.randn(s, generator=gen) * 1e-3).to(tdt) for s in shapes]

    def step():
        for p, g in zip(params, grads):
            p.grad = g.clone()  # stands in for what backward leaves; outside the reference's optimize()
        t0 = time.perf_counter()
        port.optimize()
        return time.perf_counter() - t0

    for _ in range(args.warmup):
        step()
    # bounded sample: the reference arm must finish within minutes on any host
    budget_s, times = 120.0, []
    for _ in range(args.steps):
        times.append(step())
        if sum(times) > budget_s:
            break
    t = sum(times) / len(times)
    bpp = bytes_per_param(dtype, len(rates), clip, "keep")
    value = n_params * bpp / t / 1e9
    cores = torch.get_num_threads()
    sample = f"{len(times)} optimize() calls of the full workload ({n_params} params), torch {torch.__version__} CPU"
    line = {
        "impl": "reference", "metric": "step_update_hbm_GBps", "value": value, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": len(times), "warmup": args.warmup, "ms_per_step": t * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dtype, "data": "synthetic",
        "params_per_sec": n_params / t,
        "config": {"workload": args.workload if not args.flat_params else f"flat-{args.flat_params}",
                   "params": n_params, "tensors": len(shapes), "ema_rates": rates, "clip": clip,
                   "bytes_per_param": bpp, "host_threads": cores,
                   "affinity_cores": len(os.sched_getaffinity(0))},
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def cpu_baseline(args, shapes, dtype, rates, clip, wd):
    """~10-30 s of the reference port on the host cores (rank 0, N=1)."""
    import torch
    from oracle.torch_port import ReferenceUpdatePort
    from distributed_pipeline_b200.synthetic import numel
    tdt = torch.float32 if dtype == "f32" else torch.bfloat16
    gen = torch.Generator().manual_seed(102)
    params = [torch.nn.Parameter((torch.randn(s, generator=gen) * 0.02).to(tdt)) for s in shapes]
    grads = [(torch.randn(s, generator=gen) * 1e-3).to(tdt) for s in shapes]
    port = ReferenceUpdatePort(params, lr=LR, ema_rates=rates, weight_decay=wd, gradient_clipping=clip)
    n_params = sum(numel(s) for s in shapes)
    times = []
    t_start = time.perf_counter()
    i = 0
    while True:
        for p, g in zip(params, grads):
            p.grad = g.clone()
        t0 = time.perf_counter()
        port.optimize()
        dt = time.perf_counter() - t0
        if i >= 2:
            times.append(dt)
        i += 1
        if (len(times) >= 20) or (time.perf_counter() - t_start > 20.0 and len(times) >= 3):
            break
    t = sum(times) / len(times)
    bpp = bytes_per_param(dtype, len(rates), clip, "keep")
    return {"value": n_params * bpp / t / 1e9, "unit": "GB/s", "cores": torch.get_num_threads(), "kind": "port",
            "ms_per_step": t * 1e3,
            "sample": f"{len(times)} optimize() calls (after 2 warm-up) of the full workload, {n_params} params, "
                      f"torch {torch.__version__} CPU ops as the reference calls them"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from distributed_pipeline_b200 import FusedAdamW, _lib
    from distributed_pipeline_b200.synthetic import numel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.exit("launch with torch.distributed.run --nproc-per-node N for --gpus N > 1")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", init_method="env://", device_id=dev)

    shapes, dtype, rates, clip, wd = workload_shapes(args)
    tdt = torch.float32 if dtype == "f32" else torch.bfloat16
    n_params = sum(numel(s) for s in shapes)
    gm = {"keep": _lib.GRAD_KEEP, "clipped": _lib.GRAD_CLIPPED, "zero": _lib.GRAD_ZERO}[args.grad_mode]

    gen = torch.Generator(device=dev).manual_seed(102)  # params identical on every rank
    params = [torch.nn.Parameter((torch.randn(s, generator=gen, device=dev) * 0.02).to(tdt)) for s in shapes]
    opt = FusedAdamW(params, lr=LR, weight_decay=wd, ema_rates=rates, grad_mode=gm, tile_elems=args.tile_elems,
                     bucket_cap_mb=args.bucket_mb if world > 1 else 128,
                     symmetric_grads=world > 1 and args.allreduce == "nvls")
    del params
    eng, flat = opt.engine, opt.flat
    if world > 1:
        eng.init_comm()
        eng.broadcast_state(0, True)
        if args.allreduce == "nvls":
            try:
                if not flat.symmetric:
                    raise RuntimeError(f"symmetric allocation failed: {flat.symmetric_error}")
                eng.enable_nvls(flat.grads)
            except Exception as exc:  # no NVSwitch multicast on this box: NCCL carries the all-reduce
                if rank == 0:
                    print(f"[bench] symmetric-memory all-reduce unavailable ({exc}); using NCCL", file=sys.stderr)
                args.allreduce = "nccl"
    ggen = torch.Generator(device=dev).manual_seed(4321 + rank)
    sigma = 1e-3
    grads_src = (torch.randn(flat.layout.n_elems, generator=ggen, device=dev) * sigma).to(tdt)
    flat.grads.copy_(grads_src)
    n_buckets = len(eng.buckets)

    def device_step():
        if world > 1 and not args.no_overlap:
            opt.reduce_and_step(clip)      # buckets consumed as their all-reduce lands
            return
        if world > 1:
            for b in range(n_buckets):
                eng.bucket_ready(b)
            eng.wait_allreduce()
        if clip > 0:
            opt.clip_grad_norm(clip)
        opt.step()

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident timing ------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        device_step()
    sync_all()
    launches0 = eng.launch_count
    eng.timing(True)
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        start.record()
        for _ in range(args.steps):
            device_step()
        stop.record()
        sync_all()
    eng.timing(False)
    elapsed_ms = start.elapsed_time(stop)
    kernel_ms = eng.read_timings_sync(0)
    norm_ms = eng.read_timings_sync(1)
    launches = eng.launch_count - launches0
    if world > 1:
        t = torch.tensor([elapsed_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())
        if not eng.nvls:
            launches += n_buckets * args.steps  # ncclAllReduce kernels we enqueue (ours are counted by the library)
    ms_per_step = elapsed_ms / args.steps
    bpp = bytes_per_param(dtype, len(rates), clip, args.grad_mode)
    value = world * n_params * bpp / (ms_per_step * 1e-3) / 1e9

    # ---- roofline of the dominant kernel (fused update) ---------------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        with open(peaks_path) as f:
            peak, peak_src = float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    fused_bpp = bytes_per_param(dtype, len(rates), -1.0, args.grad_mode)
    pair_fused = (world == 2 and eng.nvls and not args.no_overlap and clip <= 0 and args.grad_mode == "keep")
    per_step = 1 if pair_fused else (n_buckets if (world > 1 and not args.no_overlap and clip <= 0) else 1)
    if per_step > 1:  # bucket-granular launches: one "launch" of the roofline = one full pass
        kernel_ms = [sum(kernel_ms[i:i + per_step]) for i in range(0, len(kernel_ms) - per_step + 1, per_step)]
    k_ms = sum(kernel_ms) / len(kernel_ms) if kernel_ms else float("nan")
    achieved = n_params * fused_bpp / (k_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(f"{args.workload}:{args.grad_mode}")
    roofline = {"bound": "hbm", "kernel": "fused_update_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "kernel_ms_avg": k_ms, "kernel_ms_min": min(kernel_ms) if kernel_ms else None,
                "launches_timed": len(kernel_ms), "algorithmic_bytes_per_launch": n_params * fused_bpp,
                "frac_of_nominal_8TBps": achieved / 8000.0}
    if norm_ms:
        s = 4 if dtype == "f32" else 2
        n_ms = sum(norm_ms) / len(norm_ms)
        roofline["norm_pass"] = {"kernel": "sumsq_tiles_kernel", "kernel_ms_avg": n_ms,
                                 "achieved": n_params * s / (n_ms * 1e-3) / 1e9,
                                 "frac": n_params * s / (n_ms * 1e-3) / 1e9 / peak}

    # ---- end to end: gradients from pinned host memory, result block read back -----
    e2e = None
    if not args.no_e2e:
        host_g = torch.empty(flat.layout.n_elems, dtype=tdt, pin_memory=True)
        host_g.copy_(grads_src)
        host_res = torch.empty(_lib.R_COUNT, dtype=torch.float64, pin_memory=True)
        results = eng.results
        stream = torch.cuda.current_stream()

        def e2e_step():
            opt.step_from_host(host_g, clip)     # upload, (all-reduce,) update: bucket by bucket
            host_res.copy_(results, non_blocking=True)
            stream.synchronize()
            return float(host_res[_lib.R_LOGGED_NORM])

        e_steps = min(args.steps, 50)
        for _ in range(3):
            e2e_step()
        sync_all()
        s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s2.record()
        for _ in range(e_steps):
            norm = e2e_step()
        e2.record()
        sync_all()
        e_ms = s2.elapsed_time(e2)
        if world > 1:
            t = torch.tensor([e_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e_ms = float(t.item())
        e_ms /= e_steps
        e2e = {"value": world * n_params * bpp / (e_ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": e_ms,
               "steps": e_steps, "h2d_bytes_per_step": host_g.numel() * host_g.element_size(),
               "d2h_bytes_per_step": host_res.numel() * host_res.element_size(), "last_grad_norm": norm,
               "api": "FusedAdamW.step_from_host (TrainLoop.opt): pinned host gradients, bucket-pipelined upload"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        del opt, eng, flat
        torch.cuda.empty_cache()
        cpu = cpu_baseline(args, shapes, dtype, rates, clip, wd)

    if rank == 0:
        line = {
            "metric": "step_update_hbm_GBps", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": dtype, "data": "synthetic",
            "params_per_sec": world * n_params / (ms_per_step * 1e-3),
            "config": {"workload": args.workload if not args.flat_params else f"flat-{args.flat_params}",
                       "params": n_params, "tensors": len(shapes), "ema_rates": rates, "clip": clip,
                       "weight_decay": wd, "grad_mode": args.grad_mode, "bytes_per_param": bpp,
                       "l2_policy": f"inputs larger than L2 ({n_params * bpp / 1e6:.0f} MB streamed per step vs 126 MB L2)",
                       "allreduce_buckets": n_buckets if world > 1 else 0,
                       "bucket_mb": args.bucket_mb if world > 1 else None,
                       "allreduce": (("own kernels over symmetric memory (NVLS multimem for W>=4, pair exchange for W=2)"
                                      if eng.nvls else "ncclAllReduce(avg)") if world > 1 else None),
                       "update_overlaps_allreduce": bool(world > 1 and not args.no_overlap),
                       "one_kernel_allreduce_update": bool(pair_fused),
                       "grad_sigma": sigma},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches,
            "clocks": clocks.summary(),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
