#!/usr/bin/env python
"""Benchmark of the data-parallel post-backward update path (BASELINE.json metric:
step-update ms & HBM GB/s for grad-allreduce + AdamW + EMA).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload ...]

One "step" = one post-backward update over one set of synthetic gradients:
(N>1: all-reduce of the flat gradient buffer) -> [norm+clip pass if the workload
clips] -> fused unscale/clip/AdamW/EMA pass.  At N>1 without clipping the
all-reduce and the update are ONE kernel (dpk_reduce_update).

Default workload = BASELINE.json configs[1]: DiffuSeq-base (12-layer d=768,
209 tensors, 91 225 274 params), fp32 state, AdamW + one EMA rate 0.9999, clip
off (the norm is still produced, inside the fused pass, for logging).

Prints ONE JSON line (rank 0).  `value` = algorithmic HBM bytes of the update
processed by all ranks per second with gradients resident in HBM; `e2e` = the
same through the public optimizer API with the step's gradients arriving from
pinned host memory and the result block read back, every step.  At N>1 the line
also carries `parity_check` (a checked all-reduce + update per dtype, run outside
the timed region; the process exits non-zero if it fails) and `roofline_comm`
(the collective against the NVLink roofline).

`--impl reference`: the reference's own CPU path (torch CPU ops as it calls them,
oracle/torch_port.py) on the box's host cores; at N>1 on ALL ranks over gloo with
DDP's pre-scale + bucketed all-reduce in front and threads = physical cores // W
(the reference's rule, basic_utils/dist_run.py:31).

Two more workloads time what surrounds the update:
`trainloop-bf16-ema2-clip` (BASELINE configs[2]: whole TrainLoop.run_step with
microbatches and the all-reduce overlapped with backward) and `sampler-t2000`
(configs[3]: loss-aware sampler gather + update + weights per microbatch).
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
SPECIAL = ("trainloop-bf16-ema2-clip", "sampler-t2000")
LR = 1e-4
NVLINK_NOMINAL_GBS = 900.0      # per direction per GPU
NVLINK_MEASURED_GBS = 770.0     # peer copy on this pool (B200_PROFILING.md)


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", default="diffuseq-base-fp32-ema1", choices=sorted(WORKLOADS) + list(SPECIAL))
    ap.add_argument("--flat-params", type=int, default=0,
                    help="sweep mode: one flat tensor of this many parameters instead of the model shapes")
    ap.add_argument("--grad-mode", default="keep", choices=["keep", "clipped", "zero"])
    ap.add_argument("--tile-elems", type=int, default=0)
    ap.add_argument("--bucket-mb", type=float, default=64.0,
                    help="all-reduce bucket size for N>1 (reference DDP: 128; results do not depend on it)")
    ap.add_argument("--allreduce", choices=["nvls", "nccl"], default="nvls",
                    help="N>1: the library's own symmetric-memory kernels, or ncclAllReduce")
    ap.add_argument("--no-overlap", action="store_true",
                    help="N>1: finish the whole all-reduce before the update instead of consuming buckets as they land")
    ap.add_argument("--no-fused-reduce", action="store_true",
                    help="N>1, no clipping: bucketed all-reduce + bucket-overlapped update instead of the one-kernel form")
    ap.add_argument("--sharded", action="store_true",
                    help="N>1, no clipping: the opt-in ZeRO-1 style update (each rank updates 1/W, parameters multicast)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="N>1: skip the checked step (the line then has no parity_check)")
    return ap.parse_args(argv)


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


def bind_to_gpu_numa_node(index):
    """Pin this process to the CPUs of the NUMA node the GPU hangs off, so that the pinned staging
    buffers allocated afterwards are first-touched on that node and every rank's uploads travel its
    own socket's memory controller and PCIe root instead of all sharing one (what a launcher does
    with numactl).  Best effort: returns a description, or None where the topology is not exposed."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.split(":", 1)
        path = f"/sys/bus/pci/devices/{dom[-4:].lower()}:{rest.lower()}/numa_node"
        with open(path) as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = set()
            for part in f.read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"gpu": index, "numa_node": node, "cpus": len(cpus)}
    except Exception:
        return None


# ------------------------------------------------------------------ reference arm --

def host_core_counts():
    affinity = len(os.sched_getaffinity(0))
    try:
        import psutil
        physical = psutil.cpu_count(logical=False) or affinity
    except Exception:
        physical = affinity
    return affinity, min(physical, affinity)


def ddp_style_allreduce(grads, world, dist, cap_bytes=128 << 20):
    """What torch DDP's reducer does as the reference configures it
    (utils/trainer.py:118-125: bucket_cap_mb=128): gradients, last tensor first,
    are copied into flat buckets pre-multiplied by 1/W, the bucket is all-reduced
    (SUM), and the result copied back into each .grad."""
    import torch
    inv = 1.0 / world
    bucket, size = [], 0

    def flush():
        if not bucket:
            return
        flat = torch.cat([g.reshape(-1) for g in bucket])
        flat.mul_(inv)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        off = 0
        for g in bucket:
            n = g.numel()
            g.copy_(flat[off:off + n].view_as(g))
            off += n

    for g in reversed(grads):
        bucket.append(g)
        size += g.numel() * g.element_size()
        if size >= cap_bytes:
            flush()
            bucket, size = [], 0
    flush()


def run_reference(args):
    """The reference's own CPU implementation of the path, timed on this box's
    host cores.  /root/reference (pure Python, not installable: it has no
    setup.py / pyproject) cannot travel to the GPU box, so this is its port onto
    the very torch CPU ops it calls (oracle/torch_port.py, checked against the
    real reference via tests/golden): kind = "port".  With W > 1 ranks every rank
    runs it over gloo, with DDP's bucketed all-reduce in front of optimize()."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    os.environ["CUDA_VISIBLE_DEVICES"] = ""
    if args.workload in SPECIAL:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable":
                              f"workload {args.workload} has no CPU reference arm (its baseline is the torch-CUDA "
                              "restatement reported inside the line of the default arm)"}), flush=True)
        return
    import torch
    import torch.distributed as dist
    from oracle.torch_port import ReferenceUpdatePort
    from distributed_pipeline_b200.synthetic import numel
    if world > 1:
        dist.init_process_group("gloo", init_method="env://")
    affinity, physical = host_core_counts()
    # the reference's own rule (basic_utils/dist_run.py:31): physical cores // world size;
    # set explicitly because torchrun exports OMP_NUM_THREADS=1
    threads = max(1, physical // world)
    torch.set_num_threads(threads)

    shapes, dtype, rates, clip, wd = workload_shapes(args)
    tdt = torch.float32 if dtype == "f32" else torch.bfloat16
    gen = torch.Generator().manual_seed(102)                 # params identical on every rank
    params = [torch.nn.Parameter((torch.randn(s, generator=gen) * 0.02).to(tdt)) for s in shapes]
    n_params = sum(numel(s) for s in shapes)
    port = ReferenceUpdatePort(params, lr=LR, ema_rates=rates, weight_decay=wd, gradient_clipping=clip)
    ggen = torch.Generator().manual_seed(4321 + rank)        # gradients differ per rank
    grads = [(torch.randn(s, generator=ggen) * 1e-3).to(tdt) for s in shapes]

    def step():
        for p, g in zip(params, grads):
            p.grad = g.clone()  # stands in for what backward leaves; outside the reference's update path
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        if world > 1:
            ddp_style_allreduce([p.grad for p in params], world, dist)
        port.optimize()
        return time.perf_counter() - t0

    def agree(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=op)
        return float(t.item())

    first = agree(step(), dist.ReduceOp.MAX)
    # bounded sample: the arm must finish within minutes on any host; all ranks agree on the count
    budget_s = 90.0
    warm = max(0, min(args.warmup, int(15.0 / max(first, 1e-3))) - 1)
    for _ in range(warm):
        step()
    n_steps = max(2, min(args.steps, int(budget_s / max(first, 1e-3))))
    times = [step() for _ in range(n_steps)]
    t = agree(sum(times) / len(times), dist.ReduceOp.MAX)   # the slowest rank sets the step time
    if world > 1:
        dist.barrier()
    if rank == 0:
        bpp = bytes_per_param(dtype, len(rates), clip, "keep")
        value = world * n_params * bpp / t / 1e9             # aggregate over ranks, as the GPU arm counts
        sample = (f"{len(times)} steps ({warm + 1} warm-up) of the full workload ({n_params} params) on each of "
                  f"{world} rank(s): " + ("DDP-style bucketed gloo all-reduce + " if world > 1 else "") +
                  f"optimize(); torch {torch.__version__} CPU, {threads} thread(s) per rank")
        line = {
            "impl": "reference", "metric": "step_update_hbm_GBps", "value": value, "unit": "GB/s",
            "n_gpus": args.gpus, "steps": len(times), "warmup": warm + 1, "ms_per_step": t * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dtype, "data": "synthetic",
            "params_per_sec": world * n_params / t,
            "config": {"workload": args.workload if not args.flat_params else f"flat-{args.flat_params}",
                       "params": n_params, "tensors": len(shapes), "ema_rates": rates, "clip": clip,
                       "weight_decay": wd, "bytes_per_param": bpp, "world_size": world,
                       "host_threads_per_rank": threads, "physical_cores": physical, "affinity_cores": affinity,
                       "allreduce": "gloo, DDP-style 128 MiB buckets, pre-scaled by 1/W" if world > 1 else None},
            "cpu_baseline": {"value": value, "unit": "GB/s", "cores": threads * world, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


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


# ------------------------------------------------------------------ parity at N > 1 --

def _ordered_bits(t):
    """Integer image of a float tensor in which consecutive floats differ by 1."""
    import torch
    if t.dtype == torch.float32:
        i = t.view(torch.int32).to(torch.int64)
        return torch.where(i < 0, -(i & 0x7fffffff), i)
    i = t.view(torch.int16).to(torch.int64)
    return torch.where(i < 0, -(i & 0x7fff), i)


def parity_check(world, rank, dev, use_nvls, verbose=False):
    """One checked all-reduce and three checked update steps per dtype on a small
    ragged problem, through the same entry points the timed loop uses.

    all-reduce: the path selected for this run (pair exchange / NVSwitch multimem
      / ncclAllReduce) against (a) ncclAvg through torch and (b) the mean computed
      in float64 from an all_gather of every rank's input and rounded once.
      fp32: every element must BE that rounded mean (W = 2, 4, 8); at W = 2 also
      bit-equal to ncclAvg.  bf16: bit-equal at W = 2; for W >= 4 the switch's
      bf16 add is not one IEEE rounding, so the bound is one bf16 ulp of the
      largest addend: |ours - mean| <= 2^-8 * max_r |g_r|.
    update: (1) the one-kernel all-reduce + update (dpk_reduce_update) must leave
      parameters, both moments and every EMA copy bit-identical to all-reduce
      followed by the fused update; (2) that state against torch.optim.AdamW
      (foreach, CUDA) + the reference's eager EMA ops (utils/trainer.py:370) on the
      same averaged gradient: fp32 relative 1e-6, bf16 one ulp (`north_star`).
    """
    import torch
    import torch.distributed as dist
    from distributed_pipeline_b200 import FusedAdamW, _lib

    shapes = [(1000, 33), (7,), (300000,), (128, 128), (5,), (1200003,)]
    rates = [0.999, 0.9]
    report = {"allreduce": {}, "update": {}, "world_size": world, "path": None}
    ok = True
    worst_ulp = 0
    for dtype, name in ((torch.float32, "f32"), (torch.bfloat16, "bf16")):
        def make():
            torch.manual_seed(7)
            ps = [torch.nn.Parameter((torch.randn(s, device=dev) * 0.02).to(dtype)) for s in shapes]
            o = FusedAdamW(ps, lr=1e-3, weight_decay=0.01, ema_rates=rates, bucket_cap_mb=1.0,
                           grad_mode=_lib.GRAD_KEEP, symmetric_grads=use_nvls)
            o.engine.init_comm()
            if use_nvls:
                o.engine.enable_nvls(o.flat.grads)
            return o
        fused, staged = make(), make()
        forced = os.environ.get("DPK_FORCE_MULTIMEM") == "1"
        report["path"] = ("pair exchange" if world == 2 and not forced else "NVSwitch multimem") if staged.engine.nvls else "ncclAllReduce"
        # the torch side: same initial parameters, stock AdamW, eager EMA
        ref_p = [torch.nn.Parameter(p.detach().clone()) for p in staged.flat.param_list]
        ref_opt = torch.optim.AdamW(ref_p, lr=1e-3, weight_decay=0.01, foreach=True)
        ref_ema = [[p.detach().clone() for p in ref_p] for _ in rates]
        gen = torch.Generator(device=dev).manual_seed(100 + rank)
        ar = {"elements": 0, "off_rounded_mean": 0, "nccl_off_rounded_mean": 0, "bit_equal_ncclavg": True,
              "max_abs_diff_vs_ncclavg": 0.0, "max_err_in_spacings_of_largest_addend": 0.0}
        up = {"fused_equals_staged": True, "max_rel_err_vs_torch_adamw": 0.0, "max_ulp_vs_torch_adamw": 0,
              "logged_norm_rel_err": 0.0}
        n = staged.flat.layout.n_elems
        mask = torch.zeros(n, dtype=torch.bool, device=dev)
        for o_, c_ in zip(staged.flat.layout.offsets, staged.flat.layout.numels):
            mask[o_:o_ + c_] = True
        for it in range(3):
            g = (torch.randn(n, device=dev, generator=gen) * 0.05).to(dtype)
            nccl = g.clone()
            dist.all_reduce(nccl, op=dist.ReduceOp.AVG)
            gathered = [torch.empty_like(g) for _ in range(world)]
            dist.all_gather(gathered, g)
            mean64 = sum(x.double() for x in gathered) / world
            exact = mean64.to(dtype)
            biggest = torch.stack([x.abs().float() for x in gathered]).max(dim=0).values
            fused.flat.grads.copy_(g)
            staged.flat.grads.copy_(g)
            # the selected all-reduce path, then the update, as two steps
            if it == 2:
                staged.engine.all_buckets_ready()
            else:
                for b in range(len(staged.engine.buckets)):
                    staged.engine.bucket_ready(b)
            staged.engine.wait_allreduce()
            torch.cuda.synchronize()
            ours = staged.flat.grads.clone()
            err = (ours.double() - mean64).abs()
            # spacing of the dtype at the magnitude of the largest addend: 2^(floor(log2|x|) - mantissa bits)
            mant = 23 if dtype == torch.float32 else 7
            ulp_big = torch.exp2(torch.floor(torch.log2(torch.clamp(biggest.double(), min=1e-30))) - mant)
            ar["elements"] += n
            ar["off_rounded_mean"] += int((ours != exact).sum().item())
            ar["nccl_off_rounded_mean"] += int((nccl != exact).sum().item())
            ar["bit_equal_ncclavg"] = ar["bit_equal_ncclavg"] and bool(torch.equal(ours, nccl))
            ar["max_abs_diff_vs_ncclavg"] = max(ar["max_abs_diff_vs_ncclavg"], float((ours.float() - nccl.float()).abs().max().item()))
            ar["max_err_in_spacings_of_largest_addend"] = max(ar["max_err_in_spacings_of_largest_addend"],
                                                              float((err / ulp_big).max().item()))
            # replicas agree
            chk = ours.clone()
            dist.broadcast(chk, 0)
            ok = ok and bool(torch.equal(chk, ours))
            # the update: one-kernel form vs all-reduce-then-update
            fused.reduce_and_step()
            staged.step()
            torch.cuda.synchronize()
            for a, b in ((fused.flat.params, staged.flat.params), (fused.flat.exp_avg, staged.flat.exp_avg),
                         (fused.flat.exp_avg_sq, staged.flat.exp_avg_sq), (fused.flat.emas[0], staged.flat.emas[0]),
                         (fused.flat.emas[1], staged.flat.emas[1])):
                up["fused_equals_staged"] = up["fused_equals_staged"] and bool(torch.equal(a, b))
            ra, rb = fused.engine.read_results_sync(), staged.engine.read_results_sync()
            up["fused_equals_staged"] = up["fused_equals_staged"] and ra[_lib.R_LOGGED_NORM] == rb[_lib.R_LOGGED_NORM] \
                and ra[_lib.R_COMM_TIMEOUT] == 0 and rb[_lib.R_COMM_TIMEOUT] == 0
            # torch's AdamW + eager EMA on the same averaged gradient
            for p, gv in zip(ref_p, staged.flat.layout.views(ours, staged.flat.shapes)):
                p.grad = gv.clone()
            sq = sum(float((p.grad.float() ** 2).sum().item()) for p in ref_p)
            ref_opt.step()
            for rate, lst in zip(rates, ref_ema):
                for trg, src in zip(lst, ref_p):
                    trg.mul_(rate).add_(src.detach(), alpha=1 - rate)
            up["logged_norm_rel_err"] = max(up["logged_norm_rel_err"], abs(rb[_lib.R_LOGGED_NORM] - sq ** 0.5) / max(sq ** 0.5, 1e-30))
        # one more step WITH clipping: all-reduce + norm pass as one kernel (dpk_reduce_norm), then the update,
        # against all-reduce, norm pass, update as separate launches -- same bits expected
        g = (torch.randn(n, device=dev, generator=gen) * 0.05).to(dtype)
        clip_state = []
        for o in (fused, staged):
            o.flat.grads.copy_(g)
            keep = [t.clone() for t in (o.flat.params, o.flat.exp_avg, o.flat.exp_avg_sq, *o.flat.emas)]
            if o is fused:
                o.reduce_and_step(0.05)
            else:
                o.engine.all_buckets_ready()
                o.engine.wait_allreduce()
                o.clip_grad_norm(0.05)
                o.step()
            torch.cuda.synchronize()
            res = o.engine.read_results_sync()
            clip_state.append(([t.clone() for t in (o.flat.params, o.flat.exp_avg, o.flat.exp_avg_sq, *o.flat.emas)],
                               res[_lib.R_TOTAL_NORM], res[_lib.R_CLIP_COEF], res[_lib.R_LOGGED_NORM]))
            for t, k in zip((o.flat.params, o.flat.exp_avg, o.flat.exp_avg_sq, *o.flat.emas), keep):
                t.copy_(k)          # undo: the comparison with torch below is on the three unclipped steps
            o.step_count -= 1
        up["clip_step_fused_equals_staged"] = bool(
            all(torch.equal(a, b) for a, b in zip(clip_state[0][0], clip_state[1][0]))
            and clip_state[0][1:] == clip_state[1][1:] and clip_state[0][2] < 1.0)
        pairs = list(zip(staged.flat.param_list, ref_p))
        for k in range(len(rates)):
            pairs += list(zip(staged.flat.ema_views[k], ref_ema[k]))
        for a, b in pairs:
            if a.numel() == 0:
                continue
            a, b = a.detach(), b.detach()
            rms = b.double().pow(2).mean().sqrt()
            rel = ((a.double() - b.double()).abs() / (b.double().abs() + rms + 1e-30)).max().item()
            up["max_rel_err_vs_torch_adamw"] = max(up["max_rel_err_vs_torch_adamw"], float(rel))
            up["max_ulp_vs_torch_adamw"] = max(up["max_ulp_vs_torch_adamw"],
                                               int((_ordered_bits(a.contiguous()) - _ordered_bits(b.contiguous())).abs().max().item()))
        # verdicts.  `switch`: the NVSwitch did the additions (multimem.ld_reduce); its fp32 sums have
        # so far always come out as the correctly rounded mean (reported, and asserted in tests/), its
        # bf16 sums are NOT single IEEE roundings (measured: up to one bf16 spacing of the largest
        # addend away from the exact mean) -- so the gate for that path is an absolute one that a
        # missing or doubled contribution would still break by orders of magnitude.
        switch = staged.engine.nvls and (world > 2 or os.environ.get("DPK_FORCE_MULTIMEM") == "1")
        if switch:
            ar_ok = ar["max_err_in_spacings_of_largest_addend"] <= (1.0 if dtype == torch.float32 else 2.0)
        else:
            ar_ok = ar["bit_equal_ncclavg"]          # pair exchange / ncclAllReduce itself: one rounding of (a+b)/2, or NCCL's own bits
            if staged.engine.nvls:
                ar_ok = ar_ok and ar["off_rounded_mean"] == 0
        if dtype == torch.float32:
            up_ok = up["fused_equals_staged"] and up["max_rel_err_vs_torch_adamw"] <= 1e-6 and up["logged_norm_rel_err"] <= 1e-5
        else:
            up_ok = up["fused_equals_staged"] and up["max_ulp_vs_torch_adamw"] <= 1 and up["logged_norm_rel_err"] <= 2 ** -7
        up_ok = up_ok and up["clip_step_fused_equals_staged"]
        ar["ok"], up["ok"] = bool(ar_ok), bool(up_ok)
        ok = ok and ar_ok and up_ok
        worst_ulp = max(worst_ulp, up["max_ulp_vs_torch_adamw"])
        report["allreduce"][name] = ar
        report["update"][name] = up
        fused.engine.close()
        staged.engine.close()
        del fused, staged
    # every rank must agree that every rank passed
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    report["ok"] = bool(flag.item())
    report["max_ulp"] = worst_ulp
    report["tolerances"] = ("all-reduce: W=2 pair exchange / NCCL bit-equal to ncclAvg (= the correctly rounded mean); through "
                            "the NVSwitch |ours - float64 mean| <= 1 fp32 / 2 bf16 spacings of the largest addend "
                            "(off_rounded_mean reports how many elements are not the rounded mean: 0 expected in fp32); "
                            "update: one-kernel form bit-identical to all-reduce-then-update, and vs torch AdamW + eager EMA "
                            "on the same averaged gradient rel 1e-6 (fp32) / 1 ulp (bf16); logged norm 1e-5 / 2^-7")
    return report


# ------------------------------------------------------------------ the GPU arm --

def run_ours(args):
    import torch
    import torch.distributed as dist
    from distributed_pipeline_b200 import FusedAdamW, _lib
    from distributed_pipeline_b200.synthetic import numel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # a benchmark has no eval callbacks that could legitimately hold a rank back for minutes: a peer that does
    # not show up within a minute is gone, fail instead of parking the GPUs for the library's ten-minute default
    os.environ.setdefault("DPK_PEER_TIMEOUT_MS", "60000")
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
                     symmetric_grads=world > 1 and args.allreduce == "nvls",
                     sharded=bool(args.sharded and world > 1 and args.allreduce == "nvls"))
    opt.fused_reduce = not args.no_fused_reduce
    del params
    eng, flat = opt.engine, opt.flat
    if world > 1:
        eng.init_comm()
        eng.broadcast_state(0, True)
        if args.allreduce == "nvls":
            # every rank must reach the same decision (a rank that fell back to NCCL while the
            # others spin in the symmetric-memory kernels would hang the job)
            good = torch.tensor([1 if flat.symmetric else 0], device=dev)
            dist.all_reduce(good, op=dist.ReduceOp.MIN)
            why = f"symmetric allocation failed: {flat.symmetric_error}"
            if good.item():
                try:
                    eng.enable_nvls(flat.grads)
                except Exception as exc:
                    why = str(exc)
                good = torch.tensor([1 if eng.nvls else 0], device=dev)
                dist.all_reduce(good, op=dist.ReduceOp.MIN)
            if not good.item():
                if rank == 0:
                    print(f"[bench] symmetric-memory all-reduce unavailable ({why}); using NCCL", file=sys.stderr)
                eng.nvls = False
                args.allreduce = "nccl"
    if opt.sharded:
        args.no_e2e = True     # the host-gradient entry point (step_from_host) is a replica-path API
        if not eng.nvls:
            sys.exit("--sharded needs the symmetric-memory path")
        eng.enable_shard(flat.params)
    parity = None
    if world > 1 and not args.no_parity:
        parity = parity_check(world, rank, dev, eng.nvls)
    ggen = torch.Generator(device=dev).manual_seed(4321 + rank)
    sigma = 1e-3
    grads_src = (torch.randn(flat.layout.n_elems, generator=ggen, device=dev) * sigma).to(tdt)
    flat.grads.copy_(grads_src)
    n_buckets = len(eng.buckets)
    one_kernel = bool(world > 1 and eng.nvls and (opt.fused_reduce or opt.sharded) and not args.no_overlap and clip <= 0
                      and args.grad_mode == "keep")
    switch = world > 2 or os.environ.get("DPK_FORCE_MULTIMEM") == "1"
    # with clipping: all-reduce + norm pass as one kernel (dpk_reduce_norm) where the switch reduces
    norm_kernel = bool(world > 1 and eng.nvls and opt.fused_reduce and not args.no_overlap and clip > 0 and switch)

    def device_step():
        if world > 1 and not args.no_overlap:
            opt.reduce_and_step(clip)      # one kernel, or buckets consumed as their all-reduce lands
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
    comm_ms = eng.read_timings_sync(2)
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
    bucketed_update = world > 1 and not one_kernel and not args.no_overlap and clip <= 0
    per_step = n_buckets if bucketed_update else 1
    if per_step > 1:  # bucket-granular launches: one "launch" of the roofline = one full pass
        kernel_ms = [sum(kernel_ms[i:i + per_step]) for i in range(0, len(kernel_ms) - per_step + 1, per_step)]
    k_ms = sum(kernel_ms) / len(kernel_ms) if kernel_ms else float("nan")
    achieved = n_params * fused_bpp / (k_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(f"{args.workload}:{args.grad_mode}")
    kname = "fused_update_kernel"
    if one_kernel:
        pair = world == 2 and os.environ.get("DPK_FORCE_MULTIMEM") != "1"
        kname = "fused_update_kernel<kGradPair> (all-reduce inside)" if pair else "reduce_update_kernel (all-reduce inside)"
        if opt.sharded:
            kname = "shard_update_kernel (reduce-scatter + 1/W update + all-gather)"
    roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "kernel_ms_avg": k_ms, "kernel_ms_min": min(kernel_ms) if kernel_ms else None,
                "launches_timed": len(kernel_ms), "algorithmic_bytes_per_launch": n_params * fused_bpp,
                "frac_of_nominal_8TBps": achieved / 8000.0}
    if norm_ms:
        s = 4 if dtype == "f32" else 2
        per_norm = n_buckets if (world > 1 and not args.no_overlap and clip > 0 and not norm_kernel) else 1
        if per_norm > 1:
            norm_ms = [sum(norm_ms[i:i + per_norm]) for i in range(0, len(norm_ms) - per_norm + 1, per_norm)]
        n_ms = sum(norm_ms) / len(norm_ms)
        roofline["norm_pass"] = {"kernel": "reduce_norm_kernel (all-reduce inside)" if norm_kernel else "sumsq_tiles_kernel",
                                 "kernel_ms_avg": n_ms,
                                 "achieved": n_params * s / (n_ms * 1e-3) / 1e9,
                                 "frac": n_params * s / (n_ms * 1e-3) / 1e9 / peak}

    # ---- the collective against the NVLink roofline (N > 1) --------------------------
    roofline_comm = None
    if world > 1:
        S = flat.layout.n_elems * (4 if dtype == "f32" else 2)          # message bytes per rank
        if one_kernel:
            c_ms, c_kernel, c_n = k_ms, kname, len(kernel_ms)
        elif norm_kernel:
            c_ms, c_kernel, c_n = roofline["norm_pass"]["kernel_ms_avg"], "reduce_norm_kernel (all-reduce inside)", len(norm_ms)
        else:
            per = [sum(comm_ms[i:i + n_buckets]) for i in range(0, len(comm_ms) - n_buckets + 1, n_buckets)]
            c_ms = sum(per) / len(per) if per else float("nan")
            c_n = len(per)
            pair = world == 2 and os.environ.get("DPK_FORCE_MULTIMEM") != "1"
            c_kernel = (("pair_allreduce_kernel" if pair else "nvls_allreduce_kernel") if eng.nvls
                        else "ncclAllReduce") + f" x {n_buckets} buckets (sum of their durations)"
        algbw = S / (c_ms * 1e-3) / 1e9
        busbw = algbw * 2 * (world - 1) / world
        # bytes that must cross each GPU's links, per direction: the pair forms move S/2 out and back (one-kernel
        # pair: S in); the switch forms S(1 + 1/W); a ring 2(W-1)/W * S
        if eng.nvls:
            wire = S * (1.0 if (world == 2 and os.environ.get("DPK_FORCE_MULTIMEM") != "1") else 1.0 + 1.0 / world)
        else:
            wire = S * 2 * (world - 1) / world
        roofline_comm = {"bound": "nvlink", "kernel": c_kernel, "kernel_ms_avg": c_ms, "launches_timed": c_n,
                         "bytes": S, "algbw": algbw, "busbw": busbw, "unit": "GB/s",
                         "frac": busbw / NVLINK_NOMINAL_GBS, "peak": NVLINK_NOMINAL_GBS,
                         "peak_source": "nominal NVLink 5, per direction per GPU",
                         "wire_bytes_per_direction": wire,
                         "wire_frac": wire / (c_ms * 1e-3) / 1e9 / NVLINK_NOMINAL_GBS,
                         "wire_frac_of_measured_770": wire / (c_ms * 1e-3) / 1e9 / NVLINK_MEASURED_GBS,
                         "note": ("the collective and the update are the same kernel" if one_kernel else
                                  "the collective and the norm pass are the same kernel" if norm_kernel else
                                  "timed with CUDA events around the kernel, beside the pass it overlaps")}

    # ---- end to end: gradients from pinned host memory, result block read back -----
    e2e = None
    if not args.no_e2e:
        numa = bind_to_gpu_numa_node(local_rank) if world > 1 else None
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
               "staging_numa": numa,
               "api": "FusedAdamW.step_from_host (TrainLoop.opt): pinned host gradients, bucket-pipelined upload"}

    opt_sharded = opt.sharded
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        del opt, eng, flat
        torch.cuda.empty_cache()
        cpu = cpu_baseline(args, shapes, dtype, rates, clip, wd)

    failed = parity is not None and not parity["ok"]
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
                       "allreduce": (("own kernels over symmetric memory (NVSwitch multimem for W>=4, pair exchange for W=2)"
                                      if args.allreduce == "nvls" else "ncclAllReduce(avg)") if world > 1 else None),
                       "update_overlaps_allreduce": bool(world > 1 and not args.no_overlap),
                       "one_kernel_allreduce_update": one_kernel, "one_kernel_allreduce_norm": norm_kernel,
                       "sharded_update": bool(opt_sharded),
                       "grad_sigma": sigma},
            "roofline": roofline, "roofline_comm": roofline_comm, "parity_check": parity,
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches,
            "clocks": clocks.summary(),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if failed:
        sys.exit("parity_check failed: " + json.dumps(parity))


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "trainloop-bf16-ema2-clip":
        from scripts.bench_trainloop import run as run_trainloop
        run_trainloop(args, ClockSampler)
    elif args.workload == "sampler-t2000":
        from scripts.bench_sampler import run as run_sampler
        run_sampler(args, ClockSampler)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
