"""Worker for tests/test_gpu_multi.py: one rank of a W-rank NCCL job on real GPUs (W = 2, 4 or 8).

* bench.parity_check: the library's all-reduce for this world size (pair exchange at W=2, NVSwitch multimem at
  W>=4, or -- with DPK_FORCE_MULTIMEM=1 -- the switch kernels at W=2 too) against ncclAvg and against the
  float64 mean rounded once; the one-kernel all-reduce + update against all-reduce-then-update (bit-identical)
  and against torch.optim.AdamW + eager EMA.
* the same state against the reference port (oracle/torch_port.py) driven with ncclAvg gradients.
* a second pass with few collective blocks and small units, so that every collective block walks many units
  and update blocks wait on many different arrival counters.
* the loss-aware sampler's packed all-gather: rank-major order, ragged and empty local batches, bit-exact
  history / counts / weights against the numpy restatement.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import bench  # noqa: E402
from distributed_pipeline_b200 import FusedAdamW, _lib  # noqa: E402
from distributed_pipeline_b200.sampler import LossSecondMomentResampler  # noqa: E402
from oracle.sampler_oracle import LossSecondMomentResamplerOracle  # noqa: E402
from oracle.torch_port import ReferenceUpdatePort  # noqa: E402


def port_check(world, rank, dev, dtype):
    """Three CLIPPED steps of the library (its all-reduce + norm pass + fused update, clip 1.0, gradients
    rewritten as clip_grad_norm_ leaves them) against the reference port (oracle/torch_port.py: clip_grad_norm_,
    AdamW, eager EMA on torch-CUDA).

    (1) The port is fed the SAME averaged gradient the library's all-reduce produces (a second instance runs
        the bucket all-reduce), so what is compared is everything behind the all-reduce, clipping included:
        fp32 relative 1e-6, bf16 two bf16 ulps (the bf16 clip coefficient may land one ulp apart).
    (2) fp32 only: the port fed with ncclAvg gradients instead -- an independent end-to-end check that includes
        NCCL's own rounding.  Where the gradients are the same bits (W=2 pair exchange) 1e-6.  Through the switch
        every element is the correctly rounded mean while ncclAvg is an ulp of the largest addend off it in a
        third to half of them; where the W contributions nearly cancel that is a large relative error of the mean,
        and Adam's first steps divide by |mean|: measured 1.4e-6 at W=4, 3.4e-5 at W=8.  Gate there: 1e-4.
        (In bf16 neither NCCL's ring nor the switch yields the rounded mean and the same effect reaches 1e-1:
        not a meaningful gate, hence (1).)"""
    shapes = [(257, 33), (7,), (150001,), (64, 64), (3,)]
    switch = world > 2 or os.environ.get("DPK_FORCE_MULTIMEM") == "1"

    def make(mode):
        torch.manual_seed(11)
        ps = [torch.nn.Parameter((torch.randn(s, device=dev) * 0.02).to(dtype)) for s in shapes]
        o = FusedAdamW(ps, lr=1e-3, weight_decay=0.01, ema_rates=[0.999], bucket_cap_mb=0.25, grad_mode=mode,
                       symmetric_grads=True)
        o.engine.init_comm()
        o.engine.enable_nvls(o.flat.grads)
        return ps, o
    ps, opt = make(_lib.GRAD_CLIPPED)
    _, reducer = make(_lib.GRAD_KEEP)                   # only its all-reduce is used
    ports = {}
    for feed in (("own", "nccl") if dtype == torch.float32 else ("own",)):
        rp = [torch.nn.Parameter(p.detach().clone()) for p in ps]
        ports[feed] = (rp, ReferenceUpdatePort(rp, lr=1e-3, ema_rates=[0.999], weight_decay=0.01, gradient_clipping=1.0))
    gen = torch.Generator(device=dev).manual_seed(900 + rank)
    for _ in range(3):
        g = (torch.randn(opt.flat.layout.n_elems, device=dev, generator=gen) * 0.05).to(dtype)
        opt.flat.grads.copy_(g)
        reducer.flat.grads.copy_(g)
        reducer.engine.all_buckets_ready()
        reducer.engine.wait_allreduce()
        torch.cuda.synchronize()
        avg = {"own": reducer.flat.grads.clone(), "nccl": g.clone()}
        dist.all_reduce(avg["nccl"], op=dist.ReduceOp.AVG)
        for feed, (rp, port) in ports.items():
            for p, gv in zip(rp, opt.flat.layout.views(avg[feed], opt.flat.shapes)):
                p.grad = gv.clone()
            port.optimize()
        opt.reduce_and_step(1.0)
    torch.cuda.synchronize()
    worst = {}
    for feed, (rp, port) in ports.items():
        w = 0.0
        for a_, b_ in list(zip(ps, rp)) + list(zip(opt.flat.ema_views[0], port.ema_params[0])):
            a_, b_ = a_.detach().double(), b_.detach().double()
            rms = b_.pow(2).mean().sqrt()
            w = max(w, float(((a_ - b_).abs() / (b_.abs() + rms)).max().item()))
        worst[feed] = w
    tol_own = 1e-6 if dtype == torch.float32 else 2 ** -7
    assert worst["own"] <= tol_own, f"{dtype} W={world}: differs from the reference port by {worst['own']:.3e} (> {tol_own:.1e})"
    if "nccl" in worst:
        tol = 1e-4 if switch else 1e-6
        assert worst["nccl"] <= tol, f"{dtype} W={world}: differs from the ncclAvg-fed port by {worst['nccl']:.3e} (> {tol:.1e})"
    opt.engine.close()
    reducer.engine.close()
    return worst


def shard_check(world, rank, dev, dtype):
    """Sharded (ZeRO-1 style) update against the replica path on the same per-rank gradients: parameters
    bit-identical on every rank after every step; after gather_shards() both moments and the EMA copies too."""
    shapes = [(257, 33), (7,), (150001,), (64, 64), (3,), (40000,)]

    def make(sharded):
        torch.manual_seed(21)
        ps = [torch.nn.Parameter((torch.randn(s, device=dev) * 0.02).to(dtype)) for s in shapes]
        o = FusedAdamW(ps, lr=1e-3, weight_decay=0.01, ema_rates=[0.999, 0.9], bucket_cap_mb=0.25,
                       grad_mode=_lib.GRAD_KEEP, symmetric_grads=True, sharded=sharded)
        o.engine.init_comm()
        o.engine.enable_nvls(o.flat.grads)
        if sharded:
            o.engine.enable_shard(o.flat.params)
        return o
    sh, rep = make(True), make(False)
    gen = torch.Generator(device=dev).manual_seed(300 + rank)
    for it in range(3):
        g = (torch.randn(sh.flat.layout.n_elems, device=dev, generator=gen) * 0.05).to(dtype)
        sh.flat.grads.copy_(g)
        rep.flat.grads.copy_(g)
        sh.reduce_and_step()
        rep.engine.all_buckets_ready()
        rep.engine.wait_allreduce()
        rep.step()
        torch.cuda.synchronize()
        assert torch.equal(sh.flat.grads, g), "the sharded pass leaves the local gradient untouched"
        assert torch.equal(sh.flat.params, rep.flat.params), f"{dtype} W={world} step {it}: sharded parameters differ"
        a, b = sh.engine.read_results_sync(), rep.engine.read_results_sync()
        assert abs(a[_lib.R_LOGGED_NORM] - b[_lib.R_LOGGED_NORM]) <= 1e-6 * b[_lib.R_LOGGED_NORM] and a[_lib.R_COMM_TIMEOUT] == 0
    # owner-only state before the gather, complete state after it
    assert not torch.equal(sh.flat.exp_avg, rep.flat.exp_avg)
    sd = sh.state_dict()          # gathers
    torch.cuda.synchronize()
    for name, x, y in (("exp_avg", sh.flat.exp_avg, rep.flat.exp_avg), ("exp_avg_sq", sh.flat.exp_avg_sq, rep.flat.exp_avg_sq),
                       ("ema0", sh.flat.emas[0], rep.flat.emas[0]), ("ema1", sh.flat.emas[1], rep.flat.emas[1])):
        assert torch.equal(x, y), f"{dtype} W={world}: {name} differs after gather_shards"
    assert len(sd["state"]) == len(shapes)
    # and stepping on after a gather keeps working
    g = (torch.randn(sh.flat.layout.n_elems, device=dev, generator=gen) * 0.05).to(dtype)
    sh.flat.grads.copy_(g); rep.flat.grads.copy_(g)
    sh.reduce_and_step()
    rep.engine.all_buckets_ready(); rep.engine.wait_allreduce(); rep.step()
    sh.gather_shards()
    torch.cuda.synchronize()
    assert torch.equal(sh.flat.params, rep.flat.params) and torch.equal(sh.flat.emas[1], rep.flat.emas[1])
    sh.engine.close(); rep.engine.close()


def sampler_check(world, rank, dev):
    T = 50
    holder = FusedAdamW([torch.nn.Parameter(torch.zeros(1024, device=dev))], ema_rates=[])
    holder.engine.init_comm()
    sampler = LossSecondMomentResampler(T, history_per_term=4, device=dev, engine=holder.engine, max_local_batch=16)
    ref = LossSecondMomentResamplerOracle(T, 4)
    rng = np.random.RandomState(7)                         # same stream on every rank: all ranks know all inputs
    for _ in range(40):
        sizes = [int(rng.randint(0, 17)) for _ in range(world)]   # ragged, possibly empty, per rank
        ts = [rng.randint(0, T, size=s) for s in sizes]
        ls = [(rng.rand(s) + 0.1).astype(np.float32) for s in sizes]
        sampler.update_with_local_losses(torch.from_numpy(ts[rank]).to(dev), torch.from_numpy(ls[rank]).to(dev))
        ref.update_with_gathered(ts, ls)
    hist, counts = sampler.state()
    assert np.array_equal(counts.cpu().numpy().astype(np.int64), ref._loss_counts)
    assert np.array_equal(hist.cpu().numpy().view(np.uint64), ref._loss_history.view(np.uint64))
    assert np.array_equal(sampler.weights().view(np.uint64), ref.weights().view(np.uint64))
    np.random.seed(5)
    a = sampler.sample(32, dev)[0].cpu().numpy()
    np.random.seed(5)
    b = ref.sample(32)[0]
    assert np.array_equal(a, np.asarray(b)), "sampled timesteps differ for the same RNG state"
    sampler.close()
    holder.engine.close()


def main():
    tmp = sys.argv[1]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rep = bench.parity_check(world, rank, dev, use_nvls=True)
    if rank == 0:
        print("parity_check", rep, flush=True)
    assert rep["ok"], rep
    # stricter than bench.py's gate: every fp32 element the switch produced is the correctly rounded mean
    assert rep["allreduce"]["f32"]["off_rounded_mean"] == 0, rep["allreduce"]["f32"]
    # few collective blocks, small units: every block walks many units
    os.environ["DPK_FUSED_COMM_BLOCKS"] = "3"
    os.environ["DPK_FUSED_UNIT_KB"] = "64"
    rep2 = bench.parity_check(world, rank, dev, use_nvls=True)
    assert rep2["ok"], rep2
    del os.environ["DPK_FUSED_COMM_BLOCKS"], os.environ["DPK_FUSED_UNIT_KB"]
    for dtype in (torch.float32, torch.bfloat16):
        w = port_check(world, rank, dev, dtype)
        if rank == 0:
            print(f"reference port, {dtype}, W={world}: worst relative difference {w}", flush=True)
    switch = world > 2 or os.environ.get("DPK_FORCE_MULTIMEM") == "1"
    for dtype in (torch.float32, torch.bfloat16):
        if dtype == torch.bfloat16 and not switch:
            continue      # two ranks' replica path averages bf16 pairs exactly; the switch's bf16 add differs (see bench.parity_check)
        shard_check(world, rank, dev, dtype)
        if rank == 0:
            print(f"sharded update, {dtype}, W={world}: bit-identical to the replica path", flush=True)
    sampler_check(world, rank, dev)
    dist.barrier()
    if rank == 0:
        open(os.path.join(tmp, "allreduce_ok"), "w").write("ok")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
