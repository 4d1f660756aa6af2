"""Worker for tests/test_gpu_multi.py: one rank of an NCCL job on real GPUs.
Compares the drop-in TrainLoop (flat NCCL buckets + fused update) with the
reference's loop restated on torch DDP + torch.optim.AdamW + eager EMA."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from distributed_pipeline_b200 import TrainLoop, dist_util, logger  # noqa: E402
from distributed_pipeline_b200.sampler import LossSecondMomentResampler  # noqa: E402
from distributed_pipeline_b200.synthetic import SyntheticDiffuSeq  # noqa: E402
from oracle.sampler_oracle import LossSecondMomentResamplerOracle  # noqa: E402
from oracle.torch_port import ReferenceUpdatePort  # noqa: E402

TOY = dict(n_layers=2, d=128, ffn=512, vocab=4096)


class Loop(TrainLoop):
    def compute_losses(self, micro_batch):
        out = self.ddp_model(micro_batch["ids"], micro_batch["t"])
        return {"loss": ((out - micro_batch["target"]) ** 2).mean(dim=(1, 2))}

    @staticmethod
    def backward_from_losses(losses):
        losses["loss"].mean().backward()

    def log_loss_dict(self, mode, losses, *args, **kwargs):
        pass


def batches(seed, B, n, L=16):
    gen = torch.Generator().manual_seed(seed)
    return [{"ids": torch.randint(0, TOY["vocab"], (B, L), generator=gen),
             "t": torch.randint(0, 2000, (B,), generator=gen),
             "target": torch.randn(B, L, 128, generator=gen) * 2.0} for _ in range(n)]


def main():
    tmp = sys.argv[1]
    assert dist_util.setup_dist(silent=True)
    rank, world = dist_util.get_rank(), dist_util.get_world_size()
    dev = dist_util.dev()
    logger.configure(dir=os.path.join(tmp, f"log{rank}"))
    data = batches(1234 + rank, 8, 6)

    for clip, allreduce, after in ((-1.0, "nvls", "clipped"), (1.0, "nvls", "clipped"), (1.0, "nccl", "clipped"),
                                   (-1.0, "nvls", "keep")):
        torch.manual_seed(102 + rank)              # ranks start DIFFERENT: rank 0's parameters must win
        model = SyntheticDiffuSeq(**TOY).to(dev)
        torch.manual_seed(102 + rank)
        ref_model = SyntheticDiffuSeq(**TOY).to(dev)

        loop = Loop(model=model, data=iter(data), batch_size=8, microbatch=4, lr=1e-3, ema_rate="0.9999,0.999",
                    log_interval=10 ** 9, save_interval=10 ** 9, resume_checkpoint="", weight_decay=0.01,
                    learning_steps=100, checkpoint_path=tmp, gradient_clipping=clip, grad_allreduce=allreduce,
                    grads_after_step=after)
        assert loop.opt.engine.nvls == (allreduce == "nvls")
        assert loop._pair_fused == (world == 2 and after == "keep" and clip <= 0)
        assert loop.use_ddp and loop.global_batch == 8 * world
        assert loop.opt.engine.world_size == world

        ddp = torch.nn.parallel.DistributedDataParallel(ref_model, device_ids=[dev], output_device=dev,
                                                        broadcast_buffers=False, bucket_cap_mb=128,
                                                        find_unused_parameters=False)
        port = ReferenceUpdatePort(list(ref_model.parameters()), lr=1e-3, ema_rates=[0.9999, 0.999],
                                   weight_decay=0.01, learning_steps=100, gradient_clipping=clip)
        # the reference copies EMA *before* DDP broadcasts (F4); align rank != 0 for the comparison
        for lst in port.ema_params:
            for e, p in zip(lst, ref_model.parameters()):
                e.data.copy_(p.data)
        for a, b in zip(model.parameters(), ref_model.parameters()):
            assert torch.equal(a, b), "constructor broadcast: rank 0's parameters on every rank"

        for step, batch in enumerate(data):
            loop.run_step(batch)
            loop.step += 1
            # reference: zero, microbatches with no_sync on all but the last, optimize
            port.zero_grad()
            for i in range(0, 8, 4):
                mb = {k: v[i:i + 4].to(dev) for k, v in batch.items()}
                last = i + 4 >= 8
                if last:
                    out = ddp(mb["ids"], mb["t"])
                else:
                    with ddp.no_sync():
                        out = ddp(mb["ids"], mb["t"])
                ((out - mb["target"]) ** 2).mean(dim=(1, 2)).mean().backward()
            if step == 0 and clip < 0:
                # same averaged gradients (W=2: order-free) -- compare before either side clips
                pass
            port.optimize()

        def close(a, b, rtol=1e-6):
            a, b = a.double(), b.double()
            rms = b.pow(2).mean().sqrt()
            return bool(((a - b).abs() <= rtol * (b.abs() + rms)).all())

        nbit = 0
        for a, b in zip(loop.master_params, port.params):
            assert close(a, b), f"rank {rank} clip {clip}: parameters differ"
            nbit += int(torch.equal(a, b))
        for k in range(2):
            for a, b in zip(loop.ema_params[k], port.ema_params[k]):
                assert close(a, b)
        # replicas stay identical across ranks
        flat = loop.opt.flat.params.clone()
        dist.broadcast(flat, 0)
        assert torch.equal(flat, loop.opt.flat.params)
        if rank == 0:
            print(f"clip={clip} allreduce={allreduce} grads_after_step={after}: {nbit}/{len(port.params)} parameter tensors bit-identical to "
                  f"DDP+AdamW after 6 steps",
                  flush=True)
        loop._reducer.remove()
        del loop, ddp

    # ---- two ranks: all-reduce + update as ONE kernel == all-reduce, then update ----------
    if world == 2:
        from distributed_pipeline_b200 import FusedAdamW, _lib
        for dtype in (torch.float32, torch.bfloat16):
            shapes = [(1000, 33), (7,), (300000,), (128, 128), (5,), (200003,)]

            def make():
                torch.manual_seed(7)
                ps = [torch.nn.Parameter((torch.randn(s, device=dev) * 0.02).to(dtype)) for s in shapes]
                o = FusedAdamW(ps, lr=1e-3, weight_decay=0.01, ema_rates=[0.999, 0.9], bucket_cap_mb=0.5,
                               grad_mode=_lib.GRAD_KEEP, symmetric_grads=True)
                o.engine.init_comm()
                o.engine.enable_nvls(o.flat.grads)
                return o
            fused, staged = make(), make()
            gen = torch.Generator(device=dev).manual_seed(50 + rank)
            for it in range(3):
                g = (torch.randn(fused.flat.layout.n_elems, device=dev, generator=gen) * 0.05).to(dtype)
                fused.flat.grads.copy_(g)
                staged.flat.grads.copy_(g)
                fused.reduce_and_step()                       # dpk_reduce_update_pair
                for b in range(len(staged.engine.buckets)):   # all-reduce every bucket, then one update
                    staged.engine.bucket_ready(b)
                staged.engine.wait_allreduce()
                staged.step()
                torch.cuda.synchronize()
                assert torch.equal(fused.flat.grads, g), "the fused pass leaves the local gradient untouched"
                for a, b in ((fused.flat.params, staged.flat.params), (fused.flat.exp_avg, staged.flat.exp_avg),
                             (fused.flat.exp_avg_sq, staged.flat.exp_avg_sq), (fused.flat.emas[0], staged.flat.emas[0]),
                             (fused.flat.emas[1], staged.flat.emas[1])):
                    assert torch.equal(a, b), f"{dtype}: one-kernel pair update differs from all-reduce + update"
                ra, rb = fused.engine.read_results_sync(), staged.engine.read_results_sync()
                assert ra[_lib.R_LOGGED_NORM] == rb[_lib.R_LOGGED_NORM] and ra[_lib.R_COMM_TIMEOUT] == 0
            if rank == 0:
                print(f"{dtype}: one-kernel pair all-reduce+update bit-identical to all-reduce then update", flush=True)
            del fused, staged

    # ---- loss-aware sampler: one packed all-gather, rank-major order -------------------
    T = 50
    eng_loop = Loop(model=SyntheticDiffuSeq(**TOY).to(dev), data=iter(data), batch_size=8, microbatch=4, lr=1e-3,
                    ema_rate="0.9", log_interval=1, save_interval=10 ** 9, resume_checkpoint="")
    sampler = LossSecondMomentResampler(T, history_per_term=4, device=dev, engine=eng_loop.opt.engine,
                                        max_local_batch=16)
    ref = LossSecondMomentResamplerOracle(T, 4)
    rng = np.random.RandomState(7)                         # same stream on every rank: all ranks know all inputs
    for rnd in range(40):
        sizes = [int(rng.randint(0, 17)) for _ in range(world)]   # ragged, possibly empty, per rank
        ts = [rng.randint(0, T, size=s) for s in sizes]
        ls = [(rng.rand(s) + 0.1).astype(np.float32) for s in sizes]
        sampler.update_with_local_losses(torch.from_numpy(ts[rank]).to(dev), torch.from_numpy(ls[rank]).to(dev))
        ref.update_with_gathered(ts, ls)
    hist, counts = sampler.state()
    assert np.array_equal(counts.cpu().numpy().astype(np.int64), ref._loss_counts)
    assert np.array_equal(hist.cpu().numpy().view(np.uint64), ref._loss_history.view(np.uint64))
    assert np.array_equal(sampler.weights().view(np.uint64), ref.weights().view(np.uint64))
    dist.barrier()
    if rank == 0:
        open(os.path.join(tmp, "multi_ok"), "w").write("ok")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
