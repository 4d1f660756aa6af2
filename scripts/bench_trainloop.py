"""bench.py --workload trainloop-bf16-ema2-clip  (BASELINE.json configs[2]).

Whole ``TrainLoop.run_step`` on DiffuSeq-base with every state tensor in bf16 (the reference keeps no fp32
master copy), gradient clipping 1.0, EMA rates 0.9999 and 0.999, global batch 2048 split over the
W ranks, microbatches of 64 with ``no_sync()`` on all but the last one -- so the bucketed all-reduce runs from
gradient hooks on a side stream *while the last microbatch's backward is still running*, which is the thing
``north_star`` describes.  Reported, all as max over ranks, CUDA events around whole steps:

  step_ms           drop-in loop, everything on
  fwd_bwd_ms        the same forward/backward work with no all-reduce and no update
  tail_ms           step_ms - fwd_bwd_ms: what the post-backward path costs on top of backward
  no_comm_step_ms   drop-in loop with the all-reduce switched off (each rank updates on its local gradients)
  exposed_comm_ms   step_ms - no_comm_step_ms: communication that backward did not hide
  reference_step_ms the reference's loop restated on torch-CUDA on the same box: DistributedDataParallel
                    (bucket_cap_mb=128) + clip_grad_norm_ + 209 .item() + AdamW + eager EMA
                    (/root/reference/utils/trainer.py:198-275, 360-370)
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GLOBAL_BATCH, MICRO, SEQ = 2048, 64, 128


def run(args, ClockSampler):
    from distributed_pipeline_b200 import TrainLoop, dist_util, logger
    from distributed_pipeline_b200.synthetic import SyntheticDiffuSeq

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("LOCAL_RANK", "0")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        assert dist_util.setup_dist(silent=True)
    per_rank = GLOBAL_BATCH // world
    n_micro = per_rank // MICRO

    class Loop(TrainLoop):
        def compute_losses(self, mb):
            out = self.ddp_model(mb["ids"], mb["t"])
            return {"loss": ((out.float() - mb["target"]) ** 2).mean(dim=(1, 2))}

        @staticmethod
        def backward_from_losses(losses):
            losses["loss"].mean().backward()

        def log_loss_dict(self, mode, losses, *a, **k):
            pass

    gen = torch.Generator().manual_seed(1234 + rank)
    batch = {"ids": torch.randint(0, 30522, (per_rank, SEQ), generator=gen),
             "t": torch.randint(0, 2000, (per_rank,), generator=gen),
             "target": torch.randn(per_rank, SEQ, 128, generator=gen)}
    batch = {k: v.pin_memory() for k, v in batch.items()}
    logger.configure(dir=f"/tmp/dpk-bench-tl-{rank}")

    def make_loop():
        torch.manual_seed(102)
        model = SyntheticDiffuSeq().to(dev).to(torch.bfloat16)
        return Loop(model=model, data=iter([]), batch_size=per_rank, microbatch=MICRO, lr=1e-4,
                    ema_rate="0.9999,0.999", log_interval=10 ** 9, save_interval=10 ** 9, resume_checkpoint="",
                    gradient_clipping=1.0, grads_after_step="zero")

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / steps
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    steps = max(2, min(args.steps, 6 if n_micro > 8 else 20))
    warmup = max(3, min(args.warmup, 3))

    loop = make_loop()
    launches0 = loop.opt.engine.launch_count

    def ours():
        loop.run_step(batch)
        loop.step += 1

    with ClockSampler(local) as clocks:
        step_ms = timed(ours, steps, warmup)
    launches = loop.opt.engine.launch_count - launches0
    loop.flush_logs()

    def fwd_bwd_only():
        loop._zero_grad()
        for i in range(0, per_rank, MICRO):
            mb = {k: v[i:i + MICRO].to(dev, non_blocking=True) for k, v in batch.items()}
            with loop.ddp_model.no_sync() if loop.use_ddp else torch.enable_grad():
                losses = loop.compute_losses(mb)
            loop._grads_are_zero = False
            loop.backward_from_losses(losses)

    fwd_bwd_ms = timed(fwd_bwd_only, steps, warmup)

    no_comm_ms = None
    if loop.use_ddp:
        loop._reducer.enabled = False          # hooks never arm: no collective is launched
        no_comm_ms = timed(ours, steps, warmup)
        loop._reducer.enabled = True
    allreduce_kind = None
    if loop.use_ddp:
        allreduce_kind = "own symmetric-memory kernels" if loop.opt.engine.nvls else "ncclAllReduce"
    n_params = loop.opt.flat.layout.n_params
    if loop._reducer is not None:
        loop._reducer.remove()
    del loop
    torch.cuda.empty_cache()

    # ---- the reference's loop on torch-CUDA (same box, same model, same microbatching) ----------------
    torch.manual_seed(102)
    ref_model = SyntheticDiffuSeq().to(dev).to(torch.bfloat16)
    params = list(ref_model.parameters())
    ema = [[p.detach().clone() for p in params] for _ in (0.9999, 0.999)]
    ddp = ref_model
    if world > 1:
        ddp = torch.nn.parallel.DistributedDataParallel(ref_model, device_ids=[dev], output_device=dev,
                                                        broadcast_buffers=False, bucket_cap_mb=128,
                                                        find_unused_parameters=False)
    opt = torch.optim.AdamW(params, lr=1e-4, weight_decay=0.0)

    def reference():
        for p in params:                                    # _zero_grad, :203-207
            if p.grad is not None:
                p.grad.detach_()
                p.grad.zero_()
        for i in range(0, per_rank, MICRO):                  # forward_backward, :209-235
            mb = {k: v[i:i + MICRO].to(dev) for k, v in batch.items()}
            last = i + MICRO >= per_rank
            if last or world == 1:
                out = ddp(mb["ids"], mb["t"])
            else:
                with ddp.no_sync():
                    out = ddp(mb["ids"], mb["t"])
            ((out.float() - mb["target"]) ** 2).mean(dim=(1, 2)).mean().backward()
        torch.nn.utils.clip_grad_norm_(params, 1.0)           # optimize, :237-244
        sqsum = 0.0
        for p in params:
            sqsum += (p.grad ** 2).sum().item()              # _log_grad_norm, :265-271
        opt.step()
        for rate, targets in zip((0.9999, 0.999), ema):      # update_ema, :360-370
            for trg, src in zip(targets, params):
                trg.mul_(rate).add_(src.detach(), alpha=1 - rate)

    ref_ms = timed(reference, steps, warmup)

    if rank == 0:
        line = {
            "metric": "trainloop_step_ms", "value": step_ms, "unit": "ms", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": step_ms, "higher_is_better": False, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "trainloop-bf16-ema2-clip", "params": n_params, "global_batch": GLOBAL_BATCH,
                       "per_rank_batch": per_rank, "microbatch": MICRO, "microbatches_per_step": n_micro,
                       "seq_len": SEQ, "clip": 1.0, "ema_rates": [0.9999, 0.999], "allreduce": allreduce_kind,
                       "l2_policy": "inputs larger than L2 (model state 1.3 GB, activations per microbatch > 126 MB)"},
            "fwd_bwd_ms": fwd_bwd_ms, "tail_ms": step_ms - fwd_bwd_ms,
            "no_comm_step_ms": no_comm_ms,
            "exposed_comm_ms": (step_ms - no_comm_ms) if no_comm_ms is not None else None,
            "reference_step_ms": ref_ms, "reference_tail_ms": ref_ms - fwd_bwd_ms,
            "reference": "the reference's loop on torch-CUDA: DDP(bucket_cap_mb=128) + clip_grad_norm_ + 209 .item() + "
                         "AdamW + eager EMA, same model and microbatching",
            "gpu_launches": launches, "clocks": clocks.summary(),
            "e2e": {"value": step_ms, "unit": "ms", "h2d_bytes_per_step": sum(v.numel() * v.element_size() for v in batch.values()),
                    "d2h_bytes_per_step": 64,
                    "api": "TrainLoop.run_step: host batches, per-microbatch H2D copies, grad-norm read-back ring"},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
