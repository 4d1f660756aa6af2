"""N>1 host logic on CPU: two gloo ranks.

The product's all-reduce is NCCL inside libdpk, which cannot run here; what CAN
be pinned without a GPU is everything around it: which buckets fire, in which
order, on which microbatch (DDP decides at forward time, the reference leaves
no_sync() before backward), and that averaging in place on flat bucket ranges
gives the gradients torch DDP gives.  The engine is replaced by a test double
that performs the bucket all-reduce with gloo on the flat CPU gradient buffer.
"""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class GlooEngine:
    """Stand-in with the Engine surface BucketReducer uses."""

    def __init__(self, flat, ranges):
        self.flat, self._ranges, self.world_size = flat, ranges, dist.get_world_size()
        self.fired, self.waits = [], 0

    @property
    def buckets(self):
        return self._ranges

    def bucket_ready(self, b, stream=None):
        s, c = self._ranges[b]
        view = self.flat.grads[s:s + c]
        view.mul_(1.0 / self.world_size)          # reducer semantics: pre-multiply by 1/W, then SUM (F8)
        dist.all_reduce(view)
        self.fired.append(b)

    def wait_allreduce(self, stream=None):
        self.waits += 1


def _worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank), CUDA_VISIBLE_DEVICES="")
    from distributed_pipeline_b200 import dist_util
    from distributed_pipeline_b200.engine import plan_buckets
    from distributed_pipeline_b200.flat import FlatState
    from distributed_pipeline_b200.reducer import BucketReducer, FlatDDP
    from distributed_pipeline_b200.synthetic import SyntheticDiffuSeq
    from torch.nn.parallel import DistributedDataParallel

    assert dist_util.setup_dist(backend="gloo", silent=True) is True
    assert dist_util.is_initialized() and dist_util.get_world_size() == world and dist_util.get_rank() == rank
    assert dist_util.setup_dist() is True  # idempotent

    kw = dict(n_layers=2, d=32, ffn=64, vocab=40, emb=16, t_hidden=24, max_pos=16, heads=2)
    torch.manual_seed(102 + rank)             # the reference seeds seed+rank (utils/initialization.py:11)
    ours = SyntheticDiffuSeq(**kw)
    torch.manual_seed(102 + rank)
    theirs = SyntheticDiffuSeq(**kw)

    # coalesced sync_params == per-tensor broadcast (reference dist_util.py:141-152)
    flat = FlatState(list(ours.parameters()), n_ema=1)
    dist_util.sync_params(flat.param_views)                # one broadcast of the flat range
    ddp = DistributedDataParallel(theirs, broadcast_buffers=False, bucket_cap_mb=128, find_unused_parameters=False)
    for a, b in zip(ours.parameters(), theirs.parameters()):
        assert torch.equal(a, b), "rank-0 parameters must win on every rank"
    loose = [torch.full((3,), float(rank)), torch.full((2, 2), float(rank))]
    dist_util.sync_params(loose)                            # per-tensor path
    assert all(float(t.max()) == 0.0 for t in loose)

    ranges = plan_buckets(flat.layout, torch.float32, bucket_cap_bytes=8 * 1024)
    assert len(ranges) > 2
    eng = GlooEngine(flat, ranges)
    red = BucketReducer(flat, eng)
    wrapped = FlatDDP(ours, red)
    assert wrapped.module is ours

    gen = torch.Generator().manual_seed(1234 + rank)
    for step in range(3):
        flat.grads.zero_()
        for p in theirs.parameters():
            if p.grad is not None:
                p.grad.detach_()
                p.grad.zero_()
        n_micro = 3
        for i in range(n_micro):
            ids, t = torch.randint(0, 40, (2, 8), generator=gen), torch.randint(0, 100, (2,), generator=gen)
            last = i == n_micro - 1
            # exactly the reference's pattern (utils/trainer.py:216-235): forward inside
            # no_sync() for all but the last microbatch, backward OUTSIDE the context
            if last:
                lo, lt = wrapped(ids, t).pow(2).mean(), ddp(ids, t).pow(2).mean()
            else:
                with wrapped.no_sync():
                    lo = wrapped(ids, t).pow(2).mean()
                with ddp.no_sync():
                    lt = ddp(ids, t).pow(2).mean()
            fired_before = len(eng.fired)
            lo.backward()
            lt.backward()
            if not last:
                assert len(eng.fired) == fired_before, "no all-reduce on a no_sync microbatch"
        # one pass over every bucket, tail first, one wait
        assert eng.fired[-len(ranges):] == list(range(len(ranges))) and len(eng.fired) == (step + 1) * len(ranges)
        assert eng.waits == step + 1
        for (name, a), b in zip(ours.named_parameters(), theirs.parameters()):
            assert torch.equal(a.grad, b.grad), f"{name}: flat-bucket average differs from DDP (W=2 is order-free)"
    # identical on both ranks
    g = flat.grads.clone()
    dist.broadcast(g, 0)
    assert torch.equal(g, flat.grads)
    red.remove()
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(tmp, f"ok{rank}"), "w").write("ok")


def test_two_rank_bucket_reducer_matches_ddp(tmp_path):
    from distributed_pipeline_b200 import dist_util
    port = dist_util.find_free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()
