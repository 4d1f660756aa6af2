"""bench.py's reference arm (the reference path's CPU port, timed on host cores) runs without a GPU and prints
one JSON line with the keys the driver reads."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_contract_line():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    proc = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload",
                           "toy-fp32-ema2-clip", "--steps", "3", "--warmup", "1"],
                          capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert proc.returncode == 0, proc.stderr[-2000:]
    lines = [l for l in proc.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "step_update_hbm_GBps" and d["unit"] == "GB/s"
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["steps"] == 3
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"] == "toy-fp32-ema2-clip" and "model" not in d["config"]
    # (8 + 2K) * s bytes per parameter with clipping: K = 2, fp32 -> 48
    assert d["config"]["bytes_per_param"] == 48


def test_reference_arm_runs_on_every_rank_over_gloo():
    """N>1: the reference arm is launched like the GPU arm (torchrun); every rank runs the port with DDP's
    pre-scale + bucketed all-reduce over gloo in front, threads = physical cores // W, rank 0 prints the line."""
    sys.path.insert(0, ROOT)
    from distributed_pipeline_b200.dist_util import find_free_port
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    proc = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                           "--master-addr", "127.0.0.1", "--master-port", str(find_free_port()),
                           os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--workload",
                           "toy-fp32-ema2-clip", "--steps", "3", "--warmup", "1"],
                          capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert proc.returncode == 0, proc.stderr[-2000:]
    lines = [l for l in proc.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["config"]["world_size"] == 2
    assert d["config"]["host_threads_per_rank"] == max(1, d["config"]["physical_cores"] // 2)
    assert d["cpu_baseline"]["cores"] == 2 * d["config"]["host_threads_per_rank"]
    assert "gloo" in d["config"]["allreduce"] and d["value"] > 0


def test_ddp_style_allreduce_is_what_torch_ddp_computes():
    """bench.py's stand-in for the DDP reducer in the reference arm: pre-multiply by 1/W, SUM, in buckets."""
    import torch
    sys.path.insert(0, ROOT)
    import bench

    class FakeDist:
        class ReduceOp:
            SUM = "sum"
        calls = []

        @staticmethod
        def all_reduce(t, op=None):
            FakeDist.calls.append(t.numel())
            t.mul_(3)          # three identical ranks: SUM of three pre-scaled copies

    gs = [torch.randn(5, 7), torch.randn(3), torch.randn(11)]
    want = [g.clone() * (1.0 / 3) * 3 for g in gs]
    bench.ddp_style_allreduce(gs, 3, FakeDist, cap_bytes=50)
    for a, b in zip(gs, want):
        assert torch.allclose(a, b, rtol=1e-6)
    assert FakeDist.calls == [14, 35]      # last tensors first; a bucket closes once it reaches the cap


def test_ordered_bits_counts_ulps():
    """bench.parity_check's ulp distance: consecutive floats differ by one, across zero too."""
    import torch
    sys.path.insert(0, ROOT)
    import bench
    x = torch.tensor([1.0, -1.0, 0.0, 1e-30], dtype=torch.float32)
    up = torch.nextafter(x, torch.full_like(x, float("inf")))
    assert (bench._ordered_bits(up) - bench._ordered_bits(x)).tolist() == [1, 1, 1, 1]
    assert int(bench._ordered_bits(torch.tensor([-0.0])).item()) == int(bench._ordered_bits(torch.tensor([0.0])).item()) == 0
    b = torch.tensor([1.0, -2.5], dtype=torch.bfloat16)
    nb = (b.view(torch.int16) + 1).view(torch.bfloat16)          # next bf16 away from zero
    assert (bench._ordered_bits(nb) - bench._ordered_bits(b)).abs().tolist() == [1, 1]
    assert bench.bytes_per_param("f32", 1, -1.0, "keep") == 36 and bench.bytes_per_param("bf16", 2, 1.0, "keep") == 24
