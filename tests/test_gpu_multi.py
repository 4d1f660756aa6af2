"""Multi-GPU (NCCL) path: needs >= 2 GPUs on the box (gpurun --gpus 2)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpu_trainloop_matches_ddp_adamw(tmp_path):
    from distributed_pipeline_b200 import dist_util
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(dist_util.find_free_port()), os.path.join(HERE, "multi_worker.py"),
           str(tmp_path)]
    proc = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    sys.stdout.write(proc.stdout[-3000:])
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-6000:]
    assert (tmp_path / "multi_ok").exists()
