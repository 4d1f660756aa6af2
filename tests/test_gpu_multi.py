"""Multi-GPU (NCCL / symmetric-memory) paths.  Each case needs as many GPUs as its world size and is skipped
otherwise (gpurun --gpus N); bench.py --gpus N runs the same checks (`parity_check`) where the driver has N GPUs."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
N_GPUS = torch.cuda.device_count() if torch.cuda.is_available() else 0


def _torchrun(script, world, tmp_path, env=None, timeout=900):
    from distributed_pipeline_b200 import dist_util
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(dist_util.find_free_port()), os.path.join(HERE, script), str(tmp_path)]
    # a peer that never arrives must fail the test quickly instead of parking the GPUs for ten minutes
    full_env = dict(os.environ, DPK_PEER_TIMEOUT_MS="20000", **(env or {}))
    proc = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=full_env)
    sys.stdout.write(proc.stdout[-3000:])
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-6000:]


@pytest.mark.skipif(N_GPUS < 2, reason="needs 2 GPUs")
def test_two_gpu_trainloop_matches_ddp_adamw(tmp_path):
    _torchrun("multi_worker.py", 2, tmp_path)
    assert (tmp_path / "multi_ok").exists()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_allreduce_and_one_kernel_update(world, tmp_path):
    if N_GPUS < world:
        pytest.skip(f"needs {world} GPUs")
    _torchrun("allreduce_worker.py", world, tmp_path)
    assert (tmp_path / "allreduce_ok").exists()


@pytest.mark.skipif(N_GPUS < 2, reason="needs 2 GPUs")
def test_switch_kernels_with_two_ranks(tmp_path):
    """The W >= 4 kernels (multimem all-reduce, reduce_update_kernel) forced onto a two-rank job."""
    _torchrun("allreduce_worker.py", 2, tmp_path, env={"DPK_FORCE_MULTIMEM": "1"})
    assert (tmp_path / "allreduce_ok").exists()
