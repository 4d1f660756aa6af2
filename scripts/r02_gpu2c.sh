#!/bin/bash
# two GPUs: multi-GPU tests (sharded update, reduce_norm, switch kernels forced onto two ranks), sharded bench
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
export DPK_PEER_TIMEOUT_MS=20000
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_round2.py tests/test_gpu_update.py -m gpu -x -q > $O/r02h_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02h_pytest.log
tail -30 $O/r02h_pytest.log | cut -c1-300
DPK_FORCE_MULTIMEM=1 timeout 200 $TR --master-port 29531 bench.py --gpus 2 --steps 50 --warmup 5 --sharded --no-parity > $O/r02h_bench_n2_sharded_mm.log 2>&1
timeout 200 $TR --master-port 29532 bench.py --gpus 2 --steps 50 --warmup 5 --sharded --no-parity > $O/r02h_bench_n2_sharded.log 2>&1
DPK_FORCE_MULTIMEM=1 timeout 200 $TR --master-port 29533 bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e --workload diffuseq-base-bf16-ema2-clip > $O/r02h_bench_n2_bf16_mm.log 2>&1
for f in $O/r02h_bench*.log; do echo "== $f"; tail -2 $f | cut -c1-300; done
