#!/bin/bash
# Round-2 two-GPU check: all GPU tests (incl. multi-GPU at W=2 and the switch kernels forced onto 2 ranks), bench lines.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
export DPK_PEER_TIMEOUT_MS=20000
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r02b_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02b_pytest.log
tail -5 $O/r02b_pytest.log
timeout 300 $TR --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 > $O/r02b_bench_n2.log 2>&1
DPK_FORCE_MULTIMEM=1 timeout 300 $TR --master-port 29512 bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e > $O/r02b_bench_n2_mm.log 2>&1
DPK_FORCE_MULTIMEM=1 timeout 300 $TR --master-port 29513 bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e --no-fused-reduce > $O/r02b_bench_n2_mm_bucketed.log 2>&1
timeout 300 $TR --master-port 29514 bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e --workload diffuseq-base-bf16-ema2-clip > $O/r02b_bench_n2_bf16.log 2>&1
timeout 300 $TR --master-port 29515 bench.py --impl reference --gpus 2 --steps 5 --warmup 2 > $O/r02b_ref_n2.log 2>&1
timeout 200 python bench.py --steps 100 --warmup 10 --workload diffuseq-base-fp32-ema1-clip --no-cpu-baseline --no-e2e > $O/r02b_bench_clip.log 2>&1
timeout 200 python bench.py --steps 100 --warmup 10 --workload diffuseq-base-bf16-ema2-clip --no-cpu-baseline --no-e2e > $O/r02b_bench_bf16.log 2>&1
for f in $O/r02b_bench_*.log $O/r02b_ref_n2.log; do echo "== $f"; tail -2 $f | cut -c1-600; done
