#!/bin/bash
# Round-2 two-GPU check #2: all GPU tests, N=2 bench lines (default, bf16+clip, forced switch kernels), trainloop and sampler workloads.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
export DPK_PEER_TIMEOUT_MS=20000
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 1200 python -m pytest tests -m gpu -x -q > $O/r02e_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02e_pytest.log
tail -5 $O/r02e_pytest.log
timeout 200 $TR --master-port 29521 bench.py --gpus 2 --steps 50 --warmup 5 > $O/r02e_bench_n2.log 2>&1
timeout 200 $TR --master-port 29522 bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e --workload diffuseq-base-bf16-ema2-clip > $O/r02e_bench_n2_bf16.log 2>&1
DPK_FORCE_MULTIMEM=1 timeout 200 $TR --master-port 29523 bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e --workload diffuseq-base-bf16-ema2-clip > $O/r02e_bench_n2_bf16_mm.log 2>&1
timeout 300 $TR --master-port 29524 bench.py --gpus 2 --workload trainloop-bf16-ema2-clip --steps 3 --warmup 3 > $O/r02e_trainloop_n2.log 2>&1
timeout 200 $TR --master-port 29525 bench.py --gpus 2 --workload sampler-t2000 --steps 200 --warmup 10 > $O/r02e_sampler_n2.log 2>&1
timeout 200 python bench.py --workload sampler-t2000 --steps 200 --warmup 10 > $O/r02e_sampler_n1.log 2>&1
timeout 200 python bench.py --steps 100 --warmup 10 --workload diffuseq-base-fp32-ema1-clip --no-cpu-baseline --no-e2e > $O/r02e_bench_clip.log 2>&1
timeout 200 python bench.py --steps 100 --warmup 10 --workload diffuseq-base-bf16-ema2-clip --no-cpu-baseline --no-e2e > $O/r02e_bench_bf16.log 2>&1
for f in $O/r02e_*.log; do echo "== $f"; tail -2 $f | cut -c1-400; done
