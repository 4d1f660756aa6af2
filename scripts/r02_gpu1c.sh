#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02g_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02g_pytest.log
tail -3 $O/r02g_pytest.log
timeout 300 python bench.py > $O/r02g_bench_default.log 2>&1
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > $O/r02g_bench_reference.log 2>&1
timeout 200 python bench.py --steps 100 --warmup 10 --workload diffuseq-base-fp32-ema1-clip --no-cpu-baseline --no-e2e > $O/r02g_bench_clip.log 2>&1
timeout 200 python bench.py --steps 100 --warmup 10 --workload diffuseq-base-bf16-ema2-clip --no-cpu-baseline --no-e2e > $O/r02g_bench_bf16.log 2>&1
timeout 900 bash profiles/prof_r02.sh > $O/r02g_prof.log 2>&1
tail -12 $O/r02g_prof.log
