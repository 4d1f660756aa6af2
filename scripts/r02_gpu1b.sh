#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
for W in diffuseq-base-fp32-ema1 diffuseq-base-fp32-ema1-clip diffuseq-base-bf16-ema2-clip; do
  timeout 200 python bench.py --steps 100 --warmup 10 --workload $W --no-cpu-baseline --no-e2e > $O/r02f_$W.log 2>&1
done
timeout 200 python bench.py --steps 100 --warmup 10 --grad-mode zero --no-cpu-baseline --no-e2e > $O/r02f_zero.log 2>&1
