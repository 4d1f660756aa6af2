#!/bin/bash
# Round-2 single-GPU check: tests, bench lines, bf16 A/B of resident blocks, one ncu capture.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02_pytest.log
timeout 300 python bench.py --steps 100 --warmup 10 > $O/r02_bench_default.log 2>&1
timeout 200 python bench.py --steps 100 --warmup 10 --workload diffuseq-base-fp32-ema1-clip --no-cpu-baseline --no-e2e > $O/r02_bench_clip.log 2>&1
timeout 200 python bench.py --steps 100 --warmup 10 --workload diffuseq-base-bf16-ema2-clip --no-cpu-baseline --no-e2e > $O/r02_bench_bf16_mb4.log 2>&1
DPK_UPDATE_BLOCKS_PER_SM=3 timeout 200 python bench.py --steps 100 --warmup 10 --workload diffuseq-base-bf16-ema2-clip --no-cpu-baseline --no-e2e > $O/r02_bench_bf16_mb3.log 2>&1
for n in 10000000 30000000; do
  timeout 200 python bench.py --steps 100 --warmup 10 --workload diffuseq-base-fp32-ema1-clip --flat-params $n --no-cpu-baseline --no-e2e > $O/r02_flat_$n.log 2>&1
done
timeout 200 python bench.py --steps 5 --warmup 3 --workload diffuseq-base-bf16-ema2-clip --no-cpu-baseline --no-e2e > $O/r02_plain_bf16.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'fused_update_kernel|sumsq_tiles_kernel' -s 6 -c 4 -o $O/r02_prof_bf16 -f python bench.py --steps 5 --warmup 3 --workload diffuseq-base-bf16-ema2-clip --no-cpu-baseline --no-e2e > $O/r02_ncu_bf16.log 2>&1
tail -3 $O/r02_pytest.log
for f in $O/r02_bench_*.log $O/r02_flat_*.log; do echo "== $f"; tail -1 $f | cut -c1-1500; done
