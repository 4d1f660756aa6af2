#!/bin/bash
# Round-2 final eight-GPU run: bench lines as the driver launches them, tests at W=8, workloads, sweep points.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
N=${1:-8}
export DPK_PEER_TIMEOUT_MS=30000
P=29900
tr() { P=$((P+1)); timeout ${T:-200} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; }
tr bench.py --gpus $N --steps 100 --warmup 10 > $O/r02i_n${N}_default.log 2>&1; echo "default rc=$?"
tr bench.py --gpus $N --steps 100 --warmup 10 --no-e2e --workload diffuseq-base-bf16-ema2-clip > $O/r02i_n${N}_bf16clip.log 2>&1; echo "bf16clip rc=$?"
tr bench.py --gpus $N --steps 100 --warmup 10 --sharded --no-parity > $O/r02i_n${N}_sharded.log 2>&1; echo "sharded rc=$?"
T=300 tr bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $O/r02i_n${N}_reference.log 2>&1; echo "reference rc=$?"
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "allreduce and $N" > $O/r02i_n${N}_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02i_n${N}_pytest.log; tail -3 $O/r02i_n${N}_pytest.log
T=400 tr bench.py --gpus $N --workload trainloop-bf16-ema2-clip --steps 4 --warmup 3 > $O/r02i_n${N}_trainloop.log 2>&1; echo "trainloop rc=$?"
tr bench.py --gpus $N --workload sampler-t2000 --steps 200 --warmup 10 > $O/r02i_n${N}_sampler.log 2>&1; echo "sampler rc=$?"
bash scripts/flat_sweep.sh $N 10000000 1000000000 > $O/r02i_n${N}_sweep.log 2>&1; cat $O/r02i_n${N}_sweep.log
for f in default bf16clip sharded reference trainloop sampler; do echo "== $f"; grep '^{' $O/r02i_n${N}_$f.log | cut -c1-260; done
