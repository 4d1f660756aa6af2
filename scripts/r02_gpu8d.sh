#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
N=${1:-8}
export DPK_PEER_TIMEOUT_MS=30000
P=29950
tr() { P=$((P+1)); timeout ${T:-200} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; }
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "allreduce and $N" > $O/r02j_n${N}_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02j_n${N}_pytest.log; tail -3 $O/r02j_n${N}_pytest.log
if [ "$N" = 8 ]; then
  T=300 tr bench.py --gpus $N --workload trainloop-bf16-ema2-clip --steps 20 --warmup 3 > $O/r02j_n${N}_trainloop.log 2>&1; echo "trainloop rc=$?"
else
  tr bench.py --gpus $N --steps 100 --warmup 10 > $O/r02j_n${N}_default.log 2>&1; echo "default rc=$?"
  tr bench.py --gpus $N --steps 100 --warmup 10 --no-e2e --no-parity --workload diffuseq-base-bf16-ema2-clip > $O/r02j_n${N}_bf16clip.log 2>&1; echo "bf16clip rc=$?"
fi
grep -h '^{' $O/r02j_n${N}_*.log | cut -c1-200
