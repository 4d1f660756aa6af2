#!/bin/bash
# Round-2 eight-GPU run: multi-GPU parity tests at W=4/8, the one-kernel all-reduce + update against the bucketed form, tunables.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
export DPK_PEER_TIMEOUT_MS=20000
P=29600
run() {  # name, nproc, extra env..., then bench args after --
  local name=$1 n=$2; shift 2
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  P=$((P+1))
  env "${envs[@]}" timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $P \
     bench.py --gpus $n --steps 50 --warmup 5 --no-e2e "$@" > $O/r02c_$name.log 2>&1
  echo "== $name rc=$?"; grep '^{' $O/r02c_$name.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); r=d['roofline']; c=d.get('roofline_comm') or {}; p=d.get('parity_check') or {}
    print('   ms/step %.4f kernel_ms %.4f comm_ms %.4f busbw %.1f parity %s' % (d['ms_per_step'], r['kernel_ms_avg'], c.get('kernel_ms_avg', 0), c.get('busbw', 0), p.get('ok')))
"
}
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "allreduce and (4 or 8)" > $O/r02c_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02c_pytest.log
tail -4 $O/r02c_pytest.log
run n8_fused 8 -- 
run n8_bucketed 8 -- --no-fused-reduce
run n8_fused_c32 8 DPK_FUSED_COMM_BLOCKS=32 -- --no-parity
run n8_fused_c96 8 DPK_FUSED_COMM_BLOCKS=96 -- --no-parity
run n8_fused_u1m 8 DPK_FUSED_UNIT_KB=1024 -- --no-parity
run n8_fused_u4m 8 DPK_FUSED_UNIT_KB=4096 -- --no-parity
run n8_fused_upd296 8 DPK_FUSED_UPDATE_BLOCKS=296 -- --no-parity
run n4_fused 4 --
run n4_bucketed 4 -- --no-fused-reduce --no-parity
run n8_bf16clip 8 -- --workload diffuseq-base-bf16-ema2-clip
