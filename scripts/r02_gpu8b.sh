#!/bin/bash
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd "$(dirname "$0")/.."
O=gpurun_out
export DPK_PEER_TIMEOUT_MS=20000
P=29700
run() {
  local name=$1 n=$2; shift 2
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  P=$((P+1))
  env "${envs[@]}" timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $P \
     bench.py --gpus $n --steps 50 --warmup 5 --no-e2e --no-parity "$@" > $O/r02d_$name.log 2>&1
  echo "== $name rc=$?"; grep '^{' $O/r02d_$name.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); r=d['roofline']; c=d.get('roofline_comm') or {}
    print('   ms/step %.4f kernel_ms %.4f comm_ms %.4f busbw %.1f' % (d['ms_per_step'], r['kernel_ms_avg'], c.get('kernel_ms_avg', 0), c.get('busbw', 0)))
"
}
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29700 scripts/nvls_headroom.py > $O/r02d_headroom8.log 2>&1; tail -2 $O/r02d_headroom8.log
run n8_c16 8 DPK_FUSED_COMM_BLOCKS=16 --
run n8_c24 8 DPK_FUSED_COMM_BLOCKS=24 --
run n8_c32_u128 8 DPK_FUSED_UNIT_KB=128 --
run n8_c32_u512 8 DPK_FUSED_UNIT_KB=512 --
