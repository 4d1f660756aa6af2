run() { # label, env, extra args
  env $2 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $3 bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e $4 > gpurun_out/n2g.log 2>&1
  echo -n "$1: rc=$? "; tail -1 gpurun_out/n2g.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['config']['allreduce_buckets'], d['roofline']['kernel_ms_avg'])"
}
run "persist mb128" "X=1" 29701 "--bucket-mb 128"
run "flat mb128" "DPK_UPDATE_GRID=flat" 29702 "--bucket-mb 128"
run "flat mb32" "DPK_UPDATE_GRID=flat" 29703 "--bucket-mb 32"
run "flat mb16" "DPK_UPDATE_GRID=flat" 29704 "--bucket-mb 16"
run "mult8 mb32" "DPK_UPDATE_GRID=mult:8" 29705 "--bucket-mb 32"
run "reserve16 mb32" "DPK_UPDATE_GRID=reserve:16" 29706 "--bucket-mb 32"
run "reserve32 mb32" "DPK_UPDATE_GRID=reserve:32" 29707 "--bucket-mb 32"
run "reserve16 mb64" "DPK_UPDATE_GRID=reserve:16" 29708 "--bucket-mb 64"
run "flat mb64" "DPK_UPDATE_GRID=flat" 29709 "--bucket-mb 64"
python - <<'PY'
import os,subprocess
PY
