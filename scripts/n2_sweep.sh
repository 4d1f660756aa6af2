python -m pytest tests/test_gpu_update.py -m gpu -q -x -k "bucket_granular" > gpurun_out/pt_bucket.log 2>&1; echo "bucket test rc=$?"; tail -3 gpurun_out/pt_bucket.log
for mb in 16 32 64 128; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29600+mb)) bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e --bucket-mb $mb > gpurun_out/bench_n2_mb$mb.log 2>&1; echo "mb=$mb rc=$?"
  tail -1 gpurun_out/bench_n2_mb$mb.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['config']['allreduce_buckets'], d['roofline']['kernel_ms_avg'])"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29599 bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e --bucket-mb 128 --no-overlap > gpurun_out/bench_n2_noov.log 2>&1; echo "noov rc=$?"
tail -1 gpurun_out/bench_n2_noov.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29598 bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e --bucket-mb 32 --workload diffuseq-base-bf16-ema2-clip > gpurun_out/bench_n2_bf16.log 2>&1; echo "bf16 rc=$?"
tail -1 gpurun_out/bench_n2_bf16.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline'])"
