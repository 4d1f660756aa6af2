# BASELINE.json configs[4]: flat-parameter update sweep (grad-norm + clip 1.0 + AdamW + EMA 0.9999, fp32, synthetic
# gradients).   bash scripts/flat_sweep.sh [N_GPUS] [sizes...]      one JSON line per size
N=${1:-1}; shift
SIZES=${@:-10000000 30000000 91225274 100000000 300000000 1000000000 2000000000}
P=29800
for n in $SIZES; do
  P=$((P+1))
  if [ "$N" = 1 ]; then RUN="python"; else RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P"; fi
  $RUN bench.py --gpus $N --flat-params $n --workload diffuseq-base-fp32-ema1-clip --steps 30 --warmup 5 --no-e2e --no-cpu-baseline --no-parity 2>/dev/null | grep '^{' | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; c=d.get('roofline_comm') or {}
o={'n_gpus': d['n_gpus'], 'params': d['config']['params'], 'ms_per_step': round(d['ms_per_step'],4), 'GBps': round(d['value'],1), 'Gparam_per_s': round(d['params_per_sec']/1e9,1), 'fused_frac_of_hbm_peak': round(r['frac'],3), 'norm_pass_ms': round(r['norm_pass']['kernel_ms_avg'],4)}
if c: o.update({'comm_kernel': c['kernel'].split(' ')[0], 'busbw_GBps': round(c['busbw'],1), 'busbw_frac_of_900': round(c['frac'],3)})
print(json.dumps(o))"
done
