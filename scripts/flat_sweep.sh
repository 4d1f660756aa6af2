# BASELINE.json configs[4]: flat-parameter update sweep (grad-norm + clip + AdamW + EMA 0.9999, fp32, synthetic grads), one GPU
for n in 10000000 30000000 91225274 100000000 300000000 1000000000 2000000000; do
  python bench.py --flat-params $n --workload diffuseq-base-fp32-ema1-clip --steps 30 --warmup 5 --no-e2e --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print(json.dumps({'params': d['config']['params'], 'ms_per_step': round(d['ms_per_step'],4), 'GBps': round(d['value'],1), 'Gparam_per_s': round(d['params_per_sec']/1e9,1), 'fused_frac_of_hbm_peak': round(r['frac'],3), 'norm_pass_frac': round(r['norm_pass']['frac'],3)}))"
done
