# Round-2 profiling recipe (one gpurun call on ONE GPU; every ncu run is preceded by the same command exiting 0
# without ncu).  Writes launch lists, full captures, and -- via scripts/ncu_summary.py -- the kernel table and
# traffic.json that bench.py's roofline.traffic reads, so neither is typed in by hand.
set -x
O=gpurun_out
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e"
K='regex:fused_update_kernel|sumsq_tiles_kernel|norm_tail_kernel|logged_tail_kernel'
for W in diffuseq-base-fp32-ema1 diffuseq-base-fp32-ema1-clip diffuseq-base-bf16-ema2-clip; do
  $B --workload $W > $O/r02_plain_$W.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 12 -c 40 --csv --log-file $O/r02_launches_$W.csv $B --workload $W > $O/r02_ncul_$W.log 2>&1
done
for W in diffuseq-base-fp32-ema1 diffuseq-base-fp32-ema1-clip diffuseq-base-bf16-ema2-clip; do
  $B --workload $W > $O/r02_plain2_$W.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k 'regex:fused_update_kernel|sumsq_tiles_kernel' -s 8 -c 4 -f -o $O/r02_full_$W $B --workload $W > $O/r02_ncuf_$W.log 2>&1
done
python scripts/ncu_summary.py $O/r02_ncu_table.md $O/traffic.json \
  diffuseq-base-fp32-ema1:keep=$O/r02_full_diffuseq-base-fp32-ema1.ncu-rep \
  diffuseq-base-fp32-ema1-clip:keep=$O/r02_full_diffuseq-base-fp32-ema1-clip.ncu-rep \
  diffuseq-base-bf16-ema2-clip:keep=$O/r02_full_diffuseq-base-bf16-ema2-clip.ncu-rep
