# Round-1 profiling recipe (one gpurun call; every ncu run is preceded by the same command exiting 0 without ncu)
set -x
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e"
K='regex:fused_update_kernel|sumsq_tiles_kernel|segment_reduce_kernel|finalize_'
$B > gpurun_out/plain_noclip.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file gpurun_out/launches_noclip.csv $B > gpurun_out/ncu_noclip.log 2>&1
$B --workload diffuseq-base-fp32-ema1-clip > gpurun_out/plain_clip.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file gpurun_out/launches_clip.csv $B --workload diffuseq-base-fp32-ema1-clip > gpurun_out/ncu_clip.log 2>&1
$B --workload diffuseq-base-bf16-ema2-clip > gpurun_out/plain_bf16.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file gpurun_out/launches_bf16.csv $B --workload diffuseq-base-bf16-ema2-clip > gpurun_out/ncu_bf16.log 2>&1
$B > gpurun_out/plain_noclip2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fused_update_kernel -s 5 -c 2 -o gpurun_out/prof_fused $B > gpurun_out/ncu_full.log 2>&1
$B --workload diffuseq-base-bf16-ema2-clip > gpurun_out/plain_bf16b.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k "regex:fused_update_kernel|sumsq_tiles_kernel" -s 10 -c 2 -o gpurun_out/prof_bf16 $B --workload diffuseq-base-bf16-ema2-clip > gpurun_out/ncu_full2.log 2>&1
ls -la gpurun_out | head -30
