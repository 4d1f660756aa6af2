python scripts/one_kernel.py bf16 2 0 > gpurun_out/one_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"fused_update_kernel|sumsq_tiles_kernel" -s 6 -c 2 -o gpurun_out/prof_bf16 python scripts/one_kernel.py bf16 2 0 > gpurun_out/one_ncu.log 2>&1
tail -2 gpurun_out/one_plain.log gpurun_out/one_ncu.log
