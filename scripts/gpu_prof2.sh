mkdir -p gpurun_out
python scripts/prof_kernel.py batched 8192 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pcg_ -c 1 -o gpurun_out/prof_v2 python scripts/prof_kernel.py batched 8192 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
