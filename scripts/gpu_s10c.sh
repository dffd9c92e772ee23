mkdir -p gpurun_out
timeout 300 python scripts/time_pinv.py 2>&1 | tee gpurun_out/s10c_time_pinv.log | tail -12
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pinv_build -c 1 -o gpurun_out/prof_pinv python scripts/prof_pinv.py > gpurun_out/s10c_ncu.log 2>&1; tail -2 gpurun_out/s10c_ncu.log
