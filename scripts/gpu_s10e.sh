mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/s10e_bench.json 2> gpurun_out/s10e_bench.err; echo "bench rc=$?"; cat gpurun_out/s10e_bench.json; tail -3 gpurun_out/s10e_bench.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pinv_build -c 1 -o gpurun_out/prof_pinv2 python scripts/prof_pinv.py > gpurun_out/s10e_ncu.log 2>&1; tail -2 gpurun_out/s10e_ncu.log
