mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_pinv_gpu.py -q -x --timeout=180 -k "unaligned or zero_batch" 2>&1 | tail -3
timeout 600 ncu --set full --clock-control none --import-source on -k regex:schur_form -c 1 -o gpurun_out/prof_schur python scripts/prof_schur.py > gpurun_out/s10l_ncu.log 2>&1; tail -2 gpurun_out/s10l_ncu.log
