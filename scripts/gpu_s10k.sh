mkdir -p gpurun_out
timeout 300 python scripts/time_e2e.py 2>&1 | tee gpurun_out/s10k_e2e.log | tail -8
