mkdir -p gpurun_out
timeout 300 python scripts/time_chain.py 8192 2>&1 | tee gpurun_out/s10i_chain.log | tail -8
timeout 300 python scripts/time_chain.py 16 2>&1 | tee -a gpurun_out/s10i_chain.log | tail -6
