# round 2, fifth GPU pass: window stride 4*G in the four-row kernel; the column-split 8-warp experiment (GBDPCG_TMEM4_VARIANT=20)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
TMEM4_ARMS=auto,tmem4_v20 timeout 600 python scripts/time_tmem4.py 8192 all 2>&1 | tee gpurun_out/tmem4_r2e.txt
timeout 120 scripts/ubench/trace_tmem4 25 8192 296 2>&1 | tee gpurun_out/trace_tmem4_r2e.txt
timeout 900 python -m pytest tests/test_parity_gpu.py -q -m gpu -k "tmem or batched or full_size" > gpurun_out/pytest_gpu_r2e.txt 2>&1; tail -5 gpurun_out/pytest_gpu_r2e.txt
