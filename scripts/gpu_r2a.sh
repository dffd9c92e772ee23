# round 2, first GPU pass: four-rows-per-thread tensor-memory kernel -- parity against the oracle, A/B timing of every variant
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 900 python scripts/time_tmem4.py 8192 all 2>&1 | tee gpurun_out/tmem4_r2a.txt
timeout 900 python -m pytest tests/test_parity_gpu.py -x -q -m gpu -k "tmem_kernel_vs_oracle or tmem_kernel_edge" 2>&1 | tail -15 | tee gpurun_out/pytest_tmem_r2a.txt
