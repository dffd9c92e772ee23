mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -k tmem > gpurun_out/pytest_tmem.txt 2>&1; echo "pytest tmem rc=$?"; tail -15 gpurun_out/pytest_tmem.txt
timeout 200 python scripts/time_tmem.py 2>&1 | tee gpurun_out/time_tmem.txt
