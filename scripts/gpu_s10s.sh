mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_pinv_gpu.py -q -x --timeout=180 > gpurun_out/s10s_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s10s_pytest.log
timeout 300 python scripts/time_pinv.py 2>&1 | tee gpurun_out/s10s_time_pinv.log | tail -12
