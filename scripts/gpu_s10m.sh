mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_schur_gpu.py tests/test_pinv_gpu.py -q -x --timeout=180 > gpurun_out/s10m_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s10m_pytest.log
timeout 300 python scripts/time_chain.py 8192 2>&1 | tee gpurun_out/s10m_chain.log | tail -8
timeout 300 python scripts/time_pinv.py 2>&1 | tee gpurun_out/s10m_time_pinv.log | tail -12
