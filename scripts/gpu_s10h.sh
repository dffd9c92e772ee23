mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_schur_gpu.py -q -x --timeout=180 > gpurun_out/s10h_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/s10h_pytest.log
