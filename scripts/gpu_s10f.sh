mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout=180 > gpurun_out/s10f_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/s10f_pytest.log
