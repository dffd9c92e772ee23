mkdir -p gpurun_out
./scripts/ubench/trace_schur 8192 16 | tee gpurun_out/s10o_trace.log
timeout 900 python -m pytest tests/test_schur_gpu.py -q -x --timeout=180 > gpurun_out/s10o_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s10o_pytest.log
timeout 300 python scripts/time_chain.py 8192 2>&1 | tee gpurun_out/s10o_chain.log | tail -8
