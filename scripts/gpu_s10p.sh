mkdir -p gpurun_out
timeout 120 python __graft_entry__.py smoke > gpurun_out/s10p_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/s10p_smoke.log
timeout 600 python -m pytest tests/test_parity_gpu.py -q -x --timeout=60 > gpurun_out/s10p_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/s10p_pytest.log
timeout 600 python scripts/bench_configs.py gpurun_out/configs_r1c.json > gpurun_out/s10p_configs.log 2>&1; echo "configs rc=$?"; tail -16 gpurun_out/s10p_configs.log
