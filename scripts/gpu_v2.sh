mkdir -p gpurun_out
timeout 120 python __graft_entry__.py smoke > gpurun_out/v2_smoke.log 2>&1; echo "smoke rc=$?"; tail -6 gpurun_out/v2_smoke.log
timeout 900 python -m pytest tests -m gpu -q --timeout=120 2>&1 | tail -40 > gpurun_out/v2_pytest.log; tail -25 gpurun_out/v2_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/v2_bench.json 2> gpurun_out/v2_bench.err; echo "bench rc=$?"; cat gpurun_out/v2_bench.json; tail -3 gpurun_out/v2_bench.err
