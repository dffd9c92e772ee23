set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv
nproc
timeout 900 python -m pytest tests -m gpu -q --timeout=300 2>&1 | tail -40 > gpurun_out/r1_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r1_smoke.log 2>&1
timeout 600 python oracle/make_ref_fixtures.py gpurun_out/ref_fixtures > gpurun_out/r1_fixtures.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r1_bench_ours.json 2> gpurun_out/r1_bench_ours.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r1_bench_ref.json 2> gpurun_out/r1_bench_ref.err
tail -5 gpurun_out/r1_pytest.log; cat gpurun_out/r1_smoke.log; tail -20 gpurun_out/r1_fixtures.log; cat gpurun_out/r1_bench_ours.json; tail -3 gpurun_out/r1_bench_ours.err; cat gpurun_out/r1_bench_ref.json; tail -3 gpurun_out/r1_bench_ref.err
