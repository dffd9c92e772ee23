#!/bin/bash
# full GPU check: parity suite, random sweep, per-config latency table, bench (ours + reference arm)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.txt
timeout 600 python scripts/sweep_parity.py > gpurun_out/sweep.txt 2>&1; echo "sweep rc=$?"; tail -3 gpurun_out/sweep.txt
timeout 600 python scripts/bench_configs.py gpurun_out/bench_configs.json > gpurun_out/bench_configs.txt 2>&1; echo "configs rc=$?"
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json | cut -c1-600
timeout 900 python bench.py --impl reference --steps 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_ref.json
