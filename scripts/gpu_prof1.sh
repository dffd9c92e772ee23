set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout=300 2>&1 | tail -15 > gpurun_out/r1_pytest.log
tail -5 gpurun_out/r1_pytest.log
python scripts/prof_kernel.py all > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pcg_rows -c 12 -o gpurun_out/prof_r1_base python scripts/prof_kernel.py all > gpurun_out/ncu_full.log 2>&1
tail -5 gpurun_out/ncu_full.log
python bench.py --steps 2 --warmup 3 > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1_base.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
tail -3 gpurun_out/ncu_launches.log
