# final pass of a round on one GPU: whole GPU suite, smoke, both bench arms, launch list of the bench command, ncu --set full of the batched kernel
mkdir -p gpurun_out
R=${1:-r2}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu_${R}.txt 2>&1; tail -5 gpurun_out/pytest_gpu_${R}.txt
timeout 600 python -c "import __graft_entry__ as e; e.smoke()" 2>&1 | tail -6 | tee gpurun_out/smoke_${R}.txt
timeout 900 python bench.py > gpurun_out/bench_${R}.json 2> gpurun_out/bench_${R}.err; tail -3 gpurun_out/bench_${R}.err; cut -c1-300 gpurun_out/bench_${R}.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_${R}_ref.json 2> gpurun_out/bench_${R}_ref.err; tail -3 gpurun_out/bench_${R}_ref.err; cut -c1-200 gpurun_out/bench_${R}_ref.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_${R}.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches_${R}.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pcg_tmem4 -c 1 -f -o gpurun_out/ncu_tmem4_${R} python scripts/prof_kernel.py batched 8192 > gpurun_out/ncu_tmem4_${R}.log 2>&1; tail -2 gpurun_out/ncu_tmem4_${R}.log
timeout 300 python scripts/time_grid.py 2>&1 | tee gpurun_out/grid_${R}.txt
timeout 300 python scripts/stress_tmem.py 100 2>&1 | tail -3 | tee gpurun_out/stress_tmem_${R}.txt
timeout 300 python scripts/time_tmem64.py 2>&1 | grep -v tmem_b | tee gpurun_out/tmem64_${R}.txt | cut -c1-120
timeout 300 python scripts/time_tmem4_wide.py 2>&1 | tee gpurun_out/tmem4_wide_${R}.txt | cut -c1-120
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pcg_tmem4 -c 1 -f -o gpurun_out/ncu_tmem4_wide_${R} python scripts/time_tmem4_wide.py 4096 > gpurun_out/ncu_tmem4_wide_${R}.log 2>&1; tail -2 gpurun_out/ncu_tmem4_wide_${R}.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pcg_fused -c 1 -f -o gpurun_out/ncu_tmem64_${R} python scripts/prof_kernel.py batched64 2048 > gpurun_out/ncu_tmem64_${R}.log 2>&1; tail -2 gpurun_out/ncu_tmem64_${R}.log
