# round 2, fourth GPU pass: pair-wise padded staging with the bulk copies spread over the warps; whole GPU suite; both bench arms; launch list and an
# ncu --set full capture of the batched kernel from the bench command; phase timeline
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 120 scripts/ubench/trace_tmem4 25 8192 296 2>&1 | tee gpurun_out/trace_tmem4_r2d.txt
TMEM4_ARMS=auto timeout 600 python scripts/time_tmem4.py 8192 all 2>&1 | tee gpurun_out/tmem4_r2d.txt
timeout 300 python scripts/time_tail.py 2>&1 | tee gpurun_out/tail_r2d.txt
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu_r2d_full.txt 2>&1; tail -12 gpurun_out/pytest_gpu_r2d_full.txt
timeout 600 python -c "import __graft_entry__ as e; e.smoke()" 2>&1 | tail -6 | tee gpurun_out/smoke_r2d.txt
timeout 900 python bench.py > gpurun_out/bench_r2d.json 2> gpurun_out/bench_r2d.err; tail -3 gpurun_out/bench_r2d.err; cut -c1-300 gpurun_out/bench_r2d.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_r2d_ref.json 2> gpurun_out/bench_r2d_ref.err; tail -3 gpurun_out/bench_r2d_ref.err; cut -c1-200 gpurun_out/bench_r2d_ref.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r2d.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches_r2d.log 2>&1; tail -2 gpurun_out/ncu_launches_r2d.log | cut -c1-200
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pcg_tmem4 -c 1 -f -o gpurun_out/ncu_tmem4_r2d python scripts/prof_kernel.py batched 8192 > gpurun_out/ncu_tmem4_r2d.log 2>&1; tail -2 gpurun_out/ncu_tmem4_r2d.log
