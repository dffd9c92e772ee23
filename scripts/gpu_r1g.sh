# last pass of the round after the tensor-memory kernel became the batched default: launch list of the bench command, bench lines
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 > gpurun_out/bench_r1g_short.json 2> gpurun_out/bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1g.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log
python bench.py > gpurun_out/bench_r1g.json 2> gpurun_out/bench.err; cut -c1-200 gpurun_out/bench_r1g.json
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_r1g_ref.json 2> gpurun_out/bench_ref.err; cut -c1-200 gpurun_out/bench_r1g_ref.json
timeout 300 python scripts/time_batched_iters.py 2>&1 | tee gpurun_out/batched_iters_tmem.txt | tail -3
