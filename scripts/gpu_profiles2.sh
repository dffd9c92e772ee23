# evidence for profiles/ (round 1, second pass): bench lines, launch list of the bench command, ncu --set full of the three kernels
tag=${1:-r1c}
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_${tag}.json 2> gpurun_out/bench_${tag}.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_${tag}.json
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_${tag}_ref.json 2> gpurun_out/bench_${tag}_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_${tag}_ref.json
python bench.py --steps 2 --warmup 3 > gpurun_out/bench_${tag}_short.json 2> gpurun_out/bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${tag}.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
tail -1 gpurun_out/ncu_launches.log | cut -c1-300
python scripts/prof_kernel.py batched 8192 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pcg_ -c 1 -o gpurun_out/prof_${tag} python scripts/prof_kernel.py batched 8192 > gpurun_out/ncu_full.log 2>&1
tail -1 gpurun_out/ncu_full.log
ncu --set full --clock-control none --import-source on -k regex:pinv_build -c 1 -o gpurun_out/prof_pinv_${tag} python scripts/prof_pinv.py > gpurun_out/ncu_pinv.log 2>&1; tail -1 gpurun_out/ncu_pinv.log
ncu --set full --clock-control none --import-source on -k regex:schur_form -c 1 -o gpurun_out/prof_schur_${tag} python scripts/prof_schur.py > gpurun_out/ncu_schur.log 2>&1; tail -1 gpurun_out/ncu_schur.log
python scripts/time_chain.py 8192 > gpurun_out/chain_${tag}.log 2>&1; python scripts/time_chain.py 16 >> gpurun_out/chain_${tag}.log 2>&1; cat gpurun_out/chain_${tag}.log
python scripts/time_pinv.py > gpurun_out/time_pinv_${tag}.log 2>&1; cat gpurun_out/time_pinv_${tag}.log
