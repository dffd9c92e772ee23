# Produces the evidence committed under profiles/: plain bench line, ncu launch list of the same bench command,
# one ncu --set full capture of the batched kernel.   usage (GPU box): bash scripts/gpu_profiles.sh <tag>
tag=${1:-r1}
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 > gpurun_out/bench_${tag}_short.json 2> gpurun_out/bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${tag}.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log
python scripts/prof_kernel.py batched 8192 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pcg_ -c 1 -o gpurun_out/prof_${tag} python scripts/prof_kernel.py batched 8192 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_${tag}.json 2> gpurun_out/bench.err; cat gpurun_out/bench_${tag}.json
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_${tag}_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_${tag}_ref.json
