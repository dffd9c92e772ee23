#!/bin/bash
# end-of-round evidence: bench (ours + reference arm), ncu --set full of the clustered solver kernel (C4 and C3 shapes)
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_r1f.json 2> gpurun_out/bench_r1f.err; echo "bench rc=$?"; python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_r1f.json"))
print(d["value"], d["e2e"]["value"], d["roofline"]["frac"], d["other_configs"], d["mpc_step"]["steps_per_s"])
PY
timeout 900 python bench.py --impl reference --steps 5 > gpurun_out/bench_r1f_ref.json 2> gpurun_out/bench_r1f_ref.err; echo "ref rc=$?"
for c in C4 C3; do
  timeout 200 python scripts/prof_cluster.py $c && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:pcg_fused -c 1 -f -o gpurun_out/prof_cluster_$c python scripts/prof_cluster.py $c > gpurun_out/ncu_cluster_$c.log 2>&1
  tail -1 gpurun_out/ncu_cluster_$c.log
done
ls -la gpurun_out/*.ncu-rep
