mkdir -p gpurun_out
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/scale_${n}.json 2> gpurun_out/scale_${n}.err; echo "bench$n rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/scale_${n}.json')); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['strong_scaling']['value'])
PY
done
