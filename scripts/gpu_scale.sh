# bench.py under torchrun on N GPUs of one box (strong scaling headline: the same 8192 systems partitioned by problem index)
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv,noheader | head -$N
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_r2_${N}gpu.json 2> gpurun_out/bench_r2_${N}gpu.err
tail -3 gpurun_out/bench_r2_${N}gpu.err; cut -c1-500 gpurun_out/bench_r2_${N}gpu.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_r2_${N}gpu_ref.json 2> gpurun_out/bench_r2_${N}gpu_ref.err
cut -c1-200 gpurun_out/bench_r2_${N}gpu_ref.json
