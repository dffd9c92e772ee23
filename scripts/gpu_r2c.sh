# round 2, third GPU pass: phase timelines of the four-row kernel (two CTAs per SM vs one, packed vs scalar FMA), variants again after the
# reciprocal rewrite, grid kernel with its own scratch pool, symmetric-shipping debug, whole GPU suite, both bench arms
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
for exe in trace_tmem4 trace_tmem4_scalar; do
  for cfg in "25 8192 296" "25 4096 148"; do timeout 120 scripts/ubench/$exe $cfg; done
done 2>&1 | tee gpurun_out/trace_tmem4_r2c.txt
TMEM4_ARMS=auto,tmem_r2,rows timeout 600 python scripts/time_tmem4.py 8192 all 2>&1 | tee gpurun_out/tmem4_r2c.txt
GBDPCG_LIB=$PWD/gbd_pcg_b200/variants/libgbdpcg_scalar.so TMEM4_ARMS=auto timeout 600 python scripts/time_tmem4.py 8192 all 2>&1 | tee gpurun_out/tmem4_scalar_r2c.txt
timeout 300 python scripts/time_grid.py 2>&1 | tee gpurun_out/grid_r2c.txt
timeout 300 python scripts/debug_sym.py 2>&1 | tee gpurun_out/debug_sym_r2c.txt
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu_r2c_full.txt 2>&1; tail -12 gpurun_out/pytest_gpu_r2c_full.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r2c.json 2> gpurun_out/bench_r2c.err; tail -3 gpurun_out/bench_r2c.err; cut -c1-400 gpurun_out/bench_r2c.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_r2c_ref.json 2> gpurun_out/bench_r2c_ref.err; tail -3 gpurun_out/bench_r2c_ref.err; cut -c1-300 gpurun_out/bench_r2c_ref.json
