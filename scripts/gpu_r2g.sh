# round 2, pass g: the solo form of the four-row kernel (single solves and batches of up to #SMs systems); whole suite; bench
mkdir -p gpurun_out
timeout 300 python scripts/bench_variants.py C2 stair "0:0" "0:0x800" "0:0x400" 2>&1 | tee gpurun_out/single_tmem4_r2g.txt
timeout 300 python scripts/bench_variants.py C2 jacobi "0:0" "0:0x800" 2>&1 | tee -a gpurun_out/single_tmem4_r2g.txt
GBDPCG_TMEM4_NO_SOLO=1 timeout 300 python scripts/bench_variants.py C2 stair "0:0x400" 2>&1 | tee -a gpurun_out/single_tmem4_r2g.txt
timeout 300 python scripts/time_tail.py 2>&1 | tee gpurun_out/tail_r2g.txt
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu_r2g_full.txt 2>&1; tail -15 gpurun_out/pytest_gpu_r2g_full.txt
timeout 600 python -c "import __graft_entry__ as e; e.smoke()" 2>&1 | tail -6 | tee gpurun_out/smoke_r2g.txt
timeout 900 python bench.py > gpurun_out/bench_r2g.json 2> gpurun_out/bench_r2g.err; tail -3 gpurun_out/bench_r2g.err; cut -c1-300 gpurun_out/bench_r2g.json
