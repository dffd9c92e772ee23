# single-solve latency: the four-row tensor-memory kernel forced on one C2 system against the planner's register-row kernel
mkdir -p gpurun_out
timeout 300 python scripts/bench_variants.py C2 stair "0:0" "0:0x400" "0:0x1400" 2>&1 | tee gpurun_out/single_tmem4_r2f.txt
timeout 300 python scripts/bench_variants.py C2 jacobi "0:0" "0:0x400" 2>&1 | tee -a gpurun_out/single_tmem4_r2f.txt
timeout 300 python scripts/time_tail.py 2>&1 | tee gpurun_out/tail_r2f.txt
