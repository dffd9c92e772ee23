# round 2, second GPU pass: FFMA2 operand microbenchmark, every tensor-memory variant again (new defaults + small-system variants + scalar-FMA build),
# the whole GPU test suite, smoke, the C++ API-surface program, an ncu --set full capture of the four-row kernel, a short bench run
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
scripts/ubench/ffma2_banks 2>&1 | tee gpurun_out/ffma2_banks_r2b.txt
timeout 900 python scripts/time_tmem4.py 8192 all 2>&1 | tee gpurun_out/tmem4_r2b.txt
GBDPCG_LIB=$PWD/gbd_pcg_b200/variants/libgbdpcg_scalar.so TMEM4_ARMS=tmem4_v0,tmem4_v1,tmem4_v3 timeout 600 python scripts/time_tmem4.py 8192 c5 2>&1 | tee gpurun_out/tmem4_scalar_r2b.txt
GBDPCG_LIB=$PWD/gbd_pcg_b200/variants/libgbdpcg_nopad.so TMEM4_ARMS=auto timeout 600 python scripts/time_tmem4.py 8192 all 2>&1 | tee gpurun_out/tmem4_nopad_r2b.txt
timeout 300 python scripts/time_grid.py 2>&1 | tee gpurun_out/grid_r2b.txt
timeout 300 python scripts/time_tail.py 2>&1 | tee gpurun_out/tail_r2b.txt
timeout 600 python oracle/make_ref_fixtures.py gpurun_out/ref_fixtures pinv 2>&1 | tee gpurun_out/ref_glass_pinv_fixtures.log
LD_LIBRARY_PATH=$PWD/gbd_pcg_b200 timeout 300 examples/api_surface.exe 2>&1 | tee gpurun_out/api_surface_r2b.txt
timeout 600 python -c "import __graft_entry__ as e; e.smoke()" 2>&1 | tail -8 | tee gpurun_out/smoke_r2b.txt
timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -25 | tee gpurun_out/pytest_gpu_r2b.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r2b.json 2> gpurun_out/bench_r2b.err; tail -3 gpurun_out/bench_r2b.err; cut -c1-600 gpurun_out/bench_r2b.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pcg_tmem4 -c 1 -f -o gpurun_out/ncu_tmem4_r2b python scripts/prof_kernel.py batched 8192 > gpurun_out/ncu_tmem4_r2b.log 2>&1; tail -2 gpurun_out/ncu_tmem4_r2b.log
