# session-10 first call: sanity of the restored tree + scalar-FMA A/B + FMA micro-benchmarks
mkdir -p gpurun_out
timeout 200 python __graft_entry__.py smoke > gpurun_out/s10_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/s10_smoke.log
timeout 900 python -m pytest tests -m gpu -q -x --timeout=180 > gpurun_out/s10_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/s10_pytest.log
for i in 1 2; do
timeout 200 python scripts/time_batched.py 8192 2>&1 | tail -1 | tee -a gpurun_out/s10_ab.log
GBDPCG_LIB=$PWD/gbd_pcg_b200/variants/libgbdpcg_scalar.so timeout 200 python scripts/time_batched.py 8192 2>&1 | tail -1 | tee -a gpurun_out/s10_ab.log
done
cd scripts/ubench && nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fma2_operands fma2_operands.cu && ./fma2_operands | tee ../../gpurun_out/s10_fma2.log; nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fma_rate fma_rate.cu && ./fma_rate | tee ../../gpurun_out/s10_fma_rate.log; cd ../..
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/s10_bench.json 2> gpurun_out/s10_bench.err; echo "bench rc=$?"; cat gpurun_out/s10_bench.json
