# ncu --set full of one batched launch of the tensor-memory kernel (8192 x C2)
mkdir -p gpurun_out
PROF_FLAGS=0x400 python scripts/prof_kernel.py batched 8192 > gpurun_out/prof_tmem_plain.log 2>&1 && \
PROF_FLAGS=0x400 ncu --set full --clock-control none --import-source on -k regex:pcg_tmem -s 2 -c 1 -f -o gpurun_out/prof_tmem python scripts/prof_kernel.py batched 8192 > gpurun_out/ncu_tmem.log 2>&1
tail -2 gpurun_out/ncu_tmem.log
ls -la gpurun_out/prof_tmem.ncu-rep
