mkdir -p gpurun_out
timeout 900 python scripts/sweep_parity.py 80 1 > gpurun_out/s10r_sweep.log 2>&1; echo "sweep rc=$?"; grep -c " ok$" gpurun_out/s10r_sweep.log; grep "FAIL\|EXCEPTION\|Error\|ALL OK" gpurun_out/s10r_sweep.log | head -20; tail -3 gpurun_out/s10r_sweep.log
