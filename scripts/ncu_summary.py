"""Text summary of one-kernel .ncu-rep (raw metrics + hottest SASS).  usage: python scripts/ncu_summary.py rep.ncu-rep "title" > out.txt"""
import csv, subprocess, sys, tempfile, os
rep, title = sys.argv[1], sys.argv[2]
raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
rr = list(csv.reader(raw.splitlines()))
h, u, d = rr[0], rr[1], rr[2]
want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"]
hi = {k: i for i, k in enumerate(h)}
print(title)
for k in want:
    if k in hi:
        print(f"{k} [{u[hi[k]]}] = {d[hi[k]]}")
print()
with tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False) as f:
    f.write(subprocess.check_output(["ncu", "-i", rep, "--page", "source", "--csv"], text=True))
    tmp = f.name
here = os.path.dirname(os.path.abspath(__file__))
print(subprocess.check_output([sys.executable, os.path.join(here, "ncu_hot.py"), tmp, "1.0"], text=True))
os.unlink(tmp)
