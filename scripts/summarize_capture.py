"""One ncu --set full capture -> profiles/<name>.txt (key counters + stall reasons + hottest SASS) and, with --traffic,
profiles/traffic.json.   usage: python scripts/summarize_capture.py gpurun_out/prof_tmem.ncu-rep ncu_tmem_r1g "<command line>" [--traffic]"""
import csv
import json
import os
import subprocess
import sys

rep, name, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "profiles")
raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
rr = list(csv.reader(raw.splitlines()))
h, u, d = rr[0], rr[1], rr[-1]
want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__maximum_warps_avg_per_active_cycle",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active", "smsp__sass_inst_executed_op_tmem_ldt.sum", "smsp__sass_inst_executed_op_tmem_stt.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]
hi = {k: i for i, k in enumerate(h)}
out = {}
lines = [f"ncu --set full --clock-control none --import-source on : {cmd}"]
for k in want:
    if k in hi:
        lines.append(f"{k} [{u[hi[k]]}] = {d[hi[k]]}")
        out[k] = (d[hi[k]], u[hi[k]])
src = subprocess.check_output(["ncu", "-i", rep, "--page", "source", "--csv"], text=True)
open("/tmp/_src.csv", "w").write(src)
hot = subprocess.check_output([sys.executable, os.path.join(ROOT, "scripts", "ncu_hot.py"), "/tmp/_src.csv", "1.0"], text=True)
lines.append("")
lines.append("ncu --page source (warp-state sampling) of the same capture: stall reasons and SASS instructions with >= 1% of samples")
lines.append(hot)
open(os.path.join(P, name + ".txt"), "w").write("\n".join(lines))
print("\n".join(lines)[:6000])
if "--traffic" in sys.argv:
    def to_bytes(v, unit):
        return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    traffic = to_bytes(*out["dram__bytes_read.sum"]) + to_bytes(*out["dram__bytes_write.sum"])
    json.dump({"kernel": out["Kernel Name"][0].split("(")[0], "batch": 8192, "dram_bytes_per_launch": traffic,
               "algorithmic_bytes_per_launch": 8192 * 152768, "source": f"profiles/{name}.txt"},
              open(os.path.join(P, "traffic.json"), "w"), indent=1)
