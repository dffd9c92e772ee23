"""Turns gpurun_out/ ncu artefacts into the small text summaries committed under profiles/.
usage: python scripts/summarize_profiles.py <tag>    (reads gpurun_out/launches_<tag>.csv, prof_<tag>.ncu-rep)"""
import collections
import csv
import json
import os
import subprocess
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)

# ---- launch list of the bench command: per-kernel totals and shares ----
rows = [r for r in csv.reader(open(os.path.join(G, f"launches_{tag}.csv"))) if len(r) > 10]
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        v = float(r[ix["Metric Value"]])
    except (ValueError, KeyError):
        continue
    unit = r[ix["Metric Unit"]]
    us = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)
    name = r[ix["Kernel Name"]].split("(")[0][:90] + " grid=" + r[ix["Grid Size"]] if "Grid Size" in ix else r[ix["Kernel Name"]][:90]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += us
tot = sum(a[1] for a in agg.values())
with open(os.path.join(P, f"launches_{tag}.txt"), "w") as f:
    f.write(f"ncu --metrics gpu__time_duration.sum --clock-control none -c 400 : python bench.py --steps 2 --warmup 3   (first 400 launches; "
            f"cold-cache, serialised -- compare shares, not absolutes)\n")
    f.write(f"{'launches':>8} {'total us':>12} {'share':>7}  kernel\n")
    for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{n:8d} {us:12.1f} {100 * us / tot:6.1f}%  {name}\n")
print(open(os.path.join(P, f"launches_{tag}.txt")).read())

# ---- full capture of the batched kernel ----
rep = os.path.join(G, f"prof_{tag}.ncu-rep")
raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
rr = list(csv.reader(raw.splitlines()))
h, u, d = rr[0], rr[1], rr[2]
want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"]
hi = {k: i for i, k in enumerate(h)}
out = {}
with open(os.path.join(P, f"ncu_full_{tag}.txt"), "w") as f:
    f.write("ncu --set full --clock-control none --import-source on : python scripts/prof_kernel.py batched 8192  (one launch: 8192 C2 systems, 148 CTAs)\n")
    for k in want:
        if k in hi:
            f.write(f"{k} [{u[hi[k]]}] = {d[hi[k]]}\n")
            out[k] = (d[hi[k]], u[hi[k]])
print(open(os.path.join(P, f"ncu_full_{tag}.txt")).read())

def to_bytes(v, unit):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
traffic = to_bytes(*out["dram__bytes_read.sum"]) + to_bytes(*out["dram__bytes_write.sum"])
json.dump({"kernel": out["Kernel Name"][0].split("(")[0], "batch": 8192, "dram_bytes_per_launch": traffic,
           "algorithmic_bytes_per_launch": 8192 * 152768, "source": f"profiles/ncu_full_{tag}.txt"},
          open(os.path.join(P, "traffic.json"), "w"), indent=1)

# ---- stall reasons / hottest instructions of the same capture ----
src = subprocess.check_output(["ncu", "-i", rep, "--page", "source", "--csv"], text=True)
open("/tmp/_src.csv", "w").write(src)
hot = subprocess.check_output([sys.executable, os.path.join(ROOT, "scripts", "ncu_hot.py"), "/tmp/_src.csv", "1.0"], text=True)
open(os.path.join(P, f"ncu_stalls_{tag}.txt"), "w").write("ncu --page source (warp-state sampling) of the capture above: stall reasons and SASS instructions with >= 1% of samples\n" + hot)
print(hot[:1500])
