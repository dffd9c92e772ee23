"""Summarise an `ncu --page source --csv` export: stall-reason totals and the hottest SASS instructions.
usage: ncu -i rep.ncu-rep --page source --csv --launch-skip K --launch-count 1 > src.csv ; python scripts/ncu_hot.py src.csv [min_pct]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 0.4
hdr = next(r for r in rows if "# Samples" in r and "Source" in r)
ix = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows:
    if len(r) == len(hdr) and r is not hdr:
        try:
            int(r[ix["# Samples"]])
        except ValueError:
            continue
        if r[ix["Address"]].startswith("0x"):
            data.append(r)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
T = sum(int(d[ix["# Samples"]]) for d in data)
tot = {s: sum(int(d[ix[s]] or 0) for d in data) for s in stalls}
print("kernel:", rows[0][1][:100] if rows[0] else "")
print("SASS instructions", len(data), "samples", T, "warp-inst", sum(int(d[ix["Instructions Executed"]] or 0) for d in data))
for s, v in sorted(tot.items(), key=lambda x: -x[1])[:10]:
    print(f"  {s:26s} {100 * v / max(1, sum(tot.values())):5.1f}%")
for n, d in enumerate(data):
    s = int(d[ix["# Samples"]])
    if s * 100 < T * minpct:
        continue
    st = {k[6:]: int(d[ix[k]] or 0) for k in stalls}
    top = [f"{k}:{v}" for k, v in sorted(st.items(), key=lambda x: -x[1])[:2] if v]
    print(f"{n:5d} {100 * s / T:5.1f}% ex={d[ix['Instructions Executed']]:>9s} {d[ix['Source']][:56]:56s} {' '.join(top)}")
