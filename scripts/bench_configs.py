"""Single-solve latency of every BASELINE config (C1..C4) x preconditioner, ours vs the unmodified reference CUDA kernel
(oracle/_ref, when present): kernel scope, CUDA events, p50 of 300 solves after 30 warm-ups.
usage: python scripts/bench_configs.py [out.json]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gbd_pcg_b200 as g  # noqa: E402
from helpers import make_system  # noqa: E402
from oracle import ref_cuda  # noqa: E402


def p50(launch, d_l, n_warm=30, n=300):
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for _ in range(n_warm):
        d_l.zero_(); launch()
    torch.cuda.synchronize()
    for a, b in evs:
        torch.cuda._sleep(100000)
        d_l.zero_()
        a.record(); launch(); b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) * 1e3 for a, b in evs)
    return ts[len(ts) // 2], ts[int(len(ts) * 0.99)]


rows = []
for name in ("C1", "C2", "C3", "C4"):
    for pc in ("identity", "jacobi", "stair"):
        S, P, gam, nx, N, tol, max_iter = make_system(name, pc)
        dt = S.dtype
        P_full = P if P is not None else g.layout.identity_compressed(nx, N, dt)
        d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        d_S, d_P, d_g = d(S), (d(P) if P is not None else None), d(gam)
        d_Pf = d(P_full)
        d_l = torch.zeros(N * nx, dtype=d_S.dtype, device="cuda")
        d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
        cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0 if P is not None else 1)
        plan = g.plan_query(nx, N, dt, P is not None, cfg)
        ours = p50(lambda: g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None), d_l)
        it = int(d_it.item())
        row = {"config": name, "nx": nx, "N": N, "dtype": np.dtype(dt).name, "precond": pc, "iters": it, "kernel": plan["kernel"],
               "cluster": plan["cluster"], "threads": plan["threads"], "ours_p50_us": ours[0], "ours_p99_us": ours[1],
               "ours_us_per_iter": ours[0] / max(it, 1)}
        if ref_cuda.has(dt, nx, N):
            ref = ref_cuda.get(dt, nx, N)
            r = p50(lambda: ref.launch(d_S, d_Pf, d_g, d_l, max_iter, tol), d_l)
            row.update(ref_p50_us=r[0], ref_p99_us=r[1], ref_iters=int(ref.d_iters.item()), speedup=r[0] / ours[0])
        rows.append(row)
        print(json.dumps(row), flush=True)
if len(sys.argv) > 1:
    json.dump(rows, open(sys.argv[1], "w"), indent=1)
