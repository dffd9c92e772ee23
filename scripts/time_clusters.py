"""Single-solve p50 latency against the forced cluster size (stair preconditioner).
usage: python scripts/time_clusters.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gbd_pcg_b200 as g  # noqa: E402
from helpers import make_system  # noqa: E402


def p50(launch, d_l, n_warm=10, n=100):
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
    return ts[len(ts) // 2]


for shape in ("C2", "C3", "C4", (6, 3, 48, np.float32, 1e-6, 2000), (8, 4, 200, np.float64, 1e-10, 4000), (16, 8, 48, np.float32, 1e-6, 2000),
              (20, 10, 32, np.float32, 1e-6, 2000), (24, 12, 64, np.float32, 1e-6, 2000), (24, 12, 64, np.float64, 1e-10, 2000),
              (32, 16, 32, np.float32, 1e-6, 2000)):
    for pc in ("stair", "identity"):
        S, P, gam, nx, N, tol, max_iter = make_system(shape, pc)
        d = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).cuda()
        d_S, d_P, d_g = d(S), d(P), d(gam)
        d_l = torch.zeros(N * nx, dtype=d_S.dtype, device="cuda")
        d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
        out = []
        for cl in (0, 1, 2, 4, 8, 16):
            cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0 if P is not None else 1, cluster=cl)
            try:
                plan = g.plan_query(nx, N, S.dtype, P is not None, cfg)
                us = p50(lambda: g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None), d_l)
                out.append(f"{'auto=' + str(plan['cluster']) if cl == 0 else cl}:{us:.1f}us({plan['threads']}t,pm{plan['pinv_mode']})")
            except g.GbdpcgError:
                out.append(f"{cl}:-")
        print(f"nx={nx} N={N} {np.dtype(S.dtype).name} {pc:8s} iters {int(d_it.item())}: " + "  ".join(out), flush=True)
