"""Run-time-nx fallback kernel vs the specialised one: p50 single-solve latency on C2 / C3 / C4 shapes (stair).
usage: python scripts/time_generic.py"""
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


for shape in ("C2", "C3", "C4", (18, 9, 32, np.float32, 1e-6, 2000), (20, 10, 32, np.float32, 1e-6, 2000), (20, 10, 32, np.float64, 1e-10, 2000),
              (24, 12, 64, np.float32, 1e-6, 2000), (24, 12, 64, np.float64, 1e-10, 2000), (28, 14, 100, np.float32, 1e-6, 2000), (32, 16, 32, np.float32, 1e-6, 2000),
              (32, 16, 32, np.float64, 1e-10, 2000)):
    S, P, gam, nx, N, tol, max_iter = make_system(shape, "stair")
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_S, d_P, d_g = d(S), d(P), d(gam)
    d_l = torch.zeros(N * nx, dtype=d_S.dtype, device="cuda")
    d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
    for nm, fl in (("planner", 0), ("generic", g._lib.FLAG_FORCE_GENERIC)):
        cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0, flags=fl)
        try:
            plan = g.plan_query(nx, N, S.dtype, True, cfg)
            us = p50(lambda: g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None), d_l)
            print(f"nx={nx} N={N} {np.dtype(S.dtype).name} {nm:8s} {plan['kernel']:40s} cluster {plan['cluster']} x {plan['threads']}: iters {int(d_it.item())} p50 {us:.1f} us", flush=True)
        except g.GbdpcgError as e:
            print(f"nx={nx} N={N} {nm}: {e}")
