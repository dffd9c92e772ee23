"""fp64: reference-order recurrence (2 reductions / iteration) vs the single-reduction Chronopoulos-Gear form, per
preconditioner -- lambda against the host oracle at equal iteration count, iteration counts, and p50 latency.
usage: python scripts/f64_recurrence.py          prints one line per (shape, preconditioner, recurrence[, cluster])"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gbd_pcg_b200 as g  # noqa: E402
from helpers import make_system  # noqa: E402
from oracle import oracle  # noqa: E402

L = g._lib


def p50(launch, d_l, n_warm=20, n=200):
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


shapes = [("C3", 0), ("C3", 1), ("C3", 2), ((14, 7, 32, np.float64, 1e-10, 2000), 3), ((12, 4, 128, np.float64, 1e-10, 4000), 4),
          ((8, 4, 200, np.float64, 1e-12, 4000), 5), ((16, 8, 48, np.float64, 1e-10, 4000), 6), ((6, 3, 20, np.float64, 1e-12, 2000), 7),
          ((10, 5, 20, np.float64, 1e-12, 2000), 8)]
worst = {}
for shape, seed in shapes:
    for pc in ("identity", "jacobi", "stair"):
        S, P, gam, nx, N, tol, max_iter = make_system(shape, pc, seed=seed)
        lam0 = np.zeros(N * nx)
        lam_o, it_o, _ = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=0)
        d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        d_S, d_P, d_g = d(S), (d(P) if P is not None else None), d(gam)
        d_l = torch.zeros(N * nx, dtype=torch.float64, device="cuda")
        d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
        for rec, fl in (("reference-order", L.FLAG_TEXTBOOK), ("single-reduction", L.FLAG_FUSED)):
            for cl in ((0,) if shape != "C3" or seed else (0, 8)):
                cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0 if P is not None else 1, flags=fl, cluster=cl)
                try:
                    us = p50(lambda: g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None), d_l)
                except g.GbdpcgError as e:
                    print(f"{shape} {pc} {rec} cluster={cl}: {e}")
                    continue
                it = int(d_it.item())
                lam = d_l.cpu().numpy()
                lam_ref = lam_o if it == it_o else oracle.pcg(S, P, gam, lam0, nx, N, 0.0, it, order=0)[0]
                rel = np.abs(lam - lam_ref).max() / np.abs(lam_ref).max()
                plan = g.plan_query(nx, N, np.float64, P is not None, cfg)
                worst[(pc, rec)] = max(worst.get((pc, rec), 0.0), rel)
                print(f"{shape if isinstance(shape, str) else shape[:3]} seed {seed} {pc:8s} {rec:16s} cluster {plan['cluster']} x {plan['threads']}: "
                      f"iters {it} (oracle {it_o}) rel diff {rel:.2e}  p50 {us:.1f} us", flush=True)
for k, v in sorted(worst.items()):
    print("worst", k, f"{v:.2e}")
