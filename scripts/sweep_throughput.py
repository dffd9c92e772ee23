"""Randomised parity sweep of the throughput-mode kernels (four-row tensor-memory kernel incl. its wide form, two-row tensor-memory kernel,
fp64 tensor-memory form of the cluster kernel, one and two CTAs per SM): random (nx, N, dtype, batch, preconditioner), GBDPCG_FLAG_TMEM
where a tensor-memory kernel exists and the planner's own choice, every distinct system against the host oracle (iterations within +-1,
lambda at the bar at equal count), replicas bit-identical.   usage: python scripts/sweep_throughput.py [n_cases] [seed]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gbd_pcg_b200 as g  # noqa: E402
from oracle import oracle  # noqa: E402

ncase = int(sys.argv[1]) if len(sys.argv) > 1 else 100
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
FLAG_TMEM = 0x400
bad = 0
seen = {}
for case in range(ncase):
    nx = int(rng.choice([8, 10, 12, 14, 16]))
    dt = rng.choice([np.float32, np.float64])
    N = int(rng.integers(2, 129 if dt == np.float32 else 101))
    B = int(rng.integers(149, 700))
    kind = str(rng.choice(["jacobi", "stair"]))
    nseed = 3
    tol, bar = (1e-6, 1e-5) if dt == np.float32 else (1e-10, 1e-10)
    S, P, gam = g.synth.lqr_schur_batch(nx, nx // 2, N, range(case, case + nseed), kind, dt)
    idx = np.arange(B) % nseed
    d_S, d_P, d_g = (torch.from_numpy(a[idx].copy()).cuda() for a in (S, P, gam))
    res = {}
    for arm, flags in (("forced", FLAG_TMEM), ("planner", 0)):
        cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=3000, empty_pinv=0, flags=flags)
        try:
            kern = g.plan_query(nx, N, dt, True, cfg, batch=B)["kernel"]
        except g.GbdpcgError:
            continue
        d_l = torch.zeros(B, N * nx, dtype=d_S.dtype, device="cuda")
        d_it = torch.zeros(B, dtype=torch.int32, device="cuda")
        d_ex = torch.ones(B, dtype=torch.uint8, device="cuda")
        g.solve_batched_dev(nx, N, B, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
        torch.cuda.synchronize()
        res[arm] = (kern, d_l.cpu().numpy(), d_it.cpu().numpy(), d_ex.cpu().numpy())
    line = f"case {case:3d}: nx={nx:2d} N={N:3d} B={B:3d} {np.dtype(dt).name:7s} {kind:6s}"
    fail = False
    for arm, (kern, lam, it, ex) in res.items():
        short = kern.split("<")[0] + ("/wide" if "wide" in kern else "") + ("/P_TMEM" if "P_TMEM" in kern else "") + ("/2perSM" if "2 per SM" in kern and "P_TMEM" in kern else "")
        seen[short] = seen.get(short, 0) + 1
        worst = 0.0
        ok = not ex.any()
        for s_ in range(nseed):
            rows = np.nonzero(idx == s_)[0]
            ok = ok and bool((it[rows] == it[rows[0]]).all()) and all(np.array_equal(lam[r], lam[rows[0]]) for r in rows)
            lo, ito, _ = oracle.pcg(S[s_], P[s_], gam[s_], np.zeros(N * nx, dt), nx, N, tol, 3000)
            # (fp32 systems of a few dozen unknowns: the iteration at which |r.Phi^-1 r| crosses the tolerance is order-sensitive -- SURVEY App. E,
            #  tests judge C1 on the solution for the same reason; +-2 there, +-1 everywhere else)
            ok = ok and abs(int(it[rows[0]]) - ito) <= (2 if (N * nx <= 200 and dt == np.float32) else 1)
            if abs(int(it[rows[0]]) - ito) > 1: line += f" [oracle iters {ito}]"
            if int(it[rows[0]]) != ito:
                lo = oracle.pcg(S[s_], P[s_], gam[s_], np.zeros(N * nx, dt), nx, N, 0.0, int(it[rows[0]]))[0]
            e = float(np.linalg.norm(lam[rows[0]].astype(np.float64) - lo) / np.linalg.norm(lo))
            lim = bar
            if N * nx <= 200:
                # small systems: CG run to finite termination amplifies rounding, two CPU statements of the same recurrence already differ by
                # more than the bar (tests: noise_floor) -- the bar is then 3x that measured CPU-vs-CPU floor
                k = int(it[rows[0]])
                l1 = oracle.pcg(S[s_], P[s_], gam[s_], np.zeros(N * nx, dt), nx, N, 0.0, k, order=1)[0]
                ln = oracle.pcg_numpy(S[s_], P[s_], gam[s_], np.zeros(N * nx, dt), nx, N, 0.0, k)[0]
                lk = oracle.pcg(S[s_], P[s_], gam[s_], np.zeros(N * nx, dt), nx, N, 0.0, k)[0]
                rel = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64) - b) / np.linalg.norm(b))
                lim = max(bar, 3 * rel(l1, lk), 3 * rel(ln, lk))
            ok = ok and e <= lim
            worst = max(worst, e)
        fail = fail or not ok
        line += f" | {arm}: {short} err {worst:.1e} iters {[int(it[np.nonzero(idx == s_)[0][0]]) for s_ in range(nseed)]} ex {int(ex.sum())} {'ok' if ok else 'FAIL'}"
    bad += fail
    print(line, flush=True)
print("kernels exercised:", seen)
print("FAILED" if bad else "ALL OK", bad)
