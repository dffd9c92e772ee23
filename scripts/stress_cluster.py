"""Determinism stress of the clustered solver (push-mode exchanges): the same systems solved thousands of times, alone and in
batches with several clusters in flight, must give bit-identical lambda and iteration counts every time -- a lost or early
message would show up as a different result (or a hang: run under `timeout`).
usage: python scripts/stress_cluster.py [repeats]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gbd_pcg_b200 as g  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
bad = 0
for (nx, nu, N, dt, tol, cl, batch) in [(14, 7, 64, np.float64, 1e-10, 0, 1), (12, 4, 512, np.float32, 1e-6, 0, 1), (12, 4, 100, np.float32, 1e-6, 16, 1),
                                        (20, 10, 32, np.float32, 1e-6, 0, 1), (24, 12, 64, np.float64, 1e-10, 0, 1), (14, 7, 64, np.float64, 1e-10, 0, 9),
                                        (12, 4, 100, np.float32, 1e-6, 16, 40), (6, 3, 48, np.float32, 1e-6, 2, 300), (20, 10, 32, np.float32, 1e-6, 0, 64),
                                        (15, 0, 64, np.float32, 1e-6, 0, 30)]:
    nseed = min(batch, 4)
    if nx % 2:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from helpers import random_block_tridiagonal
        ss = [random_block_tridiagonal(nx, N, "stair", dt, seed=s, coupled=True) for s in range(nseed)]
        S, P, gam = (np.stack([t[i] for t in ss]) for i in range(3))
    else:
        S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", dt)
    k = (batch + nseed - 1) // nseed
    d = lambda a: torch.from_numpy(np.tile(a, (k, 1))[:batch].copy()).cuda()
    d_S, d_P, d_g = d(S), d(P), d(gam)
    d_l = torch.zeros(batch, N * nx, dtype=d_S.dtype, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=3000, empty_pinv=0, cluster=cl)
    plan = g.plan_query(nx, N, dt, True, cfg, batch=batch)
    n = reps if batch == 1 else max(reps // 10, 50)
    ref_l = ref_i = None
    diff = 0
    for r in range(n):
        d_l.zero_()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None)
        if r == 0:
            ref_l, ref_i = d_l.clone(), d_it.clone()
        elif not (torch.equal(d_l, ref_l) and torch.equal(d_it, ref_i)):
            diff += 1
    torch.cuda.synchronize()
    bad += diff
    print(f"nx={nx} N={N} {np.dtype(dt).name} batch={batch} cluster {plan['cluster']} x {plan['threads']} ({plan['kernel']}): {n} runs, "
          f"iters {ref_i[:nseed].tolist()}, {diff} differing", flush=True)
print("ALL IDENTICAL" if bad == 0 else f"NONDETERMINISM: {bad}")
sys.exit(1 if bad else 0)
