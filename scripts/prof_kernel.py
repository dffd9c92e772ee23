"""Small driver for ncu captures: a few launches of the batched kernel (C5) and of the single-system kernels (C2, C3, C4).
usage: python scripts/prof_kernel.py [batched|batched64|single|all] [batch]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
nseed = 64

if what in ("batched", "all"):
    nx, nu, N = 14, 7, 32
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", np.float32)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(batch // nseed, 1) for a in (S, P, gam))
    d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, empty_pinv=0, flags=int(os.environ.get("PROF_FLAGS", "0"), 0))
    for _ in range(3):
        d_l.zero_()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None)
    torch.cuda.synchronize()
    print("batched iters", d_it.float().mean().item())

if what == "batched64":      # fp64 throughput form (P_TMEM): C3 systems
    nx, nu, N = 14, 7, 64
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(32), "stair", np.float64)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(batch // 32, 1) for a in (S, P, gam))
    d_l = torch.zeros(batch, N * nx, dtype=torch.float64, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=1e-10, pcg_max_iter=2000, empty_pinv=0, flags=int(os.environ.get("PROF_FLAGS", "0"), 0))
    for _ in range(3):
        d_l.zero_()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None)
    torch.cuda.synchronize()
    print("batched64 iters", d_it.float().mean().item(), g.plan_query(nx, N, np.float64, True, cfg, batch=batch)["kernel"])

if what in ("single", "all"):
    for name in ("C2", "C3", "C4"):
        nx, nu, N, dt, tol, mi = g.synth.CONFIGS[name]
        S, P, gam = g.synth.lqr_schur(nx, nu, N, 0, "stair", dt)
        d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (S, P, gam))
        d_l = torch.zeros(N * nx, dtype=d_S.dtype, device="cuda")
        cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=mi, empty_pinv=0)
        for _ in range(3):
            d_l.zero_()
            it = g.solvePCG(nx, N, d_S, d_P, d_g, d_l, None, None, None, None, cfg)
        print(name, "iters", it)
