"""Debug aid: GBDPCG_FLAG_SYMMETRIC_S against the full-S host path, with the first differences printed."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g
nx, nu, N, batch = 14, 7, 32, 300
for dt, tol, mi in ((np.float32, 1e-6, 1000), (np.float64, 1e-10, 2000)):
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(6), "stair", dt)
    Sb, gb = (np.tile(a, (batch // 6, 1)).copy() for a in (S, gam))
    Sp = Sb.copy().reshape(batch, N, 3, nx * nx); Sp[:, :, 0, :] = np.nan; Sp = Sp.reshape(batch, -1)
    full = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=mi, flags=g._lib.FLAG_BUILD_STAIR)
    sym = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=mi, flags=g._lib.FLAG_BUILD_STAIR | g._lib.FLAG_SYMMETRIC_S)
    lf, ls = np.zeros((batch, N * nx), dt), np.zeros((batch, N * nx), dt)
    itf, exf = g.solve_batched_host(Sb, None, gb, lf, nx, N, batch, full)
    its, exs = g.solve_batched_host(Sp, None, gb, ls, nx, N, batch, sym)
    print(np.dtype(dt).name, "full: iters", itf[:6], "exit", exf.any(), "| sym: iters", its[:6], "exit", exs.any(), "| nan in sym lambda", np.isnan(ls).any(),
          "| max |dlam|", np.nanmax(np.abs(lf - ls)), "systems differing", int((np.abs(lf - ls).max(axis=1) > 0).sum()))
    its2, _ = g.solve_batched_host(Sb, None, gb, ls * 0, nx, N, batch, sym)
    print("   sym flag on the unpoisoned S: iters", its2[:6])
    l1 = np.zeros(N * nx, dt)
    it1 = g.solvePCG(Sp[1].copy(), gam[1], l1, nx, N, sym)
    print("   single host solve with the flag: iters", it1, "vs", itf[1], "max|d|", np.abs(l1 - lf[1]).max())
