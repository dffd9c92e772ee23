"""Determinism stress of the cooperative-grid solver (one barrier per iteration, exchange buffers alternating with the iteration's
parity): the same systems solved hundreds of times -- alone, as a batch on one grid, and from two streams at once (each launch has its
own scratch from the pool) -- must give bit-identical lambda and iteration counts every time; a buffer overwritten one iteration
early would show up as a different result (or a hang: run under `timeout`).   usage: python scripts/stress_grid.py [repeats]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gbd_pcg_b200 as g  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
bad = 0
for (nx, nu, N, dt, tol, batch) in [(14, 7, 1500, np.float32, 1e-6, 1), (14, 7, 1000, np.float64, 1e-10, 1), (14, 7, 1041, np.float32, 1e-6, 1),
                                    (8, 4, 2500, np.float64, 1e-10, 1), (14, 7, 1100, np.float32, 1e-6, 3), (6, 3, 6000, np.float32, 1e-6, 2)]:
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(batch), "stair", dt)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (S, P, gam))
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=5000, empty_pinv=0)
    plan = g.plan_query(nx, N, dt, True, cfg, batch=batch)
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = [(torch.zeros(batch, N * nx, dtype=d_S.dtype, device="cuda"), torch.zeros(batch, dtype=torch.int32, device="cuda")) for _ in streams]
    ref_l = ref_i = None
    diff = 0
    for r in range(reps):
        two = r % 3 == 2                                           # every third round: both streams at once
        for si, st in enumerate(streams if two else streams[:1]):
            d_l, d_it = outs[si]
            with torch.cuda.stream(st):
                d_l.zero_()
                g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None)
        torch.cuda.synchronize()
        for si in range(2 if two else 1):
            l, i = outs[si][0].cpu().numpy(), outs[si][1].cpu().numpy()
            if ref_l is None:
                ref_l, ref_i = l.copy(), i.copy()
            elif not (np.array_equal(l, ref_l) and np.array_equal(i, ref_i)):
                diff += 1
    bad += diff
    print(f"nx={nx} N={N} {np.dtype(dt).name} batch={batch} {plan['kernel']} G={plan['cluster']} x {plan['threads']}: {reps} rounds, iters {ref_i.tolist()}, "
          f"{diff} differing", flush=True)
print("ALL IDENTICAL" if bad == 0 else f"{bad} DIFFERENCES")
sys.exit(1 if bad else 0)
