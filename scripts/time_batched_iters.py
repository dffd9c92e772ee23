"""Batched C5 launch time against pcg_max_iter: separates the per-system fixed cost (matrices staging -> registers, prologue,
outputs) from the per-iteration cost.   usage: python scripts/time_batched_iters.py [batch]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
nseed = 64
nx, nu, N = 14, 7, 32
S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", np.float32)
d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(batch // nseed, 1) for a in (S, P, gam))
d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
rows = []
for mi in (0, 1, 2, 4, 8, 16, 24, 1000):
    cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=mi, empty_pinv=0)
    ts = []
    for i in range(13):
        d_l.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None)
        b.record()
        torch.cuda.synchronize()
        if i >= 3:
            ts.append(a.elapsed_time(b))
    ms = float(np.median(ts))
    its = d_it.float().mean().item()
    rows.append((its, ms))
    per_sys = ms * 1e3 / ((batch + 147) // 148)
    print(f"max_iter {mi:5d}: iters mean {its:6.2f}  {ms:.4f} ms/launch  {per_sys:.3f} us per system per SM", flush=True)
(i0, t0), (i1, t1) = rows[2], rows[-1]
slope = (t1 - t0) / (i1 - i0)
print(f"per-iteration {slope * 1e3 / ((batch + 147) // 148):.3f} us; fixed per system {(t0 - slope * i0) * 1e3 / ((batch + 147) // 148):.3f} us")
