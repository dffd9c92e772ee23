"""Determinism stress of the tensor-memory batched kernel: the same batch solved over and over (alone, and with a second
stream launching the same kernel concurrently so that CTAs of two launches share SMs and the 512 TMEM columns); every lambda
and iteration count must be bit-identical to the first run.   usage: python scripts/stress_tmem.py [reps]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
bad = 0
for (nx, nu, N, batch) in [(14, 7, 32, 2048), (16, 8, 32, 1500), (12, 4, 40, 1111), (8, 4, 60, 700)]:
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(16), "stair", np.float32)
    r = (batch + 15) // 16
    d_S, d_P, d_g = (torch.from_numpy(np.tile(a, (r, 1))[:batch].copy()).cuda() for a in (S, P, gam))
    cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, empty_pinv=0, flags=0x400)
    s2 = torch.cuda.Stream()
    ref = None
    for i in range(reps):
        d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
        d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
        d_l2 = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None)
        if i % 2:
            with torch.cuda.stream(s2):
                g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l2, cfg, None, None, stream=s2)
        torch.cuda.synchronize()
        cur = (d_l.clone(), d_it.clone())
        if ref is None:
            ref = cur
        else:
            same = bool(torch.equal(cur[0], ref[0]) and torch.equal(cur[1], ref[1]) and (i % 2 == 0 or torch.equal(d_l2, ref[0])))
            bad += 0 if same else 1
    print(f"nx={nx} N={N} batch={batch}: {reps} runs ({reps // 2} with a concurrent second launch), mismatches so far {bad}, iters mean {ref[1].float().mean().item():.2f}", flush=True)
print("ALL IDENTICAL" if bad == 0 else f"MISMATCHES {bad}")
sys.exit(1 if bad else 0)
