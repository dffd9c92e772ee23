"""A/B of the tensor-memory batched kernel (GBDPCG_FLAG_TMEM, two systems per SM) against the register-row kernel (GBDPCG_FLAG_NO_TMEM):
lambda and iteration counts must be bit-identical; launch time of both.   usage: python scripts/time_tmem.py [batch]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g  # noqa: E402

FLAG_TMEM, FLAG_NO_TMEM = 0x400, 0x800
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
nseed = 64
SHAPES = [(14, 7, 32), (12, 4, 40), (16, 8, 30), (8, 4, 60), (10, 5, 17)]
if len(sys.argv) > 2 and sys.argv[2] == "small":      # systems of <= 128 threads: 128 TMEM columns per CTA, four CTAs per SM
    SHAPES = [(10, 5, 17), (8, 4, 30), (12, 6, 20), (14, 7, 16), (16, 8, 16), (8, 4, 16)]
for (nx, nu, N) in SHAPES:
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", np.float32)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(batch // nseed, 1) for a in (S, P, gam))
    out = {}
    for name, flags in (("default", FLAG_NO_TMEM), ("tmem", FLAG_TMEM)):
        d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
        d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
        d_ex = torch.zeros(batch, dtype=torch.uint8, device="cuda")
        cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, empty_pinv=0, flags=flags)
        ts = []
        for i in range(9):
            d_l.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
            b.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(a.elapsed_time(b))
        out[name] = (float(np.median(ts)), d_l.cpu().numpy(), d_it.cpu().numpy(), d_ex.cpu().numpy())
    (t0, l0, i0, e0), (t1, l1, i1, e1) = out["default"], out["tmem"]
    same = bool(np.array_equal(l0.view(np.uint32), l1.view(np.uint32)) and np.array_equal(i0, i1) and np.array_equal(e0, e1))
    print(f"nx={nx} N={N} batch={batch}: default {t0:.4f} ms ({batch / t0 / 1e3:.2f} M/s)  tmem {t1:.4f} ms ({batch / t1 / 1e3:.2f} M/s)  "
          f"speed-up {t0 / t1:.3f}  iters mean {i0.mean():.2f}  bit-identical {same}  max|dlambda| {np.abs(l0 - l1).max():.3e}", flush=True)
