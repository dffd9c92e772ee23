"""Times the on-GPU preconditioner builder on the C5 batch (8192 x nx=14, N=32 fp32) and a few other shapes.
usage: python scripts/time_pinv.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g  # noqa: E402

for (nx, nu, N, dt, batch) in [(14, 7, 32, np.float32, 8192), (14, 7, 64, np.float64, 2048), (12, 4, 512, np.float32, 512), (14, 7, 32, np.float32, 1)]:
    nseed = min(batch, 16)
    S, _, _ = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "identity", dt)
    d_S = torch.from_numpy(S).cuda().repeat(batch // nseed, 1).contiguous()
    d_P = torch.empty_like(d_S)
    for kind in ("jacobi", "stair"):
        ts = []
        for i in range(13):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(200000)      # let the CPU run ahead so that the events bracket GPU time only
            a.record()
            g.build_pinv_dev(nx, N, batch, d_S, d_P, kind)
            b.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(a.elapsed_time(b))
        ms = float(np.median(ts))
        by = 2 * d_S.numel() * d_S.element_size()
        print(f"build_pinv {kind:6s} nx={nx} N={N} {np.dtype(dt).name} batch={batch}: {ms * 1e3:.1f} us, {by / ms / 1e6:.0f} GB/s (read S + write Pinv)", flush=True)
