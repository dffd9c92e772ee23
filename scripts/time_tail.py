"""Batched C2 launch time against the batch size around whole numbers of resident waves (296 tensor-memory slots on a B200):
separates the partial last wave from the per-launch fixed cost behind the strong-scaling efficiency at 8 GPUs (1024 systems per GPU).
usage: python scripts/time_tail.py [flags]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g  # noqa: E402

flags = int(sys.argv[1], 0) if len(sys.argv) > 1 else 0
nx, nu, N, nseed = 14, 7, 32, 64
S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", np.float32)
maxb = 8192
d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(maxb // nseed, 1) for a in (S, P, gam))
flush = torch.zeros(64 << 20, dtype=torch.int32, device="cuda")      # 256 MB = 2 x L2, READ between launches (no dirty lines left behind)
for batch in (148, 296, 444, 592, 740, 888, 1024, 1036, 1184, 2048, 4096, 8192):
    d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, empty_pinv=0, flags=flags)
    kern = g.plan_query(nx, N, np.float32, True, cfg, batch=batch)["kernel"]
    ts = []
    for i in range(25):
        d_l.zero_()
        flush.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None)
        b.record()
        torch.cuda.synchronize()
        if i >= 5:
            ts.append(a.elapsed_time(b) * 1e3)
    t = float(np.median(ts))
    print(f"batch {batch:5d} ({batch / 296:5.2f} waves of 296): {t:8.1f} us  {batch / t:6.2f} solves/us  {t / batch * 148:7.2f} us per system per SM | {kern}", flush=True)
