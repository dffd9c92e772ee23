"""Wide form of the four-row tensor-memory kernel (8 warps, one system per SM) against the register-row clusters it replaces in
throughput mode: launch times on fp32 batches of systems of 129..256 four-row threads.   usage: python scripts/time_tmem4_wide.py [batch]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g  # noqa: E402

FLAG_TMEM, FLAG_NO_TMEM = 0x400, 0x800
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nseed = 32
for (nx, nu, N) in [(14, 7, 64), (14, 7, 48), (16, 8, 64), (12, 4, 85), (8, 4, 128), (10, 5, 80)]:
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", np.float32)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(batch // nseed, 1) for a in (S, P, gam))
    out = {}
    for name, flags in (("regs", FLAG_NO_TMEM), ("wide", FLAG_TMEM), ("planner", 0)):
        d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
        d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
        cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=2000, empty_pinv=0, flags=flags)
        info = g.plan_query(nx, N, np.float32, True, cfg, batch=batch)
        ts = []
        for i in range(8):
            d_l.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None)
            b.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(a.elapsed_time(b))
        out[name] = (float(np.median(ts)), d_l.cpu().numpy(), d_it.cpu().numpy(), info)
    ref = out["regs"]
    for name, (t, l, it, info) in out.items():
        bytes_per = 2 * 3 * N * nx * nx * 4 + 3 * N * nx * 4
        print(f"nx={nx} N={N} batch={batch} {name:8s}: {t:8.4f} ms ({batch / t / 1e3:6.3f} M/s, {batch * bytes_per / t / 1e6:7.1f} GB/s) iters mean {it.mean():.2f} "
              f"max|d iters| {int(np.abs(it - ref[2]).max())} rel diff {np.linalg.norm(l - ref[1]) / np.linalg.norm(ref[1]):.1e} | {info['kernel']} cluster {info['cluster']} "
              f"threads {info['threads']} regs {info['regs']} resident {info['max_resident']}", flush=True)
