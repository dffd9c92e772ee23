"""Quick A/B timer: batched C5 launch time and C2 single-solve p50 for the library GBDPCG_LIB points at (default: in-tree build).
usage: [GBDPCG_LIB=...] python scripts/time_batched.py [batch] [flags]      prints one line"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
flags = int(sys.argv[2], 0) if len(sys.argv) > 2 else 0
nseed = 64
nx, nu, N = 14, 7, 32
S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", np.float32)
d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(batch // nseed, 1) for a in (S, P, gam))
d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, empty_pinv=0, flags=flags)
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
lam = d_l[:nseed].double().cpu().numpy()
# single C2 solve
d_l1 = torch.zeros(N * nx, dtype=torch.float32, device="cuda")
d_it1 = torch.zeros(1, dtype=torch.int32, device="cuda")
evs = []
for i in range(220):
    d_l1.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(50000)
    a.record()
    g.solve_dev_async(nx, N, d_S[0], d_P[0], d_g[0], d_l1, cfg, d_it1, None)
    b.record()
    evs.append((a, b))
torch.cuda.synchronize()
us = sorted(a.elapsed_time(b) * 1e3 for a, b in evs[20:])
plan = g.plan_query(nx, N, np.float32, True, cfg)
print(f"lib={os.environ.get('GBDPCG_LIB', 'default')} flags={flags:#x} batch={batch}: {ms:.4f} ms/launch -> {batch / ms / 1e3:.3f} M solves/s, "
      f"iters mean {d_it.float().mean().item():.2f}, lambda checksum {np.abs(lam).sum():.6f}; C2 single p50 {us[len(us) // 2]:.1f} us "
      f"({int(d_it1.item())} iters); {plan['kernel']} threads {plan['threads']} regs {plan['regs']}", flush=True)
