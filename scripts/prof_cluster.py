"""ncu driver: a few single solves of the C4 shape (nx=12, N=512, fp32, stair) -- the clustered solver kernel, 16 CTAs."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g
name = sys.argv[1] if len(sys.argv) > 1 else "C4"
nx, nu, N, dt, tol, max_iter = g.synth.CONFIGS[name]
S, P, gam = g.synth.lqr_schur(nx, nu, N, 0, "stair", dt)
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
d_S, d_P, d_g = d(S), d(P), d(gam)
d_l = torch.zeros(N * nx, dtype=d_S.dtype, device="cuda")
d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0)
for _ in range(3):
    d_l.zero_()
    g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None)
torch.cuda.synchronize()
print("ok", int(d_it.item()), g.plan_query(nx, N, dt, True, cfg))
