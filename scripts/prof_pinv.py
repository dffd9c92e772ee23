"""ncu driver: a few launches of the preconditioner builder on the C5 batch."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g
nx, nu, N, batch = 14, 7, 32, 8192
S, _, _ = g.synth.lqr_schur_batch(nx, nu, N, range(16), "identity", np.float32)
d_S = torch.from_numpy(S).cuda().repeat(batch // 16, 1).contiguous()
d_P = torch.empty_like(d_S)
for _ in range(3):
    g.build_pinv_dev(nx, N, batch, d_S, d_P, "stair")
torch.cuda.synchronize()
print("ok")
