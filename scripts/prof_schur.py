"""ncu driver: a few launches of the Schur-formation kernel on the C5 batch."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g
batch, nx, nu, N, nseed = 8192, 14, 7, 32, 16
ds = [g.synth.lqr_stage_data(nx, nu, N, s, dtype=np.float32) for s in range(nseed)]
dd = {k: torch.from_numpy(np.concatenate([d[k] for d in ds])).cuda().repeat(batch // nseed, 1).contiguous() for k in ds[0]}
f32 = dict(dtype=torch.float32, device="cuda")
S, gam = torch.empty(batch * 3 * N * nx * nx, **f32), torch.empty(batch * N * nx, **f32)
Qi, Ri = torch.empty(batch * N * nx * nx, **f32), torch.empty(batch * (N - 1) * nu * nu, **f32)
for _ in range(3):
    g.schur_form_dev(nx, nu, N, batch, dd["Q"], dd["R"], dd["A"], dd["B"], dd["q"], dd["r"], dd["c"], S, gam, Qi, Ri)
torch.cuda.synchronize()
print("ok")
