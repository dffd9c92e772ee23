"""Times the device-resident MPC step on the C5 batch: Schur formation -> stair Phi^-1 -> PCG -> primal recovery.
usage: python scripts/time_chain.py [batch]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
nx, nu, N, nseed = 14, 7, 32, 16
ds = [g.synth.lqr_stage_data(nx, nu, N, s, dtype=np.float32) for s in range(nseed)]
dd = {k: torch.from_numpy(np.concatenate([d[k] for d in ds])).cuda().repeat(batch // nseed, 1).contiguous() for k in ds[0]}
f32 = dict(dtype=torch.float32, device="cuda")
S, gam = torch.empty(batch * 3 * N * nx * nx, **f32), torch.empty(batch * N * nx, **f32)
P, lam = torch.empty_like(S), torch.zeros(batch * N * nx, **f32)
Qi, Ri = torch.empty(batch * N * nx * nx, **f32), torch.empty(batch * (N - 1) * nu * nu, **f32)
x, u = torch.empty(batch * N * nx, **f32), torch.empty(batch * (N - 1) * nu, **f32)
it = torch.zeros(batch, dtype=torch.int32, device="cuda")
cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, empty_pinv=0)
steps = {
    "schur_form": lambda: g.schur_form_dev(nx, nu, N, batch, dd["Q"], dd["R"], dd["A"], dd["B"], dd["q"], dd["r"], dd["c"], S, gam, Qi, Ri),
    "build_pinv(stair)": lambda: g.build_pinv_dev(nx, N, batch, S, P, "stair"),
    "pcg_solve": lambda: (lam.zero_(), g.solve_batched_dev(nx, N, batch, S, P, gam, lam, cfg, it, None)),
    "schur_recover": lambda: g.schur_recover_dev(nx, nu, N, batch, Qi, Ri, dd["A"], dd["B"], dd["q"], dd["r"], lam, x, u),
}
tot = 0.0
for name, fn in steps.items():
    ts = []
    for i in range(8):
        torch.cuda._sleep(200000)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        if i >= 3:
            ts.append(a.elapsed_time(b))
    ms = float(np.median(ts)); tot += ms
    print(f"{name:18s} batch={batch}: {ms * 1e3:9.1f} us", flush=True)
print(f"whole step: {tot * 1e3:.1f} us -> {batch / tot / 1e3:.2f} M MPC steps/s; PCG iterations mean {it.float().mean().item():.2f}")
