"""Systems beyond one 16-CTA cluster: the cooperative-grid kernel (pcg_grid.cuh) next to the unmodified reference kernel on the same
system -- kernel-scope p50 and microseconds per iteration.   usage: python scripts/time_grid.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gbd_pcg_b200 as g  # noqa: E402
from oracle import ref_cuda  # noqa: E402


def p50(fn, reset, n=60, warm=10):
    for _ in range(warm):
        reset(); fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        torch.cuda._sleep(200000)
        reset()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return float(np.median(ts))


for (nx, nu, N, dt, tol) in [(14, 7, 1500, np.float32, 1e-6), (14, 7, 1000, np.float64, 1e-10), (14, 7, 832, np.float32, 1e-6)]:
    for pc in ("stair", "jacobi"):
        S, P, gam = g.synth.lqr_schur(nx, nu, N, 1, pc, dt)
        d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (S, P, gam))
        d_l = torch.zeros(N * nx, dtype=d_S.dtype, device="cuda")
        d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
        cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=5000, empty_pinv=0)
        info = g.plan_query(nx, N, dt, True, cfg)
        t = p50(lambda: g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None), d_l.zero_)
        it = int(d_it.item())
        line = (f"nx={nx} N={N} {np.dtype(dt).name} {pc:6s}: {info['kernel']} G={info['cluster']} x {info['threads']} thr: {t:8.1f} us, {it} iterations, "
                f"{t / it:6.2f} us/iteration")
        if ref_cuda.has(dt, nx, N):
            rk = ref_cuda.get(dt, nx, N)
            tr = p50(lambda: rk.launch(d_S, d_P, d_g, d_l, 5000, tol), d_l.zero_)
            itr = int(rk.d_iters.item())
            line += f" | reference pcg<> {tr:8.1f} us, {itr} iterations, {tr / itr:6.2f} us/iteration -> x{tr / t:.2f}"
        print(line, flush=True)

# Between the register-row cluster's limit and the tensor-memory cluster's: the planner's choice (16-CTA cluster, Phi^-1 in tensor
# memory) next to what GBDPCG_FLAG_NO_TMEM leaves (the cooperative grid).
print("--- tensor-memory cluster form before the grid ---", flush=True)
for (nx, nu, N, dt, tol) in [(14, 7, 800, np.float32, 1e-6), (14, 7, 900, np.float32, 1e-6), (14, 7, 1024, np.float32, 1e-6), (12, 4, 1200, np.float32, 1e-6),
                             (14, 7, 400, np.float64, 1e-10), (14, 7, 576, np.float64, 1e-10), (12, 4, 672, np.float64, 1e-10), (8, 4, 1024, np.float64, 1e-10)]:
    S, P, gam = g.synth.lqr_schur(nx, nu, N, 1, "stair", dt)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (S, P, gam))
    d_l = torch.zeros(N * nx, dtype=d_S.dtype, device="cuda")
    d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
    line = f"nx={nx} N={N} {np.dtype(dt).name}:"
    for flags in (0, g._lib.FLAG_NO_TMEM):
        cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=5000, empty_pinv=0, flags=flags)
        info = g.plan_query(nx, N, dt, True, cfg)
        t = p50(lambda: g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None), d_l.zero_)
        it = int(d_it.item())
        line += f" {info['kernel']} G={info['cluster']} x {info['threads']}: {t:7.1f} us / {it} it = {t / it:5.2f} us/it |"
    print(line, flush=True)
