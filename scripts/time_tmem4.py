"""A/B of the batched fp32 kernels on one shape list: register-row (GBDPCG_FLAG_NO_TMEM), tensor memory with two rows per thread
(GBDPCG_FLAG_TMEM | GBDPCG_FLAG_TMEM_R2) and with four rows per thread (GBDPCG_FLAG_TMEM; GBDPCG_TMEM4_VARIANT picks where a
thread's quads live).  Every kernel is held to the host oracle on the distinct seeds (iterations within +-1, lambda <= 1e-5 at
equal count); launch time = median of 8 after 3 warm-ups; a max_iter sweep splits fixed from per-iteration cost.
usage: python scripts/time_tmem4.py [batch] [shapes: all|c5|small]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gbd_pcg_b200 as g  # noqa: E402
from oracle import oracle  # noqa: E402

F = g._lib
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
which = sys.argv[2] if len(sys.argv) > 2 else "all"
nseed = 16
SHAPES = {"c5": [(14, 7, 32)],
          "all": [(14, 7, 32), (12, 4, 40), (16, 8, 30), (16, 8, 32), (8, 4, 60), (10, 5, 17), (12, 6, 20), (14, 7, 16), (8, 4, 16)],
          "small": [(10, 5, 17), (8, 4, 30), (12, 6, 20), (14, 7, 16), (16, 8, 16), (8, 4, 16)]}[which]
VARIANTS = {8: [0, 10], 10: [0, 10], 12: [0, 1, 10], 14: [0, 1, 2, 3, 10], 16: [0, 1]}


def run(nx, N, d_S, d_P, d_g, flags, max_iter=1000, reps=11):
    d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    d_ex = torch.zeros(batch, dtype=torch.uint8, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=max_iter, empty_pinv=0, flags=flags)
    ts = []
    for i in range(reps):
        d_l.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
        b.record()
        torch.cuda.synchronize()
        if i >= 3:
            ts.append(a.elapsed_time(b))
    return float(np.median(ts)), d_l.cpu().numpy(), d_it.cpu().numpy(), d_ex.cpu().numpy()


for (nx, nu, N) in SHAPES:
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", np.float32)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(batch // nseed, 1) for a in (S, P, gam))
    ora = [oracle.pcg(S[s], P[s], gam[s], np.zeros(N * nx, np.float32), nx, N, 1e-6, 1000) for s in range(nseed)]
    arms = [("auto", 0, None), ("rows", F.FLAG_NO_TMEM, None), ("tmem_r2", F.FLAG_TMEM | F.FLAG_TMEM_R2, None)]
    arms += [(f"tmem4_v{v}", F.FLAG_TMEM, v) for v in VARIANTS.get(nx, [])]
    if os.environ.get("TMEM4_ARMS"):
        arms = [a for a in arms if a[0] in os.environ["TMEM4_ARMS"].split(",")]
    for name, flags, var in arms:
        os.environ.pop("GBDPCG_TMEM4_VARIANT", None)
        if var is not None:
            os.environ["GBDPCG_TMEM4_VARIANT"] = str(var)
        try:
            kern = g.plan_query(nx, N, np.float32, True, g.pcg_config(flags=flags), batch=batch)
            t, lam, it, ex = run(nx, N, d_S, d_P, d_g, flags)
        except Exception as e:  # noqa: BLE001
            print(f"nx={nx} N={N} {name}: {str(e)[:100]}", flush=True)
            continue
        worst, dit = 0.0, 0
        for s in range(nseed):
            lo, ito, _ = ora[s]
            dit = max(dit, abs(int(it[s]) - ito))
            if int(it[s]) == ito:
                worst = max(worst, float(np.linalg.norm(lam[s].astype(np.float64) - lo) / np.linalg.norm(lo)))
        reps_same = all(np.array_equal(lam[s::nseed], np.broadcast_to(lam[s], lam[s::nseed].shape)) for s in range(nseed))
        t0 = run(nx, N, d_S, d_P, d_g, flags, max_iter=0, reps=8)[0]
        t2 = run(nx, N, d_S, d_P, d_g, flags, max_iter=2, reps=8)[0]
        k = float(it.mean())
        ok = (dit <= 1 and worst < 1e-5 and reps_same and not ex.any())
        print(f"nx={nx} N={N} batch={batch} {name:9s}: {t:.4f} ms ({batch / t / 1e3:6.2f} M/s) iters {k:.2f} | load-only {t0:.4f} ms, per-iteration "
              f"{(t - t2) / max(k - 2, 1) * 1e3:.2f} us | oracle: d_iters {dit} rel {worst:.1e} replicas {reps_same} {'OK' if ok else 'FAIL'} | "
              f"{kern['kernel']} thr {kern['threads']} regs {kern['regs']} res {kern['max_resident']}", flush=True)
