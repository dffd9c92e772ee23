"""Times one config under several (cluster, flags) choices.  usage: python scripts/bench_variants.py C4 stair "16:0" "8:0" "16:0x10" ..."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import gbd_pcg_b200 as g
from helpers import make_system
name, pc = sys.argv[1], sys.argv[2]
S, P, gam, nx, N, tol, max_iter = make_system(name, pc)
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
d_S, d_P, d_g = d(S), (d(P) if P is not None else None), d(gam)
d_l = torch.zeros(N * nx, dtype=d_S.dtype, device="cuda"); d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
for spec in sys.argv[3:]:
    cl, fl = spec.split(":")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0 if P is not None else 1, cluster=int(cl), flags=int(fl, 0))
    try:
        plan = g.plan_query(nx, N, S.dtype, P is not None, cfg)
        launch = lambda: g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None)
        for _ in range(20): d_l.zero_(); launch()
        torch.cuda.synchronize()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(200)]
        for a, b in evs:
            torch.cuda._sleep(100000); d_l.zero_(); a.record(); launch(); b.record()
        torch.cuda.synchronize()
        ts = sorted(a.elapsed_time(b) * 1e3 for a, b in evs)
        print(f"{name} {pc} cluster={cl} flags={fl}: p50 {ts[100]:.1f} us, iters {int(d_it.item())}, {ts[100]/max(1,int(d_it.item())):.2f} us/it, {plan['kernel']} threads {plan['threads']} regs {plan['regs']} smem {plan['smem_bytes']}", flush=True)
    except Exception as e:
        print(f"{name} {pc} cluster={cl} flags={fl}: {e}")
