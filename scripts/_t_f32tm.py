import os, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import gbd_pcg_b200 as g
from oracle import oracle
def p50(launch, d_l, n=120):
    for _ in range(15): d_l.zero_(); launch()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in evs:
        torch.cuda._sleep(100000); d_l.zero_(); a.record(); launch(); b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) * 1e3 for a, b in evs)
    return ts[len(ts) // 2]
print("== single solves (p50 us)")
for (nx, nu, N) in [(12, 4, 512), (12, 4, 100), (8, 4, 200), (16, 8, 100), (12, 6, 256), (8, 4, 512), (16, 8, 64)]:
    S, P, gam = g.synth.lqr_schur(nx, nu, N, 0, "stair", np.float32)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (S, P, gam))
    d_l = torch.zeros(N * nx, dtype=torch.float32, device="cuda"); d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
    lo, ito, _ = oracle.pcg(S, P, gam, np.zeros(N * nx, np.float32), nx, N, 1e-6, 2000)
    out = []
    for cl, fl in [(0, 0), (2, 0x400), (4, 0x400), (8, 0x400), (16, 0x400)]:
        cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=2000, empty_pinv=0, flags=fl, cluster=cl)
        try:
            info = g.plan_query(nx, N, np.float32, True, cfg)
        except g.GbdpcgError:
            continue
        t = p50(lambda: g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None), d_l)
        it = int(d_it.item()); lam = d_l.cpu().numpy()
        lk = lo if it == ito else oracle.pcg(S, P, gam, np.zeros(N * nx, np.float32), nx, N, 0.0, it)[0]
        err = float(np.linalg.norm(lam - lk) / np.linalg.norm(lk))
        out.append(f"{'planner' if fl == 0 else 'TM'} cl{info['cluster']} thr{info['threads']}: {t:.1f} (it {it}/{ito} err {err:.0e})")
    print(f"nx={nx} N={N} | " + " | ".join(out), flush=True)
print("== batches of 2048 (ms)")
for (nx, nu, N) in [(12, 4, 100), (8, 4, 200), (16, 8, 100), (12, 4, 512), (8, 4, 150), (12, 6, 128)]:
    batch, nseed = 2048, 16
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", np.float32)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(batch // nseed, 1) for a in (S, P, gam))
    ref = None; out = []
    for name, fl in (("regs", 0x800), ("TM", 0x400)):
        cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=2000, empty_pinv=0, flags=fl)
        try:
            info = g.plan_query(nx, N, np.float32, True, cfg, batch=batch)
        except g.GbdpcgError as e:
            out.append(f"{name}: unsupported"); continue
        d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda"); d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
        ts = []
        for i in range(6):
            d_l.zero_(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None); b.record(); torch.cuda.synchronize()
            if i >= 2: ts.append(a.elapsed_time(b))
        l = d_l.cpu().numpy(); it = d_it.cpu().numpy()
        if ref is None: ref = (l, it)
        out.append(f"{name} {info['kernel'].split(',')[3] if 'P_TMEM' in info['kernel'] else ''} cl{info['cluster']} thr{info['threads']} res{info['max_resident']}: {np.median(ts):.3f} ms d_it {int(np.abs(it - ref[1]).max())} rel {np.linalg.norm(l - ref[0]) / np.linalg.norm(ref[0]):.0e}")
    print(f"nx={nx} N={N} | " + " | ".join(out), flush=True)
