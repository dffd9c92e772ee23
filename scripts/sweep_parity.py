"""Randomised parity sweep (GPU vs host oracle) over shapes / batch sizes / kinds for the builder, the Schur formation and the solver.
usage: python scripts/sweep_parity.py [n_cases] [seed]"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import gbd_pcg_b200 as g
from oracle import oracle
ncase = int(sys.argv[1]) if len(sys.argv) > 1 else 60
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
mr = lambda a, b: float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / max(1e-300, np.abs(np.asarray(b, np.float64)).max()))
for case in range(ncase):
    nx = int(rng.choice([2, 4, 6, 8, 10, 12, 14, 16, 3, 5, 9]))
    nu = int(rng.integers(1, nx + 1))
    N = int(rng.choice([2, 3, 5, 8, 16, 31, 32, 33, 47, 64, 100, 130]))
    B = int(rng.choice([1, 2, 3, 7, 40, 150, 300]))
    dt = rng.choice([np.float32, np.float64])
    kind = str(rng.choice(["jacobi", "stair"]))
    bar = 5e-5 if dt == np.float32 else 1e-10
    # stage data: random SPD costs, contractive dynamics
    spd = lambda n: (lambda M: M @ M.T / n + np.eye(n))(rng.standard_normal((n, n)))
    cm = lambda Ms: np.stack([np.ascontiguousarray(M.T).reshape(-1) for M in Ms]).astype(dt)
    nb = min(B, 3)
    ds = []
    for _ in range(nb):
        ds.append({"Q": cm([spd(nx) for _ in range(N)]), "R": cm([spd(nu) for _ in range(N - 1)]),
                   "A": cm([0.7 * np.eye(nx) + 0.1 * rng.standard_normal((nx, nx)) for _ in range(N - 1)]),
                   "B": cm([0.3 * rng.standard_normal((nx, nu)) for _ in range(N - 1)]),
                   "q": rng.standard_normal((N, nx)).astype(dt), "r": rng.standard_normal((N - 1, nu)).astype(dt), "c": rng.standard_normal((N, nx)).astype(dt)})
    d = {k: np.concatenate([ds[i % nb][k] for i in range(B)]) for k in ds[0]}
    dd = {k: torch.from_numpy(v).cuda() for k, v in d.items()}
    tdt = dd["Q"].dtype
    S = torch.empty(B * 3 * N * nx * nx, dtype=tdt, device="cuda"); P = torch.full_like(S, float("nan"))
    gam = torch.empty(B * N * nx, dtype=tdt, device="cuda"); lam = torch.zeros_like(gam)
    Qi = torch.empty(B * N * nx * nx, dtype=tdt, device="cuda"); Ri = torch.empty(B * (N - 1) * nu * nu, dtype=tdt, device="cuda")
    it = torch.zeros(B, dtype=torch.int32, device="cuda"); ex = torch.zeros(B, dtype=torch.uint8, device="cuda")
    tol = 1e-6 if dt == np.float32 else 1e-11
    try:
        g.schur_form_dev(nx, nu, N, B, dd["Q"], dd["R"], dd["A"], dd["B"], dd["q"], dd["r"], dd["c"], S, gam, Qi, Ri)
        g.build_pinv_dev(nx, N, B, S, P, kind)
        g.solve_batched_dev(nx, N, B, S, P, gam, lam, g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=3000, empty_pinv=0), it, ex)
        torch.cuda.synchronize()
    except Exception as e:
        print(f"case {case}: nx={nx} nu={nu} N={N} B={B} {np.dtype(dt).name} {kind}: EXCEPTION {e}"); bad += 1; continue
    Sh, Ph, gh, lh = (x.cpu().numpy().reshape(B, -1) for x in (S, P, gam, lam))
    errs = []
    for i in range(nb):
        fl = tuple(ds[i][k].reshape(-1) for k in "QRABqrc")
        So, go, _, _ = oracle.schur_form(*fl, nx, nu, N)
        e1, e2 = mr(Sh[i], So), mr(gh[i], go)
        Po = oracle.build_pinv(Sh[i], nx, N, kind)          # from the GPU's own S: isolates the builder
        e3 = mr(Ph[i], Po)
        lo, ito, exo = oracle.pcg(Sh[i], Ph[i], gh[i], np.zeros(N * nx, dt), nx, N, tol, 3000)
        ok_it = abs(int(it[i].item()) - ito) <= 1
        A = g.layout.unpack_dense(Sh[i].astype(np.float64), nx, N)
        xs = np.linalg.solve(A, gh[i].astype(np.float64))
        e4 = float(np.linalg.norm(lh[i] - xs) / np.linalg.norm(xs))
        # the same for the oracle's own lambda: an ill-conditioned tiny system leaves BOTH this far from the dense solution at the exit test
        e4o = float(np.linalg.norm(np.asarray(lo, np.float64) - xs) / np.linalg.norm(xs))
        # GPU against the oracle at equal iteration count (the north-star bar; tiny systems: 3x the CPU-vs-CPU floor, tests' noise_floor)
        k_ = int(it[i].item())
        lk = lo if k_ == ito else oracle.pcg(Sh[i], Ph[i], gh[i], np.zeros(N * nx, dt), nx, N, 0.0, k_)[0]
        e5 = float(np.linalg.norm(lh[i].astype(np.float64) - lk) / np.linalg.norm(lk))
        l1 = oracle.pcg(Sh[i], Ph[i], gh[i], np.zeros(N * nx, dt), nx, N, 0.0, k_, order=1)[0]
        floor5 = 3 * float(np.linalg.norm(np.asarray(l1, np.float64) - lk) / np.linalg.norm(lk)) if N * nx <= 200 else 0.0
        e4 = max(0.0, e4 - 3 * e4o)          # judged net of what the oracle itself leaves
        e5 = max(0.0, e5 - floor5)
        errs.append((e1, e2, e3, e4, ok_it and e5 <= (1e-5 if dt == np.float32 else 1e-10), int(it[i].item()), ito))
    same = all(np.array_equal(lh[i], lh[i % nb]) for i in range(B))
    worst = [max(e[j] for e in errs) for j in range(4)]
    fail = worst[0] > bar or worst[1] > bar or worst[2] > bar * 2 or worst[3] > (5e-3 if dt == np.float32 else 1e-6) or not all(e[4] for e in errs) or not same or int(ex.sum().item()) != 0
    bad += fail
    kern = g.plan_query(nx, N, dt, True, g.pcg_config(empty_pinv=0), batch=B)["kernel"].split("<")[0] + ("/P_TMEM" if "P_TMEM" in g.plan_query(nx, N, dt, True, g.pcg_config(empty_pinv=0), batch=B)["kernel"] else "")
    print(f"case {case:3d}: nx={nx:2d} nu={nu:2d} N={N:3d} B={B:3d} {np.dtype(dt).name:7s} {kind:6s} {kern:18s} S {worst[0]:.1e} gamma {worst[1]:.1e} Pinv {worst[2]:.1e} lambda {worst[3]:.1e} iters {[(e[5], e[6]) for e in errs]} same={same} {'FAIL' if fail else 'ok'}", flush=True)
print("FAILED" if bad else "ALL OK", bad)
