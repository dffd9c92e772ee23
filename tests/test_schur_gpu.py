"""Schur-complement formation and primal recovery on the GPU (gbdpcg_schur_form_dev_*, gbdpcg_schur_recover_dev_*; SURVEY
s8(f) rank 4) against the host oracle, and the whole resident chain form -> build Phi^-1 -> PCG -> recover against a dense
KKT solve."""
import numpy as np
import pytest

import gbd_pcg_b200 as g
from conftest import rel_err
from oracle import oracle

pytestmark = pytest.mark.gpu
BAR = {np.dtype(np.float32): 3e-5, np.dtype(np.float64): 1e-11}
FLAT = lambda d: tuple(d[k].reshape(-1) for k in "QRABqrc")


def _torch():
    import torch
    return torch


def to_dev(d):
    torch = _torch()
    return {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in d.items()}


def random_stage_data(nx, nu, N, dt, seed=11):
    """Any (nx, nu): random SPD Q_k, R_k, random A_k, B_k; blocks column-major (include/gbdpcg.h)."""
    rng = np.random.default_rng(seed)
    spd = lambda n: (lambda M: M @ M.T / n + np.eye(n))(rng.standard_normal((n, n)))
    cm = lambda Ms: np.stack([np.ascontiguousarray(M.T).reshape(-1) for M in Ms]).astype(dt)
    return {"Q": cm([spd(nx) for _ in range(N)]), "R": cm([spd(nu) for _ in range(N - 1)]),
            "A": cm([0.7 * np.eye(nx) + 0.1 * rng.standard_normal((nx, nx)) for _ in range(N - 1)]),
            "B": cm([0.3 * rng.standard_normal((nx, nu)) for _ in range(N - 1)]),
            "q": rng.standard_normal((N, nx)).astype(dt), "r": rng.standard_normal((N - 1, nu)).astype(dt),
            "c": rng.standard_normal((N, nx)).astype(dt)}


def max_rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / np.abs(b).max())


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("nx,nu,N,dense", [(2, 1, 3, False), (14, 7, 32, False), (14, 7, 32, True), (12, 4, 512, False), (5, 3, 9, True),
                                            (16, 16, 6, True), (7, 1, 4, True)])
def test_form_and_recover_vs_oracle(nx, nu, N, dense, dt):
    torch = _torch()
    if dense:
        d = random_stage_data(nx, nu, N, dt)
    else:
        d = g.synth.lqr_stage_data(nx, nu, N, 3, dtype=dt)
    dd = to_dev(d)
    S = torch.full((3 * N * nx * nx,), float("nan"), dtype=dd["Q"].dtype, device="cuda")
    gam = torch.full((N * nx,), float("nan"), dtype=S.dtype, device="cuda")
    Qi = torch.empty(N * nx * nx, dtype=S.dtype, device="cuda")
    Ri = torch.empty((N - 1) * nu * nu, dtype=S.dtype, device="cuda")
    before = g.kernel_launches()
    g.schur_form_dev(nx, nu, N, 1, dd["Q"], dd["R"], dd["A"], dd["B"], dd["q"], dd["r"], dd["c"], S, gam, Qi, Ri)
    torch.cuda.synchronize()
    assert g.kernel_launches() == before + 1
    So, go, Qio, Rio = oracle.schur_form(*FLAT(d), nx, nu, N)
    bar = BAR[np.dtype(dt)]
    assert max_rel(S.cpu().numpy(), So) < bar and max_rel(gam.cpu().numpy(), go) < bar
    assert max_rel(Qi.cpu().numpy(), Qio) < bar and max_rel(Ri.cpu().numpy(), Rio) < bar
    Sb = S.cpu().numpy().reshape(N, 3, nx * nx)
    assert not Sb[0, 0].any() and not Sb[N - 1, 2].any()
    lam = np.random.default_rng(0).standard_normal(N * nx).astype(dt)
    x = torch.empty(N * nx, dtype=S.dtype, device="cuda")
    u = torch.empty((N - 1) * nu, dtype=S.dtype, device="cuda")
    g.schur_recover_dev(nx, nu, N, 1, Qi, Ri, dd["A"], dd["B"], dd["q"], dd["r"], torch.from_numpy(lam).cuda(), x, u)
    xo, uo = oracle.schur_recover(Qio, Rio, d["A"].reshape(-1), d["B"].reshape(-1), d["q"].reshape(-1), d["r"].reshape(-1), lam, nx, nu, N)
    assert max_rel(x.cpu().numpy(), xo) < bar and max_rel(u.cpu().numpy(), uo) < bar


@pytest.mark.parametrize("dt,tol,zbar", [(np.float32, 1e-6, 5e-4), (np.float64, 1e-12, 1e-6)])
def test_resident_chain_against_the_kkt_solution(dt, tol, zbar):
    """The whole device-resident MPC step for a batch: stage data -> S, gamma -> stair Phi^-1 -> PCG -> (x, u)."""
    torch = _torch()
    nx, nu, N, B = 14, 7, 32, 40
    ds = [g.synth.lqr_stage_data(nx, nu, N, s, dtype=dt, dense_costs=True) for s in range(4)]
    d = {k: np.concatenate([ds[i % 4][k] for i in range(B)]) for k in ds[0]}
    dd = to_dev(d)
    tdt = dd["Q"].dtype
    S = torch.empty(B * 3 * N * nx * nx, dtype=tdt, device="cuda")
    P = torch.empty_like(S)
    gam = torch.empty(B * N * nx, dtype=tdt, device="cuda")
    lam = torch.zeros(B * N * nx, dtype=tdt, device="cuda")
    Qi = torch.empty(B * N * nx * nx, dtype=tdt, device="cuda")
    Ri = torch.empty(B * (N - 1) * nu * nu, dtype=tdt, device="cuda")
    x = torch.empty(B * N * nx, dtype=tdt, device="cuda")
    u = torch.empty(B * (N - 1) * nu, dtype=tdt, device="cuda")
    it = torch.zeros(B, dtype=torch.int32, device="cuda")
    ex = torch.zeros(B, dtype=torch.uint8, device="cuda")
    g.schur_form_dev(nx, nu, N, B, dd["Q"], dd["R"], dd["A"], dd["B"], dd["q"], dd["r"], dd["c"], S, gam, Qi, Ri)
    g.build_pinv_dev(nx, N, B, S, P, "stair")
    g.solve_batched_dev(nx, N, B, S, P, gam, lam, g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=2000, empty_pinv=0), it, ex)
    g.schur_recover_dev(nx, nu, N, B, Qi, Ri, dd["A"], dd["B"], dd["q"], dd["r"], lam, x, u)
    torch.cuda.synchronize()
    assert int(ex.sum().item()) == 0
    xs, us = x.cpu().numpy().reshape(B, N, nx), u.cpu().numpy().reshape(B, N - 1, nu)
    for i in range(4):
        G, C, gg, c = g.synth.kkt_dense(ds[i], nx, nu, N)
        K = np.block([[G, C.T], [C, np.zeros((N * nx, N * nx))]])
        z = np.linalg.solve(K, np.concatenate([-gg, -c]))[:G.shape[0]]
        zz = np.concatenate([np.concatenate([xs[i, k], us[i, k] if k + 1 < N else []]) for k in range(N)])
        assert rel_err(zz, z) < zbar
    for i in range(4, B):
        assert np.array_equal(xs[i], xs[i % 4]) and np.array_equal(us[i], us[i % 4])


def test_reference_example_matrix_from_its_stage_data():
    """GPU formation of the pendulum LQR the reference example decodes to reproduces examples/pcg_solve.cu:14-22 (fp32 and fp64)."""
    torch = _torch()
    from test_schur_cpu import example_stage_data
    for dt in (np.float32, np.float64):
        d, nx, nu, N = example_stage_data(dt)
        dd = to_dev(d)
        S = torch.empty(3 * N * nx * nx, dtype=dd["Q"].dtype, device="cuda")
        gam = torch.empty(N * nx, dtype=S.dtype, device="cuda")
        g.schur_form_dev(nx, nu, N, 1, dd["Q"], dd["R"], dd["A"], dd["B"], dd["q"], dd["r"], dd["c"], S, gam)
        torch.cuda.synchronize()
        assert np.abs(S.cpu().numpy().astype(np.float64) - g.synth.EXAMPLE_S).max() < 6e-5


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("nx,nu,N,batch", [(14, 7, 32, 40), (12, 4, 33, 7), (2, 1, 3, 5), (7, 3, 9, 3), (16, 8, 17, 2), (16, 16, 6, 3)])
def test_staged_recovery_equals_the_global_memory_form_bit_for_bit(nx, nu, N, batch, dt, monkeypatch):
    """The staged recovery kernel (matrices through per-warp shared-memory slices, next knot in flight; sizes at compile time where
    nu = nx/2) keeps the arithmetic and its order: x, u equal those of the global-memory warp kernel bit for bit
    (GBDPCG_RECOVER_GLOBAL=1 selects that one).  Odd pointers' alignment classes are all met: knot strides of 196 / 98 / 49 elements."""
    torch = _torch()
    rng = np.random.default_rng(5)
    mk = lambda *shape: torch.from_numpy(rng.standard_normal(shape).astype(dt)).cuda()
    Qi, Ri = mk(batch * N * nx * nx), mk(batch * (N - 1) * nu * nu)
    A, B = mk(batch * (N - 1) * nx * nx), mk(batch * (N - 1) * nx * nu)
    q, r, lam = mk(batch * N * nx), mk(batch * (N - 1) * nu), mk(batch * N * nx)
    outs = []
    for env in ("0", "1"):
        monkeypatch.setenv("GBDPCG_RECOVER_GLOBAL", env)
        x = torch.full((batch * N * nx,), float("nan"), dtype=Qi.dtype, device="cuda")
        u = torch.full((batch * (N - 1) * nu,), float("nan"), dtype=Qi.dtype, device="cuda")
        g.schur_recover_dev(nx, nu, N, batch, Qi, Ri, A, B, q, r, lam, x, u)
        torch.cuda.synchronize()
        outs.append((x.cpu().numpy(), u.cpu().numpy()))
    assert np.isfinite(outs[0][0]).all() and np.isfinite(outs[0][1]).all()
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
