"""Parity of the CUDA path (through the C ABI) against the host oracle, the reference CUDA kernel
and the committed golden fixtures.  Everything here needs a GPU."""
import os

import numpy as np
import pytest

import gbd_pcg_b200 as g
from conftest import GOLDEN, LAMBDA_TOL, rel_err
from helpers import random_block_tridiagonal, PRECONDS, make_system, oracle_at_iters
from oracle import oracle

pytestmark = pytest.mark.gpu


def _torch():
    import torch
    return torch


def gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter, **cfg_kw):
    """One solve through the device-pointer overload; returns (lambda, iters, max_iter_exit)."""
    torch = _torch()
    d = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_S, d_P, d_g, d_l = d(S), d(P), d(gam), d(np.array(lam0, dtype=S.dtype))
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, **cfg_kw)
    before = g.kernel_launches()
    it, ex = g.api._solve_dev(nx, N, d_S, d_P, d_g, d_l, config=cfg, return_flag=True)
    assert g.kernel_launches() == before + 1, "the solve did not go through the CUDA kernel"
    return d_l.cpu().numpy(), it, ex


def check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex, noise_floor=False):
    """North-star bar: iterations within +-1, lambda within 1e-5 (fp32) / 1e-10 (fp64) at equal iteration count.
    noise_floor=True (small / unpreconditioned systems only): CG run to finite termination amplifies rounding, so two
    CPU statements of the same recurrence (plain vs reference reduction order vs dense numpy) already differ by more
    than the bar on some of these shapes; the bar is then widened to 3x that measured CPU-vs-CPU floor."""
    lam_o, it_o, ex_o = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=0)
    assert abs(it - it_o) <= 1, (it, it_o)
    if it == it_o:
        assert ex == ex_o
    lam_eq = lam_o if it == it_o else oracle_at_iters(S, P, gam, lam0, nx, N, it)
    err = rel_err(lam, lam_eq)
    bar = LAMBDA_TOL[S.dtype]
    if noise_floor:
        lam_1, it_1, _ = oracle.pcg(S, P, gam, lam0, nx, N, 0.0, it, order=1)
        lam_n, _, _ = oracle.pcg_numpy(S, P, gam, lam0, nx, N, 0.0, it)
        bar = max(bar, 3 * rel_err(lam_1, lam_eq), 3 * rel_err(lam_n, lam_eq))
    assert err <= bar, (err, bar, it, it_o)
    return err


@pytest.mark.parametrize("precond", PRECONDS)
@pytest.mark.parametrize("name", ["C1", "C2", "C3", "C4"])
def test_baseline_configs_vs_oracle(name, precond):
    S, P, gam, nx, N, tol, max_iter = make_system(name, precond)
    if name == "C1":
        # C1's iteration count is legitimately order-sensitive in fp32 (cond 1.6e3, 6 unknowns):
        # judge the fp32 run on the solution and the fp64 run on the hard KAT below.
        lam, it, ex = gpu_solve(S, P, gam, np.zeros(N * nx), nx, N, tol, max_iter)
        A = g.layout.unpack_dense(S.astype(np.float64), nx, N)
        x = np.linalg.solve(A, gam.astype(np.float64))
        assert not ex and rel_err(lam, x) < 2e-4
        return
    lam0 = np.zeros(N * nx, dtype=S.dtype)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    assert not ex
    check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex)


@pytest.mark.parametrize("precond,iters", [("identity", 6), ("jacobi", 5), ("stair", 3)])
def test_c1_fp64_known_answer(precond, iters):
    """SURVEY App. E hard KAT: the reference example's system in fp64."""
    S, P, gam, nx, N, tol, max_iter = make_system("C1", precond, dtype=np.float64)
    lam, it, ex = gpu_solve(S, P, gam, np.zeros(6), nx, N, 1e-6, 25)
    lam_star = np.array([-303.7029860857, -46.4159396812, -315.1763026319, -14.8983094180, -298.7908619196, 13.5037826883])
    assert it == iters and not ex
    assert rel_err(lam, lam_star) < 1e-9


@pytest.mark.parametrize("shape", [
    (2, 1, 2, np.float32, 1e-6, 100), (2, 1, 2, np.float64, 1e-12, 100),      # minimum N
    (4, 2, 8, np.float32, 1e-6, 200), (4, 2, 8, np.float64, 1e-10, 200),
    (6, 3, 16, np.float32, 1e-6, 300), (8, 4, 33, np.float32, 1e-6, 300),      # ragged: rows not a multiple of 32
    (10, 5, 20, np.float64, 1e-10, 500), (16, 8, 24, np.float32, 1e-6, 500),
    (14, 7, 40, np.float32, 1e-6, 500),                                          # one CTA, Phi^-1 kept in shared memory
    (14, 7, 96, np.float32, 1e-6, 500), (12, 4, 100, np.float64, 1e-10, 800),   # clusters, N not a multiple of the cluster
])
@pytest.mark.parametrize("precond", ["identity", "stair"])
def test_other_shapes_vs_oracle(shape, precond):
    S, P, gam, nx, N, tol, max_iter = make_system(shape, precond, seed=3)
    lam0 = np.zeros(N * nx, dtype=S.dtype)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    assert not ex
    check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex, noise_floor=True)


EXTRA_SHAPES = [(3, 3), (3, 700), (5, 12), (7, 30), (9, 33), (11, 100), (13, 64), (15, 64), (15, 250), (17, 20), (18, 32), (19, 40), (20, 32),
                (21, 16), (22, 50), (23, 33), (24, 64), (25, 24), (26, 20), (27, 48), (28, 100), (29, 10), (30, 32), (31, 17), (32, 32)]


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("precond", ["identity", "jacobi", "stair"])
@pytest.mark.parametrize("nx,N", EXTRA_SHAPES)
def test_odd_and_large_state_sizes_vs_oracle(nx, N, precond, dt):
    """State sizes beyond the even 2..16 set: odd ones (one row per thread, no bulk copies: a block-row is not a multiple
    of 16 bytes) and 18..32 (Phi^-1 resident in shared memory, clusters) run the specialised kernel too."""
    if nx % 2:
        S, P, gam = random_block_tridiagonal(nx, N, precond, dt, seed=nx + N, coupled=True)
    else:
        S, P, gam = g.synth.lqr_schur(nx, nx // 2, N, 2, precond, dt)
        P = None if precond == "identity" else P
    tol, max_iter = (1e-6 if dt == np.float32 else 1e-10), 3000
    info = g.plan_query(nx, N, dt, P is not None)
    assert info["specialised"] == (0 if (nx > 25 and dt == np.float64) else 1), info
    lam0 = np.zeros(N * nx, dtype=dt)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    assert not ex
    check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex, noise_floor=True)


@pytest.mark.parametrize("shape", [(3, 1, 3, np.float32, 1e-6, 100), (5, 2, 12, np.float64, 1e-10, 300), (33, 0, 20, np.float32, 1e-6, 300),
                                   (7, 3, 30, np.float32, 1e-6, 300), (14, 7, 32, np.float32, 1e-6, 1000), (35, 0, 12, np.float64, 1e-10, 300)])
@pytest.mark.parametrize("precond", ["identity", "stair"])
def test_generic_fallback_vs_oracle(shape, precond):
    """Runtime-nx kernel (state sizes without an instantiation, e.g. 33, take it by themselves): forced here by flag."""
    nx, nu, N, dt, tol, max_iter = shape
    if nx % 2:
        S, P, gam = random_block_tridiagonal(nx, N, precond, dt)
    else:
        S, P, gam, nx, N, tol, max_iter = make_system(shape, precond)
    flags = g._lib.FLAG_FORCE_GENERIC
    assert g.plan_query(nx, N, dt, P is not None, g.pcg_config(flags=flags))["specialised"] == 0
    lam0 = np.zeros(N * nx, dtype=dt)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter, flags=flags)
    assert not ex
    check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex, noise_floor=True)


@pytest.mark.parametrize("cluster", [1, 2, 4, 8, 16])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_cluster_sizes_agree(cluster, dtype):
    """The same system solved by 1..16 CTAs: identical iteration count, lambda equal to rounding."""
    shape = (6, 3, 48, dtype, 1e-6 if dtype == np.float32 else 1e-10, 500)
    S, P, gam, nx, N, tol, max_iter = make_system(shape, "stair", seed=5)
    lam0 = np.zeros(N * nx, dtype=dtype)
    assert g.plan_query(nx, N, dtype, True, g.pcg_config(cluster=cluster))["cluster"] == cluster
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter, cluster=cluster)
    assert not ex
    check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex)


@pytest.mark.parametrize("flag", ["reg", "smem"])
def test_pinv_placement_variants_agree(flag):
    S, P, gam, nx, N, tol, max_iter = make_system("C2", "stair")
    flags = g._lib.FLAG_PINV_REG if flag == "reg" else g._lib.FLAG_PINV_SMEM
    lam0 = np.zeros(N * nx, dtype=S.dtype)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter, flags=flags)
    check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex)


def test_c4_cluster16_registers_and_cluster8_smem():
    S, P, gam, nx, N, tol, max_iter = make_system("C4", "stair")
    lam0 = np.zeros(N * nx, dtype=S.dtype)
    for cl in (8, 16):
        lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter, cluster=cl)
        check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex)


def _largest_plannable_N(nx, dt):
    lo, hi = 2, 1 << 18
    while lo < hi:
        mid = (lo + hi + 1) // 2
        try:
            g.plan_query(nx, mid, dt, True)
            lo = mid
        except g.GbdpcgError:
            hi = mid - 1
    return lo


@pytest.mark.parametrize("nx,nu,dt", [(14, 7, np.float32), (14, 7, np.float64), (16, 8, np.float32), (2, 1, np.float32), (8, 4, np.float64)])
def test_maximum_knot_points(nx, nu, dt):
    """The largest system the planner accepts for this state size (a cooperative grid of register-row CTAs, one per SM): solved,
    judged by its residual (the dense cross-checks of check_against_oracle do not scale to it) and against the oracle's iteration
    count; one knot point more is refused with GBDPCG_ERR_UNSUPPORTED, never mis-solved."""
    N = _largest_plannable_N(nx, dt)
    assert N >= 1024                                   # beyond what one 16-CTA cluster holds (1024 for nx=14 fp32, Phi^-1 in tensor memory)
    tol = 1e-6 if dt == np.float32 else 1e-10
    S, P, gam, nx, N, tol, max_iter = make_system((nx, nu, N, dt, tol, 4000), "stair", seed=5)
    lam0 = np.zeros(N * nx, dtype=dt)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    assert not ex
    lam_o, it_o, _ = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=0)
    assert abs(it - it_o) <= 1
    if it == it_o:
        assert rel_err(lam, lam_o) <= 3 * LAMBDA_TOL[np.dtype(dt)]
    assert rel_err(oracle.bdmv(S, lam, nx, N), gam) < (5e-3 if dt == np.float32 else 1e-6)
    S2, P2, gam2, _, N2, _, _ = make_system((nx, nu, N + 1, dt, tol, 4000), "stair", seed=5)
    with pytest.raises(g.GbdpcgError) as e:
        gpu_solve(S2, P2, gam2, np.zeros(N2 * nx, dtype=dt), nx, N2, tol, max_iter)
    assert e.value.status == g._lib.ERR_UNSUPPORTED


GRID_CASES = [(14, 7, 1500, np.float32), (14, 7, 1000, np.float64), (12, 4, 1200, np.float32), (14, 7, 833, np.float32), (8, 4, 2500, np.float64),
              (16, 8, 1000, np.float32), (6, 3, 6000, np.float32)]


@pytest.mark.parametrize("nx,nu,N,dt", GRID_CASES)
@pytest.mark.parametrize("precond", PRECONDS)
def test_systems_beyond_one_cluster(nx, nu, N, dt, precond):
    """Systems no 16-CTA cluster of register rows can hold (the reference solves them on its cooperative grid, include/pcg.cuh:23-49):
    the planner takes the cooperative-grid kernel (pcg_grid.cuh), or first the tensor-memory cluster form where that reaches; iterations within +-1 of the oracle, lambda at the bar at equal count."""
    tol = 1e-6 if dt == np.float32 else 1e-10
    S, P, gam, nx, N, tol, max_iter = make_system((nx, nu, N, dt, tol, 5000), precond, seed=3)
    info = g.plan_query(nx, N, dt, P is not None, g.pcg_config(empty_pinv=0 if P is not None else 1))
    assert info["cluster"] * info["block_rows"] >= N
    if P is not None and (nx, N) in ((14, 833), (12, 1200)):
        # past the register-row cluster (nx 14 fp32: N <= 752) a 16-CTA cluster with Phi^-1 in tensor memory still holds the system
        assert "P_TMEM" in info["kernel"] and info["cluster"] == 16
    elif P is not None or N * nx > 20000:         # (without Phi^-1 a 16-CTA cluster reaches further: S alone in registers)
        assert "pcg_grid" in info["kernel"] and info["max_resident"] == 1
    lam0 = np.zeros(N * nx, dtype=dt)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    assert not ex
    lam_o, it_o, _ = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=0)
    assert abs(it - it_o) <= 1, (it, it_o)
    lam_eq = lam_o if it == it_o else oracle.pcg(S, P, gam, lam0, nx, N, 0.0, it, order=0)[0]
    assert rel_err(lam, lam_eq) <= LAMBDA_TOL[np.dtype(dt)], (rel_err(lam, lam_eq), it, it_o)
    # deterministic: the exit decision is taken from bit-identical sums in every CTA
    lam2, it2, _ = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    assert it2 == it and np.array_equal(lam2, lam)


@pytest.mark.parametrize("scale,nx,nu,N,dt", [(1, 14, 7, 1031, np.float32), (1, 14, 7, 1032, np.float32), (1, 14, 7, 1033, np.float32),
                                               (1, 10, 5, 820, np.float64), (1, 10, 5, 821, np.float64), (1, 8, 4, 3004, np.float32),
                                               (2, 14, 7, 1041, np.float32), (2, 14, 7, 1042, np.float32), (2, 14, 7, 1043, np.float32),
                                               (2, 10, 5, 827, np.float64), (2, 10, 5, 828, np.float64), (2, 8, 4, 3025, np.float32)])
@pytest.mark.parametrize("precond", ["stair", "jacobi"])
def test_one_barrier_grid_kernel_short_last_cta(scale, nx, nu, N, dt, precond, monkeypatch):
    """pcg_grid1_kernel (one grid barrier per iteration: ghost rows of Phi^-1, ghosts of r and s two block-rows deep) where the last CTA
    owns only 1, 2 or 3 block-rows -- the second-level ghost of its left neighbour then does not exist or is the system's last row:
    against the oracle at the suite's bar, and bit for bit against the two-barrier form (GBDPCG_GRID_TWO_BARRIERS=1) where that runs the
    same decomposition: scale 1 = one CTA per SM (GBDPCG_GRID_BPC_SCALE=1); scale 2 = the planner's default for this form, half as many
    CTAs of twice the size, which the two-barrier form does not take."""
    monkeypatch.setenv("GBDPCG_GRID_BPC_SCALE", str(scale))
    tol = 1e-6 if dt == np.float32 else 1e-10
    S, P, gam, nx, N, tol, max_iter = make_system((nx, nu, N, dt, tol, 5000), precond, seed=4)
    cfg = g.pcg_config(empty_pinv=0 if P is not None else 1)
    info = g.plan_query(nx, N, dt, P is not None, cfg)
    assert "pcg_grid1" in info["kernel"] and N - (info["cluster"] - 1) * info["block_rows"] in (1, 2, 3)
    lam0 = np.random.default_rng(1).standard_normal(N * nx).astype(dt) * 0.01
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    assert not ex
    lam_o, it_o, _ = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=0)
    assert abs(it - it_o) <= 1, (it, it_o)
    lam_eq = lam_o if it == it_o else oracle.pcg(S, P, gam, lam0, nx, N, 0.0, it, order=0)[0]
    assert rel_err(lam, lam_eq) <= LAMBDA_TOL[np.dtype(dt)], (rel_err(lam, lam_eq), it, it_o)
    monkeypatch.setenv("GBDPCG_GRID_TWO_BARRIERS", "1")
    assert "pcg_grid<" in g.plan_query(nx, N, dt, P is not None, cfg)["kernel"]
    lam2, it2, ex2 = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    assert not ex2 and abs(it2 - it) <= (0 if scale == 1 else 1)
    if scale == 1:
        assert np.array_equal(lam2, lam)


@pytest.mark.parametrize("nx,nu,N,dt", [(14, 7, 1500, np.float32), (16, 8, 1000, np.float64), (10, 5, 2000, np.float64)])
def test_one_barrier_grid_kernel_pinv_in_shared_memory(nx, nu, N, dt, monkeypatch):
    """pcg_grid1_kernel with Phi^-1 in shared memory (GBDPCG_FLAG_PINV_SMEM; the planner's own choice where two matrices do not fit the
    register file, e.g. fp64 nx = 16): the ghost rows read their neighbour's block-row from the CTA's own image.  Oracle parity and
    bit-identity with the two-barrier form."""
    monkeypatch.setenv("GBDPCG_GRID_BPC_SCALE", "1")       # (the two-barrier form keeps one CTA per SM: compare the same decomposition)
    tol = 1e-6 if dt == np.float32 else 1e-10
    S, P, gam, nx, N, tol, max_iter = make_system((nx, nu, N, dt, tol, 5000), "stair", seed=6)
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0, flags=g._lib.FLAG_PINV_SMEM)
    lam0 = np.zeros(N * nx, dtype=dt)

    def run():
        torch = _torch()
        d_S, d_P, d_g, d_l = (torch.from_numpy(a).cuda() for a in (S, P, gam, lam0.copy()))
        d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
        g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None)
        torch.cuda.synchronize()
        return d_l.cpu().numpy(), int(d_it.item())

    k1 = g.plan_query(nx, N, dt, True, cfg)["kernel"]
    assert "pcg_grid1" in k1 and "P_SMEM" in k1
    lam, it = run()
    lam_o, it_o, _ = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=0)
    assert abs(it - it_o) <= 1
    lam_eq = lam_o if it == it_o else oracle.pcg(S, P, gam, lam0, nx, N, 0.0, it, order=0)[0]
    assert rel_err(lam, lam_eq) <= LAMBDA_TOL[np.dtype(dt)]
    monkeypatch.setenv("GBDPCG_GRID_TWO_BARRIERS", "1")
    k2 = g.plan_query(nx, N, dt, True, cfg)["kernel"]
    assert "pcg_grid<" in k2 and "P_SMEM" in k2
    lam2, it2 = run()
    assert it2 == it and np.array_equal(lam2, lam)


def test_one_barrier_grid_kernel_deterministic_from_two_streams():
    """The exchange buffers of pcg_grid1_kernel alternate with the iteration's parity and every launch has its own scratch from the pool:
    the same system solved 40 times, every other round from two streams at once, gives the same bits every time
    (scripts/stress_grid.py is the long form)."""
    torch = _torch()
    nx, nu, N, dt, tol = 14, 7, 1041, np.float32, 1e-6
    S, P, gam = g.synth.lqr_schur(nx, nu, N, 3, "stair", dt)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (S, P, gam))
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=5000, empty_pinv=0)
    assert "pcg_grid1" in g.plan_query(nx, N, dt, True, cfg)["kernel"]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = [(torch.zeros(N * nx, dtype=torch.float32, device="cuda"), torch.zeros(1, dtype=torch.int32, device="cuda")) for _ in streams]
    ref = None
    for r in range(40):
        n = 2 if r % 2 else 1
        for (d_l, d_it), st in list(zip(outs, streams))[:n]:
            with torch.cuda.stream(st):
                d_l.zero_()
                g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None)
        torch.cuda.synchronize()
        for d_l, d_it in outs[:n]:
            cur = (d_l.cpu().numpy(), int(d_it.item()))
            if ref is None:
                ref = cur
            assert cur[1] == ref[1] and np.array_equal(cur[0], ref[0])
    assert ref[1] > 5


def test_grid_kernel_semantics_and_batch():
    """max_iter = 0 leaves lambda; a low cap reports the flag; warm start; a batch of large systems (solved one after the other on
    the same cooperative grid) equals the single solves; NaN-poisoned padding blocks never reach the arithmetic."""
    torch = _torch()
    nx, nu, N, dt, tol = 14, 7, 1100, np.float32, 1e-6
    Ss, Ps, gs_ = g.synth.lqr_schur_batch(nx, nu, N, range(3), "stair", dt)
    for M in (Ss, Ps):
        M[:, : nx * nx] = np.nan
        M[:, (N - 1) * 3 * nx * nx + 2 * nx * nx:] = np.nan
    rng = np.random.default_rng(2)
    lam0 = rng.standard_normal((3, N * nx)).astype(dt)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (Ss, Ps, gs_))
    for mi, want_ex in ((0, 1), (3, 1), (4000, 0)):
        d_l = torch.from_numpy(lam0.copy()).cuda()
        d_it = torch.full((3,), 77, dtype=torch.int32, device="cuda")
        d_ex = torch.full((3,), 7, dtype=torch.uint8, device="cuda")
        cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=mi, empty_pinv=0)
        assert "pcg_grid" in g.plan_query(nx, N, dt, True, cfg, batch=3)["kernel"]
        g.solve_batched_dev(nx, N, 3, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
        torch.cuda.synchronize()
        lam, it, ex = d_l.cpu().numpy(), d_it.cpu().numpy(), d_ex.cpu().numpy()
        assert (ex == want_ex).all()
        if mi == 0:
            assert np.array_equal(lam, lam0) and (it == 0).all()
        elif mi == 3:
            assert (it == 3).all() and np.isfinite(lam).all()
        else:
            Sc, Pc = Ss.copy(), Ps.copy()
            for M in (Sc, Pc):
                M[np.isnan(M)] = 0
            for i in range(3):
                check_against_oracle(Sc[i], Pc[i], gs_[i], lam0[i], nx, N, tol, 4000, lam[i], int(it[i]), bool(ex[i]))
                l1, it1, _ = gpu_solve(Sc[i], Pc[i], gs_[i], lam0[i], nx, N, tol, 4000)
                assert it1 == it[i] and np.array_equal(l1, lam[i])


# ---------------------------------------------------------------------------------------------
# semantics the reference defines
# ---------------------------------------------------------------------------------------------
def test_max_iter_zero_leaves_lambda():
    S, P, gam, nx, N, tol, _ = make_system("C2", "stair")
    lam0 = np.linspace(-1, 1, N * nx).astype(np.float32)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, 0)
    assert it == 0 and ex and np.array_equal(lam, lam0)


def test_max_iter_cap_reports_flag():
    S, P, gam, nx, N, tol, _ = make_system("C2", "identity")
    lam0 = np.zeros(N * nx, dtype=np.float32)
    lam, it, ex = gpu_solve(S, None, gam, lam0, nx, N, tol, 7)
    assert it == 7 and ex
    assert rel_err(lam, oracle_at_iters(S, None, gam, lam0, nx, N, 7)) < 1e-5


def test_warm_start():
    S, P, gam, nx, N, tol, max_iter = make_system("C2", "stair")
    zero = np.zeros(N * nx, dtype=np.float32)
    lam_cold, it_cold, _ = gpu_solve(S, P, gam, zero, nx, N, tol, max_iter)
    rng = np.random.default_rng(1)
    lam0 = (lam_cold * (1 + 1e-3 * rng.standard_normal(N * nx))).astype(np.float32)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    assert it < it_cold
    check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex)


def test_padding_blocks_are_never_read():
    """The left block of row 0 and right block of row N-1 are padding (pcg.cuh:105-106): NaNs there must not matter."""
    S, P, gam, nx, N, tol, max_iter = make_system("C2", "stair")
    lam0 = np.zeros(N * nx, dtype=np.float32)
    lam_a, it_a, _ = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    S2, P2 = S.copy(), P.copy()
    for M in (S2, P2):
        M[:nx * nx] = np.nan
        M[-nx * nx:] = np.nan
    lam_b, it_b, _ = gpu_solve(S2, P2, gam, lam0, nx, N, tol, max_iter)
    assert it_a == it_b and np.array_equal(lam_a, lam_b)


def test_nan_never_converges():
    S, P, gam, nx, N, tol, _ = make_system("C2", "stair")
    gam = gam.copy()
    gam[5] = np.nan
    lam, it, ex = gpu_solve(S, P, gam, np.zeros(N * nx, dtype=np.float32), nx, N, tol, 12)
    assert it == 12 and ex and np.isnan(lam).any()


def test_empty_pinv_and_null_pinv_mean_identity():
    torch = _torch()
    S, P, gam, nx, N, tol, max_iter = make_system("C2", "stair")
    lam0 = np.zeros(N * nx, dtype=np.float32)
    ref, it_ref, _ = gpu_solve(S, None, gam, lam0, nx, N, tol, max_iter)
    # host overload with empty_pinv=1 ignores h_Pinv (the C ABI's documented rule)
    lam = lam0.copy()
    it = g.solvePCG(S, gam, lam, nx, N, g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=1), h_Pinv=P)
    assert it == it_ref and np.array_equal(lam, ref)
    del torch


def test_host_overload_matches_device_overload():
    for name, pc in (("C2", "stair"), ("C3", "jacobi")):
        S, P, gam, nx, N, tol, max_iter = make_system(name, pc)
        lam0 = np.zeros(N * nx, dtype=S.dtype)
        lam_d, it_d, _ = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
        lam_h = lam0.copy()
        it_h = g.pcg_solve(S, gam, lam_h, nx, N, g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0), h_Pinv=P)
        assert it_h == it_d and np.array_equal(lam_h, lam_d)


def test_invalid_arguments():
    S, P, gam, nx, N, tol, max_iter = make_system("C1", "identity")
    lam = np.zeros(6, dtype=np.float32)
    with pytest.raises(g.GbdpcgError):
        g.solvePCG(S, gam, lam, 2, 1, g.pcg_config())          # N < 2
    with pytest.raises(g.GbdpcgError):
        g.solvePCG(S, gam, lam, 0, 3, g.pcg_config())          # nx == 0
    with pytest.raises(ValueError):
        g.solvePCG(S, gam, lam[:3], 2, 3, g.pcg_config())      # short lambda


# ---------------------------------------------------------------------------------------------
# batched mode
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape,batch", [((14, 7, 32, np.float32, 1e-6, 1000), 333), ((2, 1, 3, np.float32, 1e-6, 50), 1000),
                                         ((14, 7, 64, np.float64, 1e-10, 2000), 40), ((6, 3, 16, np.float32, 1e-6, 300), 500)])
def test_batched_equals_single_and_oracle(shape, batch):
    torch = _torch()
    nx, nu, N, dt, tol, max_iter = shape
    nseed = 12
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", dt)
    reps = (batch + nseed - 1) // nseed
    Sb, Pb, gb = (np.tile(a, (reps, 1))[:batch].copy() for a in (S, P, gam))
    d = lambda a: torch.from_numpy(a).cuda()
    d_S, d_P, d_g = d(Sb), d(Pb), d(gb)
    d_l = torch.zeros(batch, N * nx, dtype=d_S.dtype, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    d_ex = torch.ones(batch, dtype=torch.uint8, device="cuda")
    # the register-row kernel (what a single solve runs too): batched == single, bit for bit.  (The planner's own choice for a
    # batch that fills the tensor-memory slots is the four-rows-per-thread kernel, whose accumulation order differs: that one is
    # held to the oracle in test_tmem_kernel_vs_oracle_and_register_kernel and, below, to the single solves at the parity bar.)
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0, flags=g._lib.FLAG_NO_TMEM)
    g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
    torch.cuda.synchronize()
    lam, its, exs = d_l.cpu().numpy(), d_it.cpu().numpy(), d_ex.cpu().numpy()
    assert not exs.any()
    d_l2 = torch.zeros_like(d_l)
    d_it2 = torch.zeros_like(d_it)
    g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l2, g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0), d_it2, d_ex)
    torch.cuda.synchronize()
    lam_auto, its_auto = d_l2.cpu().numpy(), d_it2.cpu().numpy()
    assert np.abs(its_auto.astype(int) - its.astype(int)).max() <= 1
    same = its_auto == its
    assert np.abs(lam_auto[same] - lam[same]).max() <= 1e-5 * np.abs(lam).max()
    # a lone system may be planned on a larger cluster than a batch (latency mode): same cluster for a bit-exact comparison
    cl = g.plan_query(nx, N, dt, True, cfg, batch=batch)["cluster"]
    for s in range(nseed):
        lam1, it1, _ = gpu_solve(S[s], P[s], gam[s], np.zeros(N * nx, dtype=dt), nx, N, tol, max_iter, cluster=cl)
        idx = np.arange(s, batch, nseed)
        assert (its[idx] == it1).all()
        assert all(np.array_equal(lam[i], lam1) for i in idx), "batched result differs from the single solve"
        check_against_oracle(S[s], P[s], gam[s], np.zeros(N * nx, dtype=dt), nx, N, tol, max_iter, lam1, it1, False)


@pytest.mark.parametrize("nx,N,dt,batch", [(7, 30, np.float32, 400), (9, 33, np.float64, 150), (15, 64, np.float32, 90), (20, 32, np.float32, 200),
                                           (24, 40, np.float64, 60)])
def test_batched_odd_and_large_state_sizes(nx, N, dt, batch):
    """The persistent batch loop on the paths the even 2..16 sizes never take: cooperative staging (odd nx: no bulk copies),
    clusters with Phi^-1 resident in shared memory (nx > 16)."""
    torch = _torch()
    nseed = 5
    tol, max_iter = (1e-6 if dt == np.float32 else 1e-10), 3000
    if nx % 2:
        sys_ = [random_block_tridiagonal(nx, N, "stair", dt, seed=100 + s, coupled=True) for s in range(nseed)]
        S, P, gam = (np.stack([t[i] for t in sys_]) for i in range(3))
    else:
        S, P, gam = g.synth.lqr_schur_batch(nx, nx // 2, N, range(nseed), "stair", dt)
    reps = (batch + nseed - 1) // nseed
    d = lambda a: torch.from_numpy(np.tile(a, (reps, 1))[:batch].copy()).cuda()
    d_S, d_P, d_g = d(S), d(P), d(gam)
    d_l = torch.zeros(batch, N * nx, dtype=d_S.dtype, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    d_ex = torch.ones(batch, dtype=torch.uint8, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0)
    plan = g.plan_query(nx, N, dt, True, cfg, batch=batch)
    assert plan["specialised"] == 1
    g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
    torch.cuda.synchronize()
    lam, its, exs = d_l.cpu().numpy(), d_it.cpu().numpy(), d_ex.cpu().numpy()
    assert not exs.any()
    for s in range(nseed):
        lam1, it1, ex1 = gpu_solve(S[s], P[s], gam[s], np.zeros(N * nx, dtype=dt), nx, N, tol, max_iter, cluster=plan["cluster"])
        idx = np.arange(s, batch, nseed)
        assert (its[idx] == it1).all()
        assert all(np.array_equal(lam[i], lam1) for i in idx), "batched result differs from the single solve"
        check_against_oracle(S[s], P[s], gam[s], np.zeros(N * nx, dtype=dt), nx, N, tol, max_iter, lam1, it1, ex1, noise_floor=True)


@pytest.mark.parametrize("dt,cluster", [(np.float32, 16), (np.float32, 8), (np.float64, 16)])
def test_batched_ragged_cluster_stays_in_step(dt, cluster):
    """A persistent ragged cluster over a batch (N = 100 on 16 CTAs: 14 full, one short, one EMPTY; on 8: the last one short):
    the push-mode exchanges carry their phase bits from one system to the next, systems of different iteration counts
    alternate, and every result equals the single solve."""
    torch = _torch()
    nx, nu, N, batch, nseed = 12, 4, 100, 130, 4
    tol, max_iter = (1e-6 if dt == np.float32 else 1e-10), 3000
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", dt)
    Pj = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "jacobi", dt)[1]
    P[1::2] = Pj[1::2]                                        # odd seeds: block-Jacobi -> about twice the iterations
    reps = (batch + nseed - 1) // nseed
    d = lambda a: torch.from_numpy(np.tile(a, (reps, 1))[:batch].copy()).cuda()
    d_S, d_P, d_g = d(S), d(P), d(gam)
    d_l = torch.zeros(batch, N * nx, dtype=d_S.dtype, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    d_ex = torch.ones(batch, dtype=torch.uint8, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0, cluster=cluster)
    assert g.plan_query(nx, N, dt, True, cfg)["cluster"] == cluster
    g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
    torch.cuda.synchronize()
    lam, its, exs = d_l.cpu().numpy(), d_it.cpu().numpy(), d_ex.cpu().numpy()
    assert not exs.any()
    for s in range(nseed):
        lam1, it1, ex1 = gpu_solve(S[s], P[s], gam[s], np.zeros(N * nx, dtype=dt), nx, N, tol, max_iter, cluster=cluster)
        idx = np.arange(s, batch, nseed)
        assert (its[idx] == it1).all()
        assert all(np.array_equal(lam[i], lam1) for i in idx), "batched result differs from the single solve"
        check_against_oracle(S[s], P[s], gam[s], np.zeros(N * nx, dtype=dt), nx, N, tol, max_iter, lam1, it1, ex1, noise_floor=True)
    assert its[0] != its[1]


@pytest.mark.parametrize("nx,nu,N,dt,cluster,batch", [(14, 7, 64, np.float64, 0, 1), (12, 4, 100, np.float32, 16, 40), (20, 10, 32, np.float32, 0, 64)])
def test_clustered_solver_is_deterministic(nx, nu, N, dt, cluster, batch):
    """Push-mode exchanges: a message that arrived early, late or in the wrong phase would change a reduction or a halo row.
    300 repeats of the same launch (several clusters in flight when batch > 1) must be bit-identical (scripts/stress_cluster.py
    is the long form)."""
    torch = _torch()
    nseed = min(batch, 4)
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", dt)
    k = (batch + nseed - 1) // nseed
    d = lambda a: torch.from_numpy(np.tile(a, (k, 1))[:batch].copy()).cuda()
    d_S, d_P, d_g = d(S), d(P), d(gam)
    d_l = torch.zeros(batch, N * nx, dtype=d_S.dtype, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=1e-6 if dt == np.float32 else 1e-10, pcg_max_iter=3000, empty_pinv=0, cluster=cluster)
    assert g.plan_query(nx, N, dt, True, cfg, batch=batch)["cluster"] > 1
    ref = None
    for r in range(300):
        d_l.zero_()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None)
        if ref is None:
            ref = (d_l.clone(), d_it.clone())
        else:
            assert torch.equal(d_l, ref[0]) and torch.equal(d_it, ref[1]), f"run {r} differs"
    assert int(d_it.min().item()) > 5


def test_batched_host_entry_point():
    nx, nu, N, dt, tol, max_iter = 14, 7, 32, np.float32, 1e-6, 1000
    batch = 700
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(10), "stair", dt)
    Sb, Pb, gb = (np.tile(a, (70, 1)).copy() for a in (S, P, gam))
    lam = np.zeros((batch, N * nx), dtype=dt)
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0)
    its, exs = g.solve_batched_host(Sb, Pb, gb, lam, nx, N, batch, cfg)
    assert not exs.any()
    for s in range(10):
        lo, ito, _ = oracle.pcg(S[s], P[s], gam[s], np.zeros(N * nx, dtype=dt), nx, N, tol, max_iter)
        for i in range(s, batch, 10):
            assert abs(int(its[i]) - ito) <= 1
            if its[i] == ito:
                assert rel_err(lam[i], lo) < 1e-5


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_symmetric_s_shipping(dt):
    """GBDPCG_FLAG_SYMMETRIC_S: only the diagonal and right blocks of S cross PCIe, the left blocks are rebuilt on the device by
    transposition.  With NaN in every left block of the host array the result must equal the full-S result bit for bit (the
    generator's S is exactly symmetric), on the single and the batched host entry points, with a device-built preconditioner."""
    nx, nu, N, batch = 14, 7, 32, 300
    tol, max_iter = (1e-6, 1000) if dt == np.float32 else (1e-10, 2000)
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(6), "stair", dt)
    Sb, gb = (np.tile(a, (batch // 6, 1)).copy() for a in (S, gam))
    for i in range(1, N):      # exact symmetry of the generator: left(k) == right(k-1)^T
        L = S[0, i * 3 * nx * nx: i * 3 * nx * nx + nx * nx].reshape(nx, nx)
        Rp = S[0, (i - 1) * 3 * nx * nx + 2 * nx * nx: i * 3 * nx * nx].reshape(nx, nx)
        assert np.array_equal(L, Rp.T)
    S_poison = Sb.copy().reshape(batch, N, 3, nx * nx)
    S_poison[:, :, 0, :] = np.nan
    S_poison = S_poison.reshape(batch, -1)
    full = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, flags=g._lib.FLAG_BUILD_STAIR)
    sym = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, flags=g._lib.FLAG_BUILD_STAIR | g._lib.FLAG_SYMMETRIC_S)
    lam_f, lam_s = np.zeros((batch, N * nx), dt), np.zeros((batch, N * nx), dt)
    it_f, ex_f = g.solve_batched_host(Sb, None, gb, lam_f, nx, N, batch, full)
    it_s, ex_s = g.solve_batched_host(S_poison, None, gb, lam_s, nx, N, batch, sym)
    assert not ex_f.any() and not ex_s.any() and np.array_equal(it_f, it_s) and np.array_equal(lam_f, lam_s)
    check_against_oracle(S[0], P[0], gam[0], np.zeros(N * nx, dt), nx, N, tol, max_iter, lam_s[0], int(it_s[0]), False)
    l1, l2 = np.zeros(N * nx, dt), np.zeros(N * nx, dt)         # single-system host entry point: flag on poisoned S == no flag on S
    it1 = g.solvePCG(S_poison[1].copy(), gam[1], l1, nx, N, sym)
    it2 = g.solvePCG(Sb[1].copy(), gam[1], l2, nx, N, full)
    assert it1 == it2 and np.array_equal(l1, l2)


def test_fp64_batch_at_bench_size():
    """The fp64 leg of bench.py (other_configs.C3_batched): 2048 x C3 (nx 14, N 64, fp64, tol 1e-10, stair), 64 distinct seeds tiled;
    every distinct system against the oracle at the fp64 bar, replicas bit-identical wherever they ran."""
    torch = _torch()
    nx, nu, N, dt, tol, max_iter = g.synth.CONFIGS["C3"]
    batch, nseed = 2048, 64
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", dt)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(batch // nseed, 1) for a in (S, P, gam))
    d_l = torch.zeros(batch, N * nx, dtype=torch.float64, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    d_ex = torch.ones(batch, dtype=torch.uint8, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0)
    g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
    torch.cuda.synchronize()
    assert int(d_ex.sum().item()) == 0
    lam, its = d_l.cpu().numpy(), d_it.cpu().numpy()
    for s_ in range(nseed):
        block = lam[s_::nseed]
        assert (block == block[0]).all() and (its[s_::nseed] == its[s_]).all()
        check_against_oracle(S[s_], P[s_], gam[s_], np.zeros(N * nx, dt), nx, N, tol, max_iter, lam[s_], int(its[s_]), False)


def test_full_size_batch_properties():
    """BASELINE C5 size (8192 x C2) through size-independent properties: the residual of every returned
    lambda satisfies the exit criterion's neighbourhood and S*lambda reproduces gamma (checked by the oracle's
    block-row mat-vec on a sample, and by a device-side norm for all)."""
    torch = _torch()
    nx, nu, N, dt, tol, max_iter = 14, 7, 32, np.float32, 1e-6, 1000
    batch, nseed = 8192, 64
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", dt)
    d = lambda a: torch.from_numpy(a).cuda()
    d_S, d_P, d_g = (d(a).repeat(batch // nseed, 1) for a in (S, P, gam))
    d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
    d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
    d_ex = torch.ones(batch, dtype=torch.uint8, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0)
    g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
    torch.cuda.synchronize()
    assert int(d_ex.sum().item()) == 0
    its = d_it.cpu().numpy()
    assert its.min() >= 20 and its.max() <= 32
    lam = d_l.cpu().numpy()
    # replicas of the same seed are bit-identical wherever they ran
    for s in range(nseed):
        block = lam[s::nseed]
        assert (block == block[0]).all()
    # S * lambda ~= gamma
    for s in range(0, nseed, 7):
        y = oracle.bdmv(S[s], lam[s], nx, N)
        assert rel_err(y, gam[s]) < 5e-3


# ---------------------------------------------------------------------------------------------
# the reference CUDA kernel itself, when its prebuilt harness travelled with the repo
# ---------------------------------------------------------------------------------------------
REF_CASES = [("C1", np.float32), ("C1", np.float64), ("C2", np.float32), ("C3", np.float64), ("C4", np.float32)]


@pytest.mark.parametrize("name,dtype", REF_CASES)
@pytest.mark.parametrize("precond", PRECONDS)
def test_against_reference_cuda(name, dtype, precond):
    from oracle import ref_cuda
    S, P, gam, nx, N, tol, max_iter = make_system(name, precond, dtype=dtype)
    if not ref_cuda.has(dtype, nx, N):
        pytest.skip("reference harness not built for this config")
    P_ref = P if P is not None else g.layout.identity_compressed(nx, N, dtype)
    lam0 = np.zeros(N * nx, dtype=dtype)
    lam_r, it_r, ex_r = ref_cuda.get(dtype, nx, N).solve(S, P_ref, gam, lam0, max_iter, tol)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    if name == "C1" and dtype == np.float32:
        A = g.layout.unpack_dense(S.astype(np.float64), nx, N)
        x = np.linalg.solve(A, gam.astype(np.float64))
        assert rel_err(lam, x) < 2e-4 and rel_err(lam_r, x) < 2e-4
        return
    assert abs(it - it_r) <= 1, (it, it_r)
    if it == it_r:
        assert ex == ex_r
    else:
        # equal iteration count: both kernels again, capped at the smaller count (max_iter exit, the test can no longer fire first)
        k = min(it, it_r)
        lam_r, it_r2, _ = ref_cuda.get(dtype, nx, N).solve(S, P_ref, gam, lam0, k, tol)
        lam, it2, _ = gpu_solve(S, P, gam, lam0, nx, N, tol, k)
        assert it2 == k and it_r2 == k
    assert rel_err(lam, lam_r) <= LAMBDA_TOL[np.dtype(dtype)]


@pytest.mark.parametrize("nx,nu,N,dtype", [(14, 7, 1500, np.float32), (14, 7, 1000, np.float64)])
@pytest.mark.parametrize("precond", ["stair", "jacobi"])
def test_large_systems_against_reference_cuda(nx, nu, N, dtype, precond):
    """Beyond one cluster, next to the LIVE reference kernel (its cooperative grid of N CTAs is co-resident up to 12 x SMs knot
    points here): iterations within +-1, lambda at the bar at equal count."""
    from oracle import ref_cuda
    if not ref_cuda.has(dtype, nx, N):
        pytest.skip("reference harness not built for this config")
    tol = 1e-6 if dtype == np.float32 else 1e-10
    S, P, gam, nx, N, tol, max_iter = make_system((nx, nu, N, dtype, tol, 5000), precond, seed=1)
    lam0 = np.zeros(N * nx, dtype=dtype)
    lam_r, it_r, ex_r = ref_cuda.get(dtype, nx, N).solve(S, P, gam, lam0, max_iter, tol)
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
    assert "pcg_grid" in g.plan_query(nx, N, dtype, True)["kernel"]
    assert not ex and not ex_r and abs(it - it_r) <= 1, (it, it_r)
    if it != it_r:
        k = min(it, it_r)
        lam_r, _, _ = ref_cuda.get(dtype, nx, N).solve(S, P, gam, lam0, k, tol)
        lam, _, _ = gpu_solve(S, P, gam, lam0, nx, N, tol, k)
    assert rel_err(lam, lam_r) <= LAMBDA_TOL[np.dtype(dtype)]


def test_golden_reference_fixtures():
    """Outputs of the reference CUDA kernel recorded on a B200 (tests/golden/ref_cuda_*.npz)."""
    import glob
    files = sorted(glob.glob(os.path.join(GOLDEN, "ref_cuda_*.npz")))
    if not files:
        pytest.skip("no reference fixtures committed yet")
    for f in files:
        z = np.load(f)
        name, precond, dtype = str(z["config"]), str(z["precond"]), np.dtype(str(z["dtype"]))
        S, P, gam, nx, N, tol, max_iter = make_system(name, precond, dtype=dtype)
        lam0 = np.zeros(N * nx, dtype=dtype)
        lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter)
        if name == "C1" and dtype == np.float32:
            continue
        it_ref = int(z["iters"])
        assert abs(it - it_ref) <= 1
        if it > it_ref:        # equal iteration count: this kernel again, capped at the recorded count
            lam, it, _ = gpu_solve(S, P, gam, lam0, nx, N, tol, it_ref)
        elif it < it_ref:      # the recorded run went one further: the oracle (held to the same fixture on the CPU) bridges the step
            lam_o = oracle.pcg(S, P, gam, lam0, nx, N, 0.0, it, order=0)[0]
            assert rel_err(lam, lam_o) <= LAMBDA_TOL[dtype], f
            continue
        assert it == it_ref and rel_err(lam, z["lam"]) <= LAMBDA_TOL[dtype], f


# ---------------------------------------------------------------------------------------------
# the C++ header shim (include/gpu_pcg.cuh -> libgbdpcg.so): reference-style programs, run for real
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("example,iters", [("pcg_solve.cu", None), ("pcg_solve_dp.cu", 6)])
def test_cpp_examples_run(example, iters, tmp_path):
    import re
    import shutil
    import subprocess
    from conftest import ROOT
    if shutil.which("nvcc") is None:
        pytest.skip("needs nvcc")
    libdir = os.path.join(ROOT, "gbd_pcg_b200")
    exe = tmp_path / "a.out"
    subprocess.check_call(["nvcc", "-std=c++17", "-Wno-deprecated-gpu-targets", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", example), "-L", libdir, "-lgbdpcg", "-o", str(exe)])
    env = dict(os.environ, LD_LIBRARY_PATH=libdir + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run([str(exe)], env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"(\d+) iterations", r.stdout)
    assert m and (iters is None or int(m.group(1)) == iters), r.stdout


def test_cpp_api_surface(tmp_path):
    """examples/api_surface.cu, compiled and RUN: overload 3 with real scratch buffers (reference include/interface.cuh:92-103),
    pcg_solve<T> with the config defaulted (README.md:42), pcgSharedMemSize<T> / checkPcgOccupancy<T> (include/pcg.cuh:13-49), and
    two host threads with their own gbdpcg_workspace and stream solving different systems concurrently -- all against known answers."""
    import shutil
    import subprocess
    from conftest import ROOT
    if shutil.which("nvcc") is None:
        pytest.skip("needs nvcc")
    libdir = os.path.join(ROOT, "gbd_pcg_b200")
    exe = tmp_path / "api_surface"
    subprocess.check_call(["nvcc", "-std=c++17", "-O2", "-Wno-deprecated-gpu-targets", "-I", os.path.join(ROOT, "include"), "-I",
                           os.path.join(ROOT, "examples"), os.path.join(ROOT, "examples", "api_surface.cu"), "-L", libdir, "-lgbdpcg",
                           "-Xcompiler", "-pthread", "-o", str(exe)])
    env = dict(os.environ, LD_LIBRARY_PATH=libdir + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run([str(exe)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "all checks passed" in r.stdout, r.stdout + r.stderr
    assert "two threads, own workspace + stream each: 80 + 80 solves, failures 0 + 0" in r.stdout


def test_python_mirror_workspace_handle():
    """gbdpcg_workspace_create / _destroy through the Python mirror: two workspaces on two host threads, host and device entry points."""
    import threading
    torch = _torch()
    out = {}

    def work(tag, name, precond):
        S, P, gam, nx, N, tol, max_iter = make_system(name, precond)
        cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0)
        with g.workspace() as ws:
            res = []
            for _ in range(20):
                lam = np.zeros(N * nx, dtype=S.dtype)
                it = g.solvePCG(S, gam, lam, nx, N, cfg, h_Pinv=P, workspace=ws)
                res.append((it, lam.copy()))
            st = torch.cuda.Stream()
            d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (S, P, gam))
            d_l = torch.zeros(N * nx, dtype=d_S.dtype, device="cuda")
            torch.cuda.synchronize()
            it_d = g.api._solve_dev(nx, N, d_S, d_P, d_g, d_l, config=cfg, workspace=ws, stream=st)
            out[tag] = (res, it_d, d_l.cpu().numpy(), (S, P, gam, nx, N, tol, max_iter))

    ts = [threading.Thread(target=work, args=("a", "C2", "stair")), threading.Thread(target=work, args=("b", "C3", "jacobi"))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    for tag in ("a", "b"):
        res, it_d, lam_d, (S, P, gam, nx, N, tol, max_iter) = out[tag]
        lam0 = np.zeros(N * nx, dtype=S.dtype)
        for it, lam in res:
            assert it == res[0][0] and np.array_equal(lam, res[0][1])
        check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, res[0][1], res[0][0], False)
        assert it_d == res[0][0] and np.array_equal(lam_d, res[0][1])


# ---------------------------------------------------------------------------------------------
# kernel-selection flags: every recurrence / staging / generation variant solves the same systems
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("flag", ["TEXTBOOK", "FUSED", "NO_TMA", "LEGACY_ROWS"])
@pytest.mark.parametrize("name,precond", [("C2", "stair"), ("C2", "identity"), ("C3", "stair"), ("C3", "identity"), ("C4", "jacobi")])
def test_kernel_variant_flags(flag, name, precond):
    S, P, gam, nx, N, tol, max_iter = make_system(name, precond)
    lam0 = np.zeros(N * nx, dtype=S.dtype)
    flags = getattr(g._lib, "FLAG_" + flag)
    info = g.plan_query(nx, N, S.dtype, P is not None, g.pcg_config(flags=flags))
    assert ("pcg_rows" in info["kernel"]) == (flag == "LEGACY_ROWS")
    lam, it, ex = gpu_solve(S, P, gam, lam0, nx, N, tol, max_iter, flags=flags)
    assert not ex
    check_against_oracle(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex)


def test_unaligned_device_pointers_fall_back_from_tma():
    """S / Phi^-1 at addresses that are only 4-byte aligned: the bulk-copy path must not be used, results unchanged."""
    torch = _torch()
    S, P, gam, nx, N, tol, max_iter = make_system("C2", "stair")
    lam_a, it_a, _ = gpu_solve(S, P, gam, np.zeros(N * nx, np.float32), nx, N, tol, max_iter, flags=g._lib.FLAG_NO_TMEM)
    pad = lambda a: torch.from_numpy(np.concatenate([np.zeros(1, a.dtype), a])).cuda()[1:]
    d_S, d_P, d_g = pad(S), pad(P), pad(gam)
    assert d_S.data_ptr() % 16 == 4
    d_l = torch.zeros(N * nx, dtype=torch.float32, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter)
    it = g.api._solve_dev(nx, N, d_S, d_P, d_g, d_l, config=cfg)
    assert it == it_a and np.array_equal(d_l.cpu().numpy(), lam_a)


# ---------------------------------------------------------------------------------------------------------------
# tensor-memory batched kernel (pcg_tmem.cuh): two systems per SM, Phi^-1 and the right blocks of S in TMEM
# ---------------------------------------------------------------------------------------------------------------
def _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter, flags):
    torch = _torch()
    d_l = torch.from_numpy(np.ascontiguousarray(lam0)).cuda()
    d_it = torch.full((batch,), 77, dtype=torch.int32, device="cuda")
    d_ex = torch.full((batch,), 7, dtype=torch.uint8, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0, flags=flags)
    g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
    torch.cuda.synchronize()
    return d_l.cpu().numpy(), d_it.cpu().numpy(), d_ex.cpu().numpy()


TMEM_ARMS = {"r4": lambda: g._lib.FLAG_TMEM, "r2": lambda: g._lib.FLAG_TMEM | g._lib.FLAG_TMEM_R2}


@pytest.mark.parametrize("nx,nu,N,batch", [(14, 7, 32, 700), (14, 7, 32, 5), (12, 4, 40, 301), (16, 8, 30, 297), (8, 4, 60, 330), (10, 5, 17, 1000),
                                           (14, 7, 2, 64), (16, 8, 32, 40), (16, 8, 16, 1300), (8, 4, 30, 2000), (12, 6, 20, 777)])
@pytest.mark.parametrize("precond", ["stair", "jacobi"])
@pytest.mark.parametrize("arm", ["r4", "r2"])
def test_tmem_kernel_vs_oracle_and_register_kernel(nx, nu, N, batch, precond, arm):
    """GBDPCG_FLAG_TMEM on ragged batch sizes (fewer systems than CTAs, not a multiple of the grid), the smallest and the
    largest supported system: every system against the host oracle.  Arm r4 = four rows per thread (pcg_tmem4.cuh) where the
    system fits 128 such threads; arm r2 = two rows per thread (pcg_tmem.cuh), bit for bit equal to the register-row kernel
    where both reduce over the same number of warp slots."""
    torch = _torch()
    nseed = 6
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), precond, np.float32)
    reps = (batch + nseed - 1) // nseed
    Sb, Pb, gb = (np.tile(a, (reps, 1))[:batch].copy() for a in (S, P, gam))
    d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (Sb, Pb, gb))
    lam0 = np.zeros((batch, N * nx), np.float32)
    tol, max_iter = 1e-6, 1000
    flags = TMEM_ARMS[arm]()
    kern = g.plan_query(nx, N, np.float32, True, g.pcg_config(flags=flags), batch=batch)["kernel"]
    four = N * ((nx + 3) // 4) <= 128 and arm == "r4" and not (nx == 16 and N * 4 <= 64)     # (nx = 16 has no <= 128-column variant for 2-warp systems)
    assert ("pcg_tmem4" in kern) if four else ("pcg_tmem<" in kern)
    lam_t, it_t, ex_t = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter, flags)
    lam_r, it_r, ex_r = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter, g._lib.FLAG_NO_TMEM)
    assert not ex_t.any()
    assert np.abs(it_t.astype(int) - it_r.astype(int)).max() <= 1
    if not four and nx in (12, 14, 16) and N * nx // 2 > 128:
        assert np.array_equal(it_t, it_r) and np.array_equal(lam_t.view(np.uint32), lam_r.view(np.uint32))
    for s in range(min(nseed, batch)):
        idx = np.arange(s, batch, nseed)
        assert (it_t[idx] == it_t[s]).all() and all(np.array_equal(lam_t[i], lam_t[s]) for i in idx)
        check_against_oracle(S[s], P[s], gam[s], np.zeros(N * nx, np.float32), nx, N, tol, max_iter, lam_t[s], int(it_t[s]), bool(ex_t[s]))


@pytest.mark.parametrize("arm", ["r4", "r2"])
def test_tmem_kernel_edge_semantics(arm):
    """max_iter = 0 leaves lambda, a low cap reports the flag, warm starts, padding blocks never reach the arithmetic."""
    torch = _torch()
    fl = TMEM_ARMS[arm]()
    nx, nu, N, batch = 14, 7, 32, 300
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(4), "stair", np.float32)
    Sb, Pb, gb = (np.tile(a, (batch // 4, 1)).copy() for a in (S, P, gam))
    rng = np.random.default_rng(5)
    # poison the padding blocks (left of row 0, right of row N-1): the reference never reads them (include/pcg.cuh:105-106)
    for M in (Sb, Pb):
        M[:, : nx * nx] = np.nan
        M[:, (N - 1) * 3 * nx * nx + 2 * nx * nx:] = np.nan
    d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (Sb, Pb, gb))
    lam0 = rng.standard_normal((batch, N * nx)).astype(np.float32)
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, 1e-6, 0, fl)
    assert np.array_equal(lam, lam0) and (it == 0).all() and (ex == 1).all()
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, 1e-6, 3, fl)
    assert (it == 3).all() and (ex == 1).all() and np.isfinite(lam).all()
    lam_w, it_w, ex_w = _batched(nx, N, batch, d_S, d_P, d_g, lam0, 1e-6, 1000, fl)
    lam_r, it_r, ex_r = _batched(nx, N, batch, d_S, d_P, d_g, lam0, 1e-6, 1000, g._lib.FLAG_NO_TMEM)
    assert not ex_w.any() and np.abs(it_w.astype(int) - it_r.astype(int)).max() <= 1
    if arm == "r2":
        assert np.array_equal(it_w, it_r) and np.array_equal(lam_w.view(np.uint32), lam_r.view(np.uint32))
    for s in range(4):
        check_against_oracle(S[s], P[s], gam[s], lam0[s], nx, N, 1e-6, 1000, lam_w[s], int(it_w[s]), bool(ex_w[s]))


def test_tmem_kernel_selection_and_refusals():
    """The planner takes the tensor-memory kernel for fp32 batches that fill the GPU twice; it is refused (flag forced) or
    avoided (planner) for the identity preconditioner, odd / large state sizes and unaligned pointers."""
    torch = _torch()
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, empty_pinv=0)
    assert "pcg_tmem" in g.plan_query(14, 32, np.float32, True, cfg, batch=2 * sms)["kernel"]
    assert g.plan_query(14, 32, np.float32, True, cfg, batch=2 * sms)["max_resident"] == 2 * sms
    # systems of 3-4 four-row warps take it at EVERY batch size: two CTAs per SM above #SMs systems, the solo form (one per SM, S with a
    # staging image of its own) up to #SMs -- including the single solve
    for b_, res in ((2 * sms - 1, 2 * sms), (sms + 1, 2 * sms), (sms, sms), (1, sms)):
        info = g.plan_query(14, 32, np.float32, True, cfg, batch=b_)
        assert "pcg_tmem4" in info["kernel"] and info["max_resident"] == res, (b_, info)
    assert g.plan_query(14, 32, np.float32, True, cfg, batch=1)["smem_bytes"] > 2 * g.plan_query(14, 32, np.float32, True, cfg, batch=8192)["smem_bytes"] - 8192
    assert "pcg_tmem" not in g.plan_query(14, 32, np.float64, True, cfg, batch=8192)["kernel"]
    assert "pcg_tmem" not in g.plan_query(14, 32, np.float32, False, cfg, batch=8192)["kernel"]
    # the four-rows-per-thread kernel is the planner's choice where the system fits 128 such threads and an SM holds >= 8 warps of it
    assert "pcg_tmem4" in g.plan_query(14, 32, np.float32, True, cfg, batch=2 * sms)["kernel"]
    assert "pcg_tmem4" in g.plan_query(12, 40, np.float32, True, cfg, batch=8192)["kernel"]
    assert "pcg_tmem<" in g.plan_query(14, 32, np.float32, True, g.pcg_config(flags=g._lib.FLAG_TMEM_R2), batch=8192)["kernel"]
    # systems of <= 2 warps take at most 128 tensor-memory columns: four CTAs per SM, chosen once the batch fills them
    assert "pcg_tmem4" in g.plan_query(10, 17, np.float32, True, cfg, batch=4 * sms)["kernel"]
    assert g.plan_query(10, 17, np.float32, True, cfg, batch=4 * sms)["max_resident"] == 4 * sms
    assert "pcg_tmem" not in g.plan_query(10, 17, np.float32, True, cfg, batch=4 * sms - 1)["kernel"]
    assert "pcg_tmem" not in g.plan_query(8, 16, np.float32, True, cfg, batch=8192)["kernel"]       # one warp per system: the register-row kernel wins
    assert "pcg_tmem" not in g.plan_query(14, 32, np.float32, True, g.pcg_config(flags=g._lib.FLAG_NO_TMEM), batch=8192)["kernel"]
    # (fp64 + GBDPCG_FLAG_TMEM is the tensor-memory form of the cluster kernel: tests/test_tmem64_gpu.py)
    assert "P_TMEM" in g.plan_query(14, 32, np.float64, True, g.pcg_config(flags=g._lib.FLAG_TMEM), batch=8192)["kernel"]
    for (nx, N, dt, has_p) in [(14, 32, np.float64, False), (14, 32, np.float32, False), (15, 20, np.float32, True), (18, 20, np.float32, True),
                               (14, 1100, np.float32, True)]:      # (beyond 16 CTAs x 256 four-row threads; up to there: wide form / cluster form, tests/test_tmem4_wide_gpu.py, test_tmem64_gpu.py)
        with pytest.raises(g.GbdpcgError):
            g.plan_query(nx, N, dt, has_p, g.pcg_config(flags=g._lib.FLAG_TMEM), batch=8192)
    # unaligned S / Pinv: the planner's choice falls back to the register-row kernel, same results
    nx, nu, N, batch = 14, 7, 32, 2 * sms
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(4), "stair", np.float32)
    Sb, Pb, gb = (np.tile(a, (batch // 4, 1)).copy() for a in (S, P, gam))
    pad = lambda a: torch.from_numpy(np.concatenate([np.zeros(1, a.dtype), a.ravel()])).cuda()[1:]
    d_S, d_P, d_g = pad(Sb), pad(Pb), torch.from_numpy(gb).cuda()
    assert d_S.data_ptr() % 16 == 4
    lam0 = np.zeros((batch, N * nx), np.float32)
    lam_u, it_u, _ = _batched(nx, N, batch, d_S, d_P, d_g, lam0, 1e-6, 1000, 0)
    lam_a, it_a, _ = _batched(nx, N, batch, torch.from_numpy(Sb).cuda(), torch.from_numpy(Pb).cuda(), d_g, lam0, 1e-6, 1000, g._lib.FLAG_NO_TMEM)
    assert np.array_equal(it_u, it_a) and np.array_equal(lam_u, lam_a)        # = the register-row kernel, bit for bit
    with pytest.raises(g.GbdpcgError):
        _batched(nx, N, batch, d_S, d_P, d_g, lam0, 1e-6, 1000, g._lib.FLAG_TMEM)


@pytest.mark.parametrize("arm", ["r4", "r2"])
@pytest.mark.parametrize("nx,nu,N,batch", [(14, 7, 32, 2048), (16, 8, 32, 1500), (12, 4, 40, 1111)])
def test_tmem_kernels_are_deterministic_under_concurrent_launches(nx, nu, N, batch, arm):
    """The -m gpu form of scripts/stress_tmem.py (compute-sanitizer is closed on this pool, so repeat-and-compare is the race
    evidence): the same batch 50 times, every second time with a SECOND launch of the same kernel on another stream so that CTAs
    of two launches share SMs and the 512 tensor-memory columns (tcgen05.alloc of the late CTAs blocks until columns are free).
    Every lambda and iteration count must be bit-identical to the first run, in both launches."""
    torch = _torch()
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(16), "stair", np.float32)
    r = (batch + 15) // 16
    d_S, d_P, d_g = (torch.from_numpy(np.tile(a, (r, 1))[:batch].copy()).cuda() for a in (S, P, gam))
    cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, empty_pinv=0, flags=TMEM_ARMS[arm]())
    s2 = torch.cuda.Stream()
    ref = None
    for i in range(50):
        d_l = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
        d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
        d_l2 = torch.zeros(batch, N * nx, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, None)
        if i % 2:
            with torch.cuda.stream(s2):
                g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l2, cfg, None, None, stream=s2)
        torch.cuda.synchronize()
        if ref is None:
            ref = (d_l.clone(), d_it.clone())
            assert int(d_it.min().item()) > 5
        else:
            assert torch.equal(d_l, ref[0]) and torch.equal(d_it, ref[1]), f"run {i} differs"
            assert i % 2 == 0 or torch.equal(d_l2, ref[0]), f"run {i}: the concurrent launch differs"
