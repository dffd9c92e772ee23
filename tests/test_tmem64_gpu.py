"""fp64 throughput form of the cluster kernel (pcg_fused.cuh, PMODE = P_TMEM): two rows per thread, the rows of Phi^-1 and the
last columns of the rows of S in tensor memory.  Forced with GBDPCG_FLAG_TMEM, and as the planner's own choice for fp64 batches
that cannot be resident at once.  Everything against the host oracle (north-star bar: iterations within +-1, lambda <= 1e-10
at equal iteration count) and against the register-row kernel it replaces."""
import os

import numpy as np
import pytest

import gbd_pcg_b200 as g
from conftest import LAMBDA_TOL, rel_err
from helpers import make_system, oracle_at_iters
from oracle import oracle

pytestmark = pytest.mark.gpu
F64 = np.float64


def _torch():
    import torch
    return torch


def _solve1(S, P, gam, lam0, nx, N, tol, max_iter, **kw):
    torch = _torch()
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_S, d_P, d_g, d_l = d(S), d(P), d(gam), d(np.array(lam0, dtype=S.dtype))
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, **kw)
    before = g.kernel_launches()
    it, ex = g.api._solve_dev(nx, N, d_S, d_P, d_g, d_l, config=cfg, return_flag=True)
    assert g.kernel_launches() == before + 1
    return d_l.cpu().numpy(), it, ex


def _check(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex, noise_floor=False):
    """The suite's bar (tests/test_parity_gpu.py::check_against_oracle).  noise_floor=True (small systems only): CG run to finite
    termination amplifies rounding, so two CPU statements of the same recurrence already differ by more than 1e-10 on some of these
    shapes; the bar is then 3x that measured CPU-vs-CPU floor."""
    lam_o, it_o, ex_o = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=0)
    assert abs(int(it) - it_o) <= 1, (it, it_o)
    if it == it_o:
        assert bool(ex) == bool(ex_o)
    lam_eq = lam_o if it == it_o else oracle_at_iters(S, P, gam, lam0, nx, N, int(it))
    err = rel_err(lam, lam_eq)
    bar = LAMBDA_TOL[np.dtype(F64)]
    if noise_floor:
        lam_1, _, _ = oracle.pcg(S, P, gam, lam0, nx, N, 0.0, int(it), order=1)
        lam_n, _, _ = oracle.pcg_numpy(S, P, gam, lam0, nx, N, 0.0, int(it))
        bar = max(bar, 3 * rel_err(lam_1, lam_eq), 3 * rel_err(lam_n, lam_eq))
    assert err <= bar, (err, bar, it, it_o)
    return err


SHAPES = [(14, 7, 64), (14, 7, 32), (14, 7, 65), (14, 7, 2), (14, 7, 3), (14, 7, 200), (12, 4, 100), (12, 6, 37), (8, 4, 200), (8, 4, 33),
          (10, 5, 20), (10, 5, 90), (16, 8, 24), (16, 8, 16)]


@pytest.mark.parametrize("nx,nu,N", SHAPES)
@pytest.mark.parametrize("precond", ["stair", "jacobi"])
def test_tmem64_single_solves_vs_oracle(nx, nu, N, precond):
    """One CTA and every cluster size the planner comes up with (N not a multiple of the cluster, the minimum system, ragged last CTA)."""
    S, P, gam, nx, N, tol, max_iter = make_system((nx, nu, N, F64, 1e-10, 3000), precond, seed=7)
    flags = g._lib.FLAG_TMEM
    info = g.plan_query(nx, N, F64, True, g.pcg_config(flags=flags, empty_pinv=0))
    assert "P_TMEM" in info["kernel"] and info["pinv_mode"] == 4 and info["cluster"] * info["block_rows"] >= N
    lam0 = np.zeros(N * nx)
    lam, it, ex = _solve1(S, P, gam, lam0, nx, N, tol, max_iter, flags=flags)
    assert not ex
    _check(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex, noise_floor=N * nx <= 64)
    lam2, it2, _ = _solve1(S, P, gam, lam0, nx, N, tol, max_iter, flags=flags)
    assert it2 == it and np.array_equal(lam2, lam)              # deterministic


@pytest.mark.parametrize("cluster", [1, 2, 4, 8, 16])
def test_tmem64_forced_cluster_sizes(cluster):
    shape = (12, 6, 40, F64, 1e-10, 1000)
    S, P, gam, nx, N, tol, max_iter = make_system(shape, "stair", seed=5)
    cfg_kw = dict(flags=g._lib.FLAG_TMEM, cluster=cluster)
    assert g.plan_query(nx, N, F64, True, g.pcg_config(empty_pinv=0, **cfg_kw))["cluster"] == cluster
    lam0 = np.zeros(N * nx)
    lam, it, ex = _solve1(S, P, gam, lam0, nx, N, tol, max_iter, **cfg_kw)
    _check(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex)


def _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter, flags=0):
    torch = _torch()
    d_l = torch.from_numpy(np.ascontiguousarray(lam0)).cuda()
    d_it = torch.full((batch,), 77, dtype=torch.int32, device="cuda")
    d_ex = torch.full((batch,), 7, dtype=torch.uint8, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0, flags=flags)
    g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
    torch.cuda.synchronize()
    return d_l.cpu().numpy(), d_it.cpu().numpy(), d_ex.cpu().numpy()


@pytest.mark.parametrize("nx,nu,N,batch", [(14, 7, 64, 300), (14, 7, 64, 75), (14, 7, 32, 500), (12, 4, 100, 200), (8, 4, 200, 333), (14, 7, 100, 97),
                                           (14, 7, 16, 700), (8, 4, 32, 901), (16, 8, 32, 310), (10, 5, 24, 333)])   # CTAs of <= 128 threads: two per SM
@pytest.mark.parametrize("precond", ["stair", "jacobi"])
def test_tmem64_batches_vs_oracle_and_register_kernel(nx, nu, N, batch, precond):
    """The planner's choice for fp64 batches that cannot be resident at once (more clusters than the GPU holds): ragged batch sizes,
    every distinct system against the oracle, replicas bit-identical, and agreement with the register-row plan."""
    torch = _torch()
    nseed, tol, max_iter = 5, 1e-10, 3000
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), precond, F64)
    idx = np.arange(batch) % nseed
    d_S, d_P, d_g = (torch.from_numpy(a[idx].copy()).cuda() for a in (S, P, gam))
    info = g.plan_query(nx, N, F64, True, g.pcg_config(empty_pinv=0), batch=batch)
    info_reg = g.plan_query(nx, N, F64, True, g.pcg_config(empty_pinv=0, flags=g._lib.FLAG_NO_TMEM), batch=batch)
    assert "P_TMEM" not in info_reg["kernel"]
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    if batch * info_reg["cluster"] > sms:
        assert "P_TMEM" in info["kernel"] and info["cluster"] <= info_reg["cluster"], (info, info_reg)
    forced = g.plan_query(nx, N, F64, True, g.pcg_config(empty_pinv=0, flags=g._lib.FLAG_TMEM), batch=batch)
    assert ("2 per SM" in forced["kernel"]) == (forced["threads"] <= 128), forced
    if forced["threads"] <= 128:
        assert forced["max_resident"] * forced["cluster"] > sms                       # more CTAs resident than there are SMs
    lam0 = np.zeros((batch, N * nx))
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter, flags=g._lib.FLAG_TMEM)
    lam_r, it_r, ex_r = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter, flags=g._lib.FLAG_NO_TMEM)
    lam_p, it_p, _ = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter)
    assert not ex.any() and not ex_r.any()
    if "P_TMEM" in info["kernel"]:
        assert np.array_equal(lam_p, lam) and np.array_equal(it_p, it)
    for s_ in range(nseed):
        rows = np.nonzero(idx == s_)[0]
        assert (it[rows] == it[rows[0]]).all() and all(np.array_equal(lam[r], lam[rows[0]]) for r in rows)
        _check(S[s_], P[s_], gam[s_], lam0[0], nx, N, tol, max_iter, lam[rows[0]], it[rows[0]], ex[rows[0]])
        assert abs(int(it[rows[0]]) - int(it_r[rows[0]])) <= 1
        if it[rows[0]] == it_r[rows[0]]:
            assert rel_err(lam[rows[0]], lam_r[rows[0]]) <= LAMBDA_TOL[np.dtype(F64)]


def test_tmem64_edge_semantics():
    """max_iter = 0 leaves lambda and sets the flag; a low cap reports it; warm start; NaN-poisoned padding blocks never reach the
    arithmetic; NaN in gamma never converges (reference semantics)."""
    torch = _torch()
    nx, nu, N, batch, tol = 14, 7, 64, 160, 1e-10
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(4), "stair", F64)
    Sn, Pn = S.copy(), P.copy()
    for M in (Sn, Pn):
        M[:, : nx * nx] = np.nan
        M[:, (N - 1) * 3 * nx * nx + 2 * nx * nx:] = np.nan
    idx = np.arange(batch) % 4
    d_S, d_P, d_g = (torch.from_numpy(a[idx].copy()).cuda() for a in (Sn, Pn, gam))
    rng = np.random.default_rng(3)
    lam0 = rng.standard_normal((batch, N * nx))
    F = g._lib.FLAG_TMEM
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 0, F)
    assert np.array_equal(lam, lam0) and (it == 0).all() and (ex == 1).all()
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 3, F)
    assert (it == 3).all() and (ex == 1).all() and np.isfinite(lam).all()
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 3000, F)
    assert (ex == 0).all()
    for i in (0, 1, 2, 3, batch - 1):
        _check(S[idx[i]], P[idx[i]], gam[idx[i]], lam0[i], nx, N, tol, 3000, lam[i], it[i], ex[i])
    gam_n = gam[idx].copy()
    gam_n[5, 17] = np.nan
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, torch.from_numpy(gam_n).cuda(), np.zeros((batch, N * nx)), tol, 40, F)
    assert it[5] == 40 and ex[5] == 1 and (ex[np.arange(batch) != 5] == 0).all()


def test_tmem64_selection_and_refusals():
    F, NO = g._lib.FLAG_TMEM, g._lib.FLAG_NO_TMEM
    q = lambda nx, N, batch, **kw: g.plan_query(nx, N, F64, kw.pop("has_p", True), g.pcg_config(empty_pinv=0, **kw), batch=batch)
    assert "P_TMEM" in q(14, 64, 2048)["kernel"] and q(14, 64, 2048)["cluster"] == 2
    # latency mode (the batch gets one cluster per system in one wave): the tensor-memory form on the largest cluster whose CTAs keep two warps;
    # nx 10 and 16 keep the register rows there (measured 2-4 % behind)
    one = q(14, 64, 1)
    assert "P_TMEM" in one["kernel"] and one["cluster"] == 8 and one["threads"] == 64 and one["max_resident"] >= 1, one
    assert q(14, 64, 9)["cluster"] == 8 and "P_TMEM" in q(14, 64, 9)["kernel"]
    assert "P_TMEM" not in q(10, 60, 1)["kernel"] and "P_TMEM" not in q(16, 32, 1)["kernel"]
    assert "P_TMEM" not in q(14, 64, 2048, flags=NO)["kernel"]
    assert "P_TMEM" not in q(14, 64, 2048, flags=g._lib.FLAG_TEXTBOOK)["kernel"]
    assert "P_TMEM" not in q(14, 64, 2048, has_p=False)["kernel"]
    for kw in (dict(flags=F | g._lib.FLAG_TEXTBOOK), dict(flags=F, has_p=False)):
        with pytest.raises(g.GbdpcgError) as e:
            q(14, 64, 8, **kw)
        assert e.value.status == g._lib.ERR_UNSUPPORTED
    with pytest.raises(g.GbdpcgError):
        q(7, 64, 8, flags=F)                                             # odd state size: no instantiation
    with pytest.raises(g.GbdpcgError):
        q(14, 16 * 40, 8, flags=F)                                       # beyond 16 CTAs x 256 threads


def test_tmem64_unaligned_pointers_fall_back():
    torch = _torch()
    nx, nu, N, batch, tol = 14, 7, 64, 200, 1e-10
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(2), "stair", F64)
    idx = np.arange(batch) % 2
    pad = lambda a: torch.from_numpy(np.concatenate([np.zeros(1, a.dtype), a.reshape(-1)])).cuda()[1:]
    d_S, d_P = pad(S[idx]), pad(P[idx])
    assert d_S.data_ptr() % 16 == 8
    d_g = torch.from_numpy(gam[idx].copy()).cuda()
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, np.zeros((batch, N * nx)), tol, 3000)
    assert not ex.any()
    _check(S[0], P[0], gam[0], np.zeros(N * nx), nx, N, tol, 3000, lam[0], it[0], ex[0])
    with pytest.raises(g.GbdpcgError):
        _batched(nx, N, batch, d_S, d_P, d_g, np.zeros((batch, N * nx)), tol, 3000, flags=g._lib.FLAG_TMEM)


def test_tmem64_variant_b_and_concurrent_launches():
    """The A/B alternative (more columns of S in tensor memory) solves the same systems; two launches on two streams share SMs and
    tensor memory and stay bit-identical over 30 repeats."""
    torch = _torch()
    nx, nu, N, batch, tol = 14, 7, 64, 150, 1e-10
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(3), "stair", F64)
    idx = np.arange(batch) % 3
    d_S, d_P, d_g = (torch.from_numpy(a[idx].copy()).cuda() for a in (S, P, gam))
    lam0 = np.zeros((batch, N * nx))
    lam_a, it_a, _ = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 3000, g._lib.FLAG_TMEM)
    os.environ["GBDPCG_TMEM64_VARIANT"] = "1"
    try:
        assert "S_tmem=b" in g.plan_query(nx, N, F64, True, g.pcg_config(empty_pinv=0, flags=g._lib.FLAG_TMEM), batch=batch)["kernel"]
        lam_b, it_b, ex_b = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 3000, g._lib.FLAG_TMEM)
    finally:
        del os.environ["GBDPCG_TMEM64_VARIANT"]
    assert not ex_b.any() and np.array_equal(it_a, it_b) and rel_err(lam_b, lam_a) <= LAMBDA_TOL[np.dtype(F64)]
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=3000, empty_pinv=0, flags=g._lib.FLAG_TMEM)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(30):
        l1 = torch.zeros(batch, N * nx, dtype=torch.float64, device="cuda")
        l2 = torch.zeros(batch, N * nx, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, l1, cfg, stream=s1)
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, l2, cfg, stream=s2)
        torch.cuda.synchronize()
        assert np.array_equal(l1.cpu().numpy(), lam_a) and np.array_equal(l2.cpu().numpy(), lam_a)


# ---------------------------------------------------------------------------------------------------------------
# the same form in fp32: FOUR rows per thread (state sizes that are multiples of 4), for systems beyond the single-CTA
# tensor-memory kernels (more than 256 four-row threads) -- C4-sized systems on 8 CTAs instead of 16
# ---------------------------------------------------------------------------------------------------------------
F32 = np.float32


def _check32(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex):
    lam_o, it_o, ex_o = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=0)
    assert abs(int(it) - it_o) <= 1, (it, it_o)
    lam_eq = lam_o if it == it_o else oracle_at_iters(S, P, gam, lam0, nx, N, int(it))
    assert rel_err(lam, lam_eq) <= LAMBDA_TOL[np.dtype(F32)], (rel_err(lam, lam_eq), it, it_o)


@pytest.mark.parametrize("nx,nu,N,cluster", [(12, 4, 512, 16), (12, 4, 512, 8), (8, 4, 200, 4), (8, 4, 200, 16), (16, 8, 100, 8), (16, 8, 100, 4),
                                             (12, 6, 70, 2), (8, 4, 33, 1), (16, 8, 7, 2), (12, 4, 2, 1),
                                             # nx = 2 mod 4: the last thread of a block-row owns two padding rows
                                             (14, 7, 100, 4), (14, 7, 100, 2), (14, 7, 512, 8), (14, 7, 60, 1), (14, 7, 3, 1), (14, 7, 2, 1), (14, 7, 130, 16),
                                             (10, 5, 150, 4), (10, 5, 33, 2), (10, 5, 85, 1)])
@pytest.mark.parametrize("precond", ["stair", "jacobi"])
def test_tmem32_cluster_form_single_solves(nx, nu, N, cluster, precond):
    S, P, gam, nx, N, tol, max_iter = make_system((nx, nu, N, F32, 1e-6, 3000), precond, seed=9)
    kw = dict(flags=g._lib.FLAG_TMEM, cluster=cluster)
    info = g.plan_query(nx, N, F32, True, g.pcg_config(empty_pinv=0, **kw))
    assert "P_TMEM" in info["kernel"] and "R=4" in info["kernel"] and info["cluster"] == cluster, info
    lam0 = np.zeros(N * nx, F32)
    lam, it, ex = _solve1(S, P, gam, lam0, nx, N, tol, max_iter, **kw)
    assert not ex
    if N * nx > 64:
        _check32(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex)
    lam2, it2, _ = _solve1(S, P, gam, lam0, nx, N, tol, max_iter, **kw)
    assert it2 == it and np.array_equal(lam2, lam)


@pytest.mark.parametrize("nx,nu,N,batch", [(12, 4, 100, 301), (8, 4, 200, 222), (16, 8, 100, 160), (12, 4, 300, 60), (14, 7, 100, 300), (14, 7, 128, 150)])
@pytest.mark.parametrize("precond", ["stair", "jacobi"])
def test_tmem32_cluster_form_batches(nx, nu, N, batch, precond):
    torch = _torch()
    nseed, tol, max_iter = 4, 1e-6, 3000
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), precond, F32)
    idx = np.arange(batch) % nseed
    d_S, d_P, d_g = (torch.from_numpy(a[idx].copy()).cuda() for a in (S, P, gam))
    auto = g.plan_query(nx, N, F32, True, g.pcg_config(empty_pinv=0), batch=batch)
    reg = g.plan_query(nx, N, F32, True, g.pcg_config(empty_pinv=0, flags=g._lib.FLAG_NO_TMEM), batch=batch)
    # the planner's choice in throughput mode: fewer SMs per system (a smaller cluster, or the same with two CTAs per SM)
    assert "P_TMEM" in auto["kernel"] and (auto["cluster"] < reg["cluster"] or "2 per SM" in auto["kernel"]), (auto, reg)
    lam0 = np.zeros((batch, N * nx), F32)
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter)
    lam_r, it_r, ex_r = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter, flags=g._lib.FLAG_NO_TMEM)
    assert not ex.any() and not ex_r.any()
    for s_ in range(nseed):
        rows = np.nonzero(idx == s_)[0]
        assert (it[rows] == it[rows[0]]).all() and all(np.array_equal(lam[r], lam[rows[0]]) for r in rows)
        _check32(S[s_], P[s_], gam[s_], lam0[0], nx, N, tol, max_iter, lam[rows[0]], it[rows[0]], ex[rows[0]])
        assert abs(int(it[rows[0]]) - int(it_r[rows[0]])) <= 1


def test_tmem32_selection():
    q = lambda nx, N, batch, **kw: g.plan_query(nx, N, F32, True, g.pcg_config(empty_pinv=0, **kw), batch=batch)
    # latency mode: nx 8 and 16 take the tensor-memory form, nx 12 keeps the register rows (measured); single-CTA shapes keep their kernels
    assert "P_TMEM" in q(8, 200, 1)["kernel"] and "P_TMEM" in q(16, 100, 1)["kernel"] and "P_TMEM" in q(16, 64, 1)["kernel"]
    assert "P_TMEM" not in q(12, 512, 1)["kernel"] and "P_TMEM" not in q(12, 100, 1)["kernel"]
    assert "pcg_tmem4" in q(16, 32, 1)["kernel"] and "pcg_tmem4" in q(12, 40, 8192)["kernel"] and "wide" in q(16, 64, 4096)["kernel"]
    assert "P_TMEM" in q(12, 512, 2048)["kernel"] and q(12, 512, 2048)["cluster"] == 8
    assert "P_TMEM" not in q(12, 512, 2048, flags=g._lib.FLAG_NO_TMEM)["kernel"]
    assert "P_TMEM" in q(14, 100, 2048)["kernel"] and q(14, 100, 2048)["cluster"] == 2      # nx 14: padded row groups, 2 CTAs instead of 4
    assert "P_TMEM" not in q(10, 100, 2048)["kernel"]                    # nx 10: no fewer SMs than the register rows -- not taken
    assert "P_TMEM" not in q(6, 300, 2048)["kernel"]                     # no instantiation below nx 8
    with pytest.raises(g.GbdpcgError):
        q(6, 300, 2048, flags=g._lib.FLAG_TMEM)


def test_tmem32_padded_rows_edge_semantics():
    """nx 14, four rows per thread, N beyond the single-CTA kernels: max_iter 0 / cap / warm start, NaN-poisoned padding blocks, NaN in gamma."""
    torch = _torch()
    nx, nu, N, batch, tol = 14, 7, 100, 160, 1e-6
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(4), "stair", F32)
    Sn, Pn = S.copy(), P.copy()
    for M in (Sn, Pn):
        M[:, : nx * nx] = np.nan
        M[:, (N - 1) * 3 * nx * nx + 2 * nx * nx:] = np.nan
    idx = np.arange(batch) % 4
    d_S, d_P, d_g = (torch.from_numpy(a[idx].copy()).cuda() for a in (Sn, Pn, gam))
    lam0 = np.random.default_rng(5).standard_normal((batch, N * nx)).astype(F32)
    F = g._lib.FLAG_TMEM
    assert "P_TMEM" in g.plan_query(nx, N, F32, True, g.pcg_config(empty_pinv=0, flags=F), batch=batch)["kernel"]
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 0, F)
    assert np.array_equal(lam, lam0) and (it == 0).all() and (ex == 1).all()
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 3, F)
    assert (it == 3).all() and (ex == 1).all() and np.isfinite(lam).all()
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 3000, F)
    assert (ex == 0).all()
    for i in (0, 1, 2, 3, batch - 1):
        _check32(S[idx[i]], P[idx[i]], gam[idx[i]], lam0[i], nx, N, tol, 3000, lam[i], it[i], ex[i])
    gam_n = gam[idx].copy()
    gam_n[7, 13] = np.nan
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, torch.from_numpy(gam_n).cuda(), np.zeros((batch, N * nx), F32), tol, 40, F)
    assert it[7] == 40 and ex[7] == 1 and (ex[np.arange(batch) != 7] == 0).all()
