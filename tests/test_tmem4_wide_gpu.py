"""Wide form of the four-rows-per-thread tensor-memory kernel (pcg_tmem4.cuh, NWQ = 2: 8 warps, two per lane quarter, one CTA per
SM) -- fp32 systems of 129..256 four-row threads (nx 14: 37 <= N <= 64), the planner's choice for batches that the register-row
clusters cannot hold at once.  Against the host oracle and the register-row plan."""
import numpy as np
import pytest

import gbd_pcg_b200 as g
from conftest import LAMBDA_TOL, rel_err
from helpers import oracle_at_iters
from oracle import oracle

pytestmark = pytest.mark.gpu
F32 = np.float32


def _torch():
    import torch
    return torch


def _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter, flags=0):
    torch = _torch()
    d_l = torch.from_numpy(np.ascontiguousarray(lam0)).cuda()
    d_it = torch.full((batch,), 77, dtype=torch.int32, device="cuda")
    d_ex = torch.full((batch,), 7, dtype=torch.uint8, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0, flags=flags)
    g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
    torch.cuda.synchronize()
    return d_l.cpu().numpy(), d_it.cpu().numpy(), d_ex.cpu().numpy()


def _check(S, P, gam, lam0, nx, N, tol, max_iter, lam, it, ex):
    lam_o, it_o, ex_o = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=0)
    assert abs(int(it) - it_o) <= 1, (it, it_o)
    if it == it_o:
        assert bool(ex) == bool(ex_o)
    lam_eq = lam_o if it == it_o else oracle_at_iters(S, P, gam, lam0, nx, N, int(it))
    assert rel_err(lam, lam_eq) <= LAMBDA_TOL[np.dtype(F32)], (rel_err(lam, lam_eq), it, it_o)


@pytest.mark.parametrize("nx,nu,N,batch", [(14, 7, 64, 300), (14, 7, 40, 77), (14, 7, 37, 160), (16, 8, 64, 149), (16, 8, 33, 200), (12, 4, 85, 150),
                                           (12, 6, 43, 333), (8, 4, 128, 151), (8, 4, 65, 400), (10, 5, 70, 222)])
@pytest.mark.parametrize("precond", ["stair", "jacobi"])
def test_wide_tmem4_vs_oracle_and_register_plan(nx, nu, N, batch, precond):
    torch = _torch()
    nseed, tol, max_iter = 5, 1e-6, 2000
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), precond, F32)
    idx = np.arange(batch) % nseed
    d_S, d_P, d_g = (torch.from_numpy(a[idx].copy()).cuda() for a in (S, P, gam))
    forced = g.plan_query(nx, N, F32, True, g.pcg_config(empty_pinv=0, flags=g._lib.FLAG_TMEM), batch=batch)
    assert "wide" in forced["kernel"] and forced["threads"] > 128 and forced["max_resident"] == torch.cuda.get_device_properties(0).multi_processor_count
    reg = g.plan_query(nx, N, F32, True, g.pcg_config(empty_pinv=0, flags=g._lib.FLAG_NO_TMEM), batch=batch)
    auto = g.plan_query(nx, N, F32, True, g.pcg_config(empty_pinv=0), batch=batch)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    assert ("wide" in auto["kernel"]) == (batch * reg["cluster"] > sms if reg["cluster"] > 1 else batch >= sms), (auto, reg)
    lam0 = np.zeros((batch, N * nx), F32)
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter, g._lib.FLAG_TMEM)
    lam_r, it_r, ex_r = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, max_iter, g._lib.FLAG_NO_TMEM)
    assert not ex.any() and not ex_r.any()
    for s_ in range(nseed):
        rows = np.nonzero(idx == s_)[0]
        assert (it[rows] == it[rows[0]]).all() and all(np.array_equal(lam[r], lam[rows[0]]) for r in rows)
        _check(S[s_], P[s_], gam[s_], lam0[0], nx, N, tol, max_iter, lam[rows[0]], it[rows[0]], ex[rows[0]])
        assert abs(int(it[rows[0]]) - int(it_r[rows[0]])) <= 1


def test_wide_tmem4_edge_semantics_and_concurrency():
    torch = _torch()
    nx, nu, N, batch, tol = 14, 7, 64, 170, 1e-6
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(4), "stair", F32)
    Sn, Pn = S.copy(), P.copy()
    for M in (Sn, Pn):
        M[:, : nx * nx] = np.nan
        M[:, (N - 1) * 3 * nx * nx + 2 * nx * nx:] = np.nan
    idx = np.arange(batch) % 4
    d_S, d_P, d_g = (torch.from_numpy(a[idx].copy()).cuda() for a in (Sn, Pn, gam))
    lam0 = np.random.default_rng(3).standard_normal((batch, N * nx)).astype(F32)
    F = g._lib.FLAG_TMEM
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 0, F)
    assert np.array_equal(lam, lam0) and (it == 0).all() and (ex == 1).all()
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 3, F)
    assert (it == 3).all() and (ex == 1).all() and np.isfinite(lam).all()
    lam, it, ex = _batched(nx, N, batch, d_S, d_P, d_g, lam0, tol, 2000, F)
    assert (ex == 0).all()
    for i in (0, 1, 2, 3, batch - 1):
        _check(S[idx[i]], P[idx[i]], gam[idx[i]], lam0[i], nx, N, tol, 2000, lam[i], it[i], ex[i])
    # two launches on two streams (they queue for the SMs: one wide CTA fills an SM's registers), 20 repeats, bit-identical
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=2000, empty_pinv=0, flags=F)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(20):
        l1, l2 = torch.from_numpy(lam0).cuda(), torch.from_numpy(lam0).cuda()
        torch.cuda.synchronize()
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, l1, cfg, stream=s1)
        g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, l2, cfg, stream=s2)
        torch.cuda.synchronize()
        assert np.array_equal(l1.cpu().numpy(), lam) and np.array_equal(l2.cpu().numpy(), lam)
    # beyond 256 four-row threads the tensor-memory form of the CLUSTER kernel takes over (tests/test_tmem64_gpu.py); the two-row arm has no wide form
    assert "P_TMEM" in g.plan_query(nx, 65, F32, True, g.pcg_config(empty_pinv=0, flags=F), batch=batch)["kernel"]
    with pytest.raises(g.GbdpcgError) as e:
        g.plan_query(nx, 64, F32, True, g.pcg_config(empty_pinv=0, flags=F | g._lib.FLAG_TMEM_R2), batch=batch)
    assert e.value.status == g._lib.ERR_UNSUPPORTED
