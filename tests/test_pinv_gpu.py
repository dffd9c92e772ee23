"""On-GPU preconditioner construction (gbdpcg_build_pinv_*) against the host oracle
(oracle_build_pinv_*: glass::invertMatrix order, reference GLASS/src/L3/inv.cuh:6-37) and dense-inverse properties,
and the host solve entry points with GBDPCG_FLAG_BUILD_* against solves that were handed Phi^-1."""
import numpy as np
import pytest

import gbd_pcg_b200 as g
from gbd_pcg_b200 import _lib
from conftest import rel_err
from helpers import make_system
from oracle import oracle

pytestmark = pytest.mark.gpu

# same arithmetic order on both sides; the GPU contracts a - (c*pvInv)*b into an FMA, the x86 oracle does not
BAR = {np.dtype(np.float32): 2e-5, np.dtype(np.float64): 1e-11}


def _torch():
    import torch
    return torch


def gpu_build(S, nx, N, batch, kind):
    torch = _torch()
    d_S = torch.from_numpy(np.ascontiguousarray(S)).cuda()
    d_P = torch.full_like(d_S, float("nan"))
    before = g.kernel_launches()
    g.build_pinv_dev(nx, N, batch, d_S, d_P, kind)
    torch.cuda.synchronize()
    assert g.kernel_launches() == before + 1
    return d_P.cpu().numpy()


def max_rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / np.abs(b).max())


@pytest.mark.parametrize("kind", ["identity", "jacobi", "stair"])
@pytest.mark.parametrize("name", ["C1", "C2", "C3", "C4"])
def test_baseline_configs_vs_oracle(name, kind):
    S, _, _, nx, N, _, _ = make_system(name, "identity")
    P = gpu_build(S, nx, N, 1, kind)
    Po = oracle.build_pinv(S, nx, N, kind)
    assert np.isfinite(P).all()
    assert max_rel(P, Po) < BAR[S.dtype]
    # exact zeros where the layout has padding / the kind has no coupling
    Pb, Pob = P.reshape(N, 3, nx * nx), Po.reshape(N, 3, nx * nx)
    assert np.array_equal(Pb == 0, Pob == 0) or kind == "stair"
    assert not Pb[0, 0].any() and not Pb[N - 1, 2].any()


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("nx,nu,N", [(2, 1, 2), (4, 2, 9), (6, 3, 33), (8, 4, 70), (10, 5, 17), (12, 6, 300), (16, 8, 129), (14, 7, 1000)])
def test_specialised_shapes_and_chunking(nx, nu, N, dt):
    """Every compiled state size; N chosen so that systems are cut into several chunks with halo blocks (and ragged tails)."""
    S, _, _ = g.synth.lqr_schur(nx, nu, N, 7, "identity", dt)
    for kind in ("jacobi", "stair"):
        P = gpu_build(S, nx, N, 1, kind)
        assert max_rel(P, oracle.build_pinv(S, nx, N, kind)) < BAR[np.dtype(dt)]


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("nx,N", [(3, 5), (5, 12), (7, 8), (18, 6), (25, 4)])
def test_runtime_nx_fallback(nx, N, dt):
    """Odd / large state sizes go through the runtime-nx kernel; systems here are random SPD block-tridiagonal."""
    rng = np.random.default_rng(nx * 100 + N)
    A = np.zeros((N * nx, N * nx))
    for k in range(N):
        M = rng.standard_normal((nx, nx))
        A[k * nx:(k + 1) * nx, k * nx:(k + 1) * nx] = M @ M.T + nx * np.eye(nx)
        if k + 1 < N:
            L = 0.3 * rng.standard_normal((nx, nx))
            A[(k + 1) * nx:(k + 2) * nx, k * nx:(k + 1) * nx] = L
            A[k * nx:(k + 1) * nx, (k + 1) * nx:(k + 2) * nx] = L.T
    S = g.layout.pack_dense(A, nx, N).astype(dt)
    for kind in ("identity", "jacobi", "stair"):
        P = gpu_build(S, nx, N, 1, kind)
        assert max_rel(P, oracle.build_pinv(S, nx, N, kind)) < BAR[np.dtype(dt)]


def test_batched_build_matches_per_problem_and_ignores_padding():
    nx, nu, N, B = 14, 7, 32, 300
    S, Pref, _ = g.synth.lqr_schur_batch(nx, nu, N, range(8), "stair", np.float32)
    S = np.tile(S, (B // 8 + 1, 1))[:B].copy()
    Sn = S.reshape(B, N, 3, nx * nx).copy()
    Sn[:, 0, 0] = np.nan            # padding blocks of S are never read (reference include/pcg.cuh:105-106)
    Sn[:, N - 1, 2] = np.nan
    P = gpu_build(Sn.reshape(B, -1), nx, N, B, "stair").reshape(B, -1)
    assert np.isfinite(P).all()
    for i in range(8):
        assert max_rel(P[i], oracle.build_pinv(S[i], nx, N, "stair")) < BAR[np.dtype(np.float32)]
        assert max_rel(P[i], Pref[i]) < 1e-5            # vs numpy.linalg.inv in fp64, cast
    for i in range(8, B):
        assert np.array_equal(P[i], P[i % 8])           # deterministic: same input, same bits, whichever CTA ran it


def test_host_form_and_property():
    nx, nu, N = 12, 4, 64
    S, _, _ = g.synth.lqr_schur(nx, nu, N, 2, "identity", np.float64)
    P = g.build_pinv_host(S, nx, N, 1, "jacobi")
    Sb = S.reshape(N, 3, nx, nx).transpose(0, 1, 3, 2)
    Pb = P.reshape(N, 3, nx, nx).transpose(0, 1, 3, 2)
    for k in range(N):
        assert np.abs(Pb[k, 1] @ Sb[k, 1] - np.eye(nx)).max() < 1e-12


def test_bad_arguments():
    torch = _torch()
    x = torch.zeros(3 * 4 * 4 * 2, device="cuda")
    with pytest.raises(g.GbdpcgError) as e:
        g.build_pinv_dev(4, 2, 1, x, x.clone(), 7)
    assert e.value.status == _lib.ERR_INVALID
    with pytest.raises(g.GbdpcgError):
        g.build_pinv_dev(4, 1, 1, x, x.clone(), "stair")     # N < 2


@pytest.mark.parametrize("name,kind", [("C2", "stair"), ("C2", "jacobi"), ("C3", "stair"), ("C4", "stair")])
def test_host_solve_with_device_built_preconditioner(name, kind):
    """solvePCG(h_S, h_gamma, h_lambda, ...) -- the reference's own host signature, no Pinv argument -- with
    GBDPCG_FLAG_BUILD_*: same iteration count (+-1) and solution as the solve that was handed the fp64-built Phi^-1."""
    S, P, gam, nx, N, tol, max_iter = make_system(name, kind)
    lam_a = np.zeros(N * nx, dtype=S.dtype)
    it_a = g.pcg_solve(S, gam, lam_a, nx, N, g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0), h_Pinv=P)
    lam_b = np.zeros(N * nx, dtype=S.dtype)
    before = g.kernel_launches()
    it_b = g.solvePCG(S, gam, lam_b, nx, N, g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, flags=g.api.BUILD_FLAGS[kind]))
    assert g.kernel_launches() == before + 2                 # builder + solver
    assert abs(it_a - it_b) <= 1
    # both are converged solutions of the same system: compare against the dense solve at the accuracy tol allows
    A = g.layout.unpack_dense(S.astype(np.float64), nx, N)
    x = np.linalg.solve(A, gam.astype(np.float64))
    bar = 2e-4 if S.dtype == np.float32 else 2e-6        # what |r.Phi^-1 r| < tol leaves in lambda
    assert rel_err(lam_a, x) < bar and rel_err(lam_b, x) < bar
    # and against the oracle run with the oracle-built preconditioner, at equal iteration count
    Po = oracle.build_pinv(S, nx, N, kind)
    lam_o, it_o, _ = oracle.pcg(S, Po, gam, np.zeros(N * nx, S.dtype), nx, N, tol, max_iter)
    assert abs(it_b - it_o) <= 1
    if it_b == it_o:
        assert rel_err(lam_b, lam_o) < (1e-5 if S.dtype == np.float32 else 1e-10)


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("nx,N", [(7, 30), (15, 64), (20, 32), (27, 20), (32, 32), (33, 12), (40, 8)])
def test_host_signature_on_odd_large_and_uninstantiated_state_sizes(nx, N, dt):
    """The reference's host signature end to end (run-time-nx builder, then whichever solver kernel the planner picks) on
    state sizes outside the even 2..16 set, against the oracle with the oracle-built stair preconditioner."""
    from helpers import random_block_tridiagonal
    if nx % 2:
        S, _, gam = random_block_tridiagonal(nx, N, "identity", dt, seed=nx, coupled=True)
    else:
        S, _, gam = g.synth.lqr_schur(nx, nx // 2, N, 1, "identity", dt)
    tol, max_iter = (1e-6 if dt == np.float32 else 1e-10), 3000
    lam = np.zeros(N * nx, dtype=dt)
    before = g.kernel_launches()
    it = g.solvePCG(S, gam, lam, nx, N, g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0))
    assert g.kernel_launches() == before + 2                 # builder + solver
    Po = oracle.build_pinv(S, nx, N, "stair")
    lam_o, it_o, ex_o = oracle.pcg(S, Po, gam, np.zeros(N * nx, dt), nx, N, tol, max_iter)
    assert not ex_o and abs(it - it_o) <= 1
    if it != it_o:
        lam_o = oracle.pcg(S, Po, gam, np.zeros(N * nx, dt), nx, N, 0.0, it)[0]
    lam_1 = oracle.pcg(S, Po, gam, np.zeros(N * nx, dt), nx, N, 0.0, it, order=1)[0]
    bar = max(1e-5 if dt == np.float32 else 1e-10, 3 * rel_err(lam_1, lam_o))
    assert rel_err(lam, lam_o) < bar


def test_batched_host_solve_with_device_built_preconditioner():
    nx, nu, N, B = 14, 7, 32, 600
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(16), "stair", np.float32)
    rep = lambda a: np.tile(a, (B // 16 + 1, 1))[:B].copy()
    S, P, gam = rep(S), rep(P), rep(gam)
    cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, empty_pinv=0)
    lam_a = np.zeros((B, N * nx), np.float32)
    it_a, ex_a = g.solve_batched_host(S, P, gam, lam_a, nx, N, B, cfg)
    lam_b = np.zeros((B, N * nx), np.float32)
    cfg_b = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, flags=_lib.FLAG_BUILD_STAIR)
    it_b, ex_b = g.solve_batched_host(S, None, gam, lam_b, nx, N, B, cfg_b)
    assert not ex_a.any() and not ex_b.any()
    assert np.abs(it_a.astype(int) - it_b.astype(int)).max() <= 1
    for i in range(16):
        A = g.layout.unpack_dense(S[i].astype(np.float64), nx, N)
        x = np.linalg.solve(A, gam[i].astype(np.float64))
        assert rel_err(lam_b[i], x) < 2e-4
    for i in range(16, B):
        assert np.array_equal(lam_b[i], lam_b[i % 16])


def test_reference_signature_builds_the_preconditioner_when_empty_pinv_is_zero():
    """solvePCG(h_S, h_gamma, h_lambda, stateSize, knotPoints, &config) with config.empty_pinv = 0: the reference warns
    and solves with uninitialised memory (include/interface.cuh:33-34,45-46); here Phi^-1 (config.pinv_build) is formed
    on the device, and the iteration counts are those of SURVEY s8(d): 103 / 50 / 25 for identity / Jacobi / stair."""
    S, _, gam, nx, N, tol, max_iter = make_system("C2", "identity")
    for kw, want in ((dict(empty_pinv=1), 103), (dict(empty_pinv=0, pinv_build="jacobi"), 50), (dict(empty_pinv=0), 25)):
        lam = np.zeros(N * nx, dtype=S.dtype)
        it = g.solvePCG(S, gam, lam, nx, N, g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, **kw))
        assert abs(it - want) <= 1, (kw, it)


def test_build_and_solve_are_capturable_in_a_cuda_graph():
    """resident MPC loop -- build Phi^-1 + solve captured once, replayed with new S / gamma."""
    torch = _torch()
    nx, nu, N = 14, 7, 32
    cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, empty_pinv=0)
    S0, _, g0 = g.synth.lqr_schur(nx, nu, N, 0, "identity", np.float32)
    d_S, d_g = torch.from_numpy(S0).cuda(), torch.from_numpy(g0).cuda()
    d_P, d_l = torch.empty_like(d_S), torch.zeros(N * nx, device="cuda")
    d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(2):                        # warm-up outside capture (function attributes, occupancy queries)
            d_l.zero_()
            g.build_pinv_dev(nx, N, 1, d_S, d_P, "stair", stream=side)
            g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None, stream=side)
    side.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        d_l.zero_()
        g.build_pinv_dev(nx, N, 1, d_S, d_P, "stair", stream=side)
        g.solve_dev_async(nx, N, d_S, d_P, d_g, d_l, cfg, d_it, None, stream=side)
    for seed in (3, 4, 5):
        S1, P1, g1 = g.synth.lqr_schur(nx, nu, N, seed, "stair", np.float32)
        d_S.copy_(torch.from_numpy(S1).cuda())
        d_g.copy_(torch.from_numpy(g1).cuda())
        graph.replay()
        torch.cuda.synchronize()
        lam_o, it_o, _ = oracle.pcg(S1, P1, g1, np.zeros(N * nx, np.float32), nx, N, 1e-6, 1000)
        assert abs(int(d_it.item()) - it_o) <= 1
        A = g.layout.unpack_dense(S1.astype(np.float64), nx, N)
        assert rel_err(d_l.cpu().numpy(), np.linalg.solve(A, g1.astype(np.float64))) < 2e-4


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_unaligned_device_pointers_fall_back_from_tma(dt):
    """S / Pinv not 16-byte aligned: the TMA bulk copies are replaced by ordinary loads and stores, same bits."""
    torch = _torch()
    nx, nu, N, B = 14, 7, 32, 5
    S, _, _ = g.synth.lqr_schur_batch(nx, nu, N, range(B), "identity", dt)
    ref = gpu_build(S, nx, N, B, "stair")
    n = S.size
    buf_s = torch.zeros(n + 3, dtype=torch.float32 if dt == np.float32 else torch.float64, device="cuda")
    buf_p = torch.full((n + 3,), float("nan"), dtype=buf_s.dtype, device="cuda")
    off = 1                                                   # 4- or 8-byte aligned only
    buf_s[off:off + n].copy_(torch.from_numpy(S.reshape(-1)))
    d_S, d_P = buf_s[off:off + n], buf_p[off:off + n]
    assert d_S.data_ptr() % 16 != 0
    g.build_pinv_dev(nx, N, B, d_S, d_P, "stair")
    torch.cuda.synchronize()
    assert np.array_equal(d_P.cpu().numpy().reshape(ref.shape), ref)
    assert torch.isnan(buf_p[:off]).all() and torch.isnan(buf_p[off + n:]).all()     # nothing written outside


def test_zero_batch_is_a_no_op():
    torch = _torch()
    x = torch.zeros(3 * 4 * 4 * 2, device="cuda")
    before = g.kernel_launches()
    g.build_pinv_dev(4, 2, 0, x, x.clone(), "stair")
    assert g.kernel_launches() == before


# ---------------------------------------------------------------------------------------------------------------
# the product's builder kernel held to the reference's own primitives (glass::invertMatrix / glass::gemm): recorded fixtures, and
# the live harness when it travelled with the repo
# ---------------------------------------------------------------------------------------------------------------
import glob  # noqa: E402
import os  # noqa: E402

from conftest import GOLDEN  # noqa: E402

GLASS_FIXTURES = sorted(glob.glob(os.path.join(GOLDEN, "ref_glass_pinv_*.npz")))


@pytest.mark.skipif(not GLASS_FIXTURES, reason="no glass::invertMatrix fixtures recorded yet")
@pytest.mark.parametrize("path", GLASS_FIXTURES, ids=lambda p: os.path.basename(p)[len("ref_glass_pinv_"):-len(".npz")])
def test_builder_kernel_matches_recorded_glass_outputs(path):
    z = np.load(path)
    name, kind, dtype = str(z["config"]), str(z["kind"]), np.dtype(str(z["dtype"]))
    S, _, _, nx, N, _, _ = make_system(name, "identity", dtype=dtype.type)
    P = gpu_build(S, nx, N, 1, kind).reshape(N, 3 * nx * nx)[z["rows"]]
    assert max_rel(P, z["pinv_rows"]) < BAR[dtype]


@pytest.mark.parametrize("name,dtype", [("C1", np.float32), ("C1", np.float64), ("C2", np.float32), ("C3", np.float64), ("C4", np.float32)])
@pytest.mark.parametrize("kind", ["jacobi", "stair"])
def test_builder_kernel_against_live_glass_primitives(name, dtype, kind):
    from oracle import ref_cuda
    S, _, _, nx, N, _, _ = make_system(name, "identity", dtype=dtype)
    if not ref_cuda.has(dtype, nx, N) or not hasattr(ref_cuda.get(dtype, nx, N).L, "ref_glass_build_pinv"):
        pytest.skip("reference harness not built for this config")
    P_ref = ref_cuda.get(dtype, nx, N).glass_build_pinv(S, kind)
    P = gpu_build(S, nx, N, 1, kind)
    assert max_rel(P, P_ref) < BAR[np.dtype(dtype)]
    assert max_rel(oracle.build_pinv(S, nx, N, kind), P_ref) < BAR[np.dtype(dtype)]
    # a solve preconditioned with the reference-built Phi^-1 and one with ours take the same number of iterations
    if name != "C1":
        _, _, gam, _, _, tol, max_iter = make_system(name, "identity", dtype=dtype)
        cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0)
        l1, l2 = np.zeros(N * nx, dtype), np.zeros(N * nx, dtype)
        it1 = g.solvePCG(S, gam, l1, nx, N, cfg, h_Pinv=P_ref)
        it2 = g.solvePCG(S, gam, l2, nx, N, cfg, h_Pinv=P)
        assert abs(it1 - it2) <= 1
