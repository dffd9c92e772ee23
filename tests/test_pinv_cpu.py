"""Host oracle of the preconditioner construction (oracle_build_pinv_*: restates glass::invertMatrix, reference
GLASS/src/L3/inv.cuh:6-37, plus the stair products) pinned against dense numpy inverses.  The reference never
calls its inversion primitive on this path and ships no vectors for it (parity unpinned by the reference), so the
pins are mathematical: P_kk S_kk = I, the stair formula in fp64, zero padding blocks, symmetry for symmetric S."""
import numpy as np
import pytest

import gbd_pcg_b200 as g
from oracle import oracle

SHAPES = [(2, 1, 3), (4, 2, 6), (14, 7, 32), (12, 4, 40), (16, 8, 5)]


def blocks(M, nx, N):
    return np.asarray(M).reshape(N, 3, nx, nx).transpose(0, 1, 3, 2)      # [k, left/diag/right, row, col]


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("nx,nu,N", SHAPES)
def test_oracle_jacobi_is_the_block_inverse(nx, nu, N, dt):
    S, _, _ = g.synth.lqr_schur(nx, nu, N, 3, "identity", dt)
    P = oracle.build_pinv(S, nx, N, "jacobi")
    Sb, Pb = blocks(S.astype(np.float64), nx, N), blocks(P.astype(np.float64), nx, N)
    tol = 2e-5 if dt == np.float32 else 1e-12
    for k in range(N):
        assert np.abs(Pb[k, 1] @ Sb[k, 1] - np.eye(nx)).max() < tol
    assert not Pb[:, 0].any() and not Pb[:, 2].any()


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("nx,nu,N", SHAPES)
def test_oracle_stair_matches_the_formula(nx, nu, N, dt):
    S, _, _ = g.synth.lqr_schur(nx, nu, N, 1, "identity", dt)
    P = oracle.build_pinv(S, nx, N, "stair")
    Sb, Pb = blocks(S.astype(np.float64), nx, N), blocks(P.astype(np.float64), nx, N)
    Dinv = np.linalg.inv(Sb[:, 1])
    tol = 5e-6 if dt == np.float32 else 1e-12
    scale = np.abs(Dinv).max()
    for k in range(N):
        assert np.abs(Pb[k, 1] - Dinv[k]).max() < tol * scale
        if k + 1 < N:
            assert np.abs(Pb[k, 2] + Dinv[k] @ Sb[k, 2] @ Dinv[k + 1]).max() < tol * scale
        if k > 0:
            assert np.abs(Pb[k, 0] + Dinv[k] @ Sb[k, 0] @ Dinv[k - 1]).max() < tol * scale
    assert not Pb[0, 0].any() and not Pb[N - 1, 2].any()          # padding blocks are zero
    # S symmetric => Phi^-1 symmetric: right block of row k is the transpose of the left block of row k+1
    for k in range(N - 1):
        assert np.abs(Pb[k, 2] - Pb[k + 1, 0].T).max() < 10 * tol * scale


def test_oracle_matches_the_generator_preconditioners():
    """synth.py builds Phi^-1 with numpy.linalg.inv in fp64; the oracle's Gauss-Jordan agrees to rounding."""
    for kind in ("jacobi", "stair"):
        S, P, _ = g.synth.lqr_schur(14, 7, 32, 0, kind, np.float64)
        Po = oracle.build_pinv(S, 14, 32, kind)
        assert np.abs(Po - P).max() < 1e-13 * np.abs(P).max()


def test_oracle_identity_and_padding_inputs():
    S, _, _ = g.synth.lqr_schur(6, 3, 5, 0, "identity", np.float64)
    P = oracle.build_pinv(S, 6, 5, "identity")
    Pb = blocks(P, 6, 5)
    assert all(np.array_equal(Pb[k, 1], np.eye(6)) for k in range(5)) and not Pb[:, 0].any() and not Pb[:, 2].any()
    # padding blocks of S (left of row 0, right of row N-1) are never read: NaN there must not reach the output
    S2 = S.copy().reshape(5, 3, 36)
    S2[0, 0] = np.nan
    S2[4, 2] = np.nan
    P2 = oracle.build_pinv(S2.reshape(-1), 6, 5, "stair")
    assert np.array_equal(P2, oracle.build_pinv(S, 6, 5, "stair"))


def test_stair_preconditioner_from_the_oracle_drives_the_solve():
    """End to end on the host: oracle-built Phi^-1 gives the documented C2 iteration counts (50 / 25)."""
    nx, nu, N = 14, 7, 32
    S, _, gam = g.synth.lqr_schur(nx, nu, N, 0, "identity", np.float32)
    for kind, iters in (("jacobi", 50), ("stair", 25)):
        P = oracle.build_pinv(S, nx, N, kind)
        _, it, ex = oracle.pcg(S, P, gam, np.zeros(N * nx, np.float32), nx, N, 1e-6, 1000)
        assert not ex and abs(it - iters) <= 1
