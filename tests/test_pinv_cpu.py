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


# ---------------------------------------------------------------------------------------------------------------
# the builder oracle held to the reference's OWN primitives: tests/golden/ref_glass_pinv_*.npz were recorded by running
# glass::invertMatrix / glass::gemm (reference GLASS/src/L3/inv.cuh:6-37, gemm.cuh) on the blocks of S on a B200
# (oracle/ref_harness.cu: ref_glass_build_pinv; oracle/make_ref_fixtures.py <dir> pinv)
# ---------------------------------------------------------------------------------------------------------------
import glob  # noqa: E402
import os  # noqa: E402

from conftest import GOLDEN  # noqa: E402
from helpers import make_system  # noqa: E402

GLASS_FIXTURES = sorted(glob.glob(os.path.join(GOLDEN, "ref_glass_pinv_*.npz")))
GLASS_BAR = {np.dtype(np.float32): 2e-5, np.dtype(np.float64): 1e-11}      # same order of operations; FMA contraction differs (GPU vs x86)


@pytest.mark.skipif(not GLASS_FIXTURES, reason="no glass::invertMatrix fixtures recorded yet")
@pytest.mark.parametrize("path", GLASS_FIXTURES, ids=lambda p: os.path.basename(p)[len("ref_glass_pinv_"):-len(".npz")])
def test_oracle_builder_matches_recorded_glass_outputs(path):
    z = np.load(path)
    name, kind, dtype = str(z["config"]), str(z["kind"]), np.dtype(str(z["dtype"]))
    S, _, _, nx, N, _, _ = make_system(name, "identity", dtype=dtype.type)
    assert (nx, N) == (int(z["nx"]), int(z["N"]))
    P = oracle.build_pinv(S, nx, N, kind).reshape(N, 3 * nx * nx)[z["rows"]]
    ref = z["pinv_rows"]
    assert ref.dtype == dtype and np.isfinite(ref).all()
    assert float(np.abs(P.astype(np.float64) - ref).max() / np.abs(ref).max()) < GLASS_BAR[dtype]
    # the recorded preconditioner really is one: diagonal blocks invert those of S
    Sb = S.reshape(N, 3, nx, nx)
    for i, k in enumerate(z["rows"][:3]):
        D = Sb[k, 1].T.astype(np.float64)                                  # column-major block
        Dinv = ref[i].reshape(3, nx, nx)[1].T.astype(np.float64)
        assert np.abs(D @ Dinv - np.eye(nx)).max() < (5e-4 if dtype == np.float32 else 1e-10)
