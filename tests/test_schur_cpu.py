"""Host oracle of the Schur-complement formation / primal recovery (oracle_schur_form_*, oracle_schur_recover_*; SURVEY
s8(f) rank 4) pinned against a dense fp64 KKT solve and against the benchmark generator (synth.lqr_blocks), whose S it
must reproduce.  Not in the reference tree (parity unpinned by the reference): the pins are mathematical."""
import numpy as np
import pytest

import gbd_pcg_b200 as g
from oracle import oracle

FLAT = lambda d: tuple(d[k].reshape(-1) for k in "QRABqrc")


@pytest.mark.parametrize("nx,nu,N", [(2, 1, 3), (14, 7, 32), (12, 4, 40), (6, 6, 5), (5, 2, 7)])
def test_diagonal_costs_reproduce_the_benchmark_generator(nx, nu, N):
    if nx % 2:
        pytest.skip("synth family needs an even state size")
    d = g.synth.lqr_stage_data(nx, nu, N, 0)
    S, gam, _, _ = oracle.schur_form(*FLAT(d), nx, nu, N)
    S0, _, _ = g.synth.lqr_schur(nx, nu, N, 0, "identity", np.float64)
    assert np.abs(S - S0).max() < 1e-14 * np.abs(S0).max()


@pytest.mark.parametrize("dt,tol", [(np.float64, 1e-12), (np.float32, 3e-5)])
@pytest.mark.parametrize("nx,nu,N", [(4, 2, 6), (14, 7, 32), (12, 4, 20)])
def test_dense_costs_against_the_kkt_system(nx, nu, N, dt, tol):
    d = g.synth.lqr_stage_data(nx, nu, N, 5, dtype=dt, dense_costs=True)
    S, gam, Qi, Ri = oracle.schur_form(*FLAT(d), nx, nu, N)
    G, C, gg, c = g.synth.kkt_dense(d, nx, nu, N)
    Sd = -C @ np.linalg.inv(G) @ C.T
    gd = C @ np.linalg.solve(G, gg) - c
    assert np.abs(g.layout.unpack_dense(S.astype(np.float64), nx, N) - Sd).max() < tol * np.abs(Sd).max()
    assert np.abs(gam - gd).max() < tol * np.abs(gd).max()
    Sb = S.reshape(N, 3, nx * nx)
    assert not Sb[0, 0].any() and not Sb[N - 1, 2].any()
    # recovery: z from the exact multipliers equals the KKT solution
    K = np.block([[G, C.T], [C, np.zeros((N * nx, N * nx))]])
    sol = np.linalg.solve(K, np.concatenate([-gg, -c]))
    z, lam = sol[:G.shape[0]], sol[G.shape[0]:]
    x, u = oracle.schur_recover(Qi, Ri, d["A"].reshape(-1), d["B"].reshape(-1), d["q"].reshape(-1), d["r"].reshape(-1),
                                lam.astype(dt), nx, nu, N)
    zz = np.concatenate([np.concatenate([x[k * nx:(k + 1) * nx], u[k * nu:(k + 1) * nu]]) for k in range(N)])
    assert np.abs(zz - z).max() < 20 * tol * np.abs(z).max()


def test_whole_chain_on_the_host():
    """form -> stair preconditioner -> PCG -> recover, all oracle, against the dense KKT solution."""
    nx, nu, N = 14, 7, 32
    d = g.synth.lqr_stage_data(nx, nu, N, 2, dense_costs=True)
    S, gam, Qi, Ri = oracle.schur_form(*FLAT(d), nx, nu, N)
    P = oracle.build_pinv(S, nx, N, "stair")
    lam, it, ex = oracle.pcg(S, P, gam, np.zeros(N * nx), nx, N, 1e-12, 500)
    assert not ex
    x, u = oracle.schur_recover(Qi, Ri, d["A"].reshape(-1), d["B"].reshape(-1), d["q"].reshape(-1), d["r"].reshape(-1), lam, nx, nu, N)
    G, C, gg, c = g.synth.kkt_dense(d, nx, nu, N)
    K = np.block([[G, C.T], [C, np.zeros((N * nx, N * nx))]])
    z = np.linalg.solve(K, np.concatenate([-gg, -c]))[:G.shape[0]]
    zz = np.concatenate([np.concatenate([x[k * nx:(k + 1) * nx], u[k * nu:(k + 1) * nu]]) for k in range(N)])
    assert np.abs(zz - z).max() < 1e-6 * np.abs(z).max()


def example_stage_data(dt=np.float64):
    """The pendulum LQR the reference example's literal S decodes to (examples/pcg_solve.cu:14-22):
    A = [[1, 0.1], [-0.981, 1]], B = [0, 0.1]^T, Q^-1 = 0.999 I, R^-1 = 9.9, Q_N^-1 = 0.01 I, N = 3."""
    nx, nu, N = 2, 1, 3
    A = np.array([[1.0, 0.1], [-0.981, 1.0]])
    B = np.array([[0.0], [0.1]])
    cm = lambda Ms: np.stack([np.ascontiguousarray(M.T).reshape(-1) for M in Ms]).astype(dt)
    Q = [np.eye(2) / 0.999, np.eye(2) / 0.999, np.eye(2) / 0.01]
    return {"Q": cm(Q), "R": cm([np.eye(1) / 9.9] * 2), "A": cm([A, A]), "B": cm([B, B]),
            "q": np.zeros((N, nx), dt), "r": np.zeros((N - 1, nu), dt), "c": np.zeros((N, nx), dt)}, nx, nu, N


def test_reference_example_matrix_is_the_schur_complement_of_its_pendulum():
    """Known answer from the reference's own input: forming S from the decoded stage data reproduces the 36 literal
    numbers of examples/pcg_solve.cu:14-22 to the 4 digits they are printed with."""
    d, nx, nu, N = example_stage_data()
    S, gam, _, _ = oracle.schur_form(*FLAT(d), nx, nu, N)
    assert np.abs(S - g.synth.EXAMPLE_S).max() < 6e-5
    assert not gam.any()                                        # q = r = c = 0


def test_example_preconditioners_known_answer():
    """Phi^-1 of the reference example's S: oracle Gauss-Jordan against numpy.linalg.inv (synth.example_system)."""
    for kind in ("jacobi", "stair"):
        S, P, _ = g.synth.example_system(np.float64, kind)
        assert np.abs(oracle.build_pinv(S, 2, 3, kind) - P).max() < 1e-13
