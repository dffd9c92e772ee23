"""Synthetic block-tridiagonal MPC systems (host, fp64) for parity tests and the bench.

The reference ships one 6x6 input (examples/pcg_solve.cu:14-25) and nothing else; upstream
(MPCGPU, cited at reference README.md:66-76) forms S, gamma and Phi^-1 on the GPU from LQR stage
data.  ``lqr_schur`` is a CPU restatement of that Schur-complement step with
the same sign convention as the example (S negative definite):

    S_00 = -Q^-1,  S_{k+1,k+1} = -(A_k Q^-1 A_k^T + B_k R^-1 B_k^T + Q_{k+1}^-1),  S_{k+1,k} = A_k Q^-1

Preconditioners (formulae from the cited paper, validated only through dense-inverse properties):
    identity | block-Jacobi  Phi^-1_kk = S_kk^-1
             | symmetric stair  + Phi^-1_{k,k+1} = -S_kk^-1 S_{k,k+1} S_{k+1,k+1}^-1 (and its transpose)
"""
from __future__ import annotations

import numpy as np

from . import layout

EXAMPLE_NX, EXAMPLE_N = 2, 3
# literal input of reference examples/pcg_solve.cu:14-25 (same numbers in pcg_solve_dp.cu)
EXAMPLE_S = np.array([0, 0, 0, 0,
                      -.999, 0, 0, -.999,
                      .999, .0999, -.98, .999,
                      .999, -.98, .0999, .999,
                      -2.008, .8801, .8801, -3.0584,
                      .999, .0999, -.98, .999,
                      .999, -.98, .0999, .999,
                      -1.019, .8801, .8801, -2.0694,
                      0, 0, 0, 0], dtype=np.float64)
EXAMPLE_GAMMA = np.array([3.1385, 0, 0, 3.0788, .0031, 3.0788], dtype=np.float64)


def lqr_blocks(nx: int, nu: int, N: int, seed: int, *, a=0.8, dt=0.05, qpos=2.0, qvel=0.5, r=0.1,
               qN=10.0, kscale=2.0, dscale=0.5):
    """Returns (diag [N,nx,nx], lower [N-1,nx,nx], gamma [N*nx]) in fp64.

    Draw order from numpy ``default_rng(seed)``: for k = 0..N-2: K (nq x nq), D (nq x nq), B (nq x nu);
    then gamma (N*nx).  (One flat ``standard_normal`` call split in that order is the same stream.)
    """
    assert nx % 2 == 0 and N >= 2
    nq = nx // 2
    rng = np.random.default_rng(seed)
    per_k = 2 * nq * nq + nq * nu
    z = rng.standard_normal((N - 1) * per_k + N * nx)
    zk = z[:(N - 1) * per_k].reshape(N - 1, per_k)
    K = zk[:, :nq * nq].reshape(N - 1, nq, nq) * kscale
    D = zk[:, nq * nq:2 * nq * nq].reshape(N - 1, nq, nq) * dscale
    Bq = zk[:, 2 * nq * nq:].reshape(N - 1, nq, nu) * dt
    gamma = z[(N - 1) * per_k:].copy()

    A = np.zeros((N - 1, nx, nx))
    A[:] = a * np.eye(nx)
    A[:, :nq, nq:] += dt * np.eye(nq)
    A[:, nq:, :nq] += dt * K
    A[:, nq:, nq:] += dt * D
    B = np.zeros((N - 1, nx, nu))
    B[:, nq:, :] = Bq

    q = np.concatenate([np.full(nq, qpos), np.full(nx - nq, qvel)])
    Qinv = np.diag(1.0 / q)
    QNinv = np.diag(1.0 / (qN * q))
    Rinv = np.eye(nu) / r

    diag = np.zeros((N, nx, nx))
    diag[0] = -Qinv
    AQ = A @ Qinv                                    # [N-1, nx, nx] = S_{k+1,k}
    core = AQ @ np.swapaxes(A, 1, 2) + B @ Rinv @ np.swapaxes(B, 1, 2)
    diag[1:] = -(core + Qinv)
    diag[N - 1] = -(core[N - 2] + QNinv)
    lower = AQ
    return diag, lower, gamma


def lqr_stage_data(nx: int, nu: int, N: int, seed: int, dtype=np.float64, dense_costs: bool = False, **kw):
    """Stage-wise LQR data of the same family as ``lqr_blocks`` in the layout of gbdpcg_schur_form_dev_*
    (include/gbdpcg.h): Q [N, nx*nx], R [N-1, nu*nu], A [N-1, nx*nx], B [N-1, nx*nu] (each block column-major, flattened),
    q, c [N, nx], r [N-1, nu].  With ``dense_costs=False`` the Schur complement of these data IS ``lqr_blocks``' S
    (diagonal Q, R; terminal weight qN); ``dense_costs=True`` rotates Q_k, R_k into dense SPD matrices.
    The linear terms q, r, c are extra draws from a second stream (seed + 2**31)."""
    a, dt, qpos, qvel, rr, qN = (kw.get(k, d) for k, d in (("a", 0.8), ("dt", 0.05), ("qpos", 2.0), ("qvel", 0.5), ("r", 0.1), ("qN", 10.0)))
    kscale, dscale = kw.get("kscale", 2.0), kw.get("dscale", 0.5)
    nq = nx // 2
    rng = np.random.default_rng(seed)
    per_k = 2 * nq * nq + nq * nu
    z = rng.standard_normal((N - 1) * per_k + N * nx)
    zk = z[:(N - 1) * per_k].reshape(N - 1, per_k)
    K = zk[:, :nq * nq].reshape(N - 1, nq, nq) * kscale
    D = zk[:, nq * nq:2 * nq * nq].reshape(N - 1, nq, nq) * dscale
    Bq = zk[:, 2 * nq * nq:].reshape(N - 1, nq, nu) * dt
    A = np.zeros((N - 1, nx, nx))
    A[:] = a * np.eye(nx)
    A[:, :nq, nq:] += dt * np.eye(nq)
    A[:, nq:, :nq] += dt * K
    A[:, nq:, nq:] += dt * D
    B = np.zeros((N - 1, nx, nu))
    B[:, nq:, :] = Bq
    qd = np.concatenate([np.full(nq, qpos), np.full(nx - nq, qvel)])
    Q = np.stack([np.diag(qd)] * (N - 1) + [np.diag(qN * qd)])
    R = np.stack([np.eye(nu) * rr] * (N - 1))
    rng2 = np.random.default_rng(seed + 2 ** 31)
    if dense_costs:
        for k in range(N):
            U, _ = np.linalg.qr(rng2.standard_normal((nx, nx)))
            Q[k] = U @ Q[k] @ U.T
            Q[k] = 0.5 * (Q[k] + Q[k].T)
        for k in range(N - 1):
            U, _ = np.linalg.qr(rng2.standard_normal((nu, nu)))
            R[k] = U @ (R[k] * (1.0 + np.arange(nu))) @ U.T
            R[k] = 0.5 * (R[k] + R[k].T)
    q = rng2.standard_normal((N, nx))
    r = rng2.standard_normal((N - 1, nu))
    c = rng2.standard_normal((N, nx))
    cm = lambda M: np.ascontiguousarray(np.swapaxes(M, 1, 2)).reshape(M.shape[0], -1).astype(dtype)     # column-major blocks
    return {"Q": cm(Q), "R": cm(R), "A": cm(A), "B": cm(B), "q": q.astype(dtype), "r": r.astype(dtype), "c": c.astype(dtype)}


def kkt_dense(data, nx: int, nu: int, N: int):
    """Dense fp64 KKT matrix pieces (G, C, g, c) of the stage data, z ordered (x_0, u_0, x_1, ..., x_{N-1})."""
    Q = data["Q"].astype(np.float64).reshape(N, nx, nx).swapaxes(1, 2)
    R = data["R"].astype(np.float64).reshape(N - 1, nu, nu).swapaxes(1, 2)
    A = data["A"].astype(np.float64).reshape(N - 1, nx, nx).swapaxes(1, 2)
    B = data["B"].astype(np.float64).reshape(N - 1, nu, nx).swapaxes(1, 2)
    nz = N * nx + (N - 1) * nu
    G = np.zeros((nz, nz))
    g = np.zeros(nz)
    C = np.zeros((N * nx, nz))
    xo = lambda k: k * (nx + nu)
    uo = lambda k: k * (nx + nu) + nx
    for k in range(N):
        G[xo(k):xo(k) + nx, xo(k):xo(k) + nx] = Q[k]
        g[xo(k):xo(k) + nx] = data["q"][k]
        C[k * nx:(k + 1) * nx, xo(k):xo(k) + nx] = np.eye(nx)
        if k + 1 < N:
            G[uo(k):uo(k) + nu, uo(k):uo(k) + nu] = R[k]
            g[uo(k):uo(k) + nu] = data["r"][k]
            C[(k + 1) * nx:(k + 2) * nx, xo(k):xo(k) + nx] = -A[k]
            C[(k + 1) * nx:(k + 2) * nx, uo(k):uo(k) + nu] = -B[k]
    return G, C, g, data["c"].astype(np.float64).reshape(-1)


def preconditioner_blocks(diag: np.ndarray, lower: np.ndarray, kind: str):
    """Phi^-1 blocks (diag, lower) in fp64 for kind in {'identity','jacobi','stair'}."""
    N, nx, _ = diag.shape
    if kind == "identity":
        return np.broadcast_to(np.eye(nx), (N, nx, nx)).copy(), np.zeros((N - 1, nx, nx))
    Dinv = np.linalg.inv(diag)
    if kind == "jacobi":
        return Dinv, np.zeros((N - 1, nx, nx))
    if kind == "stair":
        # Phi^-1_{k+1,k} = -S_{k+1,k+1}^-1 S_{k+1,k} S_kk^-1 ; upper = transpose (S symmetric)
        low = -(Dinv[1:] @ lower @ Dinv[:-1])
        return Dinv, low
    raise ValueError(kind)


def lqr_schur(nx: int, nu: int, N: int, seed: int, precond: str = "stair", dtype=np.float32, **kw):
    """One system in the compressed layout: (S, Pinv, gamma) cast to ``dtype`` (built in fp64)."""
    diag, lower, gamma = lqr_blocks(nx, nu, N, seed, **kw)
    S = layout.pack_blocks(diag, lower)
    pd, pl = preconditioner_blocks(diag, lower, precond)
    P = layout.pack_blocks(pd, pl)
    return S.astype(dtype), P.astype(dtype), gamma.astype(dtype)


def lqr_schur_batch(nx: int, nu: int, N: int, seeds, precond: str = "stair", dtype=np.float32, **kw):
    """Batch of systems: S [B, 3*N*nx^2], Pinv [B, 3*N*nx^2], gamma [B, N*nx] (problem-major slabs)."""
    seeds = list(seeds)
    S = np.empty((len(seeds), layout.mat_elems(nx, N)), dtype=dtype)
    P = np.empty_like(S)
    g = np.empty((len(seeds), layout.vec_elems(nx, N)), dtype=dtype)
    for i, s in enumerate(seeds):
        S[i], P[i], g[i] = lqr_schur(nx, nu, N, s, precond, dtype, **kw)
    return S, P, g


def example_system(dtype=np.float32, precond: str = "identity"):
    """The reference example's system (C1) with a *defined* preconditioner."""
    nx, N = EXAMPLE_NX, EXAMPLE_N
    A = layout.unpack_dense(EXAMPLE_S, nx, N)
    diag = np.stack([A[k * nx:(k + 1) * nx, k * nx:(k + 1) * nx] for k in range(N)])
    lower = np.stack([A[(k + 1) * nx:(k + 2) * nx, k * nx:(k + 1) * nx] for k in range(N - 1)])
    pd, pl = preconditioner_blocks(diag, lower, precond)
    P = layout.pack_blocks(pd, pl)
    return EXAMPLE_S.astype(dtype), P.astype(dtype), EXAMPLE_GAMMA.astype(dtype)


# BASELINE.json configs: name -> (nx, nu, N, dtype, exit_tol, max_iter)
CONFIGS = {
    "C1": (2, 1, 3, np.float32, 1e-6, 25),
    "C2": (14, 7, 32, np.float32, 1e-6, 1000),
    "C3": (14, 7, 64, np.float64, 1e-10, 2000),
    "C4": (12, 4, 512, np.float32, 1e-6, 5000),
}
