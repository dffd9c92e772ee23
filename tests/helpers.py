"""Shared helpers for the parity tests (test infrastructure: may import oracle/)."""
import numpy as np

from gbd_pcg_b200 import synth
from oracle import oracle

PRECONDS = ("identity", "jacobi", "stair")


def make_system(name_or_shape, precond, seed=0, dtype=None):
    """(S, Pinv_or_None, gamma, nx, N, tol, max_iter) for a BASELINE config name or an (nx, nu, N, dtype, tol, max_iter) tuple."""
    cfg = synth.CONFIGS[name_or_shape] if isinstance(name_or_shape, str) else name_or_shape
    nx, nu, N, dt, tol, max_iter = cfg
    dt = dtype or dt
    if name_or_shape == "C1":
        S, P, g = synth.example_system(dt, precond)
    else:
        S, P, g = synth.lqr_schur(nx, nu, N, seed, precond, dt)
    return S, (None if precond == "identity" else P), g, nx, N, tol, max_iter


def oracle_at_iters(S, P, g, lam0, nx, N, iters):
    """lambda after exactly `iters` oracle iterations (exit test disabled) -- 'equal iteration count' comparisons."""
    lam, it, _ = oracle.pcg(S, P, g, lam0, nx, N, 0.0, iters, order=0)
    assert it == iters
    return lam


def random_block_tridiagonal(nx, N, precond, dtype, seed=11, coupled=False):
    """(S, Pinv_or_None, gamma) for ANY state size (the LQR generator needs an even one), packed into the compressed layout;
    preconditioner from the fp64 blocks.  Default: dense random negative-definite diagonal blocks with weak random
    couplings (a few iterations).  coupled=True: A = -(G G^T + 0.05 I) with G random block-bidiagonal -- negative definite
    by construction, strongly coupled neighbours, tens to hundreds of iterations without a preconditioner."""
    from gbd_pcg_b200 import layout
    rng = np.random.default_rng(seed)
    n = nx * N
    A = np.zeros((n, n))
    if coupled:
        G = np.zeros((n, n))
        for k in range(N):
            G[k*nx:(k+1)*nx, k*nx:(k+1)*nx] = np.eye(nx) + 0.3 * rng.standard_normal((nx, nx)) / np.sqrt(nx)
            if k + 1 < N:
                G[(k+1)*nx:(k+2)*nx, k*nx:(k+1)*nx] = -0.9 * np.eye(nx) + 0.3 * rng.standard_normal((nx, nx)) / np.sqrt(nx)
        A = -(G @ G.T + 0.05 * np.eye(n))
    for k in range(N if not coupled else 0):
        M = rng.standard_normal((nx, nx))
        A[k*nx:(k+1)*nx, k*nx:(k+1)*nx] = -(M @ M.T + 4 * nx * np.eye(nx))
        if k + 1 < N:
            L = rng.standard_normal((nx, nx))
            A[(k+1)*nx:(k+2)*nx, k*nx:(k+1)*nx] = L
            A[k*nx:(k+1)*nx, (k+1)*nx:(k+2)*nx] = L.T
    S = layout.pack_dense(A, nx, N).astype(dtype)
    diag = np.stack([A[k*nx:(k+1)*nx, k*nx:(k+1)*nx] for k in range(N)])
    lower = np.stack([A[(k+1)*nx:(k+2)*nx, k*nx:(k+1)*nx] for k in range(N - 1)])
    pd, pl = synth.preconditioner_blocks(diag, lower, precond)
    P = None if precond == "identity" else layout.pack_blocks(pd, pl).astype(dtype)
    return S, P, rng.standard_normal(n).astype(dtype)
