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
