"""Compressed block-tridiagonal layout used for S and Phi^-1.

The reference never names the type (README's ``cbtd_t`` does not exist); the layout is defined by
how its kernel indexes the arrays: reference include/pcg.cuh:104-110 (block-row ``k`` at offset
``k*3*nx^2``), include/utils.cuh:58-84 (``mat[nx*c + r]``: the block-row is one nx x 3nx
column-major panel [left | diag | right]) and the literal arrays of examples/pcg_solve.cu:14-22.
The left block of row 0 and the right block of row N-1 are padding and are never read.

Dense equivalent:  A[k*nx + r, (k+j-1)*nx + c] = M[k*3nx^2 + j*nx^2 + c*nx + r],  j in {0,1,2}.
"""
from __future__ import annotations

import numpy as np


def mat_elems(nx: int, N: int) -> int:
    return 3 * N * nx * nx


def vec_elems(nx: int, N: int) -> int:
    return N * nx


def pack_dense(A: np.ndarray, nx: int, N: int, pad_value: float = 0.0) -> np.ndarray:
    """Dense (N*nx)^2 block-tridiagonal matrix -> flat compressed array (dtype preserved)."""
    A = np.asarray(A)
    assert A.shape == (N * nx, N * nx)
    out = np.full((N, 3, nx, nx), pad_value, dtype=A.dtype)  # [k, j, c, r]  (column-major blocks)
    for k in range(N):
        for j in range(3):
            kc = k + j - 1
            if 0 <= kc < N:
                out[k, j] = A[k * nx:(k + 1) * nx, kc * nx:(kc + 1) * nx].T
    return out.reshape(-1)


def pack_blocks(diag: np.ndarray, lower: np.ndarray, upper: np.ndarray | None = None,
                pad_value: float = 0.0) -> np.ndarray:
    """Blocks -> flat compressed array.

    diag  [N, nx, nx]    block (k, k)
    lower [N-1, nx, nx]  block (k+1, k)
    upper [N-1, nx, nx]  block (k, k+1)   (default: lower^T, the symmetric case)
    """
    N, nx, _ = diag.shape
    if upper is None:
        upper = np.swapaxes(lower, 1, 2)
    out = np.full((N, 3, nx, nx), pad_value, dtype=diag.dtype)
    out[:, 1] = np.swapaxes(diag, 1, 2)
    out[1:, 0] = np.swapaxes(lower, 1, 2)
    out[:-1, 2] = np.swapaxes(upper, 1, 2)
    return out.reshape(-1)


def unpack_dense(M: np.ndarray, nx: int, N: int) -> np.ndarray:
    """Flat compressed array -> dense matrix (padding blocks ignored)."""
    M = np.asarray(M).reshape(N, 3, nx, nx)
    A = np.zeros((N * nx, N * nx), dtype=M.dtype)
    for k in range(N):
        for j in range(3):
            kc = k + j - 1
            if 0 <= kc < N:
                A[k * nx:(k + 1) * nx, kc * nx:(kc + 1) * nx] = M[k, j].T
    return A


def identity_compressed(nx: int, N: int, dtype=np.float32) -> np.ndarray:
    """Identity matrix in the compressed layout (what ``empty_pinv`` means for Phi^-1)."""
    out = np.zeros((N, 3, nx, nx), dtype=dtype)
    out[:, 1] = np.eye(nx, dtype=dtype)
    return out.reshape(-1)


def algorithmic_bytes(nx: int, N: int, itemsize: int, with_pinv: bool = True) -> int:
    """Bytes one solve must move: non-padding blocks of S (and Phi^-1) read once, gamma and
    lambda_0 read, lambda written."""
    mats = 2 if with_pinv else 1
    return itemsize * (mats * (3 * N - 2) * nx * nx + 3 * N * nx)
