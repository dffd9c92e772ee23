"""ctypes door onto the host oracle (oracle/pcg_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs may import
this module; the product package (gbd_pcg_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_pcg.so")
_lib = None


def build(force: bool = False) -> str:
    src = [os.path.join(_HERE, f) for f in ("pcg_oracle.c", "pcg_oracle_impl.h")]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "_build/liboracle_pcg.so"], stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        for suf, ct in (("f32", C.c_float), ("f64", C.c_double)):
            f = getattr(_lib, f"oracle_pcg_{suf}")
            f.restype = C.c_int
            f.argtypes = [C.c_void_p] * 4 + [C.c_uint32, C.c_uint32, ct, C.c_uint32, C.c_int,
                                             C.POINTER(C.c_uint32), C.POINTER(C.c_int), C.c_void_p]
            g = getattr(_lib, f"oracle_pcg_batched_{suf}")
            g.restype = C.c_int
            g.argtypes = [C.c_void_p] * 4 + [C.c_uint32, C.c_uint32, C.c_uint32, ct, C.c_uint32, C.c_int,
                                             C.c_void_p, C.c_void_p, C.c_int]
            h = getattr(_lib, f"oracle_bdmv_{suf}")
            h.restype = None
            h.argtypes = [C.c_void_p] * 3 + [C.c_uint32, C.c_uint32]
            sf = getattr(_lib, f"oracle_schur_form_{suf}")
            sf.restype = C.c_int
            sf.argtypes = [C.c_void_p] * 7 + [C.c_uint32] * 3 + [C.c_void_p] * 4
            sr = getattr(_lib, f"oracle_schur_recover_{suf}")
            sr.restype = C.c_int
            sr.argtypes = [C.c_void_p] * 7 + [C.c_uint32] * 3 + [C.c_void_p] * 2
            b = getattr(_lib, f"oracle_build_pinv_{suf}")
            b.restype = C.c_int
            b.argtypes = [C.c_void_p] * 2 + [C.c_uint32, C.c_uint32, C.c_int]
    return _lib


def _suf(dtype):
    dtype = np.dtype(dtype)
    if dtype == np.float32:
        return "f32"
    if dtype == np.float64:
        return "f64"
    raise TypeError(dtype)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def bdmv(M, x, nx, N):
    M = np.ascontiguousarray(M)
    x = np.ascontiguousarray(x, dtype=M.dtype)
    y = np.empty_like(x)
    getattr(lib(), f"oracle_bdmv_{_suf(M.dtype)}")(_ptr(M), _ptr(x), _ptr(y), nx, N)
    return y


def pcg(S, Pinv, gamma, lambda0, nx, N, exit_tol, max_iter, order=0, trace=False):
    """Returns (lambda, iters, max_iter_exit[, eta_trace]).  Pinv=None -> identity."""
    S = np.ascontiguousarray(S)
    dt = S.dtype
    Pinv = None if Pinv is None else np.ascontiguousarray(Pinv, dtype=dt)
    gamma = np.ascontiguousarray(gamma, dtype=dt)
    lam = np.array(lambda0, dtype=dt, copy=True)
    it, ex = C.c_uint32(0), C.c_int(0)
    tr = np.full(max_iter + 1, np.nan, dtype=dt) if trace else None
    rc = getattr(lib(), f"oracle_pcg_{_suf(dt)}")(_ptr(S), _ptr(Pinv), _ptr(gamma), _ptr(lam), nx, N,
                                                  exit_tol, max_iter, order, C.byref(it), C.byref(ex), _ptr(tr))
    if rc != 0:
        raise ValueError("oracle_pcg: bad arguments")
    out = (lam, int(it.value), bool(ex.value))
    return out + (tr[:it.value + 1],) if trace else out


def pcg_batched(S, Pinv, gamma, lambda0, nx, N, exit_tol, max_iter, order=0, threads=1):
    """Problem-major slabs [B, ...].  Returns (lambda [B, N*nx], iters [B], exits [B])."""
    S = np.ascontiguousarray(S)
    dt = S.dtype
    B = S.shape[0]
    Pinv = None if Pinv is None else np.ascontiguousarray(Pinv, dtype=dt)
    gamma = np.ascontiguousarray(gamma, dtype=dt)
    lam = np.array(lambda0, dtype=dt, copy=True)
    iters = np.zeros(B, dtype=np.uint32)
    exits = np.zeros(B, dtype=np.uint8)
    getattr(lib(), f"oracle_pcg_batched_{_suf(dt)}")(_ptr(S), _ptr(Pinv), _ptr(gamma), _ptr(lam), nx, N, B,
                                                     exit_tol, max_iter, order, _ptr(iters), _ptr(exits), threads)
    return lam, iters, exits.astype(bool)


def build_pinv(S, nx, N, kind):
    """Phi^-1 from S in the compressed layout; kind 0/1/2 or 'identity'/'jacobi'/'stair'
    (restates glass::invertMatrix, reference GLASS/src/L3/inv.cuh:6-37, plus the stair products)."""
    kind = {"identity": 0, "jacobi": 1, "stair": 2}.get(kind, kind)
    S = np.ascontiguousarray(S)
    P = np.empty_like(S)
    per = 3 * N * nx * nx
    Sf, Pf = S.reshape(-1, per), P.reshape(-1, per)
    for i in range(Sf.shape[0]):
        rc = getattr(lib(), f"oracle_build_pinv_{_suf(S.dtype)}")(_ptr(Sf[i]), _ptr(Pf[i]), nx, N, int(kind))
        if rc != 0:
            raise ValueError("oracle_build_pinv: bad arguments")
    return P


def schur_form(Q, R, A, B, q, r, c, nx, nu, N):
    """One system: stage data (column-major matrices, see include/gbdpcg.h) -> (S, gamma, Qinv, Rinv)."""
    dt = np.asarray(Q).dtype
    arrs = [np.ascontiguousarray(a, dtype=dt) for a in (Q, R, A, B, q, r, c)]
    S = np.empty(3 * N * nx * nx, dt)
    gamma = np.empty(N * nx, dt)
    Qinv = np.empty(N * nx * nx, dt)
    Rinv = np.empty((N - 1) * nu * nu, dt)
    rc = getattr(lib(), f"oracle_schur_form_{_suf(dt)}")(*[_ptr(a) for a in arrs], nx, nu, N, _ptr(S), _ptr(gamma), _ptr(Qinv), _ptr(Rinv))
    if rc != 0:
        raise ValueError("oracle_schur_form: bad arguments")
    return S, gamma, Qinv, Rinv


def schur_recover(Qinv, Rinv, A, B, q, r, lam, nx, nu, N):
    dt = np.asarray(Qinv).dtype
    arrs = [np.ascontiguousarray(a, dtype=dt) for a in (Qinv, Rinv, A, B, q, r, lam)]
    x = np.empty(N * nx, dt)
    u = np.empty((N - 1) * nu, dt)
    rc = getattr(lib(), f"oracle_schur_recover_{_suf(dt)}")(*[_ptr(a) for a in arrs], nx, nu, N, _ptr(x), _ptr(u))
    if rc != 0:
        raise ValueError("oracle_schur_recover: bad arguments")
    return x, u


def pcg_numpy(S, Pinv, gamma, lambda0, nx, N, exit_tol, max_iter):
    """Independent dense-numpy statement of the same recurrence, used to
    cross-check the C oracle.  Arithmetic in the dtype of S."""
    from gbd_pcg_b200 import layout
    dt = np.asarray(S).dtype
    A = layout.unpack_dense(S, nx, N)
    P = np.eye(N * nx, dtype=dt) if Pinv is None else layout.unpack_dense(np.asarray(Pinv, dtype=dt), nx, N)
    lam = np.array(lambda0, dtype=dt, copy=True)
    g = np.asarray(gamma, dtype=dt)
    r = g - A @ lam
    rt = P @ r
    p = rt.copy()
    eta = dt.type(r @ rt)
    it, ex = 0, True
    for it in range(max_iter):
        ups = A @ p
        alpha = dt.type(eta / dt.type(p @ ups))
        lam = lam + alpha * p
        r = r - alpha * ups
        rt = P @ r
        eta_new = dt.type(r @ rt)
        if abs(eta_new) < exit_tol:
            it += 1
            ex = False
            break
        beta = dt.type(eta_new / eta)
        eta = eta_new
        p = rt + beta * p
    else:
        it = max_iter
    return lam, it, ex
