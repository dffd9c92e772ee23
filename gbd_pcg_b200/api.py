"""Python mirror of the reference host interface, over the C ABI (include/gbdpcg.h).

Same names and argument meaning as the reference's templates so that tests read like calls into
the reference:

    solvePCG(h_S, h_gamma, h_lambda, stateSize, knotPoints, config)        reference include/interface.cuh:24-89
    solvePCG(state_size, knot_points, d_S, d_Pinv, d_gamma, d_lambda,
             d_r, d_p, d_v_temp, d_eta_new_temp, config)                    reference include/interface.cuh:92-144
    pcg_solve(...)                                                          reference README.md:42 (same as the first form)
    pcg_config(exit_tol, max_iter, grid, block, empty_pinv)                 reference include/types.cuh:18-35

Host arrays are numpy; device arrays are anything with ``data_ptr()`` (torch CUDA tensors).
PyTorch is used only as a device-memory / stream provider; all arithmetic is in libgbdpcg.so.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import Config, GbdpcgError, PlanInfo  # noqa: F401  (re-exported)

DEFAULT_MAX_PCG_ITER = 25      # reference include/constants.cuh:15
DEFAULT_EPSILON = 1e-6         # reference include/constants.cuh:17
DEFAULT_GRID = (128, 1, 1)     # reference include/constants.cuh:18 (carried, ignored -- as in the reference)
DEFAULT_BLOCK = (64, 1, 1)     # reference include/constants.cuh:19 (carried, ignored -- as in the reference)


@dataclass
class pcg_config:
    """Field-for-field mirror of reference ``pcg_config<T>`` (include/types.cuh:18-35)."""
    pcg_exit_tol: float = DEFAULT_EPSILON
    pcg_max_iter: int = DEFAULT_MAX_PCG_ITER
    pcg_grid: tuple = DEFAULT_GRID
    pcg_block: tuple = DEFAULT_BLOCK
    empty_pinv: int = 1
    # extensions (not in the reference): what the host overload without a Phi^-1 argument builds on the device when
    # empty_pinv == 0 ("stair" | "jacobi" | "identity"), and the launch tuning knobs of the C ABI
    pinv_build: str = field(default="stair", repr=False)
    flags: int = field(default=0, repr=False)
    cluster: int = field(default=0, repr=False)

    def to_c(self) -> Config:
        return Config(float(self.pcg_exit_tol), int(self.pcg_max_iter), int(self.empty_pinv), int(self.flags), int(self.cluster))


@dataclass
class csr_t:
    """Mirror of reference ``csr_t<T>`` (include/types.cuh:7-15): uint32 row_ptr[rows+1], col_ind[nnz], T val[nnz]."""
    row_ptr: np.ndarray
    col_ind: np.ndarray
    val: np.ndarray
    rows: int
    cols: int
    nnz: int

    @staticmethod
    def from_dense(A, dtype=None, keep_zeros=False):
        A = np.asarray(A)
        mask = np.ones_like(A, dtype=bool) if keep_zeros else (A != 0)
        counts = mask.sum(axis=1)
        row_ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
        rr, cc = np.nonzero(mask)
        return csr_t(row_ptr, cc.astype(np.uint32), np.ascontiguousarray(A[rr, cc], dtype=dtype or A.dtype), A.shape[0], A.shape[1], int(rr.size))


def _suffix(dtype) -> str:
    dtype = np.dtype(dtype)
    if dtype == np.float32:
        return "f32"
    if dtype == np.float64:
        return "f64"
    raise TypeError(f"T must be float or double (reference templates), got {dtype}")


def _host(a, dtype, n, name, writable=False):
    if not isinstance(a, np.ndarray) or a.dtype != dtype or not a.flags.c_contiguous or a.size < n:
        raise ValueError(f"{name}: need a C-contiguous numpy array of {np.dtype(dtype)} with at least {n} elements")
    if writable and not a.flags.writeable:
        raise ValueError(f"{name}: must be writable")
    return a.ctypes.data_as(C.c_void_p)


def _dev_ptr(t, name, nullable=False):
    if t is None:
        if nullable:
            return None
        raise ValueError(f"{name} must not be None")
    if isinstance(t, int):
        return C.c_void_p(t)
    if hasattr(t, "is_cuda"):
        if not t.is_cuda or not t.is_contiguous():
            raise ValueError(f"{name}: need a contiguous CUDA tensor")
        return C.c_void_p(t.data_ptr())
    raise TypeError(f"{name}: need a CUDA tensor or an integer device address")


def _torch_dtype_suffix(t):
    import torch
    if t.dtype == torch.float32:
        return "f32"
    if t.dtype == torch.float64:
        return "f64"
    raise TypeError("T must be float32 or float64")


def _stream_ptr(stream):
    if stream is None:
        import torch
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if isinstance(stream, int):
        return C.c_void_p(stream)
    return C.c_void_p(stream.cuda_stream)


class workspace:
    """Per-caller workspace handle (``gbdpcg_workspace_create`` / ``_destroy``, include/gbdpcg.h): cached device buffers,
    pinned scalars, two streams.  Replaces the 20 allocator calls per solve of the reference host path
    (include/interface.cuh:39-53,79-86,105-108,140-141).  One workspace per host thread lets callers solve concurrently;
    ``workspace=None`` in any call selects the process-wide default workspace of the current device (internally locked).
    Use as a context manager or call ``close()``."""

    def __init__(self):
        self._h = C.c_void_p(None)
        _lib.check(_lib.lib().gbdpcg_workspace_create(C.byref(self._h)))

    @property
    def handle(self):
        if not self._h:
            raise ValueError("workspace is closed")
        return self._h

    def close(self):
        if self._h:
            _lib.lib().gbdpcg_workspace_destroy(self._h)
            self._h = C.c_void_p(None)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001  (interpreter shutdown)
            pass


def _ws(w):
    return None if w is None else w.handle


# ------------------------------------------------------------------------------------------------
# reference-shaped entry points
# ------------------------------------------------------------------------------------------------
def solvePCG(*args, **kw):
    """Both reference overloads, selected like C++ would: 6 arguments = host pointers, 11 = device pointers."""
    if args and isinstance(args[0], csr_t):
        return _solve_csr(*args, **kw)
    if args and isinstance(args[0], np.ndarray):
        return _solve_host(*args, **kw)
    if args and isinstance(args[0], (int, np.integer)):
        return _solve_dev(*args, **kw)
    raise TypeError("solvePCG: unrecognised overload (first argument must be h_S or state_size)")


def pcg_solve(h_S, h_gamma, h_lambda, stateSize, knotPoints, config=None, h_Pinv=None, workspace=None):
    """README.md:42 spelling; config defaults like the README's ``default_config``."""
    return _solve_host(h_S, h_gamma, h_lambda, stateSize, knotPoints, config, h_Pinv=h_Pinv, workspace=workspace)


def _solve_host(h_S, h_gamma, h_lambda, stateSize, knotPoints, config=None, h_Pinv=None, return_flag=False, workspace=None):
    """Host-pointer overload.  ``h_lambda`` is updated in place; returns the iteration count
    (the reference returns 1 unconditionally, include/interface.cuh:88 -- a defect this build does not keep).
    ``h_Pinv`` (extension) supplies Phi^-1; it is used only when ``config.empty_pinv == 0``.  Without it,
    ``empty_pinv == 0`` makes the library form ``config.pinv_build`` from S on the device."""
    config = config or pcg_config()
    suf = _suffix(h_S.dtype)
    nx, N = int(stateSize), int(knotPoints)
    ms, vs = 3 * N * nx * nx, N * nx
    cfg = config.to_c()
    if h_Pinv is None:
        if not config.empty_pinv and config.pinv_build in BUILD_FLAGS:
            # no Phi^-1 argument (the reference's own signature): it is formed from S on the device.  The reference
            # prints a warning here and solves with an uninitialised d_Pinv (interface.cuh:33-34,45-46).
            cfg.flags |= BUILD_FLAGS[config.pinv_build]
        elif not (cfg.flags & (_lib.FLAG_BUILD_JACOBI | _lib.FLAG_BUILD_STAIR)):
            cfg.empty_pinv = 1
    it, ex = C.c_uint32(0), C.c_int32(0)
    f = getattr(_lib.lib(), f"gbdpcg_solve_host_{suf}")
    _lib.check(f(_ws(workspace), _host(h_S, h_S.dtype, ms, "h_S"),
                 None if h_Pinv is None else _host(h_Pinv, h_S.dtype, ms, "h_Pinv"),
                 _host(h_gamma, h_S.dtype, vs, "h_gamma"), _host(h_lambda, h_S.dtype, vs, "h_lambda", True),
                 nx, N, C.byref(cfg), C.byref(it), C.byref(ex)))
    return (it.value, bool(ex.value)) if return_flag else it.value


def _solve_csr(h_S, h_Pinv, h_gamma, h_lambda, stateSize, knotPoints, config=None, return_flag=False):
    """CSR overload (reference include/interface.cuh:8-20 is a stub: "NOT IMPLEMENTED", exit 12).  ``h_S`` / ``h_Pinv`` are
    ``csr_t``; ``h_Pinv`` may be None (then ``config.empty_pinv == 0`` forms ``config.pinv_build`` on the device)."""
    config = config or pcg_config()
    dt = h_S.val.dtype
    suf = _suffix(dt)
    nx, N = int(stateSize), int(knotPoints)
    if h_S.rows != nx * N or h_S.cols != nx * N:
        raise ValueError("CSR S must be square with stateSize*knotPoints rows")
    cfg = config.to_c()
    if h_Pinv is None:
        if not config.empty_pinv and config.pinv_build in BUILD_FLAGS:
            cfg.flags |= BUILD_FLAGS[config.pinv_build]
        elif not (cfg.flags & (_lib.FLAG_BUILD_JACOBI | _lib.FLAG_BUILD_STAIR)):
            cfg.empty_pinv = 1

    def arrs(m):
        if m is None:
            return [None, None, None, 0]
        rp = np.ascontiguousarray(m.row_ptr, dtype=np.uint32)
        ci = np.ascontiguousarray(m.col_ind, dtype=np.uint32)
        v = np.ascontiguousarray(m.val, dtype=dt)
        if rp.size != m.rows + 1 or ci.size < m.nnz or v.size < m.nnz:
            raise ValueError("inconsistent CSR arrays")
        keep.extend([rp, ci, v])
        return [rp.ctypes.data_as(C.c_void_p), ci.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p), int(m.nnz)]
    keep = []
    it, ex = C.c_uint32(0), C.c_int32(0)
    f = getattr(_lib.lib(), f"gbdpcg_solve_csr_host_{suf}")
    _lib.check(f(None, *arrs(h_S), *arrs(h_Pinv), _host(h_gamma, dt, nx * N, "h_gamma"), _host(h_lambda, dt, nx * N, "h_lambda", True),
                 nx, N, C.byref(cfg), C.byref(it), C.byref(ex)))
    return (it.value, bool(ex.value)) if return_flag else it.value


def pack_csr_dev(d_row_ptr, d_col_ind, d_val, rows, state_size, knot_points, d_out, d_status=None, stream=None, nnz=None):
    """Device CSR -> compressed block-tridiagonal layout (asynchronous).  ``d_status`` (int32[1]) gets bit 0 set when an
    entry lies outside the band, bit 1 when row_ptr decreases or runs past ``nnz`` (default: the length of ``d_val``);
    such rows are clamped, never read out of bounds."""
    suf = _torch_dtype_suffix(d_val)
    if nnz is None:
        nnz = int(d_val.numel())
    f = getattr(_lib.lib(), f"gbdpcg_pack_csr_dev_{suf}")
    _lib.check(f(None, _dev_ptr(d_row_ptr, "d_row_ptr"), _dev_ptr(d_col_ind, "d_col_ind"), _dev_ptr(d_val, "d_val"), int(nnz), int(rows),
                 int(state_size), int(knot_points), _dev_ptr(d_out, "d_out"), _dev_ptr(d_status, "d_status", True), _stream_ptr(stream)))


def _solve_dev(state_size, knot_points, d_S, d_Pinv, d_gamma, d_lambda, d_r=None, d_p=None, d_v_temp=None,
               d_eta_new_temp=None, config=None, stream=None, return_flag=False, workspace=None):
    """Device-pointer overload (blocking, returns iterations like include/interface.cuh:136,143).
    The four scratch buffers of the reference are accepted and ignored.  ``d_Pinv`` is always applied,
    whatever ``config.empty_pinv`` says, because the reference kernel has no notion of that flag
    (it is passed as a 13th launch argument to a 12-parameter kernel, interface.cuh:126 vs pcg.cuh:56-68);
    ``d_Pinv=None`` (extension) selects the identity."""
    config = config or pcg_config()
    suf = _torch_dtype_suffix(d_S)
    cfg = config.to_c()
    cfg.empty_pinv = 1 if d_Pinv is None else 0
    it, ex = C.c_uint32(0), C.c_int32(0)
    f = getattr(_lib.lib(), f"gbdpcg_solve_dev_sync_{suf}")
    _lib.check(f(_ws(workspace), _dev_ptr(d_S, "d_S"), _dev_ptr(d_Pinv, "d_Pinv", True), _dev_ptr(d_gamma, "d_gamma"),
                 _dev_ptr(d_lambda, "d_lambda"), int(state_size), int(knot_points), C.byref(cfg),
                 C.byref(it), C.byref(ex), _stream_ptr(stream)))
    return (it.value, bool(ex.value)) if return_flag else it.value


# ------------------------------------------------------------------------------------------------
# asynchronous / batched forms (extensions)
# ------------------------------------------------------------------------------------------------
def solve_dev_async(state_size, knot_points, d_S, d_Pinv, d_gamma, d_lambda, config, d_iters=None, d_exit=None, stream=None):
    """Enqueue one solve on ``stream`` and return immediately (d_iters: uint32[1], d_exit: uint8[1], optional)."""
    suf = _torch_dtype_suffix(d_S)
    cfg = config.to_c()
    if d_Pinv is None:
        cfg.empty_pinv = 1
    f = getattr(_lib.lib(), f"gbdpcg_solve_dev_{suf}")
    _lib.check(f(None, _dev_ptr(d_S, "d_S"), _dev_ptr(d_Pinv, "d_Pinv", True), _dev_ptr(d_gamma, "d_gamma"),
                 _dev_ptr(d_lambda, "d_lambda"), int(state_size), int(knot_points), C.byref(cfg),
                 _dev_ptr(d_iters, "d_iters", True), _dev_ptr(d_exit, "d_exit", True), _stream_ptr(stream)))


def solve_dev_static_scratch(state_size, knot_points, d_S, d_Pinv, d_gamma, d_lambda, config, d_iters=None, d_exit=None, stream=None):
    """``gbdpcg_solve_dev_static_scratch``: one solve on the cooperative-grid kernel with its exchange buffers in the library's
    file-scope device array -- the call allocates nothing (no counterpart in the reference; include/gbdpcg.h).  Calls that
    overlap on the GPU race on that array: keep them on one stream."""
    cfg = config.to_c()
    if d_Pinv is None:
        cfg.empty_pinv = 1
    _lib.check(_lib.lib().gbdpcg_solve_dev_static_scratch(
        1 if _torch_dtype_suffix(d_S) == "f64" else 0, _dev_ptr(d_S, "d_S"), _dev_ptr(d_Pinv, "d_Pinv", True), _dev_ptr(d_gamma, "d_gamma"),
        _dev_ptr(d_lambda, "d_lambda"), int(state_size), int(knot_points), C.byref(cfg), _dev_ptr(d_iters, "d_iters", True),
        _dev_ptr(d_exit, "d_exit", True), _stream_ptr(stream)))


def solve_batched_dev(state_size, knot_points, batch, d_S, d_Pinv, d_gamma, d_lambda, config, d_iters=None, d_exit=None,
                      stream=None):
    """Enqueue ``batch`` independent solves (problem-major slabs) on ``stream``."""
    suf = _torch_dtype_suffix(d_S)
    cfg = config.to_c()
    if d_Pinv is None:
        cfg.empty_pinv = 1
    f = getattr(_lib.lib(), f"gbdpcg_solve_batched_dev_{suf}")
    _lib.check(f(None, _dev_ptr(d_S, "d_S"), _dev_ptr(d_Pinv, "d_Pinv", True), _dev_ptr(d_gamma, "d_gamma"),
                 _dev_ptr(d_lambda, "d_lambda"), int(state_size), int(knot_points), int(batch), C.byref(cfg),
                 _dev_ptr(d_iters, "d_iters", True), _dev_ptr(d_exit, "d_exit", True), _stream_ptr(stream)))


def solve_batched_host(h_S, h_Pinv, h_gamma, h_lambda, state_size, knot_points, batch, config, workspace=None):
    """Blocking batched solve on host arrays (pageable or pinned).  Returns (iters[batch], max_iter_exit[batch])."""
    suf = _suffix(h_S.dtype)
    nx, N, B = int(state_size), int(knot_points), int(batch)
    ms, vs = 3 * N * nx * nx * B, N * nx * B
    cfg = config.to_c()
    if h_Pinv is None:
        cfg.empty_pinv = 1
    iters = np.zeros(B, dtype=np.uint32)
    exits = np.zeros(B, dtype=np.uint8)
    f = getattr(_lib.lib(), f"gbdpcg_solve_batched_host_{suf}")
    _lib.check(f(_ws(workspace), _host(h_S, h_S.dtype, ms, "h_S"),
                 None if h_Pinv is None else _host(h_Pinv, h_S.dtype, ms, "h_Pinv"),
                 _host(h_gamma, h_S.dtype, vs, "h_gamma"), _host(h_lambda, h_S.dtype, vs, "h_lambda", True),
                 nx, N, B, C.byref(cfg), iters.ctypes.data_as(C.c_void_p), exits.ctypes.data_as(C.c_void_p)))
    return iters, exits.astype(bool)


PINV_KINDS = {"identity": _lib.PINV_IDENTITY, "jacobi": _lib.PINV_JACOBI, "stair": _lib.PINV_STAIR}
BUILD_FLAGS = {"jacobi": _lib.FLAG_BUILD_JACOBI, "stair": _lib.FLAG_BUILD_STAIR}


def _kind(kind) -> int:
    if isinstance(kind, str):
        if kind not in PINV_KINDS:
            raise ValueError(f"preconditioner kind must be one of {sorted(PINV_KINDS)}")
        return PINV_KINDS[kind]
    return int(kind)


def build_pinv_dev(state_size, knot_points, batch, d_S, d_Pinv, kind="stair", stream=None):
    """Enqueue the construction of Phi^-1 (identity / block-Jacobi / symmetric stair) from ``d_S`` into ``d_Pinv``
    for ``batch`` problem-major systems (the reference never builds one,
    include/interface.cuh:33-34,45-46)."""
    suf = _torch_dtype_suffix(d_S)
    f = getattr(_lib.lib(), f"gbdpcg_build_pinv_dev_{suf}")
    _lib.check(f(None, _dev_ptr(d_S, "d_S"), _dev_ptr(d_Pinv, "d_Pinv"), int(state_size), int(knot_points), int(batch),
                 _kind(kind), _stream_ptr(stream)))


def build_pinv_host(h_S, state_size, knot_points, batch=1, kind="stair"):
    """Blocking host form: returns Phi^-1 (numpy, same dtype and layout as ``h_S``), computed on the GPU."""
    suf = _suffix(h_S.dtype)
    nx, N, B = int(state_size), int(knot_points), int(batch)
    ms = 3 * N * nx * nx * B
    out = np.empty(ms, dtype=h_S.dtype)
    f = getattr(_lib.lib(), f"gbdpcg_build_pinv_host_{suf}")
    _lib.check(f(None, _host(h_S, h_S.dtype, ms, "h_S"), out.ctypes.data_as(C.c_void_p), nx, N, B, _kind(kind)))
    return out.reshape(h_S.shape) if h_S.size == ms else out


def schur_form_dev(state_size, control_size, knot_points, batch, d_Q, d_R, d_A, d_B, d_q, d_r, d_c, d_S, d_gamma,
                   d_Qinv=None, d_Rinv=None, stream=None):
    """Enqueue the Schur-complement formation S = -C G^-1 C^T, gamma = C G^-1 g - c from stage-wise LQR data
    (include/gbdpcg.h; SURVEY.md s8(f) rank 4: the step before the path, upstream of the reference)."""
    suf = _torch_dtype_suffix(d_Q)
    f = getattr(_lib.lib(), f"gbdpcg_schur_form_dev_{suf}")
    _lib.check(f(None, *[_dev_ptr(t, n) for t, n in ((d_Q, "d_Q"), (d_R, "d_R"), (d_A, "d_A"), (d_B, "d_B"), (d_q, "d_q"),
                                                     (d_r, "d_r"), (d_c, "d_c"))],
                 int(state_size), int(control_size), int(knot_points), int(batch), _dev_ptr(d_S, "d_S"), _dev_ptr(d_gamma, "d_gamma"),
                 _dev_ptr(d_Qinv, "d_Qinv", True), _dev_ptr(d_Rinv, "d_Rinv", True), _stream_ptr(stream)))


def schur_recover_dev(state_size, control_size, knot_points, batch, d_Qinv, d_Rinv, d_A, d_B, d_q, d_r, d_lambda, d_x, d_u,
                      stream=None):
    """Enqueue the primal recovery x_k = -Q_k^-1 (q_k + lambda_k - A_k^T lambda_{k+1}), u_k = -R_k^-1 (r_k - B_k^T lambda_{k+1})."""
    suf = _torch_dtype_suffix(d_Qinv)
    f = getattr(_lib.lib(), f"gbdpcg_schur_recover_dev_{suf}")
    _lib.check(f(None, *[_dev_ptr(t, n) for t, n in ((d_Qinv, "d_Qinv"), (d_Rinv, "d_Rinv"), (d_A, "d_A"), (d_B, "d_B"),
                                                     (d_q, "d_q"), (d_r, "d_r"), (d_lambda, "d_lambda"))],
                 int(state_size), int(control_size), int(knot_points), int(batch), _dev_ptr(d_x, "d_x"), _dev_ptr(d_u, "d_u"),
                 _stream_ptr(stream)))


def plan_query(state_size, knot_points, dtype=np.float32, has_pinv=True, config=None, batch=1) -> dict:
    """What the library would launch for this shape (replaces pcgSharedMemSize / checkPcgOccupancy,
    reference include/pcg.cuh:13-49).  batch: systems per launch -- a system that needs a thread-block cluster gets a
    larger one while the whole batch is resident in one wave anyway."""
    info = PlanInfo()
    cfg = (config or pcg_config()).to_c()
    _lib.check(_lib.lib().gbdpcg_plan_query_batched(int(state_size), int(knot_points), int(np.dtype(dtype) == np.float64),
                                                    int(bool(has_pinv)), int(batch), C.byref(cfg), C.byref(info)))
    return {k: (getattr(info, k).decode() if k == "kernel" else getattr(info, k)) for k, _ in PlanInfo._fields_}


def kernel_launches() -> int:
    return int(_lib.lib().gbdpcg_kernel_launches())
