"""ctypes door onto the UNMODIFIED reference CUDA kernel (oracle/_ref/libref_pcg_<dtype>_<nx>_<N>.so,
built by oracle/Makefile from /root/reference).  TEST / BASELINE INFRASTRUCTURE ONLY.

Used by tests (-m gpu) as the second parity witness, by oracle/make_ref_fixtures.py to pin the host
oracle, and by bench.py --impl reference.  Never imported by the product package.
"""
from __future__ import annotations

import ctypes as C
import glob
import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")
_cache = {}


def available():
    """[(dtype_name, nx, N)] of the prebuilt reference instantiations."""
    out = []
    for p in sorted(glob.glob(os.path.join(_REF_DIR, "libref_pcg_*.so"))):
        m = re.match(r"libref_pcg_(f32|f64)_(\d+)_(\d+)\.so", os.path.basename(p))
        if m:
            out.append((m.group(1), int(m.group(2)), int(m.group(3))))
    return out


def has(dtype, nx, N) -> bool:
    return (("f64" if np.dtype(dtype) == np.float64 else "f32"), nx, N) in available()


class RefKernel:
    """The reference instantiated for one (T, STATE_SIZE, KNOT_POINTS)."""

    def __init__(self, dtype, nx, N):
        import torch
        self.torch = torch
        self.dtype = np.dtype(dtype)
        self.tdtype = torch.float64 if self.dtype == np.float64 else torch.float32
        self.nx, self.N = nx, N
        suf = "f64" if self.dtype == np.float64 else "f32"
        path = os.path.join(_REF_DIR, f"libref_pcg_{suf}_{nx}_{N}.so")
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path}: build with `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(path)
        vp = C.c_void_p
        L.ref_kernel_launch.restype = C.c_int
        L.ref_kernel_launch.argtypes = [vp] * 10 + [C.c_uint32, C.c_double, C.c_uint64]
        if hasattr(L, "ref_kernel_launch_stream"):
            L.ref_kernel_launch_stream.restype = C.c_int
            L.ref_kernel_launch_stream.argtypes = [vp] * 10 + [C.c_uint32, C.c_double, C.c_uint64, vp]
        L.ref_solve_dev.restype = C.c_uint32
        L.ref_solve_dev.argtypes = [vp] * 8 + [C.c_uint32, C.c_double]
        L.ref_solve_host.restype = C.c_uint32
        L.ref_solve_host.argtypes = [vp] * 3 + [C.c_uint32, C.c_double]
        L.ref_smem_bytes.restype = None
        L.ref_smem_bytes.argtypes = [C.POINTER(C.c_uint64)] * 2
        if hasattr(L, "ref_glass_build_pinv"):
            L.ref_glass_build_pinv.restype = C.c_int
            L.ref_glass_build_pinv.argtypes = [vp, vp, C.c_int]
        L.ref_kernel_attrs.restype = C.c_int
        L.ref_kernel_attrs.argtypes = [C.POINTER(C.c_int)] * 2
        self.L = L
        req, used = C.c_uint64(), C.c_uint64()
        L.ref_smem_bytes(C.byref(req), C.byref(used))
        # cover the reference's 2*nx-element shared-memory under-allocation from the harness side
        self.extra_smem = max(0, used.value - req.value)
        dev = "cuda"
        self.d_r = torch.zeros(nx * N, dtype=self.tdtype, device=dev)
        self.d_p = torch.zeros(nx * N, dtype=self.tdtype, device=dev)
        self.d_v = torch.zeros(N, dtype=self.tdtype, device=dev)
        self.d_eta = torch.zeros(N, dtype=self.tdtype, device=dev)
        self.d_iters = torch.zeros(1, dtype=torch.int32, device=dev)
        self.d_exit = torch.zeros(1, dtype=torch.uint8, device=dev)

    def _p(self, t):
        return C.c_void_p(t.data_ptr())

    def launch(self, d_S, d_Pinv, d_gamma, d_lambda, max_iter, tol):
        """Scope K: asynchronous cooperative launch on the legacy default stream (include/interface.cuh:130-132)."""
        e = self.L.ref_kernel_launch(self._p(d_S), self._p(d_Pinv), self._p(d_gamma), self._p(d_lambda), self._p(self.d_r),
                                     self._p(self.d_p), self._p(self.d_v), self._p(self.d_eta), self._p(self.d_iters),
                                     self._p(self.d_exit), int(max_iter), float(tol), int(self.extra_smem))
        if e != 0:
            raise RuntimeError(f"reference cooperative launch failed: cudaError {e}")

    def make_lane(self, stream):
        """A stream with its own scratch vectors and result scalars, for launch_on()."""
        t, n = self.torch, self.nx * self.N
        z = lambda k, dt: t.zeros(k, dtype=dt, device="cuda")
        return {"stream": stream, "r": z(n, self.tdtype), "p": z(n, self.tdtype), "v": z(self.N, self.tdtype),
                "eta": z(self.N, self.tdtype), "iters": z(1, t.int32), "exit": z(1, t.uint8)}

    def launch_on(self, lane, d_S, d_Pinv, d_gamma, d_lambda, max_iter, tol):
        """The same cooperative launch on lane['stream'] (several solves in flight; not something the reference API offers)."""
        e = self.L.ref_kernel_launch_stream(self._p(d_S), self._p(d_Pinv), self._p(d_gamma), self._p(d_lambda), self._p(lane["r"]),
                                            self._p(lane["p"]), self._p(lane["v"]), self._p(lane["eta"]), self._p(lane["iters"]),
                                            self._p(lane["exit"]), int(max_iter), float(tol), int(self.extra_smem),
                                            C.c_void_p(lane["stream"].cuda_stream))
        if e != 0:
            raise RuntimeError(f"reference cooperative launch failed: cudaError {e}")

    def solve(self, S, Pinv, gamma, lambda0, max_iter, tol):
        """Run the reference kernel once on host arrays (with a DEFINED Pinv); returns (lambda, iters, max_iter_exit)."""
        torch = self.torch
        mk = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=self.dtype)).cuda()
        d_S, d_P, d_g, d_l = mk(S), mk(Pinv), mk(gamma), mk(lambda0)
        torch.cuda.synchronize()
        self.launch(d_S, d_P, d_g, d_l, max_iter, tol)
        torch.cuda.synchronize()
        return d_l.cpu().numpy(), int(self.d_iters.item()), bool(self.d_exit.item())

    def solve_dev_api(self, d_S, d_Pinv, d_gamma, d_lambda, max_iter, tol) -> int:
        """Scope D: the reference's own device-pointer solvePCG overload (blocking)."""
        return int(self.L.ref_solve_dev(self._p(d_S), self._p(d_Pinv), self._p(d_gamma), self._p(d_lambda), self._p(self.d_r),
                                        self._p(self.d_p), self._p(self.d_v), self._p(self.d_eta), int(max_iter), float(tol)))

    def solve_host_api(self, h_S, h_gamma, h_lambda, max_iter, tol) -> int:
        """Scope H: the reference's host-pointer overload.  TIME ONLY: its d_Pinv is uninitialised (interface.cuh:45-46)."""
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        return int(self.L.ref_solve_host(vp(h_S), vp(h_gamma), vp(h_lambda), int(max_iter), float(tol)))

    def glass_build_pinv(self, S, kind):
        """Phi^-1 of one system from the reference's own primitives glass::invertMatrix / glass::gemm (GLASS/src/L3/inv.cuh:6-37,
        gemm.cuh), one CTA per block-row.  kind: 'jacobi' | 'stair'."""
        torch = self.torch
        d_S = torch.from_numpy(np.ascontiguousarray(S, dtype=self.dtype)).cuda()
        d_P = torch.empty_like(d_S)
        torch.cuda.synchronize()
        e = self.L.ref_glass_build_pinv(self._p(d_S), self._p(d_P), {"jacobi": 1, "stair": 2}[kind])
        if e != 0:
            raise RuntimeError(f"ref_glass_build_pinv failed: cudaError {e}")
        return d_P.cpu().numpy()

    def attrs(self):
        r, b = C.c_int(), C.c_int()
        self.L.ref_kernel_attrs(C.byref(r), C.byref(b))
        return {"regs": r.value, "max_blocks_per_sm": b.value, "extra_smem": self.extra_smem}


def get(dtype, nx, N) -> RefKernel:
    key = (np.dtype(dtype).name, nx, N)
    if key not in _cache:
        _cache[key] = RefKernel(dtype, nx, N)
    return _cache[key]
