"""ctypes loader for libgbdpcg.so (the C ABI of include/gbdpcg.h).

The product path has no CPU fallback: if the shared library is missing, or no CUDA device is
usable, every solve raises.  Nothing in this package imports oracle/.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GBDPCG_LIB: load an experiment build of the same library (scripts/, A/B runs); the default is the in-tree build
LIB_PATH = os.environ.get("GBDPCG_LIB") or os.path.join(_HERE, "libgbdpcg.so")

OK, ERR_INVALID, ERR_CUDA, ERR_UNSUPPORTED, ERR_NOMEM = 0, -1, -2, -3, -4
FLAG_FORCE_GENERIC, FLAG_PINV_SMEM, FLAG_PINV_REG = 0x1, 0x2, 0x4
FLAG_TEXTBOOK, FLAG_FUSED, FLAG_LEGACY_ROWS, FLAG_NO_TMA = 0x8, 0x10, 0x20, 0x40
FLAG_BUILD_JACOBI, FLAG_BUILD_STAIR = 0x100, 0x200
FLAG_TMEM, FLAG_NO_TMEM, FLAG_TMEM_R2, FLAG_SYMMETRIC_S = 0x400, 0x800, 0x1000, 0x2000
PINV_IDENTITY, PINV_JACOBI, PINV_STAIR = 0, 1, 2


class GbdpcgError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"gbdpcg status {status}: {message}")
        self.status = status


class Config(C.Structure):
    """struct gbdpcg_config (include/gbdpcg.h) = POD mirror of reference pcg_config<T> (include/types.cuh:18-35)."""
    _fields_ = [("exit_tol", C.c_double), ("max_iter", C.c_uint32), ("empty_pinv", C.c_int32),
                ("flags", C.c_uint32), ("cluster", C.c_uint32)]


class PlanInfo(C.Structure):
    _fields_ = [("specialised", C.c_int32), ("cluster", C.c_int32), ("threads", C.c_int32),
                ("block_rows", C.c_int32), ("pinv_mode", C.c_int32), ("regs", C.c_int32),
                ("smem_bytes", C.c_uint64), ("max_resident", C.c_int32), ("kernel", C.c_char * 96)]


# every symbol include/gbdpcg.h declares (tests check the header against this list and the .so)
_VP, _U32, _I32 = C.c_void_p, C.c_uint32, C.c_int32
_CFG = C.POINTER(Config)
SIGNATURES = {
    "gbdpcg_config_default": (None, [_CFG]),
    "gbdpcg_workspace_create": (C.c_int, [C.POINTER(_VP)]),
    "gbdpcg_workspace_destroy": (None, [_VP]),
    "gbdpcg_plan_query": (C.c_int, [_U32, _U32, C.c_int, C.c_int, _CFG, C.POINTER(PlanInfo)]),
    "gbdpcg_plan_query_batched": (C.c_int, [_U32, _U32, C.c_int, C.c_int, _U32, _CFG, C.POINTER(PlanInfo)]),
    "gbdpcg_host_alloc": (_VP, [C.c_size_t]),
    "gbdpcg_host_free": (None, [_VP]),
    "gbdpcg_version": (C.c_int, []),
    "gbdpcg_build_arch": (C.c_char_p, []),
    "gbdpcg_last_error": (C.c_char_p, []),
    "gbdpcg_last_cuda_error": (C.c_int, []),
    "gbdpcg_kernel_launches": (C.c_uint64, []),
    "gbdpcg_solve_dev_static_scratch": (C.c_int, [C.c_int] + [_VP] * 4 + [_U32, _U32, _CFG, _VP, _VP, _VP]),
}
STATIC_SCRATCH_BYTES = 106496   # GBDPCG_STATIC_SCRATCH_BYTES
for _s in ("f32", "f64"):
    SIGNATURES[f"gbdpcg_solve_host_{_s}"] = (C.c_int, [_VP] * 5 + [_U32, _U32, _CFG, C.POINTER(_U32), C.POINTER(_I32)])
    SIGNATURES[f"gbdpcg_solve_dev_{_s}"] = (C.c_int, [_VP] * 5 + [_U32, _U32, _CFG, _VP, _VP, _VP])
    SIGNATURES[f"gbdpcg_solve_dev_sync_{_s}"] = (C.c_int, [_VP] * 5 + [_U32, _U32, _CFG, C.POINTER(_U32), C.POINTER(_I32), _VP])
    SIGNATURES[f"gbdpcg_solve_batched_dev_{_s}"] = (C.c_int, [_VP] * 5 + [_U32, _U32, _U32, _CFG, _VP, _VP, _VP])
    SIGNATURES[f"gbdpcg_solve_batched_host_{_s}"] = (C.c_int, [_VP] * 5 + [_U32, _U32, _U32, _CFG, _VP, _VP])
    SIGNATURES[f"gbdpcg_build_pinv_dev_{_s}"] = (C.c_int, [_VP] * 3 + [_U32, _U32, _U32, C.c_int, _VP])
    SIGNATURES[f"gbdpcg_build_pinv_host_{_s}"] = (C.c_int, [_VP] * 3 + [_U32, _U32, _U32, C.c_int])
    SIGNATURES[f"gbdpcg_schur_form_dev_{_s}"] = (C.c_int, [_VP] * 8 + [_U32] * 4 + [_VP] * 5)
    SIGNATURES[f"gbdpcg_schur_recover_dev_{_s}"] = (C.c_int, [_VP] * 8 + [_U32] * 4 + [_VP] * 3)
    SIGNATURES[f"gbdpcg_pack_csr_dev_{_s}"] = (C.c_int, [_VP] * 4 + [_U32, _U32, _U32, _U32, _VP, _VP, _VP])
    SIGNATURES[f"gbdpcg_solve_csr_host_{_s}"] = (C.c_int, [_VP] * 4 + [_U32] + [_VP] * 3 + [_U32, _VP, _VP, _U32, _U32, _CFG,
                                                            C.POINTER(_U32), C.POINTER(_I32)])

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `make -C gbd_pcg_b200/csrc` "
                              "(or __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def check(status: int):
    if status != OK:
        raise GbdpcgError(status, lib().gbdpcg_last_error().decode(errors="replace"))


def default_config() -> Config:
    c = Config()
    lib().gbdpcg_config_default(C.byref(c))
    return c
