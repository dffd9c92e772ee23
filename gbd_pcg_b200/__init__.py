"""gbd_pcg_b200 -- B200-native (sm_100a) block-tridiagonal PCG solve behind the GBD-PCG host API.

Only the hot path of A2R-Lab/GBD-PCG lives here: csrc/ (hand-written CUDA + the C ABI of
include/gbdpcg.h, built into libgbdpcg.so), api.py (Python mirror of the reference's
solvePCG / pcg_solve / pcg_config), layout.py (compressed block-tridiagonal layout helpers) and
synth.py (synthetic MPC systems for tests and the bench).
"""
from . import layout, synth  # noqa: F401
from .api import (GbdpcgError, build_pinv_dev, build_pinv_host, csr_t, kernel_launches, pack_csr_dev, schur_form_dev, schur_recover_dev, pcg_config, pcg_solve, plan_query, solve_batched_dev,  # noqa: F401
                  solve_batched_host, solve_dev_async, solve_dev_static_scratch, solvePCG, workspace)

__all__ = ["pcg_config", "solvePCG", "pcg_solve", "solve_dev_async", "solve_dev_static_scratch", "solve_batched_dev", "solve_batched_host",
           "build_pinv_dev", "build_pinv_host", "csr_t", "pack_csr_dev", "schur_form_dev", "schur_recover_dev", "plan_query", "kernel_launches", "GbdpcgError", "workspace", "layout", "synth"]
