"""Records outputs of the UNMODIFIED reference CUDA kernel (oracle/_ref, built from /root/reference by
oracle/Makefile) on a B200, to pin the host oracle and the new kernel against the real reference.

Run on the GPU box:   python oracle/make_ref_fixtures.py gpurun_out/ref_fixtures
then copy the .npz files into tests/golden/ and commit them.  Inputs are NOT stored: they are
regenerated from the seeded generator (gbd_pcg_b200/synth.py) or are the
reference example's literal arrays (examples/pcg_solve.cu:14-25).

The reference is driven through a direct cooperative launch of pcg<T,nx,N> (include/interface.cuh:130-132)
with a DEFINED d_Pinv (identity / block-Jacobi / symmetric stair).  Its host-pointer overload is not a valid
source of fixtures: it never initialises d_Pinv (include/interface.cuh:45-46).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from gbd_pcg_b200 import layout, synth  # noqa: E402
from helpers import PRECONDS, make_system  # noqa: E402
from oracle import oracle, ref_cuda  # noqa: E402

CASES = [("C1", np.float32), ("C1", np.float64), ("C2", np.float32), ("C3", np.float64), ("C4", np.float32)]


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    for name, dtype in CASES:
        for pc in PRECONDS:
            S, P, gam, nx, N, tol, max_iter = make_system(name, pc, dtype=dtype)
            if not ref_cuda.has(dtype, nx, N):
                print("skip", name, np.dtype(dtype).name, "(no harness)")
                continue
            P_ref = P if P is not None else layout.identity_compressed(nx, N, dtype)
            lam0 = np.zeros(N * nx, dtype=dtype)
            lam, it, ex = ref_cuda.get(dtype, nx, N).solve(S, P_ref, gam, lam0, max_iter, tol)
            lam_o, it_o, ex_o = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=0)
            lam_1, it_1, _ = oracle.pcg(S, P, gam, lam0, nx, N, tol, max_iter, order=1)
            d0 = np.linalg.norm(lam.astype(np.float64) - lam_o) / np.linalg.norm(lam_o)
            d1 = np.linalg.norm(lam.astype(np.float64) - lam_1) / np.linalg.norm(lam_1)
            print(f"{name} {np.dtype(dtype).name} {pc:8s} ref iters {it} exit {ex} | oracle iters {it_o} (ref-order {it_1}) "
                  f"| rel diff vs oracle {d0:.3e}, vs ref-order oracle {d1:.3e}", flush=True)
            np.savez(os.path.join(outdir, f"ref_cuda_{name}_{np.dtype(dtype).name}_{pc}.npz"), lam=lam, iters=it, max_iter_exit=ex,
                     config=name, precond=pc, dtype=np.dtype(dtype).name, nx=nx, N=N, tol=tol, max_iter=max_iter,
                     source="reference pcg<T,nx,N> @75d6214 on NVIDIA B200, nvcc 12.9 -O3 sm_100a")


def pinv_rows(N):
    """Block-rows of Phi^-1 kept in a fixture: all of a small system, the first and last 8 of a long one (fixtures stay small)."""
    return np.arange(N) if N <= 32 else np.concatenate([np.arange(8), np.arange(N - 8, N)])


def main_pinv(outdir):
    """Phi^-1 built by the reference's own primitives glass::invertMatrix / glass::gemm (GLASS/src/L3/inv.cuh:6-37, gemm.cuh) on the
    blocks of S -> tests/golden/ref_glass_pinv_*.npz: what pins oracle_build_pinv_* and the product's pinv_build kernel."""
    import gbd_pcg_b200 as g
    os.makedirs(outdir, exist_ok=True)
    for name, dtype in CASES:
        S, _, _, nx, N, _, _ = make_system(name, "identity", dtype=dtype)
        if not ref_cuda.has(dtype, nx, N) or not hasattr(ref_cuda.get(dtype, nx, N).L, "ref_glass_build_pinv"):
            print("skip", name, np.dtype(dtype).name, "(no harness)")
            continue
        for kind in ("jacobi", "stair"):
            P_ref = ref_cuda.get(dtype, nx, N).glass_build_pinv(S, kind)
            P_o = oracle.build_pinv(S, nx, N, kind)
            P_g = g.build_pinv_host(S, nx, N, 1, kind)
            scale = np.abs(P_ref).max()
            rows = pinv_rows(N)
            print(f"{name} {np.dtype(dtype).name} {kind:6s} glass vs oracle max|d|/max|P| {np.abs(P_ref - P_o).max() / scale:.3e} | "
                  f"glass vs pinv_build kernel {np.abs(P_ref - P_g).max() / scale:.3e} | bit-identical to oracle: {np.array_equal(P_ref, P_o)}", flush=True)
            np.savez(os.path.join(outdir, f"ref_glass_pinv_{name}_{np.dtype(dtype).name}_{kind}.npz"),
                     pinv_rows=P_ref.reshape(N, 3 * nx * nx)[rows], rows=rows, config=name, kind=kind, dtype=np.dtype(dtype).name, nx=nx, N=N,
                     source="glass::invertMatrix + glass::gemm @GLASS a3e0931 on NVIDIA B200, nvcc 12.9 -O3 sm_100a, one CTA of 64 threads per block-row")


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2] == "pinv":
        main_pinv(sys.argv[1])
        sys.exit(0)
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/ref_fixtures")
