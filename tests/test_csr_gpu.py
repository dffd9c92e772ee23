"""CSR front end: the overload the reference declares as a stub (include/interface.cuh:8-20).
Bit-exact against the host-side layout packer (gbd_pcg_b200/layout.py) -- this is index work."""
import os
import shutil
import subprocess

import numpy as np
import pytest

import gbd_pcg_b200 as g
from gbd_pcg_b200 import _lib
from helpers import make_system

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torch():
    import torch
    return torch


def pack_on_gpu(csr, nx, N, dtype):
    torch = _torch()
    d = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).astype(dt).view(np.int32 if dt == np.uint32 else dt)).cuda()
    d_rp, d_ci, d_v = d(csr.row_ptr, np.uint32), d(csr.col_ind if csr.nnz else np.zeros(1), np.uint32), d(csr.val if csr.nnz else np.zeros(1), dtype)
    out = torch.full((3 * N * nx * nx,), float("nan"), dtype=d_v.dtype, device="cuda")
    status = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    g.pack_csr_dev(d_rp, d_ci, d_v, nx * N, nx, N, out, status, nnz=csr.nnz)
    torch.cuda.synchronize()
    return out.cpu().numpy(), int(status.item())


def test_malformed_row_ptr_is_clamped_and_flagged():
    """row_ptr that runs past nnz or decreases: the device packer must not read outside col_ind / val (status bit 1, entries
    beyond nnz dropped); the host CSR solve refuses the matrix before anything is uploaded."""
    nx, N = 2, 4
    rows = nx * N
    col = np.array([0, 1, 2, 3], np.uint32)
    val = np.array([1.0, 2.0, 3.0, 4.0])
    # (a) row_ptr[rows] = 1000 > nnz = 4
    rp = np.array([0, 1, 2, 3, 4, 4, 4, 4, 1000], np.uint32)
    out, status = pack_on_gpu(g.csr_t(rp, col, val, rows, rows, 4), nx, N, np.float64)
    assert status & 2
    want = np.zeros(3 * N * nx * nx)
    for i in range(4):
        want[(i // nx) * 3 * nx * nx + nx * nx + (i % nx) * nx + (i % nx)] = val[i]       # (i, i) on the diagonal blocks
    assert np.array_equal(out, want)
    # (b) decreasing row_ptr
    rp2 = np.array([0, 3, 1, 3, 4, 4, 4, 4, 4], np.uint32)
    _, status2 = pack_on_gpu(g.csr_t(rp2, col, val, rows, rows, 4), nx, N, np.float64)
    assert status2 & 2
    for bad in (rp, rp2, np.array([1, 1, 2, 3, 4, 4, 4, 4, 4], np.uint32)):
        lam = np.zeros(rows)
        with pytest.raises(g.GbdpcgError) as e:
            g.solvePCG(g.csr_t(bad, col, val, rows, rows, 4), None, np.ones(rows), lam, nx, N, g.pcg_config())
        assert e.value.status == _lib.ERR_INVALID and not lam.any()


@pytest.mark.parametrize("precond", ["identity", "stair"])
@pytest.mark.parametrize("name", ["C1", "C2", "C3", "C4"])
def test_pack_is_bit_exact(name, precond):
    S, P, _, nx, N, _, _ = make_system(name, precond)
    M = S if P is None else P
    A = g.layout.unpack_dense(M, nx, N)
    out, status = pack_on_gpu(g.csr_t.from_dense(A), nx, N, M.dtype)
    assert status == 0
    assert np.array_equal(out, g.layout.pack_dense(A, nx, N).astype(M.dtype))


def test_duplicates_are_summed_and_out_of_band_entries_flagged():
    nx, N = 2, 4
    rows = nx * N
    # row 0: (0,0)=1 twice and (0,1)=2 ; row 3: (3,7) is outside the band ; everything else empty
    row_ptr = np.array([0, 3, 3, 3, 4, 4, 4, 4, 4], np.uint32)
    csr = g.csr_t(row_ptr, np.array([0, 0, 1, 7], np.uint32), np.array([1.0, 1.0, 2.0, 5.0]), rows, rows, 4)
    out, status = pack_on_gpu(csr, nx, N, np.float64)
    assert status == 1
    want = np.zeros(3 * N * nx * nx)
    want[nx * nx + 0] = 2.0           # block-row 0, diag block, (r=0, c=0)
    want[nx * nx + 1 * nx + 0] = 2.0  # (r=0, c=1)
    assert np.array_equal(out, want)
    lam = np.zeros(rows)
    with pytest.raises(g.GbdpcgError) as e:
        g.solvePCG(csr, None, np.ones(rows), lam, nx, N, g.pcg_config())
    assert e.value.status == _lib.ERR_INVALID and not lam.any()


@pytest.mark.parametrize("name", ["C1", "C2", "C3"])
def test_csr_solve_equals_the_compressed_layout_solve(name):
    S, P, gam, nx, N, tol, max_iter = make_system(name, "stair")
    cfg = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter, empty_pinv=0)
    lam_a = np.zeros(N * nx, S.dtype)
    it_a = g.pcg_solve(S, gam, lam_a, nx, N, cfg, h_Pinv=P)
    cS, cP = g.csr_t.from_dense(g.layout.unpack_dense(S, nx, N)), g.csr_t.from_dense(g.layout.unpack_dense(P, nx, N))
    lam_b = np.zeros(N * nx, S.dtype)
    it_b = g.solvePCG(cS, cP, gam, lam_b, nx, N, cfg)
    assert it_a == it_b and np.array_equal(lam_a, lam_b)           # same packed bytes, same kernel
    # no Phi^-1 given, empty_pinv = 0: formed on the device from the packed S
    lam_c = np.zeros(N * nx, S.dtype)
    it_c = g.solvePCG(cS, None, gam, lam_c, nx, N, cfg)
    assert abs(it_c - it_a) <= 1
    # identity
    lam_d, lam_e = np.zeros(N * nx, S.dtype), np.zeros(N * nx, S.dtype)
    cfg_i = g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=max_iter)
    assert g.solvePCG(cS, None, gam, lam_d, nx, N, cfg_i) == g.solvePCG(S, gam, lam_e, nx, N, cfg_i)
    assert np.array_equal(lam_d, lam_e)


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="needs nvcc")
def test_cpp_csr_example_runs(tmp_path):
    exe = tmp_path / "csr.out"
    subprocess.check_call(["nvcc", "-std=c++17", "-Wno-deprecated-gpu-targets", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "pcg_solve_csr.cu"), "-L", os.path.dirname(_lib.LIB_PATH), "-lgbdpcg", "-o", str(exe)])
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.dirname(_lib.LIB_PATH) + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run([str(exe)], env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
