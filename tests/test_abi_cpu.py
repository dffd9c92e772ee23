"""CPU tests of the drop-in boundary: libgbdpcg.so loads, exports every symbol include/gbdpcg.h declares,
answers the queries that need no GPU, fails loudly (no CPU fallback) when asked to solve without one, and the
reference's own example programs compile unchanged against the new headers."""
import ctypes as C
import os
import re
import shutil
import subprocess
import sys

import numpy as np
import pytest

import gbd_pcg_b200 as g
from conftest import ROOT, _cuda_ok
from gbd_pcg_b200 import _lib

HEADER = os.path.join(ROOT, "include", "gbdpcg.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gbdpcg_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 20
    L = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/gbdpcg.h but not exported by libgbdpcg.so"
    # and the ctypes table of the Python mirror covers the same set
    assert sorted(_lib.SIGNATURES) == names


def test_only_the_c_abi_is_exported():
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    syms = [l.split()[-1] for l in out.splitlines() if " T " in l]
    assert syms and all(s.startswith("gbdpcg_") for s in syms), [s for s in syms if not s.startswith("gbdpcg_")][:5]


def test_library_is_sm100a_only_and_has_no_oracle_inside():
    L = _lib.lib()
    assert L.gbdpcg_build_arch() == b"sm_100a" and L.gbdpcg_version() >= 100
    out = subprocess.check_output(["cuobjdump", "--list-elf", _lib.LIB_PATH], text=True)
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs
    nm = subprocess.check_output(["nm", "-D", _lib.LIB_PATH], text=True)
    assert "oracle" not in nm


def test_config_defaults_match_the_reference():
    c = _lib.default_config()       # reference include/constants.cuh:15-17, include/types.cuh:33
    assert c.exit_tol == 1e-6 and c.max_iter == 25 and c.empty_pinv == 1 and c.flags == 0 and c.cluster == 0
    p = g.pcg_config()
    assert (p.pcg_exit_tol, p.pcg_max_iter, p.pcg_grid, p.pcg_block, p.empty_pinv) == (1e-6, 25, (128, 1, 1), (64, 1, 1), 1)


@pytest.mark.parametrize("nx,N,dtype,cluster,spec", [
    (14, 32, np.float32, 1, 1), (2, 3, np.float32, 1, 1), (14, 64, np.float64, None, 1), (12, 512, np.float32, None, 1),
    (7, 30, np.float32, 1, 1), (3, 3, np.float64, 1, 1), (20, 32, np.float32, None, 1), (24, 64, np.float64, None, 1),
    (33, 30, np.float32, 1, 0), (32, 32, np.float64, 1, 0)])
def test_plan_query_without_gpu(nx, N, dtype, cluster, spec):
    info = g.plan_query(nx, N, dtype, True)
    assert info["specialised"] == spec
    if cluster is not None:
        assert info["cluster"] == cluster
    else:
        assert info["cluster"] > 1                                 # does not fit one SM: thread-block cluster
    assert info["block_rows"] * info["cluster"] >= N
    assert info["threads"] % 32 == 0 and info["threads"] <= 1024
    assert 0 < info["smem_bytes"] <= 227 * 1024


@pytest.mark.parametrize("nx,N,dtype", [(14, 1500, np.float32), (14, 1000, np.float64), (12, 1400, np.float32), (14, 1025, np.float32), (14, 5328, np.float32)])
def test_plan_query_beyond_one_cluster(nx, N, dtype):
    """Systems no 16-CTA cluster holds go to the cooperative-grid kernel, at most one CTA per SM (the reference accepts any N whose
    N CTAs are co-resident, include/pcg.cuh:23-49: 1776 knot points for nx=14 fp32 on a B200)."""
    info = g.plan_query(nx, N, dtype, True)
    assert "pcg_grid" in info["kernel"] and info["specialised"] == 1
    assert info["cluster"] <= 148 and info["cluster"] * info["block_rows"] >= N
    assert info["threads"] % 32 == 0 and 64 <= info["threads"] <= 1024 and 0 < info["smem_bytes"] <= 227 * 1024


@pytest.mark.parametrize("nx,N,dtype", [(14, 833, np.float32), (14, 1024, np.float32), (12, 1200, np.float32), (14, 576, np.float64), (8, 1024, np.float64)])
def test_plan_query_tensor_memory_cluster_before_the_grid(nx, N, dtype):
    """Past the register-row cluster (nx 14 fp32: N <= 752) a 16-CTA cluster with Phi^-1 in tensor memory still holds the system
    (N <= 1024; fp64 N <= 576): the planner takes it before the cooperative grid; GBDPCG_FLAG_NO_TMEM restores the grid."""
    info = g.plan_query(nx, N, dtype, True)
    assert "P_TMEM" in info["kernel"] and info["cluster"] == 16 and info["cluster"] * info["block_rows"] >= N
    assert info["threads"] % 32 == 0 and info["threads"] <= 256
    assert "pcg_grid" in g.plan_query(nx, N, dtype, True, g.pcg_config(empty_pinv=0, flags=_lib.FLAG_NO_TMEM))["kernel"]


def test_plan_query_one_barrier_form_of_the_grid_kernel(monkeypatch):
    """Beyond one cluster the planner takes the one-barrier form of the cooperative-grid kernel (ghost rows of Phi^-1: 2 x nx/R threads
    beyond the owners') wherever every CTA owns two block-rows and the CTA has room for those threads; the largest systems (CTAs
    already at their register budget's thread count) and GBDPCG_GRID_TWO_BARRIERS=1 keep the two-barrier form."""
    for nx, N, dtype in [(14, 1500, np.float32), (14, 1000, np.float64), (8, 2500, np.float64), (6, 6000, np.float32)]:
        info = g.plan_query(nx, N, dtype, True)
        assert "pcg_grid1" in info["kernel"] and "one barrier" in info["kernel"]
        rpt = 1 if dtype == np.float64 else 2
        assert info["threads"] >= info["block_rows"] * (nx // rpt) + 2 * (nx // rpt) and info["block_rows"] >= 2
    assert "pcg_grid<" in g.plan_query(14, 5328, np.float32, True)["kernel"]
    # with one barrier, half as many CTAs of twice the size where they fit (the arrivals of a barrier are serialised)
    half, full = g.plan_query(14, 1500, np.float32, True), None
    monkeypatch.setenv("GBDPCG_GRID_BPC_SCALE", "1")
    full = g.plan_query(14, 1500, np.float32, True)
    assert half["block_rows"] == 2 * full["block_rows"] and half["cluster"] <= 74 < full["cluster"]
    monkeypatch.delenv("GBDPCG_GRID_BPC_SCALE")
    monkeypatch.setenv("GBDPCG_GRID_TWO_BARRIERS", "1")
    assert "pcg_grid<" in g.plan_query(14, 1500, np.float32, True)["kernel"]


def test_plan_query_rejects_what_cannot_fit():
    with pytest.raises(g.GbdpcgError) as e:
        g.plan_query(14, 4096, np.float64, True)
    assert e.value.status == _lib.ERR_UNSUPPORTED
    with pytest.raises(g.GbdpcgError):
        g.plan_query(14, 1, np.float32, True)                      # N >= 2


@pytest.mark.skipif(_cuda_ok(), reason="only meaningful on a machine without a GPU")
def test_no_cpu_fallback():
    """Without a usable CUDA device a solve must fail loudly with GBDPCG_ERR_CUDA -- never compute on the CPU."""
    S, P, gam = g.synth.example_system(np.float32)
    lam = np.zeros(6, dtype=np.float32)
    with pytest.raises(g.GbdpcgError) as e:
        g.solvePCG(S, gam, lam, 2, 3, g.pcg_config())
    assert e.value.status == _lib.ERR_CUDA and not lam.any()
    lamb = np.zeros((4, 6), dtype=np.float32)
    with pytest.raises(g.GbdpcgError):
        g.solve_batched_host(np.tile(S, (4, 1)), None, np.tile(gam, (4, 1)), lamb, 2, 3, 4, g.pcg_config())


def test_product_package_never_imports_the_oracle():
    code = "import sys, gbd_pcg_b200; assert not any(m == 'oracle' or m.startswith('oracle.') for m in sys.modules), 'oracle imported'"
    subprocess.check_call([sys.executable, "-c", code], cwd=ROOT)
    for dirpath, _, files in os.walk(os.path.join(ROOT, "gbd_pcg_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "pcg_oracle" not in txt, f


def test_invalid_arguments_are_rejected_before_any_cuda_call():
    S, P, gam = g.synth.example_system(np.float32)
    lam = np.zeros(6, dtype=np.float32)
    with pytest.raises(g.GbdpcgError) as e:
        g.solvePCG(S, gam, lam, 2, 1, g.pcg_config())
    assert e.value.status == _lib.ERR_INVALID
    with pytest.raises(ValueError):
        g.solvePCG(S, gam, lam[:3], 2, 3, g.pcg_config())
    with pytest.raises(TypeError):
        g.solvePCG(S.astype(np.float16), gam, lam, 2, 3, g.pcg_config())


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="needs nvcc")
@pytest.mark.parametrize("example", ["pcg_solve.cu", "pcg_solve_dp.cu", "pcg_solve_csr.cu", "api_surface.cu"])
def test_reference_style_example_compiles_against_the_new_headers(example, tmp_path):
    """Acceptance of the header shim: a program written against the reference's
    gpu_pcg.cuh / solvePCG<T> builds with -I include -lgbdpcg.  Uses the reference's own example where the
    reference tree is present (this container), else the equivalent example shipped in examples/."""
    ref = os.path.join("/root/reference/examples", example)
    src = ref if os.path.exists(ref) else os.path.join(ROOT, "examples", example)
    exe = tmp_path / "a.out"
    # api_surface.cu instantiates the rest of the template surface: overload 3 with scratch buffers, pcg_solve<T> with the config
    # defaulted, pcgSharedMemSize<T>, checkPcgOccupancy<T> (uninstantiated templates are not even type-checked)
    subprocess.check_call(["nvcc", "-std=c++17", "-Wno-deprecated-gpu-targets", "-I", os.path.join(ROOT, "include"), "-I",
                           os.path.join(ROOT, "examples"), src, "-L", os.path.dirname(_lib.LIB_PATH), "-lgbdpcg", "-Xcompiler", "-pthread",
                           "-o", str(exe)])
    if not _cuda_ok():
        env = dict(os.environ, LD_LIBRARY_PATH=os.path.dirname(_lib.LIB_PATH) + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
        r = subprocess.run([str(exe)], env=env, capture_output=True, text=True)
        assert r.returncode != 0 and "GPUassert" in r.stderr       # loud failure, reference error convention


def test_csr_row_ptr_is_validated_on_the_host_before_any_upload():
    """A malformed csr_t (row_ptr[0] != 0, decreasing, or row_ptr[rows] > nnz) is refused with GBDPCG_ERR_INVALID by the host
    check alone -- it answers the same way on a machine without a GPU, i.e. before any CUDA call."""
    nx, N = 2, 4
    rows = nx * N
    col = np.array([0, 1, 2, 3], np.uint32)
    val = np.array([1.0, 2.0, 3.0, 4.0])
    for rp in ([0, 1, 2, 3, 4, 4, 4, 4, 1000], [0, 3, 1, 3, 4, 4, 4, 4, 4], [1, 1, 2, 3, 4, 4, 4, 4, 4]):
        lam = np.zeros(rows)
        with pytest.raises(g.GbdpcgError) as e:
            g.solvePCG(g.csr_t(np.array(rp, np.uint32), col, val, rows, rows, 4), None, np.ones(rows), lam, nx, N, g.pcg_config())
        assert e.value.status == _lib.ERR_INVALID and "row_ptr" in str(e.value)
