"""gbdpcg_solve_dev_static_scratch -- the allocation-free entry point (no counterpart in the reference; include/gbdpcg.h) --
called through the C ABI and held to the existing device-pointer entry point gbdpcg_solve_dev_f32 / _f64, on two calls in a
row that share the library's static scratch array.  Collected last on purpose (zz)."""
import ctypes as C

import numpy as np
import pytest

import gbd_pcg_b200 as g
from conftest import LAMBDA_TOL, rel_err
from gbd_pcg_b200 import _lib
from helpers import make_system

pytestmark = pytest.mark.gpu


def _dev(torch, a):
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _call(which, S, P, gam, lam0, nx, N, tol, max_iter, flags=0, cluster=0):
    """One solve through the C ABI: which = "static" (new entry point) or "dev" (existing gbdpcg_solve_dev_*).
    Returns (lambda, iters, max_iter_exit)."""
    import torch
    L = _lib.lib()
    d_S, d_P, d_g, d_l = _dev(torch, S), _dev(torch, P), _dev(torch, gam), _dev(torch, np.array(lam0, dtype=S.dtype))
    d_it = torch.full((1,), 12345, dtype=torch.int32, device="cuda")
    d_ex = torch.full((1,), 9, dtype=torch.uint8, device="cuda")
    cfg = _lib.default_config()
    cfg.exit_tol, cfg.max_iter, cfg.empty_pinv, cfg.flags, cfg.cluster = tol, max_iter, (1 if P is None else 0), flags, cluster
    f64 = S.dtype == np.float64
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if which == "static":
        rc = L.gbdpcg_solve_dev_static_scratch(1 if f64 else 0, _ptr(d_S), _ptr(d_P), _ptr(d_g), _ptr(d_l), nx, N, C.byref(cfg),
                                               _ptr(d_it), _ptr(d_ex), stream)
    else:
        f = L.gbdpcg_solve_dev_f64 if f64 else L.gbdpcg_solve_dev_f32
        rc = f(None, _ptr(d_S), _ptr(d_P), _ptr(d_g), _ptr(d_l), nx, N, C.byref(cfg), _ptr(d_it), _ptr(d_ex), stream)
    _lib.check(rc)
    torch.cuda.synchronize()
    return d_l.cpu().numpy(), int(d_it.item()), int(d_ex.item())


def _compare(S, P, gam, lam0, nx, N, tol, max_iter, same_kernel):
    """New entry point against the existing one on the same system.  same_kernel: the existing entry point takes the cooperative-grid
    kernel with the same plan, so the two runs execute the same arithmetic in the same order -- bit-identical; otherwise (the existing
    entry point solves this size inside one cluster) the suite's bar: iterations within +-1, lambda within 1e-5 (fp32) / 1e-10 (fp64)
    at equal iteration count."""
    before = g.kernel_launches()
    lam_n, it_n, ex_n = _call("static", S, P, gam, lam0, nx, N, tol, max_iter)
    assert g.kernel_launches() == before + 1, "the static-scratch solve did not launch the CUDA kernel"
    lam_e, it_e, ex_e = _call("dev", S, P, gam, lam0, nx, N, tol, max_iter)
    assert not ex_n and not ex_e and np.isfinite(lam_n).all()
    if same_kernel:
        assert it_n == it_e and np.array_equal(lam_n, lam_e)
        return
    assert abs(it_n - it_e) <= 1, (it_n, it_e)
    if it_n != it_e:                                    # equal iteration count: both capped at the smaller one, exit test off
        k = min(it_n, it_e)
        lam_n, it_n, _ = _call("static", S, P, gam, lam0, nx, N, 0.0, k)
        lam_e, it_e, _ = _call("dev", S, P, gam, lam0, nx, N, 0.0, k)
        assert it_n == it_e == k
    assert rel_err(lam_n, lam_e) <= LAMBDA_TOL[S.dtype], (rel_err(lam_n, lam_e), it_n)


def test_static_scratch_matches_existing_entry_point_two_calls_in_a_row():
    """Two calls in a row, different element type and size, so the second starts on the first one's leftovers in the shared array:
    fp64 nx 14, N 1000 and fp32 nx 14, N 1500 -- the two beyond-one-cluster shapes of the parity suite, where the existing entry
    point runs the same kernel from a pool allocation."""
    for shape in [(14, 7, 1000, np.float64, 1e-10, 4000), (14, 7, 1500, np.float32, 1e-6, 4000)]:
        S, P, gam, nx, N, tol, max_iter = make_system(shape, "stair", seed=3)
        assert "pcg_grid" in g.plan_query(nx, N, S.dtype, True)["kernel"]
        _compare(S, P, gam, np.zeros(N * nx, dtype=S.dtype), nx, N, tol, max_iter, same_kernel=True)


@pytest.mark.parametrize("shape,precond", [((14, 7, 32, np.float32, 1e-6, 1000), "stair"), ((8, 4, 200, np.float64, 1e-10, 2000), "stair"),
                                           ((12, 4, 1200, np.float32, 1e-6, 5000), "identity"), ((2, 1, 2, np.float64, 1e-12, 100), "jacobi")])
def test_static_scratch_any_size_vs_existing_entry_point(shape, precond):
    """Sizes the existing entry point solves inside one CTA or cluster (and the minimum system): the new entry point still runs the grid
    kernel; twice in a row with a warm start the second time."""
    S, P, gam, nx, N, tol, max_iter = make_system(shape, precond, seed=4)
    rng = np.random.default_rng(1)
    for lam0 in (np.zeros(N * nx, dtype=S.dtype), (0.1 * rng.standard_normal(N * nx)).astype(S.dtype)):
        _compare(S, P, gam, lam0, nx, N, tol, max_iter, same_kernel=False)


def test_static_scratch_semantics_and_null_outputs():
    """max_iter = 0 leaves lambda and reports the flag; a low cap reports it; d_iters / d_max_iter_exit may be NULL."""
    import torch
    S, P, gam, nx, N, tol, _ = make_system((14, 7, 900, np.float32, 1e-6, 0), "stair", seed=6)
    lam0 = np.random.default_rng(0).standard_normal(N * nx).astype(np.float32)
    lam, it, ex = _call("static", S, P, gam, lam0, nx, N, tol, 0)
    assert np.array_equal(lam, lam0) and it == 0 and ex == 1
    lam, it, ex = _call("static", S, P, gam, lam0, nx, N, tol, 3)
    assert it == 3 and ex == 1 and np.isfinite(lam).all()
    lam_full, it_full, _ = _call("static", S, P, gam, lam0, nx, N, tol, 4000)
    d_S, d_P, d_g, d_l = (_dev(torch, a) for a in (S, P, gam, lam0))
    cfg = _lib.default_config()
    cfg.exit_tol, cfg.max_iter, cfg.empty_pinv = tol, 4000, 0
    _lib.check(_lib.lib().gbdpcg_solve_dev_static_scratch(0, _ptr(d_S), _ptr(d_P), _ptr(d_g), _ptr(d_l), nx, N, C.byref(cfg), None, None,
                                                          C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    assert np.array_equal(d_l.cpu().numpy(), lam_full)
    # the Python mirror reaches the same entry point
    d_l2 = _dev(torch, lam0)
    d_it = torch.zeros(1, dtype=torch.int32, device="cuda")
    g.solve_dev_static_scratch(nx, N, d_S, d_P, d_g, d_l2, g.pcg_config(pcg_exit_tol=tol, pcg_max_iter=4000, empty_pinv=0), d_it)
    torch.cuda.synchronize()
    assert int(d_it.item()) == it_full and np.array_equal(d_l2.cpu().numpy(), lam_full)


def test_static_scratch_rejects_what_it_cannot_hold():
    """Beyond the static array, beyond the grid kernel's co-residency limit, and the options the grid kernel does not have:
    GBDPCG_ERR_UNSUPPORTED with a message, nothing launched, lambda untouched."""
    import torch
    nx, nu, dt = 16, 8, np.float64
    N = None
    for cand in range(900, 2400, 100):                   # a shape the planner accepts whose exchange buffers exceed the array
        try:
            info = g.plan_query(nx, cand, dt, True)
        except g.GbdpcgError:
            break
        if "pcg_grid" in info["kernel"] and info["cluster"] * (6 * nx + 2) * 8 > _lib.STATIC_SCRATCH_BYTES:
            N = cand
            break
    assert N is not None, "no fp64 nx=16 grid shape exceeds GBDPCG_STATIC_SCRATCH_BYTES on this GPU"
    S, P, gam, nx, N, tol, max_iter = make_system((nx, nu, N, dt, 1e-10, 4000), "stair", seed=2)
    lam0 = np.ones(N * nx, dtype=dt)
    before = g.kernel_launches()
    with pytest.raises(g.GbdpcgError) as e:
        _call("static", S, P, gam, lam0, nx, N, tol, max_iter)
    assert e.value.status == _lib.ERR_UNSUPPORTED and "GBDPCG_STATIC_SCRATCH_BYTES" in str(e.value)
    assert g.kernel_launches() == before
    lam, it, ex = _call("dev", S, P, gam, lam0, nx, N, tol, max_iter)      # the allocating entry point takes it
    assert not ex and it > 0

    S, P, gam, nx, N, tol, max_iter = make_system((14, 7, 64, np.float32, 1e-6, 500), "stair")
    lam0 = np.zeros(N * nx, dtype=np.float32)
    for kw in (dict(flags=_lib.FLAG_TEXTBOOK), dict(cluster=4)):
        with pytest.raises(g.GbdpcgError) as e:
            _call("static", S, P, gam, lam0, nx, N, tol, max_iter, **kw)
        assert e.value.status == _lib.ERR_UNSUPPORTED
    # odd state size: no grid kernel is compiled for it
    d = torch.zeros(3 * 8 * 9 + 64, dtype=torch.float32, device="cuda")
    cfg = _lib.default_config()
    rc = _lib.lib().gbdpcg_solve_dev_static_scratch(0, _ptr(d), None, _ptr(d), _ptr(d), 3, 8, C.byref(cfg), None, None, None)
    assert rc == _lib.ERR_UNSUPPORTED
    # one knot point beyond the largest system the planner accepts
    lo, hi = 1024, 1 << 18
    while lo < hi:
        mid = (lo + hi + 1) // 2
        try:
            g.plan_query(14, mid, np.float32, True)
            lo = mid
        except g.GbdpcgError:
            hi = mid - 1
    big = torch.zeros(8, dtype=torch.float32, device="cuda")   # never dereferenced: the call is refused on the host
    cfg.empty_pinv = 0
    rc = _lib.lib().gbdpcg_solve_dev_static_scratch(0, _ptr(big), _ptr(big), _ptr(big), _ptr(big), 14, lo + 1, C.byref(cfg), None, None, None)
    assert rc == _lib.ERR_UNSUPPORTED
    assert g.kernel_launches() == before + 1
