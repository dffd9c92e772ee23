"""e2e of the host batched path (S, gamma, lambda; stair Phi^-1 built on the device) for the chunk counts in GBDPCG_HOST_CHUNKS,
plus the raw pinned H2D bandwidth of the same bytes.   usage: python scripts/time_e2e.py"""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g
from gbd_pcg_b200 import _lib
import ctypes as C
nx, nu, N, batch = 14, 7, 32, 8192
S, _, gam = g.synth.lqr_schur_batch(nx, nu, N, range(64), "identity", np.float32)
S, gam = np.tile(S, (batch // 64, 1)), np.tile(gam, (batch // 64, 1))
lib = _lib.lib()
def pinned(a):
    p = lib.gbdpcg_host_alloc(a.nbytes)
    out = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(a.size,)).reshape(a.shape)
    out[:] = a
    return out
hS, hg, hl = pinned(S), pinned(gam), pinned(np.zeros_like(gam))
t = torch.from_numpy(hS)
d = torch.empty(S.shape, dtype=torch.float32, device="cuda")
for _ in range(3):
    d.copy_(t, non_blocking=True); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    d.copy_(t, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 5
print(f"raw pinned H2D of S: {S.nbytes / dt / 1e9:.1f} GB/s ({dt * 1e3:.2f} ms)")
for sym, parts in [(s_, p_) for s_ in (0, 1) for p_ in (4, 8, 16, 32)]:
    cfg = g.pcg_config(pcg_exit_tol=1e-6, pcg_max_iter=1000, flags=_lib.FLAG_BUILD_STAIR | (_lib.FLAG_SYMMETRIC_S if sym else 0))
    os.environ["GBDPCG_HOST_CHUNKS"] = str(parts)
    for _ in range(2):
        hl[:] = 0; g.solve_batched_host(hS, None, hg, hl, nx, N, batch, cfg)
    ts = []
    for _ in range(5):
        hl[:] = 0
        t0 = time.perf_counter(); g.solve_batched_host(hS, None, hg, hl, nx, N, batch, cfg); ts.append(time.perf_counter() - t0)
    print(f"symmetric shipping {sym} chunks={parts}: {np.median(ts) * 1e3:.2f} ms -> {batch / np.median(ts) / 1e3:.1f} k solves/s")
