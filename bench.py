#!/usr/bin/env python
"""bench.py -- batched block-tridiagonal PCG solves/s on B200 (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch: 8192 independent iiwa-sized systems
(knot points N=32, state size nx=14, fp32, symmetric-stair Phi^-1, tol 1e-6: BASELINE config C5 = 8192 x C2)
partitioned by problem index over the GPUs (no collective on the data path).

  value      whole-job solves/s with inputs resident in HBM, CUDA events around each launch, max over ranks; with --gpus N the
             SAME 8192 problems are partitioned by problem index ("scaling": "strong", BASELINE config 5); 8192 per GPU is
             reported beside it (weak_scaling)
  e2e        same metric through the host-pointer C ABI (gbdpcg_solve_batched_host_f32) with the reference's host
             signature (S, gamma, lambda; Phi^-1 formed on the device): pinned host buffers, H2D + build + solve + D2H
             inside the timed region; the variant that also ships Phi^-1 from the host is reported beside it
  roofline   HBM roofline of the batched kernel (algorithmic bytes / launch time vs MEASURED_PEAKS.json)
  single_solve   p50/p99 latency of ONE C2 solve (the north-star latency target) in three scopes
  cpu_baseline   the host oracle (oracle/, test infrastructure) on the box's cores, bounded sample

--impl reference runs the UNMODIFIED reference CUDA kernel (oracle/_ref, built from /root/reference by
oracle/Makefile) on the same systems: it has no batch API, so its "batch" is a loop of single cooperative
launches (include/interface.cuh:130-132), the scope most favourable to it.
"""
import argparse
import ctypes as C
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NX, NU, N_KNOT, DTYPE, TOL, MAX_ITER = 14, 7, 32, np.float32, 1e-6, 1000
PRECOND = "stair"
BATCH = 8192
# systems beyond one 16-CTA cluster, timed in both arms (oracle/Makefile builds the reference harness for exactly these)
BEYOND_CLUSTER = {"N1500_f32": (14, 7, 1500, np.float32, 1e-6, 5000), "N1000_f64": (14, 7, 1000, np.float64, 1e-10, 5000)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clock and throttle reasons during the timed region (NVML; nvidia-smi fields of the recipe)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                 "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.02)

    def start(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join()
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": []}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def measured_traffic(batch):
    """dram__bytes_read + dram__bytes_write of ONE launch of the batched kernel, from the committed ncu --set full capture
    (profiles/traffic.json, written by scripts/summarize_profiles.py) -- not from this run.  Returns (bytes or None, source)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        t = json.load(open(p))
        if int(t["batch"]) == int(batch):
            return float(t["dram_bytes_per_launch"]), f"committed ncu capture profiles/traffic.json ({t.get('kernel', 'batched kernel')}, {t.get('source', 'ncu --set full')})"
        return None, f"no committed capture for a launch of {batch} systems (profiles/traffic.json holds batch {t['batch']})"
    except Exception:
        return None, "no committed capture"


def dist_env():
    # stdout carries exactly one JSON line: keep NCCL's "NCCL version ..." banner (NCCL_DEBUG=VERSION) off it
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return rank, local, world


def pct(xs, q):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(round(q * (len(xs) - 1))))]


def make_batch(seed0, count):
    from gbd_pcg_b200 import synth
    return synth.lqr_schur_batch(NX, NU, N_KNOT, range(seed0, seed0 + count), PRECOND, DTYPE)


def pinned_array(lib, shape, dtype):
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = lib.gbdpcg_host_alloc(n)
    if not p:
        raise MemoryError("gbdpcg_host_alloc failed")
    buf = (C.c_char * n).from_address(p)
    return np.frombuffer(buf, dtype=dtype).reshape(shape), p


# ---------------------------------------------------------------------------------------------------
def time_single_solves(torch, launch, d_l, n_warm=100, n_timed=1000):
    """Kernel-scope latency of one solve: CUDA events around each launch.  A ~60 us spin kernel is queued first
    so that the host is always ahead of the device and the events bracket only the solver kernel."""
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_timed)]
    for _ in range(n_warm):
        d_l.zero_()
        launch()
    torch.cuda.synchronize()
    for a, b in evs:
        torch.cuda._sleep(120000)
        d_l.zero_()
        a.record()
        launch()
        b.record()
    torch.cuda.synchronize()
    return [a.elapsed_time(b) * 1e3 for a, b in evs]  # microseconds


def time_blocking_calls(call, n_warm=20, n_timed=200):
    for _ in range(n_warm):
        call()
    ts = []
    for _ in range(n_timed):
        t0 = time.perf_counter()
        call()
        ts.append((time.perf_counter() - t0) * 1e6)
    return ts


def bind_to_gpu_numa(local):
    """Pin this rank's host threads to the CPUs NVML calls ideal for its GPU BEFORE any pinned buffer is allocated, so the
    staging memory of the host-pointer path is first-touched on the GPU's own NUMA node (eight ranks pulling from one node is
    what capped the aggregate H2D rate at 186 GB/s in round 1).  Best effort: silently a no-op without NVML."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1 and 64 * w + b < ncpu}
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def timed_launches(torch, step, reset, steps, flush=None):
    """CUDA events on the launching stream around each step; `reset` (and the optional L2 flush) sit outside the event pair."""
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    # a ~1.5 ms spin kernel first: the host stays ahead of the device for the whole loop, so an event pair brackets the step's kernel(s) and
    # not the host's launch latency (at 1,024 systems per GPU a step is ~96 us, the same order as one Python/ctypes call; time_single_solves does the same)
    torch.cuda._sleep(3000000)
    for a, b in evs:
        reset()
        if flush is not None:
            flush.sum()                    # READ a buffer larger than L2: evicts the inputs without leaving dirty lines behind
        a.record()
        step()
        b.record()
    return evs


def cpp_device_api_latency(S0, P0, g0):
    """p50 / p99 of the blocking device-pointer call measured inside a C++ program (examples/api_surface.exe --latency, built by
    __graft_entry__.build()): solvePCG<float> overload 3 through the header shim, and gbdpcg_solve_dev_sync_f32 with its own
    workspace and stream.  None when the binary is missing."""
    import subprocess
    import tempfile
    exe = os.path.join(ROOT, "examples", "api_surface.exe")
    if not os.path.exists(exe):
        return None
    try:
        with tempfile.NamedTemporaryFile(suffix=".bin") as f:
            for a in (S0, P0, g0):
                f.write(np.ascontiguousarray(a, dtype=np.float32).tobytes())
            f.flush()
            env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(ROOT, "gbd_pcg_b200") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
            r = subprocess.run([exe, "--latency", f.name, str(NX), str(N_KNOT), repr(TOL), str(MAX_ITER), "1000"], env=env,
                               capture_output=True, text=True, timeout=300)
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)[:200]}


def run_ours(args):
    import torch
    import torch.distributed as dist

    import gbd_pcg_b200 as g
    from gbd_pcg_b200 import _lib, layout
    from gbd_pcg_b200.partition import slab

    rank, local, world = dist_env()
    numa_cpus = bind_to_gpu_numa(local)
    torch.cuda.set_device(local)
    if world > 1:
        # control plane only: the timing barrier and the max-over-ranks of the timings.  No collective touches the data path.
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _lib.lib()
    total = args.batch                        # BASELINE config 5: the SAME `total` problems partitioned by problem index
    lo, hi = slab(rank, world, total)
    nloc = hi - lo

    S, P, gam = make_batch(lo, nloc)          # seeds = global problem index
    d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (S, P, gam))
    d_l = torch.zeros(nloc, N_KNOT * NX, dtype=torch.float32, device="cuda")
    d_it = torch.zeros(nloc, dtype=torch.int32, device="cuda")
    d_ex = torch.zeros(nloc, dtype=torch.uint8, device="cuda")
    cfg = g.pcg_config(pcg_exit_tol=TOL, pcg_max_iter=MAX_ITER, empty_pinv=0)
    plan = g.plan_query(NX, N_KNOT, DTYPE, True, cfg, batch=nloc)
    # inputs of one step: 1.25 GB at one GPU (>> 126 MB of L2); a partitioned slab can come close to L2, so it is flushed
    in_bytes = S.nbytes + P.nbytes + gam.nbytes
    flush = torch.zeros(64 << 20, dtype=torch.int32, device="cuda") if in_bytes < (512 << 20) else None     # 256 MB = 2 x L2

    def step():
        g.solve_batched_dev(NX, N_KNOT, nloc, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        tt = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    for _ in range(max(args.warmup, 3)):
        d_l.zero_()
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = g.kernel_launches()
    evs = timed_launches(torch, step, d_l.zero_, args.steps, flush)
    barrier()
    launches = g.kernel_launches() - launches0
    clocks = sampler.stop()
    step_ms = [a.elapsed_time(b) for a, b in evs]
    ms_per_step = max_over_ranks(sum(step_ms)) / args.steps
    value = total / (ms_per_step * 1e-3)
    iters = d_it.cpu().numpy()
    lam_gpu = d_l.cpu().numpy()
    assert int(d_ex.sum().item()) == 0, "some problems hit max_iter"

    # ---- parity of THIS run's output, every system of the slab against the host oracle (test infrastructure used as the checker):
    #      iterations within +-1, lambda <= 1e-5 relative at equal iteration count (BASELINE north_star bar) ----
    from oracle import oracle
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lam_c = np.zeros((nloc, N_KNOT * NX), dtype=DTYPE)
    t0 = time.perf_counter()
    lam_c, it_c, _ = oracle.pcg_batched(S, P, gam, lam_c, NX, N_KNOT, TOL, MAX_ITER, order=0, threads=cores)
    t_cpu = time.perf_counter() - t0
    d_iters = np.abs(iters.astype(np.int64) - it_c.astype(np.int64))
    same = d_iters == 0
    num = np.linalg.norm(lam_gpu.astype(np.float64) - lam_c.astype(np.float64), axis=1)
    den = np.linalg.norm(lam_c.astype(np.float64), axis=1)
    rel = num / den
    # systems whose counts differ by one: the oracle again, capped at the GPU's count (exit test disabled)
    for i in np.nonzero(~same & (d_iters <= 1))[0]:
        lo_i = np.zeros(N_KNOT * NX, dtype=DTYPE)
        lo_i, _, _ = oracle.pcg(S[i], P[i], gam[i], lo_i, NX, N_KNOT, 0.0, int(iters[i]), order=0)
        rel[i] = np.linalg.norm(lam_gpu[i].astype(np.float64) - lo_i) / np.linalg.norm(lo_i)
    par_max = max_over_ranks(float(rel.max()))
    par_dit = int(max_over_ranks(float(d_iters.max())))
    par_mis = torch.tensor([int((~same).sum())], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(par_mis, op=dist.ReduceOp.SUM)
    parity = {"checked": total, "max_rel_err": par_max, "iters_mismatch": int(par_mis.item()), "max_iters_diff": par_dit,
              "bar": "iterations within +-1, lambda <= 1e-5 relative at equal iteration count, against oracle/pcg_oracle.c"}
    assert par_dit <= 1 and par_max <= 1e-5, f"GPU result fails the parity bar: {parity}"

    # ---- sustained leg: back-to-back launches for >= 2 s, own clock record ----
    sampler2 = ClockSampler(local)
    n_sus = max(10, int(2.2 / (ms_per_step * 1e-3)))
    barrier()
    sampler2.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_sus):
        d_l.zero_()                    # (a warm-started lambda would converge in one iteration: reset it, as in the timed steps)
        step()
    e1.record()
    barrier()
    sus_clocks = sampler2.stop()
    sus_ms = max_over_ranks(e0.elapsed_time(e1))
    sustained = {"value": total * n_sus / (sus_ms * 1e-3), "unit": "solves/s", "launches": n_sus, "seconds": sus_ms * 1e-3, "clocks": sus_clocks,
                 "note": "lambda reset + launch back to back for >= 2 s; inputs re-read from HBM every launch"}

    # ---- end to end through the host-pointer C ABI (pinned host buffers, first-touched after the NUMA binding) ----
    hS, pS = pinned_array(lib, S.shape, DTYPE); hS[:] = S
    hP, pP = pinned_array(lib, P.shape, DTYPE); hP[:] = P
    hg, pg = pinned_array(lib, gam.shape, DTYPE); hg[:] = gam
    hl, pl = pinned_array(lib, gam.shape, DTYPE)
    e2e_steps = max(2, min(args.steps, 5))

    def time_host_path(h_P, cfg_h):
        for _ in range(2):
            hl[:] = 0
            g.solve_batched_host(hS, h_P, hg, hl, NX, N_KNOT, nloc, cfg_h)
        barrier()
        t = 0.0
        for _ in range(e2e_steps):
            hl[:] = 0
            barrier()
            t0 = time.perf_counter()
            its, _ = g.solve_batched_host(hS, h_P, hg, hl, NX, N_KNOT, nloc, cfg_h)
            t += time.perf_counter() - t0
        return total / (max_over_ranks(t) / e2e_steps), its

    # (a) the reference's own host signature: S, gamma, lambda only (include/interface.cuh:24-31 has no Pinv parameter);
    #     the symmetric-stair Phi^-1 is formed from S on the device (GBDPCG_FLAG_BUILD_STAIR)
    cfg_build = g.pcg_config(pcg_exit_tol=TOL, pcg_max_iter=MAX_ITER, flags=_lib.FLAG_BUILD_STAIR)
    e2e_full, e_it = time_host_path(None, cfg_build)
    lam_build = hl.copy()
    assert int(np.abs(e_it.astype(np.int64) - iters.astype(np.int64)).max()) <= 1, "device-built preconditioner changed iteration counts"
    h2d_full = (S.nbytes + gam.nbytes + hl.nbytes)
    d2h = (hl.nbytes + 4 * nloc + nloc)
    # (a') the same call with GBDPCG_FLAG_SYMMETRIC_S: S is declared symmetric -> only its diagonal and right blocks cross PCIe (2/3 of
    #      the bytes), the left blocks are rebuilt on the device by transposition.  The flag is opt-in (the reference kernel itself never
    #      assumes symmetry); the bench first VERIFIES on the host that every left block is bit for bit the transpose of the right block
    #      above it, and the result must equal the full-S result bit for bit.  This is the e2e headline; the full-S number is beside it.
    Sv = S.reshape(nloc, N_KNOT, 3, NX, NX)
    symmetric = bool(np.array_equal(Sv[:, 1:, 0], np.swapaxes(Sv[:, :-1, 2], -1, -2)))
    e2e_value, h2d, e2e_api = e2e_full, h2d_full, "full S shipped"
    e2e_extra = {"full_s": {"value": e2e_full, "h2d_bytes_per_step": int(h2d_full)}, "host_s_verified_symmetric": symmetric}
    if symmetric:
        cfg_sym = g.pcg_config(pcg_exit_tol=TOL, pcg_max_iter=MAX_ITER, flags=_lib.FLAG_BUILD_STAIR | _lib.FLAG_SYMMETRIC_S)
        e2e_value, it_sym = time_host_path(None, cfg_sym)
        assert np.array_equal(it_sym, e_it) and np.array_equal(hl, lam_build), "symmetric shipping changed the result"
        h2d = S.nbytes * 2 // 3 + gam.nbytes + hl.nbytes
        e2e_api = "GBDPCG_FLAG_SYMMETRIC_S (host S verified exactly symmetric): diagonal + right blocks shipped, left blocks rebuilt on the device"
    # (b) same call with Phi^-1 handed over by the host (twice the PCIe bytes)
    e2e_hostp, e_it2 = time_host_path(hP, cfg)
    assert np.array_equal(e_it2, iters.astype(np.uint32)), "host path and device path disagree"
    lam_err = float(np.abs(lam_build - hl).max() / np.abs(hl).max())
    for p in (pS, pP, pg, pl):
        lib.gbdpcg_host_free(p)

    # ---- weak-scaling supplement: `total` problems on EVERY GPU (at one GPU it is the headline itself) ----
    weak_value = value
    if world > 1:
        del d_S, d_P, d_g, d_l, d_it, d_ex
        Sw, Pw, gw = make_batch(total * (rank + 1), total)          # fresh seeds, disjoint over ranks
        w_S, w_P, w_g = (torch.from_numpy(a).cuda() for a in (Sw, Pw, gw))
        w_l = torch.zeros(total, N_KNOT * NX, dtype=torch.float32, device="cuda")
        w_it = torch.zeros(total, dtype=torch.int32, device="cuda")

        def step_weak():
            g.solve_batched_dev(NX, N_KNOT, total, w_S, w_P, w_g, w_l, cfg, w_it, None)
        for _ in range(3):
            w_l.zero_()
            step_weak()
        barrier()
        evw = timed_launches(torch, step_weak, w_l.zero_, args.steps)
        barrier()
        weak_ms = max_over_ranks(sum(a.elapsed_time(b) for a, b in evw)) / args.steps
        weak_value = total * world / (weak_ms * 1e-3)
        del w_S, w_P, w_g, w_l, w_it

    out = None
    if rank == 0:
        peak, peak_src = peaks()
        bytes_per_solve = layout.algorithmic_bytes(NX, N_KNOT, 4, True)
        launch_s = statistics.mean(step_ms) * 1e-3
        achieved = bytes_per_solve * nloc / launch_s / 1e9
        traffic, traffic_src = measured_traffic(nloc)
        out = {
            "metric": "batched PCG solves/s (8192 x N=32,nx=14 fp32, partitioned by problem index over the GPUs)", "value": value,
            "unit": "solves/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C5: {total} independent iiwa MPC systems (knot points 32, state size 14, fp32, symmetric-stair "
                                   "preconditioner, tol 1e-6), seed = problem index, partitioned by problem index over the GPUs",
                       "total_batch": total, "batch_per_gpu": nloc, "nx": NX, "knot_points": N_KNOT, "precond": PRECOND, "exit_tol": TOL,
                       "l2": (f"inputs ({in_bytes >> 20} MB per GPU) several times larger than the 126 MB L2; no flush needed" if flush is None else
                              f"slab inputs {in_bytes >> 20} MB per GPU: a 256 MB buffer (2 x L2) is read between timed launches (L2 flush that "
                              "leaves no dirty lines to be written back under the timed launch)"),
                       "kernel": plan["kernel"], "threads": plan["threads"], "cluster": plan["cluster"], "smem_bytes": plan["smem_bytes"],
                       "collective": "none (NCCL used for the timing barrier and the max over ranks only)",
                       "host_threads_bound_to_gpu_numa_cpus": numa_cpus},
            "pcg_iters": {"min": int(iters.min()), "mean": float(iters.mean()), "max": int(iters.max())},
            "iterations_per_s": float(iters.mean()) * value,
            "parity": parity,
            "e2e": {"value": e2e_value, "unit": "solves/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "api": "gbdpcg_solve_batched_host_f32(S, gamma, lambda) with GBDPCG_FLAG_BUILD_STAIR: the "
                    "reference's host signature (no Pinv argument), Phi^-1 formed on the device; pinned host buffers, chunked "
                    "2-stream overlap; bytes are per GPU; " + e2e_api, "with_host_pinv": {"value": e2e_hostp, "h2d_bytes_per_step": int(h2d + P.nbytes)},
                    "lambda_max_diff_vs_host_pinv": lam_err, **e2e_extra},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes_per_launch": bytes_per_solve * nloc,
                         "peak_source": peak_src, "bytes_per_solve": bytes_per_solve, "kernel_ms": launch_s * 1e3},
            "clocks": clocks,
            "sustained": sustained,
            "weak_scaling": {"value": weak_value, "unit": "solves/s", "batch_per_gpu": total,
                             "note": f"supplement: {total} problems on every GPU (fixed work per GPU)"},
            "cpu_baseline": None,
        }
        if world == 1:
            out.update(single_gpu_extras(torch, g, _lib, S, P, gam, cfg, cfg_build, d_S, d_P, d_it, d_ex))
            t0 = time.perf_counter()
            for i in range(50):
                oracle.pcg(S[i], P[i], gam[i], lam_c[0] * 0, NX, N_KNOT, TOL, MAX_ITER)
            t_cpu1 = (time.perf_counter() - t0) / 50
            out["cpu_baseline"] = {"value": nloc / t_cpu, "unit": "solves/s", "cores": cores, "kind": "port",
                                   "sample": f"all {nloc} systems of the step, host oracle (oracle/pcg_oracle.c), {cores} threads -- the same run "
                                             "that checks `parity`", "single_thread_solve_us": t_cpu1 * 1e6, "iters_mean": float(it_c.mean())}
        else:
            out.pop("cpu_baseline")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def single_gpu_extras(torch, g, _lib, S, P, gam, cfg, cfg_build, d_S, d_P, d_it, d_ex):
    """What explains the headline, measured at one GPU only: single-solve latency in three scopes (plus the C++-measured API
    latency), the other BASELINE configs, an fp64 batch, the device-resident preconditioner build and the whole MPC step."""
    import statistics as st
    batch = S.shape[0]
    out = {}

    def sync():
        torch.cuda.synchronize()

    # ---- single-solve latency (C2, seed 0), kernel scope / device API / host API ----
    S1, P1, g1 = (torch.from_numpy(a[0].copy()).cuda() for a in (S, P, gam))
    l1 = torch.zeros(N_KNOT * NX, dtype=torch.float32, device="cuda")
    i1 = torch.zeros(1, dtype=torch.int32, device="cuda")
    x1 = torch.zeros(1, dtype=torch.uint8, device="cuda")
    lat_k = time_single_solves(torch, lambda: g.solve_dev_async(NX, N_KNOT, S1, P1, g1, l1, cfg, i1, x1), l1)
    it1 = int(i1.item())

    def dev_call():
        l1.zero_()
        g.api._solve_dev(NX, N_KNOT, S1, P1, g1, l1, config=cfg)
    lat_d = time_blocking_calls(dev_call)
    hS1, hP1, hg1 = S[0].copy(), P[0].copy(), gam[0].copy()
    hl1 = np.zeros(N_KNOT * NX, dtype=DTYPE)

    def host_call():
        hl1[:] = 0
        g.pcg_solve(hS1, hg1, hl1, NX, N_KNOT, cfg, h_Pinv=hP1)
    lat_h = time_blocking_calls(host_call)

    def host_call_build():
        hl1[:] = 0
        g.pcg_solve(hS1, hg1, hl1, NX, N_KNOT, cfg_build)
    lat_hb = time_blocking_calls(host_call_build)
    out["single_solve"] = {"config": "C2 N=32 nx=14 fp32 stair, seed 0", "iters": it1,
                           "kernel_p50_us": pct(lat_k, 0.5), "kernel_p99_us": pct(lat_k, 0.99),
                           "iterations_per_s": it1 / (pct(lat_k, 0.5) * 1e-6),
                           "dev_api_p50_us": pct(lat_d, 0.5), "dev_api_p99_us": pct(lat_d, 0.99),
                           "dev_api_note": "through Python/ctypes; `cpp` holds the same call timed inside a C++ program",
                           "cpp": cpp_device_api_latency(hS1, hP1, hg1),
                           "host_api_p50_us": pct(lat_h, 0.5), "host_api_p99_us": pct(lat_h, 0.99),
                           "host_api_device_built_pinv_p50_us": pct(lat_hb, 0.5), "host_api_device_built_pinv_p99_us": pct(lat_hb, 0.99)}

    # ---- the other BASELINE configs (parity cases): single-solve latency, kernel scope, stair preconditioner ----
    other = {}
    for name in ("C1", "C3", "C4"):
        onx, onu, oN, odt, otol, omax = g.synth.CONFIGS[name]
        if name == "C1":
            So, Po, go = g.synth.example_system(odt, "stair")
        else:
            So, Po, go = g.synth.lqr_schur(onx, onu, oN, 0, "stair", odt)
        dS, dP, dg = (torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (So, Po, go))
        dl = torch.zeros(oN * onx, dtype=dS.dtype, device="cuda")
        di = torch.zeros(1, dtype=torch.int32, device="cuda")
        ocfg = g.pcg_config(pcg_exit_tol=otol, pcg_max_iter=omax, empty_pinv=0)
        lat = time_single_solves(torch, lambda: g.solve_dev_async(onx, oN, dS, dP, dg, dl, ocfg, di, None), dl, n_warm=20, n_timed=200)
        oplan = g.plan_query(onx, oN, odt, True, ocfg)
        other[name] = {"nx": onx, "knot_points": oN, "dtype": np.dtype(odt).name, "iters": int(di.item()), "kernel_p50_us": pct(lat, 0.5),
                       "kernel_p99_us": pct(lat, 0.99), "cluster": oplan["cluster"], "threads": oplan["threads"]}

    # ---- beyond one 16-CTA cluster (the cooperative-grid kernel): single-solve latency, kernel scope ----
    for name, (onx, onu, oN, odt, otol, omax) in BEYOND_CLUSTER.items():
        So, Po, go = g.synth.lqr_schur(onx, onu, oN, 0, "stair", odt)
        dS, dP, dg = (torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (So, Po, go))
        dl = torch.zeros(oN * onx, dtype=dS.dtype, device="cuda")
        di = torch.zeros(1, dtype=torch.int32, device="cuda")
        ocfg = g.pcg_config(pcg_exit_tol=otol, pcg_max_iter=omax, empty_pinv=0)
        lat = time_single_solves(torch, lambda: g.solve_dev_async(onx, oN, dS, dP, dg, dl, ocfg, di, None), dl, n_warm=10, n_timed=60)
        oplan = g.plan_query(onx, oN, odt, True, ocfg)
        other[name] = {"nx": onx, "knot_points": oN, "dtype": np.dtype(odt).name, "iters": int(di.item()), "kernel_p50_us": pct(lat, 0.5),
                       "kernel_p99_us": pct(lat, 0.99), "us_per_iteration": pct(lat, 0.5) / max(int(di.item()), 1),
                       "kernel": oplan["kernel"], "ctas": oplan["cluster"], "threads": oplan["threads"]}
        del dS, dP, dg, dl

    # ---- fp64 batch: 2048 x C3 (nx 14, N 64, fp64, tol 1e-10, stair), 64 distinct seeds tiled ----
    from gbd_pcg_b200 import layout
    onx, onu, oN, odt, otol, omax = g.synth.CONFIGS["C3"]
    nb, nseed = 2048, 64
    S3, P3, g3 = g.synth.lqr_schur_batch(onx, onu, oN, range(nseed), "stair", odt)
    dS, dP, dg = (torch.from_numpy(a).cuda().repeat(nb // nseed, 1) for a in (S3, P3, g3))
    dl = torch.zeros(nb, oN * onx, dtype=torch.float64, device="cuda")
    di = torch.zeros(nb, dtype=torch.int32, device="cuda")
    ocfg = g.pcg_config(pcg_exit_tol=otol, pcg_max_iter=omax, empty_pinv=0)

    def step64():
        g.solve_batched_dev(onx, oN, nb, dS, dP, dg, dl, ocfg, di, None)
    for _ in range(3):
        dl.zero_()
        step64()
    sync()
    ev = timed_launches(torch, step64, dl.zero_, 10)
    sync()
    ms64 = st.mean(a.elapsed_time(b) for a, b in ev)
    b64 = layout.algorithmic_bytes(onx, oN, 8, True)
    peak, _ = peaks()
    p64 = g.plan_query(onx, oN, odt, True, ocfg, batch=nb)
    from oracle import oracle
    worst = 0.0
    its64 = di.cpu().numpy()
    lam64 = dl[:nseed].cpu().numpy()
    for s_ in range(0, nseed, 8):
        lo, ito, _ = oracle.pcg(S3[s_], P3[s_], g3[s_], np.zeros(oN * onx), onx, oN, otol, omax)
        assert abs(int(its64[s_]) - ito) <= 1
        if int(its64[s_]) == ito:
            worst = max(worst, float(np.linalg.norm(lam64[s_] - lo) / np.linalg.norm(lo)))
    assert worst <= 1e-10, worst
    other["C3_batched"] = {"batch": nb, "nx": onx, "knot_points": oN, "dtype": "float64", "iters_mean": float(its64.mean()), "ms_per_launch": ms64,
                           "solves_per_s": nb / (ms64 * 1e-3), "bytes_per_solve": b64, "roofline_frac": b64 * nb / (ms64 * 1e-3) / 1e9 / peak,
                           "hbm_line_solves_per_s": peak * 1e9 / b64, "kernel": p64["kernel"], "cluster": p64["cluster"], "threads": p64["threads"],
                           "parity_max_rel_err_vs_oracle": worst}
    del dS, dP, dg, dl, di

    # ---- fp32 batch of systems twice the knot points of C2: 4096 x (nx 14, N 64, fp32, tol 1e-6, stair) -- the wide form of the four-row
    #      tensor-memory kernel (8 warps, one system per SM) where the register-row plan needs a 2-CTA cluster per system ----
    wnx, wnu, wN, wtol, wmax = 14, 7, 64, 1e-6, 1000
    nbw, nseedw = 4096, 64
    Sw, Pw, gw = g.synth.lqr_schur_batch(wnx, wnu, wN, range(nseedw), "stair", np.float32)
    dS, dP, dg = (torch.from_numpy(a).cuda().repeat(nbw // nseedw, 1) for a in (Sw, Pw, gw))
    dl = torch.zeros(nbw, wN * wnx, dtype=torch.float32, device="cuda")
    di = torch.zeros(nbw, dtype=torch.int32, device="cuda")
    wcfg = g.pcg_config(pcg_exit_tol=wtol, pcg_max_iter=wmax, empty_pinv=0)

    def stepw():
        g.solve_batched_dev(wnx, wN, nbw, dS, dP, dg, dl, wcfg, di, None)
    for _ in range(3):
        dl.zero_()
        stepw()
    sync()
    ev = timed_launches(torch, stepw, dl.zero_, 10)
    sync()
    msw = st.mean(a.elapsed_time(b) for a, b in ev)
    bw = layout.algorithmic_bytes(wnx, wN, 4, True)
    pw = g.plan_query(wnx, wN, np.float32, True, wcfg, batch=nbw)
    worst = 0.0
    itsw = di.cpu().numpy()
    lamw = dl[:nseedw].cpu().numpy()
    for s_ in range(0, nseedw, 8):
        lo, ito, _ = oracle.pcg(Sw[s_], Pw[s_], gw[s_], np.zeros(wN * wnx, np.float32), wnx, wN, wtol, wmax)
        assert abs(int(itsw[s_]) - ito) <= 1
        if int(itsw[s_]) == ito:
            worst = max(worst, float(np.linalg.norm(lamw[s_].astype(np.float64) - lo) / np.linalg.norm(lo)))
    assert worst <= 1e-5, worst
    other["N64_batched"] = {"batch": nbw, "nx": wnx, "knot_points": wN, "dtype": "float32", "iters_mean": float(itsw.mean()), "ms_per_launch": msw,
                            "solves_per_s": nbw / (msw * 1e-3), "bytes_per_solve": bw, "roofline_frac": bw * nbw / (msw * 1e-3) / 1e9 / peak,
                            "hbm_line_solves_per_s": peak * 1e9 / bw, "kernel": pw["kernel"], "cluster": pw["cluster"], "threads": pw["threads"],
                            "parity_max_rel_err_vs_oracle": worst}
    del dS, dP, dg, dl, di

    # ---- fp32 batch of C4-sized systems: 1024 x (nx 12, N 512, fp32, tol 1e-6, stair) -- beyond the single-CTA tensor-memory kernels: the
    #      four-row tensor-memory form of the cluster kernel, 8 CTAs per system where the register-row plan needs 16 ----
    cnx, cnu, cN, cdt, ctol, cmax = g.synth.CONFIGS["C4"]
    nbc, nseedc = 1024, 16
    Sc, Pc, gc = g.synth.lqr_schur_batch(cnx, cnu, cN, range(nseedc), "stair", np.float32)
    dS, dP, dg = (torch.from_numpy(a).cuda().repeat(nbc // nseedc, 1) for a in (Sc, Pc, gc))
    dl = torch.zeros(nbc, cN * cnx, dtype=torch.float32, device="cuda")
    di = torch.zeros(nbc, dtype=torch.int32, device="cuda")
    ccfg = g.pcg_config(pcg_exit_tol=ctol, pcg_max_iter=cmax, empty_pinv=0)

    def stepc():
        g.solve_batched_dev(cnx, cN, nbc, dS, dP, dg, dl, ccfg, di, None)
    for _ in range(3):
        dl.zero_()
        stepc()
    sync()
    ev = timed_launches(torch, stepc, dl.zero_, 10)
    sync()
    msc = st.mean(a.elapsed_time(b) for a, b in ev)
    bc = layout.algorithmic_bytes(cnx, cN, 4, True)
    pc = g.plan_query(cnx, cN, np.float32, True, ccfg, batch=nbc)
    worst = 0.0
    itsc = di.cpu().numpy()
    lamc = dl[:nseedc].cpu().numpy()
    for s_ in range(0, nseedc, 4):
        lo, ito, _ = oracle.pcg(Sc[s_], Pc[s_], gc[s_], np.zeros(cN * cnx, np.float32), cnx, cN, ctol, cmax)
        assert abs(int(itsc[s_]) - ito) <= 1
        if int(itsc[s_]) == ito:
            worst = max(worst, float(np.linalg.norm(lamc[s_].astype(np.float64) - lo) / np.linalg.norm(lo)))
    assert worst <= 1e-5, worst
    other["C4_batched"] = {"batch": nbc, "nx": cnx, "knot_points": cN, "dtype": "float32", "iters_mean": float(itsc.mean()), "ms_per_launch": msc,
                           "solves_per_s": nbc / (msc * 1e-3), "bytes_per_solve": bc, "roofline_frac": bc * nbc / (msc * 1e-3) / 1e9 / peak,
                           "hbm_line_solves_per_s": peak * 1e9 / bc, "kernel": pc["kernel"], "cluster": pc["cluster"], "threads": pc["threads"],
                           "parity_max_rel_err_vs_oracle": worst}
    del dS, dP, dg, dl, di
    out["other_configs"] = other

    # ---- device-resident preconditioner build (events) ----
    d_Pb = torch.empty_like(d_S)
    for _ in range(3):
        g.build_pinv_dev(NX, N_KNOT, batch, d_S, d_Pb, "stair")
    sync()
    evb = timed_launches(torch, lambda: g.build_pinv_dev(NX, N_KNOT, batch, d_S, d_Pb, "stair"), lambda: None, 10)
    sync()
    build_ms = st.median(a.elapsed_time(b) for a, b in evb)
    build_err = float((d_Pb - d_P).abs().max().item() / d_P.abs().max().item())
    del d_Pb
    out["pinv_build"] = {"ms": build_ms, "gb_per_s": 2 * S.nbytes / build_ms / 1e6, "max_rel_diff_vs_fp64_host_build": build_err,
                         "note": "gbdpcg_build_pinv_dev_f32 (stair) on the resident batch: reads S, writes Phi^-1"}

    # ---- the whole device-resident MPC step: stage data -> S, gamma -> stair Phi^-1 -> PCG -> (x, u) ----
    nseed = 16
    sd = [g.synth.lqr_stage_data(NX, NU, N_KNOT, s_, dtype=DTYPE) for s_ in range(nseed)]
    dd = {k: torch.from_numpy(np.concatenate([d_[k] for d_ in sd])).cuda().repeat(batch // nseed, 1).contiguous() for k in sd[0]}
    f32 = dict(dtype=torch.float32, device="cuda")
    mS, mg = torch.empty(batch * 3 * N_KNOT * NX * NX, **f32), torch.empty(batch * N_KNOT * NX, **f32)
    mP, ml = torch.empty_like(mS), torch.zeros(batch * N_KNOT * NX, **f32)
    mQi, mRi = torch.empty(batch * N_KNOT * NX * NX, **f32), torch.empty(batch * (N_KNOT - 1) * NU * NU, **f32)
    mx, mu = torch.empty(batch * N_KNOT * NX, **f32), torch.empty(batch * (N_KNOT - 1) * NU, **f32)
    stages = [("schur_form_ms", lambda: g.schur_form_dev(NX, NU, N_KNOT, batch, dd["Q"], dd["R"], dd["A"], dd["B"], dd["q"], dd["r"], dd["c"],
                                                         mS, mg, mQi, mRi)),
              ("build_pinv_ms", lambda: g.build_pinv_dev(NX, N_KNOT, batch, mS, mP, "stair")),
              ("pcg_ms", lambda: (ml.zero_(), g.solve_batched_dev(NX, N_KNOT, batch, mS, mP, mg, ml, cfg, d_it, d_ex))),
              ("recover_ms", lambda: g.schur_recover_dev(NX, NU, N_KNOT, batch, mQi, mRi, dd["A"], dd["B"], dd["q"], dd["r"], ml, mx, mu))]
    mpc = {}
    for name, fn in stages:
        ts = []
        for i_ in range(8):
            torch.cuda._sleep(200000)
            ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ea.record(); fn(); eb.record()
            torch.cuda.synchronize()
            if i_ >= 3:
                ts.append(ea.elapsed_time(eb))
        mpc[name] = st.median(ts)
    mpc["steps_per_s"] = batch / (sum(mpc.values()) * 1e-3)
    mpc["note"] = "gbdpcg_schur_form_dev + build_pinv_dev(stair) + solve_batched_dev + schur_recover_dev on the resident batch"
    out["mpc_step"] = mpc
    return out


def run_reference(args):
    """Reference arm: the unmodified reference CUDA kernel, one cooperative launch per system."""
    rank, local, world = dist_env()
    if rank != 0:
        return None
    import torch

    from oracle import oracle, ref_cuda
    torch.cuda.set_device(0)
    if not ref_cuda.has(DTYPE, NX, N_KNOT):
        return {"impl": "reference", "unavailable": "oracle/_ref/libref_pcg_f32_14_32.so not built (needs /root/reference at build time)"}
    ref = ref_cuda.get(DTYPE, NX, N_KNOT)
    sample = min(args.batch, args.ref_sample)
    S, P, gam = make_batch(0, sample)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda() for a in (S, P, gam))
    d_l = torch.zeros(sample, N_KNOT * NX, dtype=torch.float32, device="cuda")

    # pre-sliced views keep Python overhead off the device's critical path
    rows = [(d_S[i], d_P[i], d_g[i], d_l[i]) for i in range(sample)]

    def step():
        for s, p, gg, l in rows:
            ref.launch(s, p, gg, l, MAX_ITER, TOL)

    for _ in range(max(args.warmup, 3)):
        d_l.zero_()
        step()
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for a, b in evs:
        d_l.zero_()
        a.record()
        step()
        b.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    step_ms = [a.elapsed_time(b) for a, b in evs]
    ms_per_step = statistics.mean(step_ms)
    value = sample / (ms_per_step * 1e-3)

    # supplementary: the same systems with up to 48 reference solves in flight (one stream each, own scratch).  The
    # reference API cannot do this (legacy default stream, blocking read-back); reported so that the baseline is not
    # only the serial loop.
    concurrent = None
    if hasattr(ref.L, "ref_kernel_launch_stream"):
        n_lanes = 48
        lanes = [ref.make_lane(torch.cuda.Stream()) for _ in range(n_lanes)]
        cur = torch.cuda.current_stream()

        def step_lanes():
            t0 = torch.cuda.Event(enable_timing=True)
            t1 = torch.cuda.Event(enable_timing=True)
            t0.record(cur)
            for ln in lanes:
                ln["stream"].wait_event(t0)
            for i, (s, p, gg, l) in enumerate(rows):
                ref.launch_on(lanes[i % n_lanes], s, p, gg, l, MAX_ITER, TOL)
            for ln in lanes:
                cur.wait_stream(ln["stream"])
            t1.record(cur)
            return t0, t1
        for _ in range(3):
            d_l.zero_()
            step_lanes()
        torch.cuda.synchronize()
        tl = []
        for _ in range(args.steps):
            d_l.zero_()
            tl.append(step_lanes())
        torch.cuda.synchronize()
        ms_l = statistics.mean(a.elapsed_time(b) for a, b in tl)
        concurrent = {"value": sample / (ms_l * 1e-3), "unit": "solves/s", "streams": n_lanes, "ms_per_step": ms_l,
                      "note": "same launches spread over 48 streams with per-stream scratch; not reachable through the reference API"}

    # single-solve latency, same three scopes as our arm
    l1 = d_l[0]
    lat_k = time_single_solves(torch, lambda: ref.launch(d_S[0], d_P[0], d_g[0], l1, MAX_ITER, TOL), l1)
    it1 = int(ref.d_iters.item())

    def dev_call():
        l1.zero_()
        ref.solve_dev_api(d_S[0], d_P[0], d_g[0], l1, MAX_ITER, TOL)
    lat_d = time_blocking_calls(dev_call)
    hS1, hg1, hl1 = S[0].copy(), gam[0].copy(), np.zeros(N_KNOT * NX, dtype=DTYPE)

    def host_call():  # time only: the reference never initialises d_Pinv on this path (interface.cuh:45-46)
        hl1[:] = 0
        ref.solve_host_api(hS1, hg1, hl1, 25, TOL)
    lat_h = time_blocking_calls(host_call, n_warm=5, n_timed=50)

    # ---- the other BASELINE configs through the same unmodified reference kernel (kernel scope, stair preconditioner): what the
    #      repo arm's `other_configs` is to be read against.  C4's shared-memory under-allocation is covered from
    #      the harness side (extra_smem), the reference itself is not edited. ----
    from gbd_pcg_b200 import synth
    other = {}
    for name in ("C1", "C3", "C4"):
        onx, onu, oN, odt, otol, omax = synth.CONFIGS[name]
        if not ref_cuda.has(odt, onx, oN):
            other[name] = {"unavailable": f"oracle/_ref/libref_pcg_{'f64' if np.dtype(odt) == np.float64 else 'f32'}_{onx}_{oN}.so not built"}
            continue
        if name == "C1":
            So, Po, go = synth.example_system(odt, "stair")
        else:
            So, Po, go = synth.lqr_schur(onx, onu, oN, 0, "stair", odt)
        rk = ref_cuda.get(odt, onx, oN)
        dS, dP, dg = (torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (So, Po, go))
        dl = torch.zeros(oN * onx, dtype=dS.dtype, device="cuda")
        lat = time_single_solves(torch, lambda: rk.launch(dS, dP, dg, dl, omax, otol), dl, n_warm=20, n_timed=200)
        other[name] = {"nx": onx, "knot_points": oN, "dtype": np.dtype(odt).name, "iters": int(rk.d_iters.item()), "kernel_p50_us": pct(lat, 0.5),
                       "kernel_p99_us": pct(lat, 0.99), **rk.attrs()}

    for name, (onx, onu, oN, odt, otol, omax) in BEYOND_CLUSTER.items():
        if not ref_cuda.has(odt, onx, oN):
            other[name] = {"unavailable": f"oracle/_ref/libref_pcg_{'f64' if np.dtype(odt) == np.float64 else 'f32'}_{onx}_{oN}.so not built"}
            continue
        So, Po, go = synth.lqr_schur(onx, onu, oN, 0, "stair", odt)
        rk = ref_cuda.get(odt, onx, oN)
        dS, dP, dg = (torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (So, Po, go))
        dl = torch.zeros(oN * onx, dtype=dS.dtype, device="cuda")
        lat = time_single_solves(torch, lambda: rk.launch(dS, dP, dg, dl, omax, otol), dl, n_warm=10, n_timed=60)
        its = int(rk.d_iters.item())
        other[name] = {"nx": onx, "knot_points": oN, "dtype": np.dtype(odt).name, "iters": its, "kernel_p50_us": pct(lat, 0.5),
                       "kernel_p99_us": pct(lat, 0.99), "us_per_iteration": pct(lat, 0.5) / max(its, 1), **rk.attrs()}
        del dS, dP, dg, dl

    cores = os.cpu_count() or 1
    ns = min(sample, 2048)
    lam_c = np.zeros((ns, N_KNOT * NX), dtype=DTYPE)
    t0 = time.perf_counter()
    oracle.pcg_batched(S[:ns], P[:ns], gam[:ns], lam_c, NX, N_KNOT, TOL, MAX_ITER, order=1, threads=cores)
    t_cpu = time.perf_counter() - t0
    return {
        "impl": "reference",
        "metric": "batched PCG solves/s (8192 x N=32,nx=14 fp32, partitioned by problem index over the GPUs)", "value": value, "unit": "solves/s",
        "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"C5 sample: {sample} of the 8192 iiwa MPC systems (knot points 32, state size 14, fp32, "
                               "symmetric-stair preconditioner, tol 1e-6), one cooperative launch of the reference "
                               "pcg<float,14,32> per system, back to back on one stream (it has no batch API)",
                   "batch_per_gpu": sample, "nx": NX, "knot_points": N_KNOT, "precond": PRECOND, "exit_tol": TOL},
        "e2e": {"value": value, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": sample * args.steps,
        "clocks": clocks,
        "concurrent_streams": concurrent,
        "other_configs": other,
        "single_solve": {"config": "C2 N=32 nx=14 fp32 stair, seed 0", "iters": it1,
                         "kernel_p50_us": pct(lat_k, 0.5), "kernel_p99_us": pct(lat_k, 0.99),
                         "iterations_per_s": it1 / (pct(lat_k, 0.5) * 1e-6),
                         "dev_api_p50_us": pct(lat_d, 0.5), "dev_api_p99_us": pct(lat_d, 0.99),
                         "host_api_p50_us_time_only_25_iters": pct(lat_h, 0.5), **ref.attrs()},
        "cpu_baseline": {"value": ns / t_cpu, "unit": "solves/s", "cores": cores, "kind": "reference",
                         "sample": f"reference CUDA kernel on the B200 is the line's value; this entry is the host oracle in "
                                   f"reference reduction order on {ns} systems, {cores} threads"},
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH, help="systems per GPU")
    ap.add_argument("--ref-sample", type=int, default=512, help="systems per step of the reference arm")
    args = ap.parse_args()
    # stdout carries exactly one JSON line: everything libraries print meanwhile (NCCL's version banner, ...) goes to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    out = run_reference(args) if args.impl == "reference" else run_ours(args)
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    os.close(real_stdout)
    if out is not None:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
