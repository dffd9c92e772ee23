"""fp64 batches: the tensor-memory form of the cluster kernel (pcg_fused.cuh, P_TMEM: two rows per thread, half the SMs per system;
the planner's choice for batches that cannot be resident at once) against the register-row clusters (GBDPCG_FLAG_NO_TMEM), and the
A/B alternative of the S columns kept in tensor memory.  Launch times, iteration counts, agreement of lambda.
usage: python scripts/time_tmem64.py [batch]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gbd_pcg_b200 as g  # noqa: E402

FLAG_TMEM, FLAG_NO_TMEM = 0x400, 0x800
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
nseed = 32
SHAPES = [(14, 7, 64), (14, 7, 32), (12, 4, 100), (8, 4, 200), (10, 5, 60), (16, 8, 32),
          (14, 7, 16), (8, 4, 32), (12, 4, 20), (16, 8, 16)]      # the last four: CTAs of <= 128 threads, two per SM
for (nx, nu, N) in SHAPES:
    S, P, gam = g.synth.lqr_schur_batch(nx, nu, N, range(nseed), "stair", np.float64)
    d_S, d_P, d_g = (torch.from_numpy(a).cuda().repeat(batch // nseed, 1) for a in (S, P, gam))
    out = {}
    for name, flags, env in (("regs", FLAG_NO_TMEM, None), ("tmem_a", FLAG_TMEM, None), ("tmem_b", FLAG_TMEM, "1"), ("tmem_1sm", FLAG_TMEM, "one"), ("planner", 0, None)):
        os.environ.pop("GBDPCG_TMEM64_VARIANT", None)
        os.environ.pop("GBDPCG_TMEM64_SMALL", None)
        if env == "one":
            os.environ["GBDPCG_TMEM64_SMALL"] = "0"          # one CTA per SM even where the CTAs have <= 128 threads
        elif env is not None:
            os.environ["GBDPCG_TMEM64_VARIANT"] = env
        d_l = torch.zeros(batch, N * nx, dtype=torch.float64, device="cuda")
        d_it = torch.zeros(batch, dtype=torch.int32, device="cuda")
        d_ex = torch.zeros(batch, dtype=torch.uint8, device="cuda")
        cfg = g.pcg_config(pcg_exit_tol=1e-10, pcg_max_iter=2000, empty_pinv=0, flags=flags)
        try:
            info = g.plan_query(nx, N, np.float64, True, cfg, batch=batch)
        except g.GbdpcgError as e:
            print(f"nx={nx} N={N} {name}: {e}")
            continue
        ts = []
        for i in range(8):
            d_l.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            g.solve_batched_dev(nx, N, batch, d_S, d_P, d_g, d_l, cfg, d_it, d_ex)
            b.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(a.elapsed_time(b))
        out[name] = (float(np.median(ts)), d_l.cpu().numpy(), d_it.cpu().numpy(), info)
    os.environ.pop("GBDPCG_TMEM64_VARIANT", None)
    os.environ.pop("GBDPCG_TMEM64_SMALL", None)
    ref = out.get("regs")
    for name, (t, l, it, info) in out.items():
        agree = "" if ref is None else f"  iters equal to regs {bool(np.array_equal(it, ref[2]))}  rel diff {np.linalg.norm(l - ref[1]) / np.linalg.norm(ref[1]):.2e}"
        bytes_per = 2 * 3 * N * nx * nx * 8 + 3 * N * nx * 8
        print(f"nx={nx} N={N} batch={batch} {name:8s}: {t:8.4f} ms ({batch / t / 1e3:6.3f} M/s, {batch * bytes_per / t / 1e6:7.1f} GB/s)  iters mean {it.mean():.2f}"
              f"  {info['kernel']} cluster {info['cluster']} threads {info['threads']} regs {info['regs']} resident {info['max_resident']}{agree}", flush=True)
