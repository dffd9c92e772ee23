"""world_size-2 (and 3) test of the multi-GPU host logic on CPU with the gloo backend: slabs by problem index, no
data-path collective, control-plane gather of the iteration counts.  There is no GPU here, so the per-slab solver
injected into the partition driver is the host oracle (test infrastructure, acting as the stand-in device)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

from conftest import ROOT

NX, NU, N, BATCH = 4, 2, 6, 11


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, outdir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from gbd_pcg_b200 import partition, synth
    from oracle import oracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lam_out = {}

    def solve_slab(lo, hi):
        S, P, g = synth.lqr_schur_batch(NX, NU, N, range(lo, hi), "stair", np.float64)   # seed = global problem index
        lam, its, exs = oracle.pcg_batched(S, P, g, np.zeros_like(g), NX, N, 1e-10, 200)
        lam_out["lam"] = lam
        return its, exs

    lo, hi, its, exs = partition.solve_partitioned(solve_slab, BATCH, rank, world)
    all_i, all_e = partition.gather_stats(its, exs, BATCH, rank, world)
    dist.barrier()
    np.savez(os.path.join(outdir, f"rank{rank}.npz"), lo=lo, hi=hi, lam=lam_out.get("lam", np.zeros((0, NX * N))), all_i=all_i, all_e=all_e)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_partitioned_batch_equals_single_process(world, tmp_path):
    from gbd_pcg_b200 import synth
    from oracle import oracle
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    S, P, g = synth.lqr_schur_batch(NX, NU, N, range(BATCH), "stair", np.float64)
    lam, its, exs = oracle.pcg_batched(S, P, g, np.zeros_like(g), NX, N, 1e-10, 200)
    covered = np.zeros(BATCH, bool)
    for r in range(world):
        z = np.load(tmp_path / f"rank{r}.npz")
        lo, hi = int(z["lo"]), int(z["hi"])
        assert not covered[lo:hi].any()
        covered[lo:hi] = True
        assert np.array_equal(z["lam"], lam[lo:hi])                 # each rank solved exactly its own problems
        assert np.array_equal(z["all_i"], its) and np.array_equal(z["all_e"], exs)   # every rank sees the gathered stats
    assert covered.all()
