"""Partition of a batch of independent systems over GPUs: by problem index, contiguous slabs, no collective on
the data path.  The reference has no multi-GPU path; a single system always stays on one GPU.

One process per GPU (torchrun); rank r owns problems [lo, hi) = slab(r, world, batch) and solves them with the
batched entry points on its own device.  The only inter-rank traffic is control plane: a barrier, and an optional
gather of the per-problem iteration counts / exit flags to rank 0 (a few bytes per problem) -- S, Phi^-1, gamma
and lambda never leave the GPU that owns them.
"""
from __future__ import annotations

import numpy as np


def slab(rank: int, world: int, batch: int) -> tuple[int, int]:
    """[lo, hi) of the problems owned by `rank`: problem i -> GPU floor(i*world/batch)."""
    if not (0 <= rank < world) or batch < 0:
        raise ValueError((rank, world, batch))
    lo = (rank * batch + world - 1) // world
    hi = ((rank + 1) * batch + world - 1) // world
    return lo, hi


def owner(i: int, world: int, batch: int) -> int:
    return i * world // batch


def solve_partitioned(solve_slab, batch: int, rank: int, world: int):
    """Run ``solve_slab(lo, hi) -> (iters[hi-lo], exits[hi-lo])`` on this rank's slab.  Returns (lo, hi, iters, exits)."""
    lo, hi = slab(rank, world, batch)
    if hi > lo:
        iters, exits = solve_slab(lo, hi)
    else:
        iters, exits = np.zeros(0, np.uint32), np.zeros(0, bool)
    iters = np.asarray(iters, dtype=np.uint32)
    exits = np.asarray(exits, dtype=bool)
    if iters.shape != (hi - lo,) or exits.shape != (hi - lo,):
        raise ValueError("solve_slab must return one iteration count and one flag per owned problem")
    return lo, hi, iters, exits


def gather_stats(iters: np.ndarray, exits: np.ndarray, batch: int, rank: int, world: int, device=None):
    """Control-plane gather of per-problem iteration counts and max-iter flags to every rank (torch.distributed,
    whatever backend the process group uses: nccl on GPUs, gloo in the CPU tests).  Returns (iters[batch], exits[batch])."""
    if world == 1:
        return np.asarray(iters, np.uint32), np.asarray(exits, bool)
    import torch
    import torch.distributed as dist
    per = (batch + world - 1) // world                       # slabs differ by at most one problem: pad to `per`
    mine = torch.full((2, per), -1, dtype=torch.int32, device=device)
    n = len(iters)
    mine[0, :n] = torch.from_numpy(np.asarray(iters, np.int32))
    mine[1, :n] = torch.from_numpy(np.asarray(exits, np.int32))
    allb = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allb, mine)
    out_i = np.zeros(batch, np.uint32)
    out_e = np.zeros(batch, bool)
    for r in range(world):
        lo, hi = slab(r, world, batch)
        blk = allb[r].cpu().numpy()
        out_i[lo:hi] = blk[0, :hi - lo].astype(np.uint32)
        out_e[lo:hi] = blk[1, :hi - lo] != 0
    return out_i, out_e
