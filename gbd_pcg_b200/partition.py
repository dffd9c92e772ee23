"""Partition of a batch of independent systems over GPUs: by problem index, contiguous slabs, no collective
(SURVEY.md s8(e)).  The reference has no multi-GPU path; a single system always stays on one GPU."""
from __future__ import annotations


def slab(rank: int, world: int, batch: int) -> tuple[int, int]:
    """[lo, hi) of the problems owned by `rank`: problem i -> GPU floor(i*world/batch)."""
    if not (0 <= rank < world) or batch < 0:
        raise ValueError((rank, world, batch))
    lo = (rank * batch + world - 1) // world
    hi = ((rank + 1) * batch + world - 1) // world
    return lo, hi


def owner(i: int, world: int, batch: int) -> int:
    return i * world // batch
