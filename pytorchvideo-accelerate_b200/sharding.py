"""Per-rank sharding of the preprocessing work: one process per GPU, no data-path collective.

The reference gets its multi-GPU behaviour from ``accelerator.prepare(...)`` (``run.py:196-198``):
every process iterates its own share of the batches and moves them to its own device.  Clips
are independent, so the preprocessing path inherits that sharding unchanged -- rank ``r`` of
``world`` preprocesses exactly the batches it would have received anyway.  The helpers below
state that assignment explicitly for launchers without accelerate (``torchrun``).
"""
from __future__ import annotations

import os
from typing import List, Tuple


def rank_and_world() -> Tuple[int, int, int]:
    """(rank, world_size, local_rank) from the torchrun / accelerate environment (1 process if unset)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def batches_for_rank(num_batches: int, rank: int, world: int) -> List[int]:
    """Global batch indices handled by ``rank``: round-robin, as accelerate's per-process batch
    sharding hands them out (process ``r`` takes batches ``r, r + world, ...``).  Every rank gets
    the same count; a trailing partial round is dropped, as an even split requires."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside [0, {world})")
    full_rounds = num_batches // world
    return [r * world + rank for r in range(full_rounds)]


def rank_seed(base_seed: int, rank: int) -> int:
    """Distinct, reproducible RNG stream per rank for the per-clip draws."""
    return base_seed + rank
