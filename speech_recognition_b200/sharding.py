"""Utterance sharding for multi-GPU runs.

The front-end has no cross-utterance state (the reference maps one utterance per call,
run/train.py:88), so N GPUs = N independent processes over contiguous ranges of the global
utterance index - no collective on the data path.  SpecAugment draws are keyed by the GLOBAL
utterance id, so the union of the shards is bit-identical to a single-GPU run.
"""
from typing import List, Sequence, Tuple


def shard_ranges(lengths: Sequence[int], world_size: int) -> List[Tuple[int, int]]:
    """Split utterances [0, len(lengths)) into `world_size` contiguous ranges balanced by total
    samples (greedy prefix split), not by count, so ragged batches load the GPUs evenly.
    Every rank gets a (possibly empty) [begin, end) range; ranges tile the index space."""
    n = len(lengths)
    if world_size <= 0:
        raise ValueError("world_size must be positive")
    total = float(sum(int(x) for x in lengths))
    if total <= 0.0:
        # nothing to balance by: split by count
        return [(n * r // world_size, n * (r + 1) // world_size) for r in range(world_size)]
    bounds = [0]
    acc = 0.0
    i = 0
    for r in range(1, world_size):
        target = total * r / world_size
        while i < n and acc + lengths[i] / 2.0 <= target:
            acc += lengths[i]
            i += 1
        bounds.append(i)
    bounds.append(n)
    return [(bounds[r], bounds[r + 1]) for r in range(world_size)]


def rank_range(lengths: Sequence[int], rank: int, world_size: int) -> Tuple[int, int]:
    return shard_ranges(lengths, world_size)[rank]
