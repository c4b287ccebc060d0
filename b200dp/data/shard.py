"""Deterministic contiguous per-rank dataset sharding.

Behavioural contract taken from the reference's ``SplitDataset``
(reference splitdataset.py:30-87) and the way ``run`` drives it
(reference train.py:86-93): partitions are named, sorted lexicographically,
fractional weights are converted with ``round(w * len)`` when they sum to
<= 1, otherwise they must be integer sizes; a shard is the contiguous block
``[cum[i-1], cum[i])`` of the base dataset and nothing is shuffled.

Two modes:

``compat=True``
    Reproduces the reference exactly, including its quirks:
    no upper-bound check in ``__getitem__``, ``round`` may drop or overshoot
    a few samples, at least two partitions are required.

``compat=False`` (default)
    Same block layout wherever the reference is well defined, with the
    problems fixed: a single partition is allowed (world size 1), sizes are
    computed by an exact largest-remainder split so every sample is covered
    exactly once and no shard reaches past the end, indices are bounds
    checked, and :func:`make_rank_partition` uses zero-padded names so the
    lexicographic order equals the rank order for any world size.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np
from torch.utils.data import Dataset

_UNSELECTED_MSG = "Select a partition before accessing data."


def _exact_split(weights: Sequence[float], n: int) -> List[int]:
    """Largest-remainder apportionment of ``n`` items by ``weights``.

    Weights summing to 1 cover all ``n`` items; weights summing to s < 1
    cover ``round(s * n)`` items.  Ties go to the lower index so the result
    is deterministic.
    """
    total_w = float(sum(weights))
    target = int(round(total_w * n)) if total_w < 1.0 - 1e-12 else n
    raw = [w * n for w in weights]
    base = [int(np.floor(r + 1e-9)) for r in raw]
    missing = target - sum(base)
    order = sorted(range(len(raw)), key=lambda i: (-(raw[i] - base[i]), i))
    for i in order[: max(missing, 0)]:
        base[i] += 1
    return base


class ShardedDataset(Dataset):
    """Linear (unshuffled) partition of a map-style dataset.

    Drop-in for the reference ``SplitDataset(dataset, partitions,
    initial_partition)`` with ``select(name)``, ``__len__`` and
    ``__getitem__`` (reference splitdataset.py:30-87).
    """

    def __init__(self, dataset, partitions: Dict[str, float],
                 initial_partition: Optional[str] = None, *, compat: bool = False):
        super().__init__()
        if not isinstance(partitions, dict):
            raise TypeError("partitions must be a dict")
        min_parts = 2 if compat else 1
        if len(partitions) < min_parts:
            raise ValueError(
                f"ShardedDataset needs at least {min_parts} partition(s), got {len(partitions)}")
        if min(partitions.values()) < 0:
            raise ValueError("partition sizes cannot be negative")
        if max(partitions.values()) <= 0:
            raise ValueError("all partitions cannot be empty")

        self.dataset = dataset
        self.partitions = dict(partitions)
        self.compat = compat
        self.partition_names = sorted(self.partitions.keys())
        self.partition_index = {name: i for i, name in enumerate(self.partition_names)}

        weights = [self.partitions[name] for name in self.partition_names]
        n = len(dataset)
        if sum(weights) <= 1:
            if compat:
                sizes = [round(w * n) for w in weights]
            else:
                sizes = _exact_split(weights, n)
        else:
            for w in weights:
                if w != int(w):
                    raise ValueError("partition sizes should be integers, or sum up to <= 1")
            sizes = [int(w) for w in weights]
            if not compat and sum(sizes) > n:
                raise ValueError(f"partition sizes sum to {sum(sizes)} > dataset length {n}")
        self.partition_sizes = sizes
        self.partition_cum_sizes = np.cumsum(sizes)
        self.current_partition_idx: Optional[int] = None
        if initial_partition is not None:
            self.select(initial_partition)

    # -- reference API -----------------------------------------------------
    def select(self, partition: str) -> "ShardedDataset":
        self.current_partition_idx = self.partition_index[partition]
        return self

    def __len__(self) -> int:
        if self.current_partition_idx is None:
            raise ValueError(_UNSELECTED_MSG)
        return self.partition_sizes[self.current_partition_idx]

    def __getitem__(self, idx: int):
        if self.current_partition_idx is None:
            raise ValueError(_UNSELECTED_MSG)
        if not self.compat:
            n = self.partition_sizes[self.current_partition_idx]
            if idx < 0:
                idx += n
            if not 0 <= idx < n:
                raise IndexError(f"index {idx} out of range for shard of size {n}")
        return self.dataset[self.offset + idx]

    # -- extensions --------------------------------------------------------
    @property
    def offset(self) -> int:
        """Index in the base dataset of this shard's first sample."""
        if self.current_partition_idx is None:
            raise ValueError(_UNSELECTED_MSG)
        i = self.current_partition_idx
        return 0 if i == 0 else int(self.partition_cum_sizes[i - 1])

    def bounds(self, partition: Optional[str] = None):
        """``(start, stop)`` of a shard in base-dataset indices."""
        i = self.current_partition_idx if partition is None else self.partition_index[partition]
        if i is None:
            raise ValueError(_UNSELECTED_MSG)
        stop = int(self.partition_cum_sizes[i])
        return stop - self.partition_sizes[i], stop


def make_rank_partition(world_size: int, *, compat: bool = False) -> Dict[str, float]:
    """Equal-weight partition spec keyed by rank (reference train.py:86).

    ``compat=True`` returns the reference's ``{str(i): 1/size}`` (rank 2 of 12
    then owns block 4 because '10' < '2').  The default zero-pads names so
    block index == rank.
    """
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    if compat:
        return {str(i): 1 / world_size for i in range(world_size)}
    width = len(str(world_size - 1))
    return {str(i).zfill(width): 1 / world_size for i in range(world_size)}


def rank_key(rank: int, world_size: int, *, compat: bool = False) -> str:
    """Partition name for ``rank`` under :func:`make_rank_partition`."""
    if compat:
        return str(rank)
    return str(rank).zfill(len(str(world_size - 1)))


def shard_for_rank(dataset, rank: int, world_size: int, *, compat: bool = False) -> ShardedDataset:
    """The train.py:86-93 idiom in one call: partition equally, select ``rank``."""
    ds = ShardedDataset(dataset, make_rank_partition(world_size, compat=compat), compat=compat)
    return ds.select(rank_key(rank, world_size, compat=compat))


def common_batches(base, world_size: int, batch_size: int, *, compat: bool = False, drop_last: bool = False) -> int:
    """Batches per epoch that EVERY rank can run: the minimum over ranks of its shard's batch count.  Shards may differ
    by a sample (60001 samples over 2 ranks), and then ``ceil(len/batch)`` can differ by one between ranks; each
    training step is a collective, so the rank with the extra batch would wait for gradients nobody sends (the
    reference hangs in that case, train.py:38/44).  Every rank can compute this from the shard table alone."""
    counts = []
    for r in range(world_size):
        n = len(shard_for_rank(base, r, world_size, compat=compat and world_size > 1))
        counts.append(n // batch_size if drop_last else -(-n // batch_size))
    return min(counts)
