"""Partitioning of a batch of objects over GPUs.

Objects are independent SHA-256 chains: shards exchange nothing, so there is no
collective on the data path.  Uniform batches are cut into contiguous equal
index ranges; mixed-size batches are balanced by *compressions* (the unit of
work), longest object first, and one object is never split across devices.
"""
from __future__ import annotations

import heapq
from typing import Sequence

import numpy as np


def nblocks(length):
    """Compressions a message of ``length`` bytes costs: ceil((L + 9) / 64)."""
    return (np.asarray(length, dtype=np.uint64) + np.uint64(72)) // np.uint64(64)


def contiguous_ranges(n: int, world: int) -> list[tuple[int, int]]:
    """Rank r gets objects [n*r//world, n*(r+1)//world)."""
    if world < 1:
        raise ValueError("world must be >= 1")
    return [(n * r // world, n * (r + 1) // world) for r in range(world)]


def rank_range(n: int, rank: int, world: int) -> tuple[int, int]:
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    return contiguous_ranges(n, world)[rank]


def lpt_assign(lengths: Sequence[int], world: int) -> list[np.ndarray]:
    """Longest-processing-time-first assignment balanced by compressions.

    Returns, per rank, the (ascending) original indices it hashes.
    """
    if world < 1:
        raise ValueError("world must be >= 1")
    lengths = np.asarray(lengths, dtype=np.uint64)
    cost = nblocks(lengths)
    order = np.argsort(-cost.astype(np.int64), kind="stable")
    heap = [(0, r) for r in range(world)]
    buckets: list[list[int]] = [[] for _ in range(world)]
    for i in order:
        load, r = heapq.heappop(heap)
        buckets[r].append(int(i))
        heapq.heappush(heap, (load + int(cost[i]), r))
    return [np.array(sorted(b), dtype=np.int64) for b in buckets]


def assign(lengths: Sequence[int], world: int) -> list[np.ndarray]:
    """Contiguous ranges when all lengths are equal, LPT otherwise."""
    lengths = np.asarray(lengths, dtype=np.uint64)
    n = int(lengths.shape[0])
    if n == 0 or bool((lengths == lengths[0]).all()):
        return [np.arange(lo, hi, dtype=np.int64) for lo, hi in contiguous_ranges(n, world)]
    return lpt_assign(lengths, world)


def scatter_back(parts: Sequence[np.ndarray], digests: Sequence[np.ndarray], n: int) -> np.ndarray:
    """Inverse of ``assign``: per-rank digests (k_r, 32) -> (n, 32) in input order."""
    out = np.zeros((n, 32), dtype=np.uint8)
    seen = np.zeros(n, dtype=bool)
    for idx, dig in zip(parts, digests):
        dig = np.asarray(dig, dtype=np.uint8).reshape(-1, 32)
        if dig.shape[0] != idx.shape[0]:
            raise ValueError("digest count does not match shard size")
        out[idx] = dig
        seen[idx] = True
    if not bool(seen.all()):
        raise ValueError("some objects were not assigned to any shard")
    return out
