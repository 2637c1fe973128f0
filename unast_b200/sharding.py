"""Length-bucketed sharding of utterances over ranks.

The spectral path is embarrassingly data-parallel: utterances are independent, so ranks never
exchange data on the hot path.  The only multi-GPU logic is deciding who owns what:
greedy longest-first dealing (LPT) on the frame count, then sorting each rank's share by length so
that batches handed to the GPU are length-homogeneous ("bucketed").
"""
from __future__ import annotations

import heapq

import numpy as np


def deal_longest_first(sizes, world_size):
    """Returns ``world_size`` lists of indices with near-equal total size (LPT greedy).

    Deterministic: ties break on index, so every rank computes the same partition without talking."""
    sizes = np.asarray(sizes, dtype=np.int64)
    order = np.lexsort((np.arange(sizes.size), -sizes))
    heap = [(0, r) for r in range(world_size)]
    heapq.heapify(heap)
    shards = [[] for _ in range(world_size)]
    for i in order:
        load, r = heapq.heappop(heap)
        shards[r].append(int(i))
        heapq.heappush(heap, (load + int(sizes[i]), r))
    return shards


def length_buckets(indices, sizes, batch_size):
    """Sort a rank's utterances by size (descending) and cut into batches of ``batch_size``."""
    idx = sorted(indices, key=lambda i: (-int(sizes[i]), i))
    return [idx[s:s + batch_size] for s in range(0, len(idx), batch_size)]


def shard_loads(sizes, shards):
    sizes = np.asarray(sizes, dtype=np.int64)
    return [int(sizes[s].sum()) if len(s) else 0 for s in shards]
