"""Pair schedules and their static partition over ranks.

The reference decides which pairs to match with a BoW view graph + BFS (matching.py:24-99); the
BASELINE.json configs use exhaustive and sliding-window schedules instead. Pairs are (query_img,
train_img) rows of an int32 [P,2] array, the format consumed by match_pairs and written to
output/img_pairs.npy (matching.py:188).
"""
from __future__ import annotations

import numpy as np


def exhaustive_pairs(n_img: int) -> np.ndarray:
    """All i<j pairs in row-major order: n(n-1)/2 rows."""
    i, j = np.triu_indices(n_img, k=1)
    return np.stack([i, j], 1).astype(np.int32)


def window_pairs(n_img: int, window: int) -> np.ndarray:
    """Each image against the next `window` images, clipped at the end."""
    out = [(i, j) for i in range(n_img) for j in range(i + 1, min(n_img, i + 1 + window))]
    return np.asarray(out, dtype=np.int32).reshape(-1, 2)


def blocked_order(pairs: np.ndarray, block: int = 16) -> np.ndarray:
    """Reorder pairs so that consecutive pairs share images inside block x block patches of the (i,j)
    grid (a 16 x 16 patch of 16k-keypoint images is 2 x 32 MiB of descriptors: L2-resident)."""
    pairs = np.asarray(pairs, dtype=np.int32).reshape(-1, 2)
    if len(pairs) == 0:
        return pairs
    key = np.lexsort((pairs[:, 1], pairs[:, 0], pairs[:, 1] // block, pairs[:, 0] // block))
    return pairs[key]


def pair_cost(pairs: np.ndarray, k_per_image: np.ndarray) -> np.ndarray:
    k = np.asarray(k_per_image, dtype=np.int64)
    return k[pairs[:, 0]] * k[pairs[:, 1]]


def partition(pairs: np.ndarray, world_size: int, rank: int, k_per_image=None):
    """Static, deterministic partition: contiguous chunks of the (blocked) pair list with near-equal
    total cost K_q*K_t. Returns (pairs_of_rank, indices_into_pairs). Every rank computes the same
    split, so no communication is needed and the gathered result order is well defined."""
    pairs = np.asarray(pairs, dtype=np.int32).reshape(-1, 2)
    n = len(pairs)
    if world_size <= 1:
        return pairs, np.arange(n)
    if k_per_image is None:
        cost = np.ones(n, dtype=np.float64)
    else:
        cost = np.maximum(pair_cost(pairs, k_per_image).astype(np.float64), 1.0)
    csum = np.concatenate([[0.0], np.cumsum(cost)])
    total = csum[-1]
    bounds = np.searchsorted(csum, total * np.arange(world_size + 1) / world_size, side="left")
    bounds[0], bounds[-1] = 0, n
    bounds = np.maximum.accumulate(bounds)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    idx = np.arange(lo, hi)
    return pairs[lo:hi], idx


def window_shard(n_img: int, window: int, world_size: int, rank: int):
    """Sharding of the sequential-window schedule (BASELINE.json config 4) WITHOUT a full bank broadcast:
    rank r owns a contiguous range of query images and needs only those plus a halo of `window`
    following images. Ranges are balanced on pair counts (the last `window` images have fewer pairs).

    -> (img_lo, img_hi, halo_hi, local_pairs int32 [P_r,2] indexed relative to img_lo, global_pairs int32 [P_r,2])
    The concatenation of global_pairs over the ranks is window_pairs(n_img, window) in order."""
    per_img = np.minimum(window, np.maximum(n_img - 1 - np.arange(n_img), 0)).astype(np.int64)
    csum = np.concatenate([[0], np.cumsum(per_img)])
    total = csum[-1]
    bounds = np.searchsorted(csum, total * np.arange(world_size + 1) / max(world_size, 1), side="left")
    bounds[0], bounds[-1] = 0, n_img
    bounds = np.maximum.accumulate(bounds)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    halo_hi = min(n_img, hi + window) if hi > lo else hi
    rows = [(i, j) for i in range(lo, hi) for j in range(i + 1, min(n_img, i + 1 + window))]
    gp = np.asarray(rows, dtype=np.int32).reshape(-1, 2)
    return lo, hi, halo_hi, gp - lo, gp


def partition_by_owner(pairs: np.ndarray, world_size: int, rank: int, n_loc: int, weights=None):
    """Partition for the sharded host-buffer path: rank r OWNS images [r * n_loc, (r + 1) * n_loc) (it uploads and
    prepares them), and every pair goes to the owner of one of its two images -- of the query image when
    (query + train) is even, of the train image otherwise -- so each image hands about half of its pairs to
    either side. Ranks left above the mean by that rule (the last rank usually owns fewer images) pass the tail
    of their non-local pairs on to the ranks below it. A rank's pairs come back with the pairs BETWEEN ITS OWN
    images first (in list order): they can be matched while the rest of the bank is still travelling between
    the GPUs. Deterministic, no communication. weights: optional relative throughput of every rank (identical on
    all ranks, e.g. all-gathered from a measured warm-up): rank r then gets about weights[r] / sum(weights) of the
    pairs instead of 1 / world_size -- GPUs under a power cap and pair sets of different match density do not all
    run at the same pairs/s.
    -> (pairs_of_rank, indices into `pairs`, number of leading local-only pairs)"""
    pairs = np.asarray(pairs, dtype=np.int32).reshape(-1, 2)
    if world_size <= 1:
        return pairs, np.arange(len(pairs)), len(pairs)
    idx, n_local = partition_all_by_owner(pairs, world_size, n_loc, weights)[rank]
    return pairs[idx], idx, n_local


def rank_targets(n_pairs: int, world_size: int, weights=None) -> np.ndarray:
    """Pairs per rank: ceil(n / world) each, or proportional to `weights` (largest remainders first, ties by rank)."""
    if weights is None:
        return np.full(world_size, -(-n_pairs // world_size), np.int64)
    w = np.asarray(weights, dtype=np.float64)
    if w.shape != (world_size,) or not np.all(np.isfinite(w)) or np.any(w <= 0):
        raise ValueError(f"weights must be {world_size} positive numbers")
    share = n_pairs * w / w.sum()
    t = np.floor(share).astype(np.int64)
    order = np.argsort(-(share - t), kind="stable")
    t[order[: n_pairs - int(t.sum())]] += 1
    return t


def partition_all_by_owner(pairs: np.ndarray, world_size: int, n_loc: int, weights=None):
    """partition_by_owner for every rank in one pass over the list (the sharded host path needs all of them: its
    own pairs to match, everybody's to put the gathered results back into list order).
    -> [(indices into `pairs`, number of leading local-only pairs)] per rank"""
    pairs = np.asarray(pairs, dtype=np.int32).reshape(-1, 2)
    q, t = pairs[:, 0].astype(np.int64), pairs[:, 1].astype(np.int64)
    oq, ot = q // n_loc, t // n_loc
    owner = np.where(((q + t) & 1) == 0, oq, ot)
    local = oq == ot
    target = rank_targets(len(pairs), world_size, weights)
    counts = np.bincount(owner, minlength=world_size)
    small = np.int16 if world_size < 32000 else np.int64                       # int16: numpy's stable sort is a radix sort
    by_owner = np.argsort(owner.astype(small), kind="stable")                  # grouped by owner, list order inside
    starts = np.concatenate([[0], np.cumsum(counts)])
    pool = []
    for r in np.nonzero(counts > target)[0]:         # surplus: the tail of a rank's non-local pairs
        grp = by_owner[starts[r]: starts[r + 1]]
        mine = grp[~local[grp]]
        extra = int(counts[r]) - int(target[r])
        pool.append(mine[len(mine) - min(extra, len(mine)):])
    if pool:
        pool = np.concatenate(pool)
        owner = owner.copy()
        owner[pool] = -1
        counts = np.bincount(owner[owner >= 0], minlength=world_size)
        at = 0
        for r in range(world_size):                  # deficit: filled from the pool in rank order
            room = int(target[r]) - int(counts[r])
            if room > 0 and at < len(pool):
                take = pool[at: at + room]
                owner[take] = r
                at += len(take)
        if at < len(pool):                           # rounding leftovers
            owner[pool[at:]] = world_size - 1
        counts = np.bincount(owner, minlength=world_size)
        by_owner = np.argsort(owner.astype(small), kind="stable")
        starts = np.concatenate([[0], np.cumsum(counts)])
    parts = []
    for r in range(world_size):
        mine = by_owner[starts[r]: starts[r + 1]]
        loc = local[mine] & (oq[mine] == r)
        parts.append((np.concatenate([mine[loc], mine[~loc]]), int(loc.sum())))
    return parts
