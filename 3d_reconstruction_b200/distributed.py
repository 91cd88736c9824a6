"""Multi-GPU plumbing: one process per GPU (torchrun), NCCL over NVLink for the two exchange steps.

The path shards by image pair with no collective during compute:
  1. the packed bank is broadcast once from rank 0,
  2. every rank enumerates the same pair list and takes its static, cost-balanced contiguous slice,
  3. match lists are gathered to rank 0 at the end (variable length: counts first, then payload).
Works with the gloo backend on CPU tensors too (used by the world_size-2 CPU tests).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from .scheduler import partition


def is_dist() -> bool:
    return dist.is_available() and dist.is_initialized()


def world() -> tuple[int, int]:
    return (dist.get_rank(), dist.get_world_size()) if is_dist() else (0, 1)


def broadcast_bank_tensors(desc: torch.Tensor | None, k_per_image: torch.Tensor | None, shape, device, src: int = 0):
    """Rank `src` passes (desc, k_per_image); the others pass None and the [n_img, k_stride, 128] shape.
    Returns the tensors on every rank (uint8 / int32 on `device`)."""
    rank, ws = world()
    if rank != src or desc is None:
        desc = torch.empty(tuple(shape), dtype=torch.uint8, device=device)
        k_per_image = torch.empty((shape[0],), dtype=torch.int32, device=device)
    else:
        desc = desc.to(device)
        k_per_image = k_per_image.to(device=device, dtype=torch.int32)
    if ws > 1:
        dist.broadcast(desc, src=src)
        dist.broadcast(k_per_image, src=src)
    return desc, k_per_image


def my_pairs(pairs: np.ndarray, k_per_image=None):
    """-> (pairs of this rank, their indices in the global list). Same split on every rank."""
    rank, ws = world()
    return partition(pairs, ws, rank, k_per_image)


def gather_flat(offsets: torch.Tensor, payload: list[torch.Tensor], dst: int = 0):
    """Gather per-rank flat match lists (ranks own contiguous, rank-ordered slices of the pair list).

    offsets: int64 [P_r + 1] of this rank; payload: flat tensors of length offsets[-1] (same dtypes on all ranks).
    Returns on rank `dst` (global_offsets int64 [P + 1], [concatenated payload tensors]); None elsewhere.
    The result is identical to what a single rank produces for the whole list.
    """
    rank, ws = world()
    if ws == 1:
        return offsets, payload
    dev = offsets.device
    counts = (offsets[1:] - offsets[:-1]).contiguous()
    meta = torch.tensor([counts.numel(), int(offsets[-1])], dtype=torch.int64, device=dev)
    metas = [torch.zeros_like(meta) for _ in range(ws)]
    dist.all_gather(metas, meta)
    n_pairs = [int(m[0]) for m in metas]
    n_match = [int(m[1]) for m in metas]
    pmax, mmax = max(n_pairs), max(max(n_match), 1)

    def pad(t, n):
        out = torch.zeros((n,), dtype=t.dtype, device=dev)
        out[: t.numel()] = t
        return out

    def gather(t, n):
        buf = [torch.zeros((n,), dtype=t.dtype, device=dev) for _ in range(ws)] if rank == dst else None
        dist.gather(pad(t, n), buf, dst=dst)
        return buf

    cbuf = gather(counts, max(pmax, 1))
    pbufs = [gather(t, mmax) for t in payload]
    if rank != dst:
        return None
    all_counts = torch.cat([cbuf[r][: n_pairs[r]] for r in range(ws)])
    goff = torch.zeros((all_counts.numel() + 1,), dtype=torch.int64, device=dev)
    torch.cumsum(all_counts, 0, out=goff[1:])
    out = [torch.cat([pb[r][: n_match[r]] for r in range(ws)]) for pb in pbufs]
    return goff, out


def scatter_bank_window(desc: torch.Tensor | None, k_per_image: torch.Tensor | None, n_img: int, k_stride: int, dim: int,
                        window: int, device, dtype=torch.uint8, src: int = 0):
    """Sharded bank + halo for the sequential-window schedule (config 4: 2,000 images, each against the next
    32): instead of broadcasting the whole bank, rank `src` sends every rank only the images it needs
    (its query range + `window` halo images, scheduler.window_shard). Point-to-point over NCCL / gloo.

    -> (local desc [halo_hi - lo, k_stride, dim], local k_per_image, local pairs int32 [P_r,2], global pairs)"""
    from .scheduler import window_shard
    rank, ws = world()
    lo, hi, halo_hi, local_pairs, global_pairs = window_shard(n_img, window, ws, rank)
    if ws == 1:
        return desc.to(device), k_per_image.to(device=device, dtype=torch.int32), local_pairs, global_pairs
    if rank == src:
        for r in range(ws):
            rlo, _, rhalo, _, _ = window_shard(n_img, window, ws, r)
            if r == src or rhalo <= rlo:
                continue
            dist.send(desc[rlo:rhalo].contiguous().to(device), dst=r)
            dist.send(k_per_image[rlo:rhalo].to(device=device, dtype=torch.int32).contiguous(), dst=r)
        return (desc[lo:halo_hi].contiguous().to(device), k_per_image[lo:halo_hi].to(device=device, dtype=torch.int32),
                local_pairs, global_pairs)
    n_loc = max(halo_hi - lo, 0)
    d = torch.empty((n_loc, k_stride, dim), dtype=dtype, device=device)
    k = torch.empty((n_loc,), dtype=torch.int32, device=device)
    if n_loc:
        dist.recv(d, src=src)
        dist.recv(k, src=src)
    return d, k, local_pairs, global_pairs
