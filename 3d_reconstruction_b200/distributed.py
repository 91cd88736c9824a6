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


def image_slices(n_img: int, ws: int):
    """Equal image slices for the sharded upload: rank r owns [r * n_loc, min((r + 1) * n_loc, n_img)); the bank is
    padded to n_loc * ws images (empty ones) so that every all-gather moves equal parts. -> (n_loc, n_pad)"""
    n_loc = (n_img + ws - 1) // ws
    return n_loc, n_loc * ws


def allgather_bank(bank, n_loc: int) -> None:
    """In-place all-gather of every per-image section of a bank whose images [rank * n_loc, (rank + 1) * n_loc) were
    uploaded and prepared locally (descriptors, norms, minima, norm-sorted copy, perm / inv / sorted norms / chunk
    minima / spreads): afterwards every rank holds the whole prepared bank. NCCL over NVLink (gloo on CPU tensors)."""
    rank, ws = world()
    if ws == 1:
        return
    works = []
    for t, per_img in bank.table_sections():
        n = n_loc * per_img
        works.append(dist.all_gather_into_tensor(t[: n * ws], t[rank * n: (rank + 1) * n], async_op=True))
    for w in works:
        w.wait()


def gather_packed(counts: torch.Tensor, fields: list, n_pairs_of: list, dst: int = 0):
    """Variable-length gather of per-rank match lists with ONE size exchange and ONE message per rank.
    counts: int32 [P_r] matches per pair of this rank's (contiguous, rank-ordered) slice; fields: 4-byte tensors of
    length sum(counts) each; n_pairs_of[r] = P_r of every rank (known everywhere: the partition is static).
    -> on `dst`: (global int64 offsets [P + 1], [fields concatenated in rank order]); None elsewhere."""
    rank, ws = world()
    dev = counts.device
    total = counts.sum(dtype=torch.int64).reshape(1)
    if ws == 1:
        off = torch.zeros((counts.numel() + 1,), dtype=torch.int64, device=dev)
        torch.cumsum(counts.to(torch.int64), 0, out=off[1:])
        return off, fields
    totals = torch.empty((ws,), dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(totals, total)
    totals = totals.tolist()                                   # the exchange's one host synchronisation
    nf = len(fields)
    packed = torch.empty((counts.numel() + nf * totals[rank],), dtype=torch.int32, device=dev)
    packed[: counts.numel()] = counts
    for i, f in enumerate(fields):
        packed[counts.numel() + i * totals[rank]: counts.numel() + (i + 1) * totals[rank]] = f.view(torch.int32)
    if rank != dst:
        for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, packed, dst)]):
            req.wait()
        return None
    bufs = {dst: packed}
    ops = []
    for r in range(ws):
        if r != dst:
            bufs[r] = torch.empty((n_pairs_of[r] + nf * totals[r],), dtype=torch.int32, device=dev)
            ops.append(dist.P2POp(dist.irecv, bufs[r], r))
    for req in dist.batch_isend_irecv(ops):
        req.wait()
    all_counts = torch.cat([bufs[r][: n_pairs_of[r]] for r in range(ws)])
    off = torch.zeros((all_counts.numel() + 1,), dtype=torch.int64, device=dev)
    torch.cumsum(all_counts.to(torch.int64), 0, out=off[1:])
    out = []
    for i, f in enumerate(fields):
        parts = [bufs[r][n_pairs_of[r] + i * totals[r]: n_pairs_of[r] + (i + 1) * totals[r]] for r in range(ws)]
        out.append(torch.cat(parts).view(f.dtype))
    return off, out


def match_pairs_host(desc_host_local: torch.Tensor, k_per_image, n_img: int, pairs, ratio: float = 0.75,
                     mutual: bool = True, cap: int | None = None, device="cuda", dst: int = 0, out=None, workspace=None,
                     staging: bool = False, events: dict | None = None):
    """matcher.match_pairs_host over all ranks of the job: HOST descriptors in, flat HOST match lists out (on rank
    `dst`; None elsewhere), the per-pair host -> device -> host unit of matching.py:107-128 sharded and batched.

    desc_host_local: this rank's slice of the host bank, uint8 [<= n_loc, k_stride, 128] = images
    [rank * n_loc, ...) with n_loc = ceil(n_img / world) (image_slices) -- each process holds the descriptors of the
    images it extracted or loaded, and uploads only those over its own PCIe link. k_per_image: all n_img counts
    (tiny, known everywhere). pairs: the WHOLE pair list, identical on every rank. Steps: local upload -> norms +
    norm sort of the local slice -> in-place all-gather of bank + tables over NVLink -> this rank's static,
    cost-balanced slice of the pairs (scheduler.partition) -> one packed gather to `dst` -> pinned D2H.
    With one rank this is matcher.match_pairs_host (chunked upload overlapped with the matching).
    events: optional dict, filled with (start, stop) CUDA events of the phases "upload_prepare", "allgather",
    "match", "gather" on the current stream (bench.py reports the warm collective times from them)."""
    from . import matcher as M
    from .bank import DescriptorBank
    rank, ws = world()
    if ws == 1:
        return M.match_pairs_host(desc_host_local, k_per_image, pairs, ratio, mutual, cap, device, out=out,
                                  workspace=workspace, staging=staging)
    dev = torch.device(device)
    ks = int(desc_host_local.shape[1])
    n_loc, n_pad = image_slices(n_img, ws)
    lo = rank * n_loc
    have = max(0, min(n_img, lo + n_loc) - lo)
    if int(desc_host_local.shape[0]) != have:
        raise ValueError(f"rank {rank} must hold images [{lo}, {lo + have}) of the host bank, got {desc_host_local.shape[0]}")
    kpi = np.zeros(n_pad, np.int32)
    kpi[:n_img] = np.asarray(k_per_image, dtype=np.int32)
    pairs_np = np.ascontiguousarray(np.asarray(pairs, dtype=np.int32).reshape(-1, 2))
    cap = ks if cap is None else int(cap)
    def mark(name, which):
        if events is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            events.setdefault(name, [None, None])[which] = ev

    with torch.cuda.device(dev):
        mark("upload_prepare", 0)
        bank = DescriptorBank.allocate(n_pad, ks, torch.from_numpy(kpi), dev)
        if have:
            bank.desc[lo:lo + have].copy_(desc_host_local, non_blocking=True)
        bank.prepare_range(lo, n_loc)
        mark("upload_prepare", 1)
        mark("allgather", 0)
        allgather_bank(bank, n_loc)
        mark("allgather", 1)
        mine, idx = partition(pairs_np, ws, rank, kpi[:n_img])
        n_pairs_of = [len(partition(pairs_np, ws, r, kpi[:n_img])[1]) for r in range(ws)]
        pd = torch.from_numpy(np.ascontiguousarray(mine)).to(dev, non_blocking=True)
        mark("match", 0)
        res = M.match_pairs(bank, pd, ratio=ratio, mutual=mutual, cap=cap, out=out, workspace=workspace)
        flat = res.flatten()
        mark("match", 1)
        counts = (flat.offsets[1:] - flat.offsets[:-1]).to(torch.int32)
        mark("gather", 0)
        got = gather_packed(counts, [flat.query_idx, flat.train_idx, flat.dist], n_pairs_of, dst)
        mark("gather", 1)
        if got is None:
            return None
        off, (fq, ft, fd) = got
        return M.FlatMatches(fq, ft, fd, off).cpu(staging)


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
