"""Multi-GPU plumbing: one process per GPU (torchrun), NCCL / NVLink for the exchange steps.

The path shards by image pair with no collective during compute. Two ways to run it:
  * bank resident (bench.py's `value` leg): broadcast_bank_tensors once, every rank takes its static, cost-balanced
    slice of the same pair list (my_pairs), gather_flat / gather_packed bring the variable-length lists to rank 0;
  * from HOST buffers (match_pairs_host, bench.py's `e2e` leg): every rank uploads and prepares only its own image
    slice, pairs between a rank's own images are matched while the prepared slices travel between the GPUs (copy
    engines over symmetric memory; NCCL all-gathers as the fallback -- the choice is agreed by all ranks, agree_all),
    the rest follows, one packed gather to `dst`. Count errors (overflow, bad pair) are raised on EVERY rank
    (check_counts_everywhere), never on one rank alone in front of a collective.
  * scatter_bank_window: the halo-sharded bank of a sequential sweep (BASELINE config 4).
The helpers also work with the gloo backend on CPU tensors (the world_size-2 CPU tests).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from .scheduler import partition, partition_all_by_owner


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


def agree_all(ok: bool, device) -> bool:
    """True on every rank iff `ok` is true on every rank (one MIN all-reduce of a flag). For decisions that pick which
    collectives run next: if only some ranks took the other branch, the job would hang instead of failing."""
    rank, ws = world()
    if ws == 1:
        return bool(ok)
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return bool(int(flag[0]))


def check_counts_everywhere(count: torch.Tensor, cap: int) -> None:
    """PairMatches.check_overflow as a COLLECTIVE: the smallest and the largest out_count over ALL ranks are reduced
    (one MAX all-reduce of two numbers), and the same exception is raised on every rank -- OverflowError when some
    pair on some rank found more matches than `cap`, ValueError when some rank's pair list named an image outside the
    bank. Without this only the affected rank would raise, just before gather_packed, and the others would sit in that
    collective until the backend's timeout. count: this rank's int32 out_count (may be empty)."""
    rank, ws = world()
    if count.numel():
        mn, mx = torch.aminmax(count)
        st = torch.stack([-mn, mx]).to(torch.int64)
    else:
        st = torch.tensor([0, 0], dtype=torch.int64, device=count.device)
    if ws > 1:
        dist.all_reduce(st, op=dist.ReduceOp.MAX)
    neg_mn, mx = (int(x) for x in st.tolist())
    if neg_mn > 0:
        raise ValueError("on at least one rank the pair list names an image outside the bank (out_count = -1 for that pair)")
    if mx > cap:
        raise OverflowError(f"a pair (on some rank) produced {mx} matches but cap is {cap}; matches are never truncated "
                            "silently -- raised on every rank")


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


class _SymmBank:
    """Bank tensors in SYMMETRIC memory (torch.distributed._symmetric_memory: every rank's buffer is mapped into
    every other rank's address space over NVLink), cached per shape: lets the ranks PULL each other's prepared
    image slices with plain device-to-device copies on the copy engines -- no SM is taken from the matching
    kernels that run meanwhile."""

    def __init__(self, n_pad: int, ks: int, dev: torch.device):
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        from .bank import DESC_DIM, DescriptorBank
        L = _lib.lib()
        sizes = (n_pad * ks * DESC_DIM, int(L.b200m_norm_table_elems(n_pad, ks)), n_pad * ks * DESC_DIM,
                 int(L.b200m_sort_table_elems(n_pad, ks)))
        dtypes = (torch.uint8, torch.int32, torch.uint8, torch.int32)
        self.tensors = [symm.empty(n, dtype=dt, device=dev) for n, dt in zip(sizes, dtypes)]
        self.handles = [symm.rendezvous(t, dist.group.WORLD) for t in self.tensors]
        desc, sqnorm, sdesc, stab = self.tensors
        stab[3 * n_pad * ks + n_pad * ks // 32:].zero_()
        kpi = torch.zeros((n_pad,), dtype=torch.int32, device=dev)
        self.bank = DescriptorBank(desc.view(n_pad, ks, DESC_DIM), kpi, sqnorm, sdesc.view(n_pad, ks, DESC_DIM), stab)
        self.stream = torch.cuda.Stream(device=dev)
        # (tensor index, first element, elements per image) of every per-image section, as bank.table_sections() lists them
        self.sections = []
        for view, per in self.bank.table_sections():
            for ti, t in enumerate(self.tensors):
                if view.untyped_storage().data_ptr() == t.untyped_storage().data_ptr():
                    self.sections.append((ti, int(view.storage_offset()), per))
                    break
        assert len(self.sections) == len(self.bank.table_sections())

    def pull_all(self, n_loc: int) -> None:
        """On the current stream: device barrier (every rank's slice is prepared), copy every remote slice of every
        section from its owner's buffer into the same place of the local one, device barrier (nobody overwrites a
        slice that is still being read)."""
        rank, ws = world()
        self.handles[0].barrier()
        for step in range(1, ws):
            src = (rank - step) % ws
            for ti, first, per in self.sections:
                n = n_loc * per
                t = self.tensors[ti]
                t[first + src * n: first + (src + 1) * n].copy_(
                    self.handles[ti].get_buffer(src, (n,), t.dtype, first + src * n))
        self.handles[0].barrier()


_symm_banks: dict = {}


def _symm_bank(n_pad: int, ks: int, dev: torch.device):
    """Cached symmetric-memory bank, or None when symmetric memory cannot be set up here (then NCCL moves the bank)."""
    key = (n_pad, ks, dev.index)
    if key not in _symm_banks:
        import warnings
        release_symm_banks()         # one shape at a time: a bank + its sorted copy per shape would pile up otherwise
        try:
            sb = _SymmBank(n_pad, ks, dev)
        except Exception as ex:      # no P2P mapping between these GPUs, or an older torch: fall back to the NCCL all-gather
            warnings.warn(f"symmetric memory unavailable ({type(ex).__name__}: {ex}); using the NCCL all-gather")
            sb = None
        # the transport is a JOB-wide choice: a rank that went on alone to the NCCL all-gather while the others wait in
        # the symmetric-memory barrier would hang the job. Every rank reaches this line once per shape (the cache keys
        # are the same everywhere), so one flag all-reduce settles it.
        if not agree_all(sb is not None, dev) and sb is not None:
            warnings.warn("symmetric memory failed on another rank; every rank uses the NCCL all-gather")
            sb = None
        _symm_banks[key] = sb
    return _symm_banks[key]


def release_symm_banks() -> None:
    """Drop the cached symmetric-memory bank(s) so their device memory (a whole bank, its norm-sorted copy and the
    tables, per rank) can be reused. Called automatically when match_pairs_host meets a new bank shape (the cache holds
    ONE shape); call it yourself after the last match_pairs_host of a job. Every rank must do so at the same point of
    the program, like any other step of a multi-rank job (the next call rebuilds the buffers collectively)."""
    for key in [k for k in _symm_banks if k[0] != "nccl"]:
        del _symm_banks[key]


def transport_in_use() -> str:
    """What moved the bank between the ranks in the match_pairs_host calls so far: "ce" (copy engines over
    symmetric memory), "nccl" (in-place all-gathers), or "none"."""
    if not _symm_banks:
        return "none"
    return "ce" if any(v is not None for v in _symm_banks.values()) else "nccl"


def match_pairs_host(desc_host_local: torch.Tensor, k_per_image, n_img: int, pairs, ratio: float = 0.75,
                     mutual: bool = True, cap: int | None = None, device="cuda", dst: int = 0, out=None, workspace=None,
                     staging: bool = False, events: dict | None = None, transport: str = "auto", block: int = 12,
                     rank_weights=None):
    """matcher.match_pairs_host over all ranks of the job: HOST descriptors in, flat HOST match lists out (on rank
    `dst`, in the order of `pairs`; None elsewhere) -- the per-pair host -> device -> host unit of
    matching.py:107-128, sharded and batched.

    desc_host_local: this rank's slice of the host bank, uint8 [<= n_loc, k_stride, 128] = images
    [rank * n_loc, ...) with n_loc = ceil(n_img / world) (image_slices): each process holds the descriptors of the
    images it extracted or loaded and uploads only those, over its own PCIe link. k_per_image: all n_img counts
    (tiny, known everywhere). pairs: the WHOLE pair list, identical on every rank.
      1. own images host -> device in blocks (2, 4, 8, then `block` images) on a copy stream, norms + norm sort per block;
      2. every pair belongs to the owner of one of its images (scheduler.partition_by_owner); the pairs between
         a rank's OWN images are matched first, block by block as the upload lands,
      3. while the ranks pull each other's prepared slices (bank, norm-sorted bank, tables) over NVLink with the
         copy engines (symmetric memory; transport="nccl" or a failed set-up: one in-place NCCL all-gather per
         section instead, which cannot overlap the persistent matching kernels);
      4. the remaining pairs, one packed gather to `dst`, results put back into list order, pinned D2H.
    rank_weights: optional relative throughput of the ranks (identical everywhere, see rank_throughput): the
    partition hands out pairs in that proportion. With one rank this is matcher.match_pairs_host. events: optional dict, filled with (start, stop) CUDA events of
    the phases "upload_prepare", "exchange", "match_local", "match_rest", "gather" (bench.py reports them)."""
    from . import matcher as M
    from .bank import DescriptorBank
    rank, ws = world()
    if ws == 1:
        return M.match_pairs_host(desc_host_local, k_per_image, pairs, ratio, mutual, cap, device, out=out,
                                  workspace=workspace, staging=staging)
    if transport not in ("auto", "ce", "nccl"):
        raise ValueError("transport must be 'auto', 'ce' (copy engines over symmetric memory) or 'nccl'")
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
    if pairs_np.size and (pairs_np.min() < 0 or pairs_np.max() >= n_img):
        raise ValueError(f"pairs must index images in [0, {n_img})")
    cap = ks if cap is None else int(cap)

    def mark(name, which, stream=None):
        if events is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(stream) if stream is not None else ev.record()
            events.setdefault(name, [None, None])[which] = ev

    with torch.cuda.device(dev):
        main = torch.cuda.current_stream()
        symm = _symm_bank(n_pad, ks, dev) if transport in ("auto", "ce") else None
        if transport == "nccl":
            _symm_banks.setdefault(("nccl", ks, dev.index), None)
        if symm is None and transport == "ce":
            raise RuntimeError("symmetric memory could not be set up on this machine (transport='ce')")
        if symm is not None:
            bank = symm.bank
            bank.k_per_image.copy_(torch.from_numpy(kpi), non_blocking=True)
        else:
            bank = DescriptorBank.allocate(n_pad, ks, torch.from_numpy(kpi), dev)
        # 1: chunked upload of the own images -- enqueued before any host-side planning, which then runs under it
        mark("upload_prepare", 0)
        copy = M._copy_stream(dev)
        copy.wait_stream(main)
        bounds = M.upload_blocks(n_loc, block, ramp=True)
        n_blocks = len(bounds) - 1
        landed = []
        with torch.cuda.stream(copy):
            for b in range(n_blocks):
                b0, b1 = min(have, bounds[b]), min(have, bounds[b + 1])
                if b1 > b0:
                    bank.desc[lo + b0: lo + b1].copy_(desc_host_local[b0:b1], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy)
                landed.append(ev)
        parts = partition_all_by_owner(pairs_np, ws, n_loc, rank_weights)
        for r, (ix, nl) in enumerate(parts):               # local pairs in the order their images land (stable)
            need = np.searchsorted(np.asarray(bounds[1:]), pairs_np[ix[:nl]].max(axis=1) - r * n_loc, side="right")
            ix[:nl] = ix[:nl][np.argsort(need.astype(np.int16), kind="stable")]
        idx, n_local = parts[rank]
        mine = pairs_np[idx]
        P = len(mine)
        pd = torch.from_numpy(np.ascontiguousarray(mine)).to(dev, non_blocking=True)
        if out is None:
            out = M._new_out(P, cap, dev)
        else:
            M._check_out(out, P)
        ws_t = workspace if workspace is not None else M._workspace(bank, max(P, 1))
        # 2: local pairs as their images land
        segs = M.upload_segments(mine[:n_local] - lo, n_loc, bounds) if symm is not None else []

        def match(p0, p1):
            if p1 > p0:
                M.match_pairs(bank, pd[p0:p1], ratio=ratio, mutual=mutual, cap=cap, out=M._out_rows(out, p0, p1), workspace=ws_t)

        # Every block is prepared (norms, padding, norm sort) as it lands; between two blocks only as many local pairs
        # are matched as the next block's upload lasts (~66,000 / k_stride pairs per image on a B200 behind PCIe 5), so
        # that the GPU is busy during the upload but the LAST preparation -- which the other ranks wait for -- is not
        # queued behind the bulk of the local pairs: those run while the slices travel.
        avail_end = {sg[0]: sg[2] for sg in segs}           # block -> end of the local pairs that need no later block
        done = 0
        for b in range(n_blocks):
            main.wait_event(landed[b])
            bank.prepare_range(lo + bounds[b], bounds[b + 1] - bounds[b])
            if b == 0:
                mark("match_local", 0)
            if b < n_blocks - 1:
                end_b = max([e for bb, e in avail_end.items() if bb <= b], default=0)
                budget = -(-(bounds[b + 2] - bounds[b + 1]) * 66000 // ks)
                take = min(end_b, done + budget)
                match(done, take)
                done = max(done, take)
        mark("upload_prepare", 1)
        if symm is not None:                                # 3: the exchange, on its own stream: the copy engines do it
            prepared = torch.cuda.Event()
            prepared.record(main)                           # the whole slice is ready for the other ranks
            symm.stream.wait_event(prepared)
            with torch.cuda.stream(symm.stream):
                mark("exchange", 0, symm.stream)
                symm.pull_all(n_loc)
                mark("exchange", 1, symm.stream)
                exchanged = torch.cuda.Event()
                exchanged.record(symm.stream)
            match(done, n_local)                            # the rest of the local pairs, under the exchange
        if symm is not None:
            mark("match_local", 1)
            main.wait_event(exchanged)
            rest0 = n_local
        else:
            mark("exchange", 0)
            allgather_bank(bank, n_loc)
            mark("exchange", 1)
            rest0 = 0
        # 4: everything else
        mark("match_rest", 0)
        match(rest0, P)
        if rank == dst:                                 # host work under the running kernels: gathered position -> list index
            order = torch.from_numpy(np.concatenate([pp[0] for pp in parts])).to(dev, non_blocking=True)
        check_counts_everywhere(out.count, cap)          # collective: an overflow on one rank raises on all of them
        flat = out.flatten(check=False)
        mark("match_rest", 1)
        counts = (flat.offsets[1:] - flat.offsets[:-1]).to(torch.int32)
        mark("gather", 0)
        got = gather_packed(counts, [flat.query_idx, flat.train_idx, flat.dist], [len(pp[0]) for pp in parts], dst)
        if got is None:
            mark("gather", 1)
            return None
        off, fields = got
        # rank-ordered concatenation -> the order of `pairs`
        cnt = off[1:] - off[:-1]
        cnt_list = torch.empty_like(cnt)
        cnt_list[order] = cnt
        off_list = torch.zeros_like(off)
        torch.cumsum(cnt_list, 0, out=off_list[1:])
        dest = torch.repeat_interleave(off_list[:-1][order] - off[:-1], cnt) + torch.arange(int(off[-1]), device=dev)
        outs = []
        for f in fields:
            g = torch.empty_like(f)
            g[dest] = f
            outs.append(g)
        mark("gather", 1)
        return M.FlatMatches(outs[0], outs[1], outs[2], off_list).cpu(staging)


def rank_throughput(ms_mine: float, n_pairs_mine: int, device="cuda"):
    """All-gather of every rank's measured pairs / ms (e.g. over warm-up steps), rounded so that every rank holds the
    same numbers: the `weights` of scheduler.partition_by_owner / the `rank_weights` of match_pairs_host. A rank that
    measured nothing counts as average. -> list of world_size floats (None without torch.distributed)."""
    rank, ws = world()
    if ws == 1:
        return None
    backend = dist.get_backend()
    dev = torch.device(device) if backend == "nccl" else torch.device("cpu")
    mine = torch.tensor([n_pairs_mine / ms_mine if ms_mine > 0 and n_pairs_mine > 0 else 0.0], dtype=torch.float64, device=dev)
    allv = torch.zeros((ws,), dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(allv, mine)
    v = allv.cpu().numpy()
    good = v > 0
    if not good.any():
        return None
    v[~good] = v[good].mean()
    v = np.clip(v / v.mean(), 0.8, 1.25)                    # a wild measurement must not starve a rank
    return [round(float(x), 4) for x in v]


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
