"""Host-side mirror of the reference's matching call shape over the C-ABI (include/b200m.h).

Three faces over one kernel:
  * match_pairs / knn2_pairs   -- batched: a DescriptorBank + an int32 [P,2] pair list (the scheduler's output)
  * knn_match / match          -- cv2-shaped single pair: descriptor arrays in, index/distance arrays out
                                  (cv2.BFMatcher(NORM_L2).knnMatch(k=2) + ratio + crossCheck)
  * B200Matcher.__call__       -- LightGlue-shaped: matcher({'image0': feats0, 'image1': feats1}) -> dict,
                                  so matching.py:122-128 runs unchanged on its output.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import _lib
from .bank import DescriptorBank, FloatBank


@dataclass
class MatchConfig:
    ratio: float = 0.75      # l2_u8: Lowe ratio on float32 distances (<= 0 disables)
    mutual: bool = True      # cross-check
    metric: str = "l2_u8"    # "l2_u8": exact integer path (SIFT); "ip_bf16": inner product, bf16 tensor cores (DISK / SuperPoint)
    k: int = 2
    score_threshold: float | None = None   # ip_bf16: keep matches whose similarity exceeds it (lightglue.py:243-246)


@dataclass
class PairMatches:
    """Device-side result of match_pairs: fixed-capacity slots per pair + counts."""
    query_idx: torch.Tensor   # int32 [P, cap]
    train_idx: torch.Tensor   # int32 [P, cap]
    dist: torch.Tensor        # float32 [P, cap]
    count: torch.Tensor       # int32 [P]
    cap: int

    def check_overflow(self) -> None:
        if not self.count.numel():
            return
        mn, mx = (int(x) for x in torch.aminmax(self.count))
        if mn < 0:
            raise ValueError("the pair list names an image outside the bank (out_count = -1 for that pair)")
        if mx > self.cap:
            raise OverflowError(f"a pair produced {mx} matches but cap is {self.cap}; matches are never truncated silently")

    def flatten(self, check: bool = True) -> "FlatMatches":
        """Device-side compaction into flat lists + offsets (one host sync for the total). check=False only when the
        caller has just validated the counts itself (distributed.check_counts_everywhere does so on all ranks at once)."""
        if check:
            self.check_overflow()
        P = int(self.count.numel())
        offsets = torch.zeros((P + 1,), dtype=torch.int64, device=self.count.device)
        if P:
            torch.cumsum(self.count.to(torch.int64), 0, out=offsets[1:])
        total = int(offsets[-1].item()) if P else 0
        dev = self.count.device
        fq = torch.empty((total,), dtype=torch.int32, device=dev)
        ft = torch.empty((total,), dtype=torch.int32, device=dev)
        fd = torch.empty((total,), dtype=torch.float32, device=dev)
        if total:
            with torch.cuda.device(dev):
                st = torch.cuda.current_stream().cuda_stream
                _lib.check(_lib.lib().b200m_compact_matches(
                    self.query_idx.data_ptr(), self.train_idx.data_ptr(), self.dist.data_ptr(), self.count.data_ptr(),
                    offsets.data_ptr(), P, int(self.cap), fq.data_ptr(), ft.data_ptr(), fd.data_ptr(), st))
        return FlatMatches(fq, ft, fd, offsets)

    def to_lists(self):
        """-> list of (idx0 int64 [S], idx1 int64 [S], dist float32 [S]) numpy triples, one per pair."""
        self.check_overflow()
        cnt = self.count.cpu().numpy()
        q = self.query_idx.cpu().numpy()
        t = self.train_idx.cpu().numpy()
        d = self.dist.cpu().numpy()
        return [(q[p, :c].astype(np.int64), t[p, :c].astype(np.int64), d[p, :c].copy()) for p, c in enumerate(cnt)]


@dataclass
class FlatMatches:
    """Matches of all pairs back to back; pair p owns [offsets[p], offsets[p+1])."""
    query_idx: torch.Tensor   # int32 [S_total]
    train_idx: torch.Tensor   # int32 [S_total]
    dist: torch.Tensor        # float32 [S_total]
    offsets: torch.Tensor     # int64 [P + 1]

    def cpu(self, staging: bool = False) -> "FlatMatches":
        """Host copy. staging=True copies through cached PINNED buffers with one synchronisation; the returned
        tensors then alias those buffers and are overwritten by the next staging=True call on this process."""
        if not staging or not self.offsets.is_cuda:
            return FlatMatches(self.query_idx.cpu(), self.train_idx.cpu(), self.dist.cpu(), self.offsets.cpu())
        outs = []
        for name, t in (("q", self.query_idx), ("t", self.train_idx), ("d", self.dist), ("o", self.offsets)):
            h = _pinned(name, t.numel(), t.dtype)
            h.copy_(t, non_blocking=True)
            outs.append(h)
        torch.cuda.current_stream(self.offsets.device).synchronize()
        return FlatMatches(*outs)

    def pair(self, p: int):
        a, b = int(self.offsets[p]), int(self.offsets[p + 1])
        return self.query_idx[a:b], self.train_idx[a:b], self.dist[a:b]

    def digest(self, pair_ids=None):
        """-> (per-pair int64 [P], total as an unsigned 64-bit Python int): an order-independent 64-bit digest of
        every match record (pair id, query row, train row, float32 distance bits). `pair_ids` are the GLOBAL
        indices of this list's pairs (default 0..P-1), so the per-pair values of a rank's slice are the values a
        single process computes for the same pairs and the totals of all ranks simply add (mod 2^64): the
        N-GPU == 1-GPU check of bench.py and tests/test_gpu_multirank.py (matching.py:174,188-189: one ordered
        row per pair, whoever computed it)."""
        return digest_matches(self.query_idx, self.train_idx, self.dist, self.offsets, pair_ids)


_pinned_cache: dict = {}


def _pinned(name: str, numel: int, dtype) -> torch.Tensor:
    """A pinned host tensor of `numel` elements from a grow-only cache (cudaHostAlloc costs milliseconds)."""
    buf = _pinned_cache.get((name, dtype))
    if buf is None or buf.numel() < numel:
        buf = torch.empty((max(numel, 1) * 5 // 4 + 16,), dtype=dtype).pin_memory()
        _pinned_cache[(name, dtype)] = buf
    return buf[:numel]


def _i64(c: int) -> int:
    return c - (1 << 64) if c >= (1 << 63) else c


_DIGEST_K = tuple(_i64(c) for c in (0x9E3779B97F4A7C15, 0xC2B2AE3D27D4EB4F, 0x165667B19E3779F9, 0x27D4EB2F165667C5,
                                    0xBF58476D1CE4E5B9, 0x94D049BB133111EB))


def digest_matches(query_idx: torch.Tensor, train_idx: torch.Tensor, dist: torch.Tensor, offsets: torch.Tensor,
                   pair_ids=None):
    """See FlatMatches.digest. Pure integer tensor arithmetic (wrapping int64), identical on CPU and CUDA tensors
    and restated in numpy by tests/test_digest_cpu.py."""
    dev = offsets.device
    P = int(offsets.numel()) - 1
    counts = offsets[1:] - offsets[:-1]
    if pair_ids is None:
        pair_ids = torch.arange(P, dtype=torch.int64, device=dev)
    else:
        pair_ids = torch.as_tensor(pair_ids).to(device=dev, dtype=torch.int64)
        if pair_ids.numel() != P:
            raise ValueError("pair_ids must name every pair of the list")
    slot = torch.repeat_interleave(torch.arange(P, dtype=torch.int64, device=dev), counts)
    k1, k2, k3, k4, k5, k6 = _DIGEST_K
    x = (pair_ids[slot] * k1 + query_idx.to(torch.int64) * k2 + train_idx.to(torch.int64) * k3 +
         dist.contiguous().view(torch.int32).to(torch.int64) * k4)
    x = (x ^ (x >> 30)) * k5            # splitmix64 finaliser with arithmetic shifts (same in numpy / torch)
    x = (x ^ (x >> 27)) * k6
    x = x ^ (x >> 31)
    per_pair = torch.zeros((P,), dtype=torch.int64, device=dev)
    if x.numel():
        per_pair.index_add_(0, slot, x)
    total = int(per_pair.sum().item()) & ((1 << 64) - 1) if P else 0
    return per_pair, total


def _pairs_tensor(pairs, device, n_img: int | None = None) -> torch.Tensor:
    """int32 [P,2] on `device`. Host-side lists are range-checked here (free); a pair list that already lives on
    the device is checked by the kernels themselves: an entry outside [0, n_img) is treated as a pair of empty
    images and reported through out_count = -1 (PairMatches.check_overflow raises)."""
    if isinstance(pairs, torch.Tensor):
        if pairs.dim() != 2 or pairs.shape[1] != 2:
            raise ValueError("pairs must be [P,2]")
        if not pairs.is_cuda and n_img is not None and pairs.numel() and \
                (int(pairs.min()) < 0 or int(pairs.max()) >= n_img):
            raise ValueError(f"pairs must index images in [0, {n_img})")
        t = pairs.to(device=device, dtype=torch.int32)
    else:
        a = np.asarray(pairs, dtype=np.int64).reshape(-1, 2)
        if n_img is not None and a.size and (a.min() < 0 or a.max() >= n_img):
            raise ValueError(f"pairs must index images in [0, {n_img})")
        t = torch.as_tensor(a.astype(np.int32)).to(device)
    return t.contiguous()


def _check_out(out: "PairMatches", n_pairs: int) -> None:
    for name in ("query_idx", "train_idx", "dist"):
        t = getattr(out, name)
        if tuple(t.shape) != (n_pairs, out.cap) or not t.is_contiguous():
            raise ValueError(f"out.{name} must be a contiguous [{n_pairs}, {out.cap}] tensor, got {tuple(t.shape)}")
    if tuple(out.count.shape) != (n_pairs,):
        raise ValueError(f"out.count must be [{n_pairs}], got {tuple(out.count.shape)}")


def _workspace(bank: DescriptorBank, n_pairs: int) -> torch.Tensor:
    nbytes = _lib.lib().b200m_workspace_bytes(bank.k_stride, n_pairs)
    return torch.empty(int(nbytes), dtype=torch.uint8, device=bank.device)


def match_pairs(bank: DescriptorBank, pairs, ratio: float = 0.75, mutual: bool = True, cap: int | None = None,
                out: PairMatches | None = None, workspace: torch.Tensor | None = None) -> PairMatches:
    """Full rule for every (query_img, train_img) pair; asynchronous on the current stream."""
    pairs_t = _pairs_tensor(pairs, bank.device, bank.n_img)
    P = int(pairs_t.shape[0])
    if cap is None:
        cap = bank.k_stride
    if out is None:
        dev = bank.device
        out = PairMatches(torch.empty((P, cap), dtype=torch.int32, device=dev),
                          torch.empty((P, cap), dtype=torch.int32, device=dev),
                          torch.empty((P, cap), dtype=torch.float32, device=dev),
                          torch.zeros((P,), dtype=torch.int32, device=dev), int(cap))
    else:
        _check_out(out, P)
    if P == 0:
        return out
    ws = workspace if workspace is not None else _workspace(bank, P)
    with torch.cuda.device(bank.device):
        st = torch.cuda.current_stream().cuda_stream
        sd = bank.sorted_desc.data_ptr() if bank.sorted_desc is not None else None
        stab = bank.sort_table.data_ptr() if bank.sort_table is not None else None
        _lib.check(_lib.lib().b200m_match_pairs_u8(
            bank.desc.data_ptr(), bank.sqnorm.data_ptr(), sd, stab, bank.k_per_image.data_ptr(), bank.n_img, bank.k_stride,
            pairs_t.data_ptr(), P, float(ratio if ratio is not None else 0.0), int(bool(mutual)),
            out.query_idx.data_ptr(), out.train_idx.data_ptr(), out.dist.data_ptr(), out.count.data_ptr(),
            int(out.cap), ws.data_ptr(), ws.numel(), st))
    return out


def match_pairs_float(bank: FloatBank, pairs, score_threshold: float | None = None, mutual: bool = True,
                      cap: int | None = None, out: PairMatches | None = None,
                      workspace: torch.Tensor | None = None) -> PairMatches:
    """Float path: mutual arg-max of the inner product for every (query_img, train_img) pair
    (lightglue.py:223 + :235-254 in raw-similarity form). `dist` of the result holds the similarity."""
    pairs_t = _pairs_tensor(pairs, bank.device, bank.n_img)
    P = int(pairs_t.shape[0])
    if cap is None:
        cap = bank.k_stride
    dev = bank.device
    if out is None:
        out = PairMatches(torch.empty((P, cap), dtype=torch.int32, device=dev),
                          torch.empty((P, cap), dtype=torch.int32, device=dev),
                          torch.empty((P, cap), dtype=torch.float32, device=dev),
                          torch.zeros((P,), dtype=torch.int32, device=dev), int(cap))
    else:
        _check_out(out, P)
    if P == 0:
        return out
    L = _lib.lib()
    if workspace is None:
        workspace = torch.empty(int(L.b200m_workspace_bytes_bf16(bank.k_stride, P)), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(L.b200m_match_pairs_bf16(
            bank.desc.data_ptr(), bank.k_per_image.data_ptr(), bank.n_img, bank.k_stride, bank.dim,
            pairs_t.data_ptr(), P, float(score_threshold if score_threshold is not None else 0.0),
            int(score_threshold is not None), int(bool(mutual)),
            out.query_idx.data_ptr(), out.train_idx.data_ptr(), out.dist.data_ptr(), out.count.data_ptr(),
            int(out.cap), workspace.data_ptr(), workspace.numel(), st))
    return out


def match_float(desc_q, desc_t, score_threshold: float | None = None, mutual: bool = True, device="cuda"):
    """Single pair, float path: (query_idx int32 [S], train_idx int32 [S], similarity float32 [S])."""
    bank = FloatBank.from_arrays([desc_q, desc_t], device=device)
    q, t, s = match_pairs_float(bank, [[0, 1]], score_threshold, mutual).to_lists()[0]
    return q.astype(np.int32), t.astype(np.int32), s


def knn2_pairs(bank: DescriptorBank, pairs, workspace: torch.Tensor | None = None):
    """knnMatch(k=2) for every pair -> (train_idx int32 [P,k_stride,2], dist float32 [P,k_stride,2])."""
    pairs_t = _pairs_tensor(pairs, bank.device, bank.n_img)
    P = int(pairs_t.shape[0])
    idx = torch.full((P, bank.k_stride, 2), -1, dtype=torch.int32, device=bank.device)
    dist = torch.full((P, bank.k_stride, 2), float("inf"), dtype=torch.float32, device=bank.device)
    if P == 0:
        return idx, dist
    ws = workspace if workspace is not None else _workspace(bank, P)
    with torch.cuda.device(bank.device):
        st = torch.cuda.current_stream().cuda_stream
        if bank.sorted_desc is not None and _lib.get_option("pipeline") == 2 and _lib.get_option("epilogue") == 1:
            # chunk pipeline: bounds on the norm-sorted bank + exact evaluation of two chunks per row + exact fallback
            _lib.check(_lib.lib().b200m_knn2_sorted_u8(
                bank.desc.data_ptr(), bank.sqnorm.data_ptr(), bank.sorted_desc.data_ptr(), bank.sort_table.data_ptr(),
                bank.k_per_image.data_ptr(), bank.n_img, bank.k_stride, pairs_t.data_ptr(), P, idx.data_ptr(),
                dist.data_ptr(), ws.data_ptr(), ws.numel(), st))
        else:
            _lib.check(_lib.lib().b200m_knn2_u8(
                bank.desc.data_ptr(), bank.sqnorm.data_ptr(), bank.k_per_image.data_ptr(), bank.n_img, bank.k_stride,
                pairs_t.data_ptr(), P, idx.data_ptr(), dist.data_ptr(), ws.data_ptr(), ws.numel(), st))
    return idx, dist


def _new_out(n_pairs: int, cap: int, dev) -> PairMatches:
    return PairMatches(torch.empty((n_pairs, cap), dtype=torch.int32, device=dev),
                       torch.empty((n_pairs, cap), dtype=torch.int32, device=dev),
                       torch.empty((n_pairs, cap), dtype=torch.float32, device=dev),
                       torch.zeros((n_pairs,), dtype=torch.int32, device=dev), int(cap))


def _out_rows(out: PairMatches, lo: int, hi: int) -> PairMatches:
    return PairMatches(out.query_idx[lo:hi], out.train_idx[lo:hi], out.dist[lo:hi], out.count[lo:hi], out.cap)


def upload_blocks(n_img: int, block: int, ramp: bool = False):
    """Image-block boundaries of a chunked upload: [0, b1, b2, ..., n_img]. ramp=True starts small (2, 4, 8 images,
    then `block`): the first pairs can start after two images instead of after a whole block."""
    bounds, size = [0], (2 if ramp else block)
    while bounds[-1] < n_img:
        bounds.append(min(n_img, bounds[-1] + min(size, block)))
        size *= 2
    return bounds


def upload_segments(pairs: np.ndarray, n_img: int, block):
    """Split a pair list, IN ORDER, into segments that can start as soon as a prefix of the image blocks is
    resident: -> list of (last image block needed, first pair, end pair). block: images per block, or the block
    boundaries of upload_blocks. The blocked order of the scheduler (16 x 16 patches of the pair grid) needs the
    blocks one after the other; any other order still works (its first segment simply waits for more of the bank)."""
    if len(pairs) == 0:
        return []
    hi = pairs.max(axis=1)
    blk = hi // block if np.isscalar(block) else np.searchsorted(np.asarray(block[1:]), hi, side="right")
    need = np.maximum.accumulate(blk)
    cuts = np.nonzero(np.diff(need))[0] + 1
    starts = np.concatenate([[0], cuts])
    ends = np.concatenate([cuts, [len(pairs)]])
    return [(int(need[a]), int(a), int(b)) for a, b in zip(starts, ends)]


def match_pairs_host(desc_host: torch.Tensor, k_per_image, pairs, ratio: float = 0.75, mutual: bool = True,
                     cap: int | None = None, device="cuda", block: int = 16, out: PairMatches | None = None,
                     workspace: torch.Tensor | None = None, staging: bool = False) -> FlatMatches:
    """End-to-end call with HOST buffers: descriptors (uint8 [n_img, k_stride, 128], ideally pinned) and
    the pair list go host -> device, the matches come back as flat host arrays. This is the unit the
    reference performs per pair at matching.py:107-128 (host descriptors in, idx0 / idx1 on the host out).

    The bank is uploaded in blocks of `block` images on a copy stream; every block is prepared (norms, padding,
    norm sort) as it lands, and the pairs whose images are resident are matched while the rest of the bank is
    still on its way: only the first block's copy is exposed. Results are identical to
    match_pairs(DescriptorBank.from_packed(...)) -- same kernels, same order inside every pair.
    staging=True: the result aliases cached pinned buffers (FlatMatches.cpu)."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("match_pairs_host runs on a CUDA device (there is no CPU path)")
    if desc_host.dtype != torch.uint8 or desc_host.dim() != 3 or desc_host.shape[2] != 128:
        raise ValueError("desc_host must be uint8 [n_img, k_stride, 128]")
    n_img, ks = int(desc_host.shape[0]), int(desc_host.shape[1])
    kpi = torch.as_tensor(np.asarray(k_per_image, dtype=np.int32))
    pairs_np = np.ascontiguousarray(np.asarray(pairs.cpu() if isinstance(pairs, torch.Tensor) else pairs,
                                               dtype=np.int64).reshape(-1, 2))
    if pairs_np.size and (pairs_np.min() < 0 or pairs_np.max() >= n_img):
        raise ValueError(f"pairs must index images in [0, {n_img})")
    P = len(pairs_np)
    cap = ks if cap is None else int(cap)
    with torch.cuda.device(dev):
        main = torch.cuda.current_stream()
        bank = DescriptorBank.allocate(n_img, ks, kpi, dev)
        pairs_dev = torch.from_numpy(pairs_np.astype(np.int32)).to(dev, non_blocking=True)
        if out is None:
            out = _new_out(P, cap, dev)
        else:
            _check_out(out, P)
        ws = workspace if workspace is not None else _workspace(bank, max(P, 1))
        copy = _copy_stream(dev)
        copy.wait_stream(main)                       # the bank's memory may have been in use on `main` a moment ago
        n_blocks = (n_img + block - 1) // block
        landed = []
        with torch.cuda.stream(copy):
            for b in range(n_blocks):
                lo, hi = b * block, min(n_img, (b + 1) * block)
                bank.desc[lo:hi].copy_(desc_host[lo:hi], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy)
                landed.append(ev)
        segs = upload_segments(pairs_np, n_img, block)
        si = 0
        for b in range(n_blocks):
            main.wait_event(landed[b])
            bank.prepare_range(b * block, min(n_img, (b + 1) * block) - b * block)
            while si < len(segs) and segs[si][0] <= b:
                _, p0, p1 = segs[si]
                match_pairs(bank, pairs_dev[p0:p1], ratio=ratio, mutual=mutual, cap=cap, out=_out_rows(out, p0, p1),
                            workspace=ws)
                si += 1
        return out.flatten().cpu(staging)


_copy_streams: dict = {}


def _copy_stream(dev: torch.device) -> "torch.cuda.Stream":
    key = dev.index if dev.index is not None else torch.cuda.current_device()
    if key not in _copy_streams:
        _copy_streams[key] = torch.cuda.Stream(device=dev)
    return _copy_streams[key]


# ---------------------------------------------------------------------- cv2-shaped single pair
def knn_match(desc_q, desc_t, k: int = 2, device="cuda"):
    """cv2.BFMatcher(NORM_L2).knnMatch(desc_q, desc_t, k=2) as arrays:
    (train_idx int32 [Kq,2], dist float32 [Kq,2]); -1 / inf where cv2 would return a shorter row."""
    if k != 2:
        raise ValueError("only k=2 is implemented (the hot path's knnMatch call)")
    bank = DescriptorBank.from_arrays([desc_q, desc_t], device=device)
    idx, dist = knn2_pairs(bank, [[0, 1]])
    kq = int(bank.k_host()[0])
    return idx[0, :kq].cpu().numpy(), dist[0, :kq].cpu().numpy()


def match(desc_q, desc_t, ratio: float = 0.75, mutual: bool = True, device="cuda"):
    """(query_idx int32 [S], train_idx int32 [S], dist float32 [S]), ascending query index."""
    bank = DescriptorBank.from_arrays([desc_q, desc_t], device=device)
    res = match_pairs(bank, [[0, 1]], ratio=ratio, mutual=mutual)
    q, t, d = res.to_lists()[0]
    return q.astype(np.int32), t.astype(np.int32), d


# ---------------------------------------------------------------------- LightGlue-shaped
class B200Matcher:
    """Drop-in for the callable built at matching.py:20 and called at matching.py:122.

    Input contract (lightglue/lightglue.py:326-334): data['image0'|'image1'] dicts holding
    'descriptors' [1,M,128] (+ 'keypoints', 'image_size', accepted and ignored by the kNN rule).
    Output contract (lightglue/lightglue.py:442-450): 'matches0' [1,M], 'matches1' [1,N] (-1 = unmatched),
    'matching_scores0/1', 'matches': [LongTensor[S,2]] ascending in query index, 'scores': [FloatTensor[S]].
    'scores' carries the float32 L2 distance of each match (lower is better; matching.py ignores it).
    """

    def __init__(self, config: MatchConfig | None = None, device="cuda"):
        self.conf = config or MatchConfig()
        self.device = torch.device(device)
        if self.conf.metric not in ("l2_u8", "ip_bf16"):
            raise NotImplementedError(f"unknown metric {self.conf.metric}")

    def eval(self):
        return self

    def to(self, device):
        self.device = torch.device(device)
        return self

    def __call__(self, data: dict) -> dict:
        for key in ("image0", "image1"):
            assert key in data, f"Missing key {key} in data"          # lightglue.py:347-348
        d0 = data["image0"]["descriptors"]
        d1 = data["image1"]["descriptors"]
        assert d0.dim() == 3 and d1.dim() == 3 and d0.shape[0] == 1 and d1.shape[0] == 1, "batch size must be 1"
        m, n = d0.shape[1], d1.shape[1]
        if self.conf.metric == "ip_bf16":
            assert d0.shape[-1] == d1.shape[-1] and d0.shape[-1] in (128, 256)   # lightglue.py:365-366 (DISK 128, SuperPoint 256)
            fbank = FloatBank.from_arrays([d0[0], d1[0]], device=self.device)
            res = match_pairs_float(fbank, [[0, 1]], self.conf.score_threshold, self.conf.mutual)
        else:
            assert d0.shape[-1] == 128 and d1.shape[-1] == 128        # lightglue.py:365-366
            bank = DescriptorBank.from_arrays([d0[0], d1[0]], device=self.device)
            res = match_pairs(bank, [[0, 1]], ratio=self.conf.ratio, mutual=self.conf.mutual)
        s = int(res.count[0].item())      # the call's one host sync
        if s > res.cap:
            raise OverflowError(f"the pair produced {s} matches but cap is {res.cap}")
        q = res.query_idx[0, :s].long()
        t = res.train_idx[0, :s].long()
        dist = res.dist[0, :s]
        dev = self.device
        m0 = torch.full((1, m), -1, dtype=torch.long, device=dev)
        m1 = torch.full((1, n), -1, dtype=torch.long, device=dev)
        sc0 = torch.zeros((1, m), device=dev)
        sc1 = torch.zeros((1, n), device=dev)
        m0[0, q] = t
        sc0[0, q] = dist
        if self.conf.mutual:          # one-to-one only under the cross-check
            m1[0, t] = q
            sc1[0, t] = dist
        return {
            "matches0": m0, "matches1": m1, "matching_scores0": sc0, "matching_scores1": sc1,
            "stop": 0, "matches": [torch.stack([q, t], -1)], "scores": [dist],
        }

    forward = __call__
