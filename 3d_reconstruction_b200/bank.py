"""Packed u8 descriptor bank resident in HBM.

Replaces the per-pair repacking the reference does at matching.py:107-120 (a Python list
comprehension + fresh host->device copies for every pair) with one packed tensor
uint8 [n_img, k_stride, 128] + int32 squared norms [n_img, k_stride] + int32 k_per_image [n_img].
The on-disk source format is the object array written at feature_extraction.py:50
(`output/all_descriptors.npy`, one [K_i, 128] array per image).
"""
from __future__ import annotations

from typing import Sequence

import numpy as np
import torch

from . import _lib

DESC_DIM = 128
TILE_T = 256


def round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def to_u8_descriptors(a) -> np.ndarray:
    """[K,128] array -> contiguous uint8, verifying integer values in [0,255] for non-u8 input."""
    a = np.asarray(a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else a)
    if a.ndim != 2 or a.shape[1] != DESC_DIM:
        raise ValueError(f"descriptors must be [K,{DESC_DIM}], got {a.shape}")
    if a.dtype == np.uint8:
        return np.ascontiguousarray(a)
    r = np.rint(a)
    if a.size and (not np.array_equal(r, a) or r.min() < 0 or r.max() > 255):
        raise ValueError("the exact u8 path needs integer-valued descriptors in [0,255] (OpenCV SIFT); "
                         "float descriptors (DISK/SuperPoint) are not handled by this path")
    return np.ascontiguousarray(r.astype(np.uint8))


class DescriptorBank:
    """Device-resident bank. Padding rows are zero with the sentinel norm (see include/b200m.h)."""

    def __init__(self, desc: torch.Tensor, k_per_image: torch.Tensor, sqnorm: torch.Tensor,
                 sorted_desc: torch.Tensor | None = None, sort_table: torch.Tensor | None = None):
        self.desc = desc
        self.k_per_image = k_per_image
        self.sqnorm = sqnorm
        # norm-sorted copy + perm / inv / sorted norms / chunk minima / spreads (b200m_bank_sort_u8):
        # what the default pipeline (chunk sweep) runs on. None -> the slower pipeline 1.
        self.sorted_desc = sorted_desc
        self.sort_table = sort_table
        self._k_host = None

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_packed(cls, desc: torch.Tensor, k_per_image: torch.Tensor, sort: bool = True) -> "DescriptorBank":
        """desc: uint8 CUDA [n_img, k_stride, 128], k_stride % 256 == 0. Padding is zeroed in place.
        sort=True also builds the norm-sorted copy the default pipeline needs (twice the HBM footprint)."""
        if not desc.is_cuda:
            raise RuntimeError("the descriptor bank must live on a CUDA device (there is no CPU path)")
        if desc.dtype != torch.uint8 or desc.dim() != 3 or desc.shape[2] != DESC_DIM or not desc.is_contiguous():
            raise ValueError("desc must be contiguous uint8 [n_img, k_stride, 128]")
        if desc.shape[1] % TILE_T != 0:
            raise ValueError("k_stride must be a multiple of 256")
        if k_per_image.numel() != desc.shape[0]:
            raise ValueError("k_per_image must have one entry per image")
        if not k_per_image.is_cuda and k_per_image.numel() and \
                (int(k_per_image.min()) < 0 or int(k_per_image.max()) > desc.shape[1]):
            raise ValueError("k_per_image entries must lie in [0, k_stride]")   # device-side counts are clamped by the kernels
        k_per_image = k_per_image.to(device=desc.device, dtype=torch.int32).contiguous()
        n_tab = int(_lib.lib().b200m_norm_table_elems(desc.shape[0], desc.shape[1]))
        sqnorm = torch.empty((n_tab,), dtype=torch.int32, device=desc.device)   # norms + per-32-row minima
        with torch.cuda.device(desc.device):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(_lib.lib().b200m_bank_prepare_u8(desc.data_ptr(), desc.shape[0], desc.shape[1],
                                                        k_per_image.data_ptr(), sqnorm.data_ptr(), st))
            sorted_desc = sort_table = None
            if sort:
                n_st = int(_lib.lib().b200m_sort_table_elems(desc.shape[0], desc.shape[1]))
                sorted_desc = torch.empty_like(desc)
                sort_table = torch.empty((n_st,), dtype=torch.int32, device=desc.device)
                _lib.check(_lib.lib().b200m_bank_sort_u8(desc.data_ptr(), sqnorm.data_ptr(), k_per_image.data_ptr(),
                                                         desc.shape[0], desc.shape[1], sorted_desc.data_ptr(),
                                                         sort_table.data_ptr(), st))
        return cls(desc, k_per_image, sqnorm, sorted_desc, sort_table)

    @classmethod
    def allocate(cls, n_img: int, k_stride: int, k_per_image: torch.Tensor, device="cuda") -> "DescriptorBank":
        """Uninitialised device bank (descriptors, norm table, norm-sorted copy, sort table) for `n_img` images; the
        caller fills `desc` image range by image range and calls prepare_range for each (matcher.match_pairs_host
        behind chunked host->device copies, distributed.match_pairs_host on the image slice of every rank)."""
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("the descriptor bank must live on a CUDA device (there is no CPU path)")
        if k_stride % TILE_T != 0 or k_stride <= 0 or n_img <= 0:
            raise ValueError("k_stride must be a positive multiple of 256 and n_img positive")
        k_per_image = torch.as_tensor(k_per_image)
        if k_per_image.numel() != n_img:
            raise ValueError("k_per_image must have one entry per image")
        if not k_per_image.is_cuda and (int(k_per_image.min()) < 0 or int(k_per_image.max()) > k_stride):
            raise ValueError("k_per_image entries must lie in [0, k_stride]")
        L = _lib.lib()
        desc = torch.empty((n_img, k_stride, DESC_DIM), dtype=torch.uint8, device=dev)
        sqnorm = torch.empty((int(L.b200m_norm_table_elems(n_img, k_stride)),), dtype=torch.int32, device=dev)
        sorted_desc = torch.empty_like(desc)
        sort_table = torch.empty((int(L.b200m_sort_table_elems(n_img, k_stride)),), dtype=torch.int32, device=dev)
        sort_table[3 * n_img * k_stride + n_img * k_stride // 32:].zero_()     # the two per-image spread arrays (+ padding)
        return cls(desc, k_per_image.to(device=dev, dtype=torch.int32, non_blocking=True).contiguous(), sqnorm,
                   sorted_desc, sort_table)

    def prepare_range(self, img_lo: int, img_count: int) -> None:
        """Norms, padding, norm-sorted copy and tables of the images [img_lo, img_lo + img_count) on the current
        stream (b200m_bank_prepare_range_u8 + b200m_bank_sort_range_u8); ranges are independent."""
        L = _lib.lib()
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(L.b200m_bank_prepare_range_u8(self.desc.data_ptr(), self.n_img, self.k_stride,
                                                     self.k_per_image.data_ptr(), self.sqnorm.data_ptr(),
                                                     int(img_lo), int(img_count), st))
            if self.sorted_desc is not None:
                _lib.check(L.b200m_bank_sort_range_u8(self.desc.data_ptr(), self.sqnorm.data_ptr(),
                                                      self.k_per_image.data_ptr(), self.n_img, self.k_stride,
                                                      self.sorted_desc.data_ptr(), self.sort_table.data_ptr(),
                                                      int(img_lo), int(img_count), st))

    def table_sections(self):
        """Views of every per-image section of the bank's tensors as (tensor, rows per image) -- what a multi-GPU
        all-gather of image slices has to move (distributed.allgather_bank)."""
        R, ks, n = self.n_img * self.k_stride, self.k_stride, self.n_img
        out = [(self.desc.view(-1), ks * DESC_DIM), (self.sqnorm[:R], ks), (self.sqnorm[R:R + R // 32], ks // 32)]
        if self.sorted_desc is not None:
            t = self.sort_table
            n4 = (n + 3) & ~3
            b = 3 * R + R // 32
            out += [(self.sorted_desc.view(-1), ks * DESC_DIM), (t[:R], ks), (t[R:2 * R], ks), (t[2 * R:3 * R], ks),
                    (t[3 * R:b], ks // 32), (t[b:b + n], 1), (t[b + n4:b + n4 + n], 1)]
        return out

    @classmethod
    def from_arrays(cls, arrays: Sequence, device="cuda", k_stride: int | None = None) -> "DescriptorBank":
        """arrays: one [K_i,128] array per image (u8, or integer-valued float as cv2.SIFT emits).
        CUDA tensors already on `device` are verified and packed there (no host round trip)."""
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("the descriptor bank must live on a CUDA device (there is no CPU path)")
        if len(arrays) and all(isinstance(a, torch.Tensor) and a.is_cuda and
                               (dev.index is None or a.device == dev) for a in arrays):
            return cls._from_device_tensors(list(arrays), k_stride)
        u8 = [to_u8_descriptors(a) for a in arrays]
        if not u8:
            raise ValueError("empty descriptor list")
        kmax = max(a.shape[0] for a in u8)
        ks = round_up(max(kmax, 1), TILE_T) if k_stride is None else int(k_stride)
        if ks % TILE_T or ks < kmax:
            raise ValueError("k_stride must be a multiple of 256 and >= the largest image")
        host = torch.zeros((len(u8), ks, DESC_DIM), dtype=torch.uint8)
        if torch.cuda.is_available():
            host = host.pin_memory()
        hv = host.numpy()
        for i, a in enumerate(u8):
            hv[i, :a.shape[0]] = a
        kpi = torch.tensor([a.shape[0] for a in u8], dtype=torch.int32)
        return cls.from_packed(host.to(dev, non_blocking=True), kpi.to(dev))

    @classmethod
    def _from_device_tensors(cls, ts, k_stride: int | None) -> "DescriptorBank":
        for t in ts:
            if t.dim() != 2 or t.shape[1] != DESC_DIM:
                raise ValueError(f"descriptors must be [K,{DESC_DIM}], got {tuple(t.shape)}")
        kmax = max(int(t.shape[0]) for t in ts)
        ks = round_up(max(kmax, 1), TILE_T) if k_stride is None else int(k_stride)
        if ks % TILE_T or ks < kmax:
            raise ValueError("k_stride must be a multiple of 256 and >= the largest image")
        dev = ts[0].device
        desc = torch.zeros((len(ts), ks, DESC_DIM), dtype=torch.uint8, device=dev)
        bad = []
        for i, t in enumerate(ts):
            t = t.detach()
            if t.dtype != torch.uint8:
                r = t.round()
                if t.numel():
                    bad.append(((r != t) | (r < 0) | (r > 255)).any())
                t = r.to(torch.uint8)
            desc[i, :t.shape[0]] = t
        if bad and bool(torch.stack(bad).any().item()):
            raise ValueError("the exact u8 path needs integer-valued descriptors in [0,255] (OpenCV SIFT); "
                             "float descriptors (DISK/SuperPoint) are not handled by this path")
        kpi = torch.tensor([int(t.shape[0]) for t in ts], dtype=torch.int32)
        return cls.from_packed(desc, kpi)

    # ------------------------------------------------------------------ accessors
    @property
    def n_img(self) -> int:
        return int(self.desc.shape[0])

    @property
    def k_stride(self) -> int:
        return int(self.desc.shape[1])

    @property
    def device(self) -> torch.device:
        return self.desc.device

    def k_host(self) -> np.ndarray:
        if self._k_host is None:
            self._k_host = self.k_per_image.cpu().numpy()
        return self._k_host

    def nbytes(self) -> int:
        n = self.desc.numel() + 4 * self.sqnorm.numel() + 4 * self.k_per_image.numel()
        if self.sorted_desc is not None:
            n += self.sorted_desc.numel() + 4 * self.sort_table.numel()
        return n

    def sort_views(self):
        """(perm, inv, sorted norms) as int32 [n_img, k_stride] views, chunk minima [n_img, k_stride/32], spreads [n_img]."""
        R, n, ks = self.n_img * self.k_stride, self.n_img, self.k_stride
        t = self.sort_table
        return (t[:R].view(n, ks), t[R:2 * R].view(n, ks), t[2 * R:3 * R].view(n, ks),
                t[3 * R:3 * R + R // 32].view(n, ks // 32), t[3 * R + R // 32:3 * R + R // 32 + n])

    def norms(self) -> torch.Tensor:
        """int32 [n_img, k_stride] view of the squared norms (B200M_PAD_NORM on padding rows)."""
        return self.sqnorm[: self.n_img * self.k_stride].view(self.n_img, self.k_stride)


class FloatBank:
    """Device-resident bf16 bank for the float path: bf16 [n_img, k_stride, D], D in {128, 256}
    (DISK is 128-d, lightglue/disk.py:12; SuperPoint 256-d, lightglue/superpoint.py:99).
    Rows >= k_per_image[i] are padding and never compete."""

    def __init__(self, desc: torch.Tensor, k_per_image: torch.Tensor):
        if not desc.is_cuda:
            raise RuntimeError("the descriptor bank must live on a CUDA device (there is no CPU path)")
        if desc.dtype != torch.bfloat16 or desc.dim() != 3 or desc.shape[2] not in (128, 256) or not desc.is_contiguous():
            raise ValueError("desc must be contiguous bfloat16 [n_img, k_stride, 128 or 256]")
        if desc.shape[1] % TILE_T != 0:
            raise ValueError("k_stride must be a multiple of 256")
        self.desc = desc
        if k_per_image.numel() != desc.shape[0]:
            raise ValueError("k_per_image must have one entry per image")
        if not k_per_image.is_cuda and k_per_image.numel() and \
                (int(k_per_image.min()) < 0 or int(k_per_image.max()) > desc.shape[1]):
            raise ValueError("k_per_image entries must lie in [0, k_stride]")
        self.k_per_image = k_per_image.to(device=desc.device, dtype=torch.int32).contiguous()

    @classmethod
    def from_arrays(cls, arrays: Sequence, device="cuda", k_stride: int | None = None) -> "FloatBank":
        """arrays: one float [K_i, D] array / tensor per image; rounded to bf16 on the device."""
        ts = [a.detach() if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a)) for a in arrays]
        if not ts:
            raise ValueError("empty descriptor list")
        dim = int(ts[0].shape[1])
        if dim not in (128, 256) or any(t.dim() != 2 or t.shape[1] != dim for t in ts):
            raise ValueError("descriptors must be [K, 128] or [K, 256], the same length for every image")
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("the descriptor bank must live on a CUDA device (there is no CPU path)")
        kmax = max(int(t.shape[0]) for t in ts)
        ks = round_up(max(kmax, 1), TILE_T) if k_stride is None else int(k_stride)
        desc = torch.zeros((len(ts), ks, dim), dtype=torch.bfloat16, device=dev)
        for i, t in enumerate(ts):
            desc[i, :t.shape[0]] = t.to(device=dev, dtype=torch.bfloat16)
        return cls(desc, torch.tensor([int(t.shape[0]) for t in ts], dtype=torch.int32))

    @property
    def n_img(self) -> int:
        return int(self.desc.shape[0])

    @property
    def k_stride(self) -> int:
        return int(self.desc.shape[1])

    @property
    def dim(self) -> int:
        return int(self.desc.shape[2])

    @property
    def device(self) -> torch.device:
        return self.desc.device
