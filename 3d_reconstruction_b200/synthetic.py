"""Synthetic SIFT-like descriptor banks with planted correspondences.

Uniform-random u8 descriptors give zero ratio-test survivors, so parity on them proves nothing:
images here share rows of a common "world" table (perturbed by small integer noise) with their
neighbours, which yields true matches that survive the Lowe ratio and the mutual check.
Values mimic cv2.SIFT output: non-negative integers, row L2 norm ~512, clipped at 255.
"""
from __future__ import annotations

import torch


NORM_MEAN = 512.0   # cv2.SIFT scales unit descriptors by 512 before rounding/saturating
NORM_STD = 0.56     # real SIFT rows: squared norm 260485..263667, std 577 (DSC_0001/0002) => norm std ~0.56


def _sift_quantise(x: torch.Tensor, gen: torch.Generator) -> torch.Tensor:
    """Scale rows to an L2 norm drawn around 512 (as SIFT's final normalisation does), round, saturate."""
    target = NORM_MEAN + NORM_STD * torch.randn((x.shape[0], 1), generator=gen, device=x.device)
    x = x * (target / x.norm(dim=1, keepdim=True).clamp_min(1e-6))
    return x.round_().clamp_(0, 255).to(torch.uint8)


def sift_like_rows(n: int, gen: torch.Generator, device) -> torch.Tensor:
    x = torch.randn((n, 128), generator=gen, device=device).abs_()
    return _sift_quantise(x, gen)


def make_bank_tensor(n_img: int, k: int, seed: int = 1234, p_shared: float = 0.30, noise: int = 6,
                     device="cpu", k_stride: int | None = None, k_per_image=None, dup_frac: float = 0.001):
    """-> (desc uint8 [n_img, k_stride, 128] zero-padded, k_per_image int32 [n_img]).

    Image i takes round(p_shared*K_i) rows from a sliding window of the world table (neighbouring
    images overlap => true correspondences), each perturbed by integer noise in [-noise, noise];
    the rest are fresh SIFT-like rows; `dup_frac` of the rows are exact duplicates of other rows of
    the same image so the lowest-index tie-break path is exercised; rows are shuffled. Every row
    (fresh or re-observed) goes through SIFT's final normalise-to-512 / round / saturate step, so
    squared norms have the spread of real SIFT output (std ~577 around 262144).
    """
    device = torch.device(device)
    ks = k_stride if k_stride is not None else (k + 255) // 256 * 256
    window = max(2, k // 2)
    step = max(1, window // 4)
    n_world = step * n_img + window
    gw = torch.Generator(device=device).manual_seed(seed)
    world = sift_like_rows(n_world, gw, device)
    desc = torch.zeros((n_img, ks, 128), dtype=torch.uint8, device=device)
    kpi = []
    for i in range(n_img):
        ki = int(k if k_per_image is None else k_per_image[i])
        kpi.append(ki)
        if ki == 0:
            continue
        g = torch.Generator(device=device).manual_seed(seed + 1 + i)
        n_sh = min(int(round(p_shared * ki)), window)
        sel = torch.randperm(window, generator=g, device=device)[:n_sh] + i * step
        sh = world[sel].to(torch.float32)
        sh = sh + torch.randint(-noise, noise + 1, sh.shape, generator=g, device=device).to(torch.float32)
        sh = _sift_quantise(sh.clamp_(min=0), g)        # a re-observed keypoint is normalised again by SIFT
        fresh = sift_like_rows(ki - n_sh, g, device)
        rows = torch.cat([sh, fresh], 0)
        rows = rows[torch.randperm(ki, generator=g, device=device)]
        n_dup = int(dup_frac * ki)
        if n_dup > 0 and ki > 1:
            src = torch.randint(0, ki, (n_dup,), generator=g, device=device)
            dst = torch.randint(0, ki, (n_dup,), generator=g, device=device)
            rows[dst] = rows[src]
        desc[i, :ki] = rows
    return desc, torch.tensor(kpi, dtype=torch.int32, device=device)
