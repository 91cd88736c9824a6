"""Batched geometric verification on the GPU: what matching.py:133-142 does once per visited edge with
cv2.findEssentialMat(k0, k1, K, method=cv2.RANSAC, prob=0.999, threshold=1) + cv2.recoverPose(E, k0[mask],
k1[mask], K), for every pair of a batch in one call (b200m_essential_ransac_f32, csrc/match_ransac.cuh).

Keypoints are in the reference's convention: centre-origin pixels (feature_extraction.py:34) and
K = diag(focal, focal, 1) (matching.py:133), so normalised coordinates are pixel / focal.
RANSAC is randomised, so parity with cv2 is statistical (tests/test_gpu_geometry.py); results are
deterministic for a given seed.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import _lib


@dataclass
class PoseBatch:
    """Per pair: E [P,3,3], inlier counts [P], R [P,3,3], t [P,3], pose-inlier counts [P]; per match: the
    RANSAC inlier mask and recoverPose's mask (flat, `offsets` delimits the pairs)."""
    E: np.ndarray
    inliers: np.ndarray
    R: np.ndarray
    t: np.ndarray
    pose_inliers: np.ndarray
    mask: np.ndarray
    pose_mask: np.ndarray
    offsets: np.ndarray

    def masks(self, i: int):
        a, b = int(self.offsets[i]), int(self.offsets[i + 1])
        return self.mask[a:b].astype(bool), self.pose_mask[a:b].astype(bool)


def essential_ransac(points0, points1, focal: float, threshold: float = 1.0, iters: int = 1024, seed: int = 0,
                     max_depth: float = 50.0, device="cuda") -> PoseBatch:
    """points0[i], points1[i]: float [S_i, 2] matched keypoints of pair i (centre-origin pixels).
    threshold: Sampson threshold in pixels (cv2's `threshold`); iters: hypotheses per pair."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("essential_ransac runs on a CUDA device (there is no CPU path)")
    P = len(points0)
    if P != len(points1):
        raise ValueError("points0 and points1 must list the same pairs")
    sizes = [int(np.asarray(a).shape[0]) for a in points0]
    offsets = np.zeros(P + 1, np.int64)
    np.cumsum(sizes, out=offsets[1:])
    total = int(offsets[-1])
    xy = np.zeros((max(total, 1), 4), np.float32)
    for i, (a, b) in enumerate(zip(points0, points1)):
        a = np.asarray(a, np.float64).reshape(-1, 2)
        b = np.asarray(b, np.float64).reshape(-1, 2)
        if a.shape != b.shape:
            raise ValueError(f"pair {i}: {a.shape[0]} vs {b.shape[0]} points")
        xy[offsets[i]:offsets[i + 1], :2] = a / focal
        xy[offsets[i]:offsets[i + 1], 2:] = b / focal
    out_E = torch.zeros((max(P, 1), 9), dtype=torch.float64, device=dev)
    out_R = torch.zeros((max(P, 1), 9), dtype=torch.float64, device=dev)
    out_t = torch.zeros((max(P, 1), 3), dtype=torch.float64, device=dev)
    out_in = torch.zeros((max(P, 1),), dtype=torch.int32, device=dev)
    out_pin = torch.zeros((max(P, 1),), dtype=torch.int32, device=dev)
    mask = torch.zeros((max(total, 1),), dtype=torch.uint8, device=dev)
    pmask = torch.zeros((max(total, 1),), dtype=torch.uint8, device=dev)
    if P:
        L = _lib.lib()
        xy_d = torch.from_numpy(xy).to(dev)
        off_d = torch.from_numpy(offsets).to(dev)
        ws = torch.empty(int(L.b200m_workspace_bytes_ransac(P, iters)), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.check(L.b200m_essential_ransac_f32(
                xy_d.data_ptr(), off_d.data_ptr(), P, int(iters), float(threshold / focal), int(seed) & (2 ** 64 - 1),
                float(max_depth), out_E.data_ptr(), mask.data_ptr(), out_in.data_ptr(), out_R.data_ptr(),
                out_t.data_ptr(), pmask.data_ptr(), out_pin.data_ptr(), ws.data_ptr(), ws.numel(),
                torch.cuda.current_stream().cuda_stream))
    return PoseBatch(out_E.cpu().numpy()[:P].reshape(P, 3, 3), out_in.cpu().numpy()[:P],
                     out_R.cpu().numpy()[:P].reshape(P, 3, 3), out_t.cpu().numpy()[:P], out_pin.cpu().numpy()[:P],
                     mask.cpu().numpy()[:total], pmask.cpu().numpy()[:total], offsets)


def five_point(x0, x1, device="cuda") -> np.ndarray:
    """One minimal solve: x0, x1 float [5, 2] normalised points -> [n, 3, 3] candidate essential matrices
    (n <= 10, unit Frobenius norm). Test hook for the solver inside the RANSAC kernel."""
    dev = torch.device(device)
    a = torch.as_tensor(np.ascontiguousarray(x0, np.float64).reshape(5, 2)).to(dev)
    b = torch.as_tensor(np.ascontiguousarray(x1, np.float64).reshape(5, 2)).to(dev)
    out = torch.zeros((10, 9), dtype=torch.float64, device=dev)
    n = torch.zeros((1,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().b200m_five_point_f64(a.data_ptr(), b.data_ptr(), out.data_ptr(), n.data_ptr(),
                                                   torch.cuda.current_stream().cuda_stream))
    return out.cpu().numpy()[: int(n.item())].reshape(-1, 3, 3)
