"""The reference's geometric verification calls -- TEST INFRASTRUCTURE ONLY.

matching.py:133-142 verbatim in function form: cv2.findEssentialMat(k0, k1, K, method=cv2.RANSAC, prob=0.999,
threshold=1) and cv2.recoverPose(E, k0[mask], k1[mask], K) with K = diag(focal, focal, 1). The arithmetic lives
in third-party opencv-python (requirement.txt:4, unpinned; 4.13.0 in this image), so the cv2 calls themselves
are the oracle; cv2's RANSAC is randomised, parity is statistical.
"""
from __future__ import annotations

import numpy as np


def intrinsics(focal: float) -> np.ndarray:
    return np.array([[focal, 0, 0], [0, focal, 0], [0, 0, 1]], dtype=np.float64)       # matching.py:133


def verify_pair(k0, k1, focal: float, threshold: float = 1.0):
    """-> (E, mask bool [S], R, t, pose_mask bool [S]) or None where the reference `continue`s (matching.py:130-137)."""
    import cv2
    k0 = np.asarray(k0, np.float32).reshape(-1, 2)
    k1 = np.asarray(k1, np.float32).reshape(-1, 2)
    if len(k0) <= 8:                                                                    # :130
        return None
    K = intrinsics(focal)
    cv2.setRNGSeed(0)           # cv2's RANSAC draws from the global RNG: make the oracle independent of test order
    E, mask = cv2.findEssentialMat(k0, k1, K, method=cv2.RANSAC, prob=0.999, threshold=threshold)   # :134
    if mask is None or E is None or E.shape != (3, 3):
        return None
    keep = mask.ravel() > 0
    _, R, t, inl = cv2.recoverPose(E, k0[keep], k1[keep], K)                            # :139
    pose = np.zeros(len(k0), bool)
    pose[np.nonzero(keep)[0][inl.ravel() > 0]] = True
    return E, keep, R, t.ravel(), pose


def five_point_solutions(x0, x1):
    """cv2's own minimal solver: findEssentialMat on exactly five normalised points returns every real
    solution stacked as [3 n, 3]."""
    import cv2
    E, _ = cv2.findEssentialMat(np.asarray(x0, np.float64).reshape(5, 2), np.asarray(x1, np.float64).reshape(5, 2),
                                np.eye(3), method=cv2.RANSAC, prob=0.999, threshold=1e-3)
    if E is None:
        return np.zeros((0, 3, 3))
    E = np.asarray(E).reshape(-1, 3, 3)
    return E / np.linalg.norm(E.reshape(len(E), -1), axis=1)[:, None, None]


def two_view_scene(n=800, outliers=0.3, noise_px=0.4, focal=2378.98305085, seed=0, baseline=1.0, yaw=0.15):
    """Synthetic two-view geometry in the reference's conventions: centre-origin pixels, K = diag(f, f, 1).
    -> (k0, k1, is_inlier, R_true, t_true (unit), E_true (unit Frobenius))"""
    rng = np.random.default_rng(seed)
    X = rng.uniform([-3, -2, 6], [3, 2, 12], size=(n, 3))
    c, s = np.cos(yaw), np.sin(yaw)
    R = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]]) @ np.array([[1, 0, 0], [0, np.cos(0.05), -np.sin(0.05)], [0, np.sin(0.05), np.cos(0.05)]])
    t = np.array([-baseline, 0.1, 0.2])
    X1 = X @ R.T + t
    k0 = focal * X[:, :2] / X[:, 2:3] + rng.normal(0, noise_px, (n, 2))
    k1 = focal * X1[:, :2] / X1[:, 2:3] + rng.normal(0, noise_px, (n, 2))
    good = np.ones(n, bool)
    bad = rng.permutation(n)[: int(outliers * n)]
    k1[bad] = rng.uniform(-600, 600, (len(bad), 2))
    good[bad] = False
    tn = t / np.linalg.norm(t)
    tx = np.array([[0, -tn[2], tn[1]], [tn[2], 0, -tn[0]], [-tn[1], tn[0], 0]])
    E = tx @ R
    return k0, k1, good, R, tn, E / np.linalg.norm(E)
