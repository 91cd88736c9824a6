"""The cv2 calls BASELINE.json names, as arrays -- TEST INFRASTRUCTURE / TIMED CPU BASELINE ONLY.

cv2.BFMatcher(cv2.NORM_L2).knnMatch(k=2) + Lowe ratio + cross-check. This is the authoritative
oracle (third-party opencv-python, requirement.txt:4 of the reference, unpinned; 4.13.0 here) and
the CPU implementation bench.py times as the reference arm. It feeds the same slots the reference
fills at matching.py:123-128 (idx0, idx1).
"""
from __future__ import annotations

import numpy as np


def _cv2():
    import cv2
    return cv2


def knn2(q: np.ndarray, t: np.ndarray):
    """knnMatch(q, t, k=2) -> (idx int32 [Kq,2], dist float32 [Kq,2]); -1 / inf where cv2 returns fewer."""
    cv2 = _cv2()
    q32 = np.ascontiguousarray(q, dtype=np.float32)
    t32 = np.ascontiguousarray(t, dtype=np.float32)
    kq = q32.shape[0]
    idx = np.full((kq, 2), -1, np.int32)
    dist = np.full((kq, 2), np.inf, np.float32)
    if kq == 0 or t32.shape[0] == 0:
        return idx, dist
    knn = cv2.BFMatcher(cv2.NORM_L2).knnMatch(q32, t32, k=2)
    for r, row in enumerate(knn):
        for j, m in enumerate(row[:2]):
            idx[r, j] = m.trainIdx
            dist[r, j] = np.float32(m.distance)
    return idx, dist


def match(q: np.ndarray, t: np.ndarray, ratio: float = 0.75, mutual: bool = True):
    """The rule exactly as a cv2 user writes it."""
    cv2 = _cv2()
    q32 = np.ascontiguousarray(q, dtype=np.float32)
    t32 = np.ascontiguousarray(t, dtype=np.float32)
    if q32.shape[0] == 0 or t32.shape[0] == 0:
        z = np.zeros(0, np.int32)
        return z, z.copy(), np.zeros(0, np.float32)
    bf = cv2.BFMatcher(cv2.NORM_L2)
    knn = bf.knnMatch(q32, t32, k=2)
    if ratio is not None and ratio > 0:
        good = [row[0] for row in knn if len(row) == 2 and row[0].distance < ratio * row[1].distance]
    else:
        good = [row[0] for row in knn if len(row) >= 1]
    if mutual and good:
        back = bf.knnMatch(t32, q32, k=1)           # nearest query of every train row
        nearest_q = np.array([b[0].trainIdx if b else -1 for b in back], np.int64)
        good = [m for m in good if nearest_q[m.trainIdx] == m.queryIdx]
    qi = np.array([m.queryIdx for m in good], np.int32)
    ti = np.array([m.trainIdx for m in good], np.int32)
    di = np.array([m.distance for m in good], np.float32)
    return qi, ti, di


def time_knn2(q: np.ndarray, t: np.ndarray, repeats: int = 1, full_rule: bool = True, threads: int | None = None,
              as_u8: bool = False):
    """Seconds per pair for the cv2 path (float32 inputs: cv2's fastest route for this data; as_u8: the descriptors
    as uint8, the ~4x slower route SURVEY.md section 8(d) asks to report beside it)."""
    import os
    import time
    cv2 = _cv2()
    cv2.setNumThreads(threads if threads else (os.cpu_count() or 1))
    q32 = np.ascontiguousarray(q, dtype=np.uint8 if as_u8 else np.float32)
    t32 = np.ascontiguousarray(t, dtype=np.uint8 if as_u8 else np.float32)
    bf = cv2.BFMatcher(cv2.NORM_L2)
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        knn = bf.knnMatch(q32, t32, k=2)
        if full_rule:
            good = [row[0] for row in knn if len(row) == 2 and row[0].distance < 0.75 * row[1].distance]
            back = bf.knnMatch(t32, q32, k=1)
            nearest_q = [b[0].trainIdx for b in back]
            good = [m for m in good if nearest_q[m.trainIdx] == m.queryIdx]
        best = min(best, time.perf_counter() - t0)
    return best, cv2.getNumThreads()
