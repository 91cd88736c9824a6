"""CPU restatement of the reference's matching stage around the matcher call -- TEST INFRASTRUCTURE ONLY.

Follows /root/reference/matching.py (which cannot be imported: it runs at import time, reads
output/*.npy and downloads LightGlue weights):
  * view_graph()      matching.py:24-82   BoW vq -> tf-idf -> cosine top-9 -> symmetric adjacency -> start node
  * reference_loop()  matching.py:84-189  BFS over the view graph, one matcher call per visited edge
                      (:122), >8 matches gate (:130), essential-matrix RANSAC + recoverPose (:133-142),
                      >10 inliers (:144), track overlap count (:146-158), >=500 matches / 30 % overlap
                      gate (:160), 3-D point id assignment (:161-172), [idx0, idx1, ids] rows (:174)
The matcher is a parameter (any callable with the LightGlue dict contract, lightglue.py:326-341,
442-450), so the same loop runs with the cv2 oracle matcher or with the product's B200Matcher.
Parity status: PINNED to the reference's own script. tests/golden/make_golden_stage.py runs the unmodified
/root/reference/matching.py under runpy (scratch output/ directory with real SIFT features of twelve of the
reference's photographs, `lightglue` stubbed by the cv2 matcher) and commits what it wrote
(tests/golden/stage_ref.npz); tests/test_driver_cpu.py::test_restated_loop_equals_the_reference_script_itself
checks view_graph() and reference_loop() against those files entry for entry.
"""
from __future__ import annotations

import numpy as np

FOCAL = 2378.98305085          # matching.py:92


def view_graph(all_descriptors, k: int, codebook: np.ndarray, top_k: int = 10, cos_thr: float = 0.75):
    """matching.py:24-82 -> (connection: list of neighbour lists (or None), start image)."""
    from scipy.cluster.vq import vq
    n = len(all_descriptors)
    hist = np.zeros((n, k))
    for i, d in enumerate(all_descriptors):
        words, _ = vq(np.asarray(d).astype("float"), codebook)           # :26-28
        for w in words:                                                     # :30-35
            hist[i, w] += 1
    df = np.sum(hist > 0, axis=0)                                           # :43
    with np.errstate(divide="ignore", invalid="ignore"):
        idf = np.log(n / df)                                                # :45
        tfidf = hist * idf                                                  # :47
    order, score = [], []
    for i in range(n):                                                      # :52-59
        a = tfidf[i]
        with np.errstate(divide="ignore", invalid="ignore"):
            cos = np.dot(a, tfidf.T) / (np.linalg.norm(a) * np.linalg.norm(tfidf, axis=1))
        order.append(np.argsort(-cos)[1:top_k])
        score.append(np.sort(-cos)[1:top_k])
    connection = [None] * n
    for i in range(n):                                                      # :61-73
        for j, other in enumerate(order[i]):
            if not connection[i]:
                connection[i] = []
            if not connection[other]:
                connection[other] = []
            if -score[i][j] > cos_thr:
                if other not in connection[i]:
                    connection[i].append(other)
                if i not in connection[other]:
                    connection[other].append(i)
    best, start = 0, 0
    for i, c in enumerate(connection):                                      # :77-82
        if len(c) > best:
            best, start = len(c), i
    return connection, start


def cv2_matcher(ratio: float = 0.75, mutual: bool = True):
    """A matcher callable with the LightGlue dict contract whose rule is the cv2 one BASELINE.json names."""
    import torch

    from . import oracle_cv2

    def call(data):
        q = data["image0"]["descriptors"][0].cpu().numpy()
        t = data["image1"]["descriptors"][0].cpu().numpy()
        qi, ti, di = oracle_cv2.match(q, t, ratio, mutual)
        m = torch.from_numpy(np.stack([qi, ti], -1).astype(np.int64)).reshape(-1, 2)
        return {"matches": [m], "scores": [torch.from_numpy(di)]}

    return call


def reference_loop(all_descriptors, all_points, img_size, connection, start, matcher, min_matches: int = 500,
                   min_overlap: float = 0.3, focal: float = FOCAL):
    """matching.py:84-189 -> (img_pairs list of (reference_id, id), all_matches list of [idx0, idx1, ids])."""
    import cv2
    import torch
    n = len(all_descriptors)
    next_id = 0
    all_matches = []
    tracks = [None] * n                                                     # all_points3d, :86
    queue = [(start, start)]
    visited = [False] * n
    visited[start] = True
    i = 0
    while True:                                                             # :95
        cur = queue[i][1]
        for nb in connection[cur]:
            if visited[nb]:
                continue
            ref = cur                                                       # :98-105
            for other in connection[nb]:
                if other == cur:
                    break
                if visited[other]:
                    ref = other
                    break
            feats = []
            for im in (ref, nb):                                            # :107-120
                pts = np.asarray(all_points[im], dtype=float).reshape(-1, 2)
                kp = np.stack([pts[:, 0] + img_size[im][0] / 2, -pts[:, 1] + img_size[im][1] / 2], -1)
                feats.append({"keypoints": torch.tensor(kp[None], dtype=torch.float),
                              "descriptors": torch.tensor(np.asarray(all_descriptors[im], dtype=float)[None], dtype=torch.float),
                              "image_size": torch.tensor(np.asarray(img_size[im], dtype=float)[None], dtype=torch.float)})
            res = matcher({"image0": feats[0], "image1": feats[1]})        # :122
            m = res["matches"][0] if isinstance(res["matches"], (list, tuple)) else res["matches"]
            m = m.detach().cpu().numpy().reshape(-1, 2)
            idx0, idx1 = m[:, 0], m[:, 1]                                   # :128
            k0 = np.asarray(all_points[ref])[idx0].astype(np.float32)       # :124-127
            k1 = np.asarray(all_points[nb])[idx1].astype(np.float32)
            if len(k0) <= 8:                                                # :130
                continue
            K = np.array([[focal, 0, 0], [0, focal, 0], [0, 0, 1]])
            E, mask = cv2.findEssentialMat(k0, k1, K, method=cv2.RANSAC, prob=0.999, threshold=1)   # :134
            if mask is None:
                continue
            _, _, _, inl = cv2.recoverPose(E, k0[mask.ravel() > 0], k1[mask.ravel() > 0], K)       # :139
            good = k0[mask.ravel() > 0][inl.ravel() > 0]
            if len(good) <= 10:                                             # :144
                continue
            overlap = 0
            for p1, p2 in zip(idx0, idx1):                                  # :148-158
                if not tracks[ref]:
                    tracks[ref] = [-1] * np.asarray(all_points[ref]).shape[0]
                if not tracks[nb]:
                    tracks[nb] = [-1] * np.asarray(all_points[nb]).shape[0]
                if tracks[ref][p1] == -1 and tracks[nb][p2] == -1:
                    continue
                elif tracks[ref][p1] != -1:
                    overlap += 1
                elif tracks[nb][p1] != -1:                                  # sic: p1 (the reference's own indexing)
                    overlap += 1
            if not (len(idx0) >= min_matches and (cur == start or overlap / len(idx0) >= min_overlap)):   # :160
                continue
            ids = []
            for p1, p2 in zip(idx0, idx1):                                  # :162-172
                if tracks[ref][p1] == -1 and tracks[nb][p2] == -1:
                    tracks[ref][p1] = next_id
                    tracks[nb][p2] = next_id
                    next_id += 1
                elif tracks[ref][p1] != -1:
                    tracks[nb][p2] = tracks[ref][p1]
                elif tracks[nb][p1] != -1:                                  # sic
                    tracks[ref][p2] = tracks[nb][p1]
                ids.append(tracks[ref][p1])
            all_matches.append([idx0, idx1, np.array(ids)])                 # :174
            queue.append((ref, nb))
            visited[nb] = True
        i += 1
        if i >= len(queue):
            break
    return queue[1:], all_matches                                           # :188-189
