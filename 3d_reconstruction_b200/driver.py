"""matching.py-shaped driver: the reference's matching stage with the pair matching batched on the GPU.

The reference walks its view graph breadth-first and calls the matcher once per visited edge
(matching.py:94-128), re-packing both images' features on the host every time (:107-120). Which
edge is matched next depends on earlier results, but every request is a directed edge
(reference image -> new image) of the symmetric view graph built at matching.py:61-73. So this driver
  1. packs all descriptors once into a device-resident bank (bank.py),
  2. matches ALL directed edges of the graph in one batched call (matcher.match_pairs),
  3. replays the reference's walk on the host against the cached match lists: same gates
     (matching.py:130, :144, :160), same essential-matrix verification (cv2, :133-142), same
     3-D point id assignment (:146-172),
and writes the stage files the next script reads: output/img_pairs.npy and output/all_matches.npy
(matching.py:188-189 -> sfm.py:14-19), so `python sfm.py` runs on this output unchanged.
Feature files are read in the layout feature_extraction.py:41-51 writes.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np

from .bank import DescriptorBank, FloatBank
from .matcher import MatchConfig, match_pairs, match_pairs_float

FOCAL_LENGTH = 2378.98305085      # matching.py:92 / sfm.py:24


@dataclass
class Gates:
    """The reference's acceptance constants (matching.py:130, :144, :160)."""
    min_matches_for_pose: int = 8      # "> 8" matches before findEssentialMat
    min_pose_inliers: int = 10         # "> 10" recoverPose inliers
    min_matches: int = 500             # ">= 500" raw matches to accept an edge
    min_overlap: float = 0.3           # ">= 30 %" of matches already carrying a 3-D point id (except around the start image)
    strict_indexing: bool = False      # True: raise like the reference when its [id][p1] indexing runs out of range


# ---------------------------------------------------------------------- stage files
def load_features(output_dir: str = "output"):
    """-> (all_descriptors, all_points, img_size) object arrays as feature_extraction.py:41-51 saves them."""
    d = np.load(os.path.join(output_dir, "all_descriptors.npy"), allow_pickle=True)
    p = np.load(os.path.join(output_dir, "all_points.npy"), allow_pickle=True)
    s = np.load(os.path.join(output_dir, "img_size.npy"), allow_pickle=True)
    return d, p, s


def load_codebook(output_dir: str = "output"):
    """(k, codebook) as bow.py:25 dumps it."""
    import joblib
    return joblib.load(os.path.join(output_dir, "bow_codebook.plk"))


def save_matches(img_pairs, all_matches, output_dir: str = "output") -> None:
    """output/img_pairs.npy + output/all_matches.npy exactly as matching.py:188-189 (-> sfm.py:14,17)."""
    os.makedirs(output_dir, exist_ok=True)
    np.save(os.path.join(output_dir, "img_pairs.npy"), img_pairs)
    np.save(os.path.join(output_dir, "all_matches.npy"), np.array(all_matches, dtype=object))


# ---------------------------------------------------------------------- pair scheduler (matching.py:24-82)
def vq_words(all_descriptors, codebook: np.ndarray, device="cuda"):
    """Visual word of every descriptor of every image (matching.py:26-28) in ONE GPU call:
    all images' descriptors are stacked, b200m_vq_f32 assigns words, the result is split per image."""
    import torch

    from . import _lib
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("vq_words runs on a CUDA device (there is no CPU path)")
    sizes = [int(np.asarray(d).shape[0]) for d in all_descriptors]
    stack = np.ascontiguousarray(np.concatenate([np.asarray(d, dtype=np.float32).reshape(s, -1)
                                                 for d, s in zip(all_descriptors, sizes)], 0))
    obs = torch.from_numpy(stack).to(dev)
    cb = torch.from_numpy(np.ascontiguousarray(codebook, dtype=np.float64)).to(dev)
    out = torch.empty((obs.shape[0],), dtype=torch.int32, device=dev)
    scratch = torch.empty((cb.shape[0],), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().b200m_vq_f32(obs.data_ptr(), obs.shape[0], obs.shape[1], cb.data_ptr(), cb.shape[0],
                                           out.data_ptr(), None, scratch.data_ptr(),
                                           torch.cuda.current_stream().cuda_stream))
    words = out.cpu().numpy()
    return np.split(words, np.cumsum(sizes)[:-1])


def bow_view_graph(all_descriptors, k: int, codebook: np.ndarray, top_k: int = 10, cos_thr: float = 0.75,
                   device=None):
    """BoW view graph of the reference: visual words, word histograms, idf, cosine similarity, each
    image's 9 best neighbours kept when cosine > 0.75, symmetric adjacency in insertion order,
    start = first image of maximum degree. device=None: words from scipy's vq on the host, exactly
    as the reference calls it; device="cuda": words from the GPU kernel (vq_words)."""
    n = len(all_descriptors)
    hist = np.zeros((n, k))
    if device is not None:
        for i, words in enumerate(vq_words(all_descriptors, codebook, device)):
            hist[i] = np.bincount(words, minlength=k)[:k]
    else:
        from scipy.cluster.vq import vq
        for i, d in enumerate(all_descriptors):
            words, _ = vq(np.asarray(d).astype("float"), codebook)
            hist[i] = np.bincount(words, minlength=k)[:k]
    df = np.sum(hist > 0, axis=0)
    with np.errstate(divide="ignore", invalid="ignore"):
        tfidf = hist * np.log(n / df)
        norms = np.linalg.norm(tfidf, axis=1)
    connection = [None] * n
    ranked = []
    for i in range(n):
        with np.errstate(divide="ignore", invalid="ignore"):
            cos = np.dot(tfidf[i], tfidf.T) / (norms[i] * norms)
        ranked.append((np.argsort(-cos)[1:top_k], np.sort(-cos)[1:top_k]))
    for i, (order, score) in enumerate(ranked):
        for other, s in zip(order, score):
            if not connection[i]:
                connection[i] = []
            if not connection[other]:
                connection[other] = []
            if -s > cos_thr:
                if other not in connection[i]:
                    connection[i].append(other)
                if i not in connection[other]:
                    connection[other].append(i)
    degree = [len(c) if c else 0 for c in connection]
    start = int(np.argmax(degree)) if n and max(degree) > 0 else 0      # first image of maximum degree
    return connection, start


def directed_edges(connection) -> np.ndarray:
    """Every (reference image, new image) request the walk can make: both directions of every edge."""
    out = [(a, b) for a, nbrs in enumerate(connection) if nbrs for b in nbrs]
    return np.asarray(sorted(set(out)), dtype=np.int32).reshape(-1, 2)


# ---------------------------------------------------------------------- batched matching of the graph's edges
def match_edges(all_descriptors, edges: np.ndarray, conf: MatchConfig, device="cuda") -> dict:
    """{(a, b): (idx0, idx1)} for every directed edge, ONE batched GPU call, dispatched on conf.metric:
      "l2_u8"   the exact integer rule (2-NN, Lowe ratio conf.ratio, cross-check conf.mutual) on SIFT-valued
                descriptors -- integer values in [0, 255], verified while packing (bank.to_u8_descriptors);
      "ip_bf16" mutual arg-max of the inner product with conf.score_threshold on float descriptors: what the
                reference's own stage file holds (DISK, 128-d float32, feature_extraction.py:10,28,50) and the
                raw-similarity form of its matcher head (lightglue.py:223,235-254).
    Anything else raises (no setting is silently ignored)."""
    if conf.metric not in ("l2_u8", "ip_bf16"):
        raise NotImplementedError(f"unknown metric {conf.metric!r} (l2_u8 or ip_bf16)")
    if conf.metric == "l2_u8" and conf.k != 2:
        raise NotImplementedError("the l2_u8 rule is knnMatch k=2 + ratio + cross-check; other k are not implemented")
    if conf.metric == "l2_u8" and conf.score_threshold is not None:
        raise ValueError("score_threshold belongs to metric 'ip_bf16'; the l2_u8 rule filters with `ratio`")
    out = {}
    if not len(edges):
        return out
    if conf.metric == "ip_bf16":
        bank = FloatBank.from_arrays([np.asarray(d, dtype=np.float32) for d in all_descriptors], device=device)
        res = match_pairs_float(bank, edges, conf.score_threshold, conf.mutual)
    else:
        bank = DescriptorBank.from_arrays(list(all_descriptors), device=device)
        res = match_pairs(bank, edges, ratio=conf.ratio, mutual=conf.mutual)
    for (a, b), (q, t, _) in zip(np.asarray(edges).tolist(), res.to_lists()):
        out[(a, b)] = (q, t)
    return out


# ---------------------------------------------------------------------- the walk
def verify_edges_gpu(all_points, pair_matches, gates: "Gates", focal: float = FOCAL_LENGTH, device="cuda",
                     iters: int = 1024, seed: int = 0) -> dict:
    """Pose-inlier count of every cached edge in ONE batched call (geometry.essential_ransac): what the
    reference computes per visited edge with cv2.findEssentialMat + cv2.recoverPose (matching.py:133-142).
    Edges the reference skips before the geometry (`len(kpts0) <= 8`, matching.py:130) are left out."""
    from .geometry import essential_ransac
    keys = [k for k, (q, _) in pair_matches.items() if len(q) > gates.min_matches_for_pose]
    if not keys:
        return {}
    p0 = [np.asarray(all_points[a])[np.asarray(pair_matches[(a, b)][0], dtype=np.int64)] for a, b in keys]
    p1 = [np.asarray(all_points[b])[np.asarray(pair_matches[(a, b)][1], dtype=np.int64)] for a, b in keys]
    res = essential_ransac(p0, p1, focal, threshold=1.0, iters=iters, seed=seed, device=device)
    return {k: int(c) for k, c in zip(keys, res.pose_inliers)}


def match_view_graph(all_descriptors, all_points, connection, start: int, conf: MatchConfig | None = None,
                     gates: Gates | None = None, focal: float = FOCAL_LENGTH, device="cuda", pair_matches=None,
                     geometry: str = "cv2"):
    """-> (img_pairs [(reference_id, id), ...], all_matches [[idx0, idx1, point3d_ids], ...]).

    pair_matches: optional {(a, b): (idx0, idx1)} to replay the walk on precomputed matches
    (tests inject an oracle here); by default all directed edges are matched on the GPU in one call.
    geometry: "cv2" = the reference's own calls per visited edge (matching.py:133-142); "gpu" = every edge
    verified up front in one batched launch (verify_edges_gpu), the walk reads the cached inlier counts."""
    import cv2
    if geometry not in ("cv2", "gpu"):
        raise ValueError("geometry must be 'cv2' or 'gpu'")
    conf = conf or MatchConfig()
    gates = gates or Gates()
    n = len(all_descriptors)
    if pair_matches is None:
        pair_matches = match_edges(all_descriptors, directed_edges(connection), conf, device)

    K = np.array([[focal, 0, 0], [0, focal, 0], [0, 0, 1]])
    pose_inliers = verify_edges_gpu(all_points, pair_matches, gates, focal, device) if geometry == "gpu" else None
    next_id = 0
    tracks = [None] * n                 # per image: 3-D point id of every keypoint, -1 = none
    all_matches = []
    queue = [(start, start)]
    visited = [False] * n
    visited[start] = True

    def track_of(img, p):
        t = tracks[img]
        if p >= len(t):
            if gates.strict_indexing:
                raise IndexError("list index out of range")     # what the reference's [id][p1] does
            return -1
        return t[p]

    i = 0
    while i < len(queue):
        cur = queue[i][1]
        for nb in (connection[cur] or []):
            if visited[nb]:
                continue
            # reference image: the first already-visited neighbour of nb listed before cur, else cur
            ref = cur
            for other in connection[nb]:
                if other == cur:
                    break
                if visited[other]:
                    ref = other
                    break
            idx0, idx1 = pair_matches[(ref, nb)]
            idx0 = np.asarray(idx0, dtype=np.int64)
            idx1 = np.asarray(idx1, dtype=np.int64)
            if len(idx0) <= gates.min_matches_for_pose:
                continue
            if pose_inliers is not None:
                if pose_inliers.get((ref, nb), 0) <= gates.min_pose_inliers:
                    continue
            else:
                k0 = np.asarray(all_points[ref])[idx0].astype(np.float32)
                k1 = np.asarray(all_points[nb])[idx1].astype(np.float32)
                E, mask = cv2.findEssentialMat(k0, k1, K, method=cv2.RANSAC, prob=0.999, threshold=1)
                if mask is None:
                    continue
                keep = mask.ravel() > 0
                _, _, _, inl = cv2.recoverPose(E, k0[keep], k1[keep], K)
                if int((inl.ravel() > 0).sum()) <= gates.min_pose_inliers:
                    continue
            if not tracks[ref]:
                tracks[ref] = [-1] * np.asarray(all_points[ref]).shape[0]
            if not tracks[nb]:
                tracks[nb] = [-1] * np.asarray(all_points[nb]).shape[0]
            tr, tn = tracks[ref], tracks[nb]
            overlap = 0
            for p1, p2 in zip(idx0.tolist(), idx1.tolist()):
                if tr[p1] == -1 and tn[p2] == -1:
                    continue
                # the reference tests [reference][p1] and then [id][p1] (sic), never [id][p2]
                if tr[p1] != -1 or track_of(nb, p1) != -1:
                    overlap += 1
            if not (len(idx0) >= gates.min_matches and (cur == start or overlap / len(idx0) >= gates.min_overlap)):
                continue
            ids = []
            for p1, p2 in zip(idx0.tolist(), idx1.tolist()):
                if tr[p1] == -1 and tn[p2] == -1:
                    tr[p1] = tn[p2] = next_id
                    next_id += 1
                elif tr[p1] != -1:
                    tn[p2] = tr[p1]
                elif track_of(nb, p1) != -1:
                    # sic: the reference writes [reference][p2] = [id][p1]
                    if p2 < len(tr):
                        tr[p2] = tn[p1]
                    elif gates.strict_indexing:
                        raise IndexError("list assignment index out of range")
                ids.append(tr[p1])
            all_matches.append([idx0, idx1, np.array(ids)])
            queue.append((ref, nb))
            visited[nb] = True
        i += 1
    return queue[1:], all_matches


def run_matching_stage(output_dir: str = "output", conf: MatchConfig | None = None, gates: Gates | None = None,
                       device="cuda", geometry: str = "cv2"):
    """The whole stage: stage files in -> img_pairs.npy / all_matches.npy out (matching.py as a function)."""
    all_descriptors, all_points, _ = load_features(output_dir)
    k, codebook = load_codebook(output_dir)
    connection, start = bow_view_graph(all_descriptors, k, codebook, device=device)
    pairs, matches = match_view_graph(all_descriptors, all_points, connection, start, conf, gates, device=device,
                                      geometry=geometry)
    save_matches(pairs, matches, output_dir)
    return pairs, matches
