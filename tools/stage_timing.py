"""Where the matching stage (matching.py as a function) spends its time on a synthetic scene: BoW view
graph, batched GPU matching of every directed edge, host replay of the reference's walk (cv2 RANSAC per
visited edge) -- and the reference-shaped loop (one cv2 matcher call per visited edge) beside it."""
import importlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3d_reconstruction_b200")
drv = importlib.import_module("3d_reconstruction_b200.driver")
from tests.scene import make_scene  # noqa: E402


def main(n_img=24, k_img=4000, with_reference=True):
    d, p, s, k, cb = make_scene(n_img=n_img, n_world=int(k_img * 2.5), k_img=k_img, seed=3, codebook_k=800)
    pkg.match(d[0], d[1])                       # warm up the library
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    conn, start = drv.bow_view_graph(d, k, cb, device="cuda", cos_thr=0.0)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    edges = drv.directed_edges(conn)
    bank = pkg.DescriptorBank.from_arrays(list(d), device="cuda")
    res = pkg.match_pairs(bank, edges)
    lists = res.to_lists()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    pm = {(a, b): (q, t) for (a, b), (q, t, _) in zip(edges.tolist(), lists)}
    pairs, matches = drv.match_view_graph(d, p, conn, start, gates=drv.Gates(min_matches=100), pair_matches=pm)
    t3 = time.perf_counter()
    pairs_g, matches_g = drv.match_view_graph(d, p, conn, start, gates=drv.Gates(min_matches=100), pair_matches=pm,
                                              geometry="gpu")
    torch.cuda.synchronize()
    t3b = time.perf_counter()
    out = {"n_img": n_img, "keypoints_per_image": int(d[0].shape[0]), "directed_edges": int(len(edges)),
           "accepted_pairs": len(pairs), "view_graph_s": round(t1 - t0, 3), "gpu_match_all_edges_s": round(t2 - t1, 3),
           "host_walk_cv2_ransac_s": round(t3 - t2, 3), "host_walk_gpu_ransac_s": round(t3b - t3, 3),
           "gpu_geometry_same_pairs": [tuple(map(int, x)) for x in pairs_g] == [tuple(map(int, x)) for x in pairs]}
    if with_reference:
        from oracle import oracle_driver
        t4 = time.perf_counter()
        rp, rm = oracle_driver.reference_loop(d, p, s, conn, start, oracle_driver.cv2_matcher(), min_matches=100)
        t5 = time.perf_counter()
        out["reference_shaped_loop_cv2_matcher_s"] = round(t5 - t4, 3)
        out["same_pairs"] = [tuple(map(int, x)) for x in rp] == [tuple(map(int, x)) for x in pairs]
    print(json.dumps(out), flush=True)


def ransac_rate(n_pairs=512, s=2000, iters=1024):
    """Throughput of the batched RANSAC alone: n_pairs pairs of s matches (30 % outliers)."""
    from oracle import oracle_geometry as og
    geo = importlib.import_module("3d_reconstruction_b200.geometry")
    k0, k1, good, *_ = og.two_view_scene(n=s, outliers=0.3, seed=1)
    p0, p1 = [k0] * n_pairs, [k1] * n_pairs
    geo.essential_ransac(p0[:2], p1[:2], 2378.98305085, iters=iters)
    pkg.set_option("profile", 1)
    pkg.profile_read()
    t0 = time.perf_counter()
    res = geo.essential_ransac(p0, p1, 2378.98305085, iters=iters, seed=5)
    t1 = time.perf_counter()
    prof = pkg.profile_read()
    pkg.set_option("profile", 0)
    import cv2
    t2 = time.perf_counter()
    for _ in range(4):
        og.verify_pair(k0, k1, 2378.98305085)
    t3 = time.perf_counter()
    print(json.dumps({"ransac_pairs": n_pairs, "matches_per_pair": s, "hypotheses": iters,
                      "call_s_with_host_staging": round(t1 - t0, 4), "ransac_kernel_ms": round(prof["first_sweep_ms"], 3),
                      "pose_kernel_ms": round(prof["emit_ms"], 3),
                      "pairs_per_s_kernels": round(n_pairs / ((prof["first_sweep_ms"] + prof["emit_ms"]) / 1e3), 1),
                      "mean_inliers": float(res.inliers.mean()), "planted": int(good.sum()),
                      "cv2_ms_per_pair": round((t3 - t2) / 4 * 1e3, 2), "cv2_threads": cv2.getNumThreads()}), flush=True)


if __name__ == "__main__":
    main()
    ransac_rate()
