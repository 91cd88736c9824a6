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
    conn, start = drv.bow_view_graph(d, k, cb, device="cuda")
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
    out = {"n_img": n_img, "keypoints_per_image": int(d[0].shape[0]), "directed_edges": int(len(edges)),
           "accepted_pairs": len(pairs), "view_graph_s": round(t1 - t0, 3), "gpu_match_all_edges_s": round(t2 - t1, 3),
           "host_walk_cv2_ransac_s": round(t3 - t2, 3)}
    if with_reference:
        from oracle import oracle_driver
        t4 = time.perf_counter()
        rp, rm = oracle_driver.reference_loop(d, p, s, conn, start, oracle_driver.cv2_matcher(), min_matches=100)
        t5 = time.perf_counter()
        out["reference_shaped_loop_cv2_matcher_s"] = round(t5 - t4, 3)
        out["same_pairs"] = [tuple(map(int, x)) for x in rp] == [tuple(map(int, x)) for x in pairs]
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
