"""Loader of tests/golden/stage_ref.npz: inputs and outputs of the reference's own matching.py run
(tests/golden/make_golden_stage.py)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_stage_ref():
    g = np.load(os.path.join(ROOT, "tests", "golden", "stage_ref.npz"), allow_pickle=True)
    n = len(g["names"])
    desc = np.array([None] * n, dtype=object)
    pts = np.array([None] * n, dtype=object)
    size = np.array([None] * n, dtype=object)
    for i in range(n):
        desc[i], pts[i], size[i] = g[f"desc{i}"].astype(np.float32), g[f"pts{i}"], g[f"size{i}"]
    matches = [[g[f"idx0_{p}"], g[f"idx1_{p}"], g[f"ids_{p}"]] for p in range(int(g["n_pairs"]))]
    connection = [list(map(int, c)) for c in g["connection"]]
    return {"desc": desc, "pts": pts, "size": size, "k": int(g["k"]), "codebook": g["codebook"],
            "img_pairs": [tuple(map(int, p)) for p in g["img_pairs"]], "all_matches": matches,
            "connection": connection, "start": int(g["start"])}


def assert_same_stage(ref_pairs, ref_matches, pairs, matches):
    assert [tuple(map(int, p)) for p in pairs] == list(ref_pairs)
    assert len(matches) == len(ref_matches)
    for x, y in zip(ref_matches, matches):
        for u, v in zip(x, y):
            assert np.array_equal(np.asarray(u, dtype=np.int64), np.asarray(v, dtype=np.int64))
