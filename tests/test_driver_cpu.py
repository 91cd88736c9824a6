"""CPU: the matching.py-shaped driver (host logic only) against the restated reference loop.
The GPU matcher is replaced by injected oracle matches here; tests/test_gpu_driver.py runs the real one."""
import importlib
import os

import numpy as np
import pytest

from oracle import oracle_cv2, oracle_driver
from tests.scene import make_scene


@pytest.fixture(scope="module")
def scene():
    return make_scene()


def test_bow_view_graph_matches_reference_restatement(pkg, scene):
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    desc, pts, size, k, codebook = scene
    for thr in (0.75, 0.2, 0.0):
        c0, s0 = oracle_driver.view_graph(desc, k, codebook, cos_thr=thr)
        c1, s1 = drv.bow_view_graph(desc, k, codebook, cos_thr=thr)
        assert s0 == s1
        assert [list(map(int, a)) for a in c0] == [list(map(int, a)) for a in c1]


def test_walk_replay_equals_reference_loop(pkg, scene):
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    desc, pts, size, k, codebook = scene
    connection, start = drv.bow_view_graph(desc, k, codebook, cos_thr=0.0)
    assert sum(len(c) for c in connection) > 10
    pairs0, matches0 = oracle_driver.reference_loop(desc, pts, size, connection, start, oracle_driver.cv2_matcher(),
                                                    min_matches=60)
    assert len(pairs0) >= 3, "scene too weak: the walk accepted almost nothing"
    cache = {}
    for a, b in drv.directed_edges(connection).tolist():
        q, t, _ = oracle_cv2.match(desc[a], desc[b], 0.75, True)
        cache[(a, b)] = (q, t)
    pairs1, matches1 = drv.match_view_graph(desc, pts, connection, start, gates=drv.Gates(min_matches=60),
                                            pair_matches=cache)
    assert [tuple(map(int, p)) for p in pairs0] == [tuple(map(int, p)) for p in pairs1]
    assert len(matches0) == len(matches1)
    for m0, m1 in zip(matches0, matches1):
        for a, b in zip(m0, m1):
            assert np.array_equal(np.asarray(a), np.asarray(b))


def test_stage_files_roundtrip(pkg, scene, tmp_path):
    """img_pairs.npy / all_matches.npy in the layout sfm.py:14-19 reads."""
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    desc, pts, size, k, codebook = scene
    out = str(tmp_path)
    np.save(os.path.join(out, "all_descriptors.npy"), desc)
    np.save(os.path.join(out, "all_points.npy"), pts)
    np.save(os.path.join(out, "img_size.npy"), size)
    d2, p2, s2 = drv.load_features(out)
    assert len(d2) == len(desc) and np.array_equal(d2[3], desc[3]) and np.array_equal(p2[1], pts[1])
    pairs = [(0, 1), (1, 2)]
    matches = [[np.arange(5), np.arange(5) + 1, np.arange(5) + 7], [np.arange(3), np.arange(3), np.arange(3)]]
    drv.save_matches(pairs, matches, out)
    img_pairs = np.load(os.path.join(out, "img_pairs.npy"), allow_pickle=True)           # sfm.py:14
    all_matches = np.load(os.path.join(out, "all_matches.npy"), allow_pickle=True)       # sfm.py:17
    assert img_pairs.shape == (2, 2)
    idx0, idx1, ids = all_matches[0]                                                      # sfm.py:102-103
    assert np.array_equal(idx1, np.arange(5) + 1) and np.array_equal(ids, np.arange(5) + 7)
    assert np.array_equal(pts[0][idx0], pts[0][:5])                                       # fancy indexing as sfm.py:110
