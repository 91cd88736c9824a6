"""CPU: the matching.py-shaped driver (host logic only) against the restated reference loop.
The GPU matcher is replaced by injected oracle matches here; tests/test_gpu_driver.py runs the real one."""
import importlib
import os

import numpy as np
import pytest

from oracle import oracle_cv2, oracle_driver
from tests.scene import make_scene
from tests.stage_fixture import assert_same_stage, load_stage_ref


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


def test_restated_loop_equals_the_reference_script_itself(pkg):
    """tests/golden/stage_ref.npz holds what the UNMODIFIED /root/reference/matching.py wrote (img_pairs.npy,
    all_matches.npy) when run under runpy on real SIFT features of twelve of its own photographs with its
    hard-coded gates (cos > 0.75, > 8, > 10, >= 500, >= 30 %) and the cv2 matcher behind its LightGlue import.
    oracle_driver (view graph + loop) and the product's host walk must reproduce those files entry for entry."""
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    s = load_stage_ref()
    assert len(s["img_pairs"]) >= 5
    conn, start = oracle_driver.view_graph(s["desc"], s["k"], s["codebook"])
    assert start == s["start"] and [list(map(int, c)) if c else [] for c in conn] == s["connection"]
    conn2, start2 = drv.bow_view_graph(s["desc"], s["k"], s["codebook"])
    assert start2 == s["start"] and [list(map(int, c)) if c else [] for c in conn2] == s["connection"]
    pairs, matches = oracle_driver.reference_loop(s["desc"], s["pts"], s["size"], conn, start, oracle_driver.cv2_matcher())
    assert_same_stage(s["img_pairs"], s["all_matches"], pairs, matches)
    # the product's walk on injected cv2 matches (its GPU matcher is exercised by tests/test_gpu_driver.py)
    cache = {}
    for a, b in drv.directed_edges(conn2).tolist():
        q, t, _ = oracle_cv2.match(s["desc"][a], s["desc"][b], 0.75, True)
        cache[(a, b)] = (q, t)
    pairs1, matches1 = drv.match_view_graph(s["desc"], s["pts"], conn2, start2, pair_matches=cache)
    assert_same_stage(s["img_pairs"], s["all_matches"], pairs1, matches1)
    # the gates did something on this scene: some visited edges were refused
    n_edges = len(drv.directed_edges(conn2)) // 2
    assert len(pairs) < n_edges
