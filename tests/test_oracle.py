"""CPU: pin the oracle (oracle/oracle_int.py) against the committed cv2 fixtures and against the
cv2 installed here, before it is trusted as the checker of the CUDA path."""
import os

import numpy as np
import pytest

from oracle import oracle_cv2, oracle_int


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_float32_gemm_is_exact_for_u8():
    rng = np.random.default_rng(0)
    q = rng.integers(0, 256, size=(64, 128)).astype(np.uint8)
    t = rng.integers(0, 256, size=(96, 128)).astype(np.uint8)
    q[0] = 255
    t[0] = 255
    ref = ((q.astype(np.int64)[:, None, :] - t.astype(np.int64)[None, :, :]) ** 2).sum(-1)
    assert np.array_equal(oracle_int.sq_dists_int(q, t), ref)
    assert ref[0, 0] == 0 and ref.max() <= 128 * 255 * 255


@pytest.mark.parametrize("name", ["sift_pair.npz", "ties.npz"])
def test_oracle_knn2_matches_golden(golden_dir, name):
    g = _load(golden_dir, name)
    idx, dist, d2 = oracle_int.knn2(g["q"], g["t"])
    assert np.array_equal(idx, g["knn_idx"])
    assert np.array_equal(dist.view(np.uint32), g["knn_dist"].view(np.uint32))   # bit-exact float32 sqrt


def test_oracle_match_matches_golden(golden_dir):
    g = _load(golden_dir, "sift_pair.npz")
    q, t, d = oracle_int.match(g["q"], g["t"], 0.75, True)
    assert np.array_equal(q, g["m_q"]) and np.array_equal(t, g["m_t"])
    assert np.array_equal(d.view(np.uint32), g["m_d"].view(np.uint32))
    q, t, d = oracle_int.match(g["q"], g["t"], 0.75, False)
    assert np.array_equal(q, g["r_q"]) and np.array_equal(t, g["r_t"])
    g = _load(golden_dir, "ties.npz")
    q, t, d = oracle_int.match(g["q"], g["t"], 0.0, True)
    assert np.array_equal(q, g["m_q"]) and np.array_equal(t, g["m_t"])


@pytest.mark.parametrize("kq,kt,alphabet", [(257, 300, 256), (100, 513, 3), (5, 1, 256), (3, 2, 2), (1, 7, 256)])
def test_oracle_vs_cv2_live(kq, kt, alphabet):
    rng = np.random.default_rng(kq * 1000 + kt)
    q = rng.integers(0, alphabet, size=(kq, 128)).astype(np.uint8) * (255 // max(alphabet - 1, 1))
    t = rng.integers(0, alphabet, size=(kt, 128)).astype(np.uint8) * (255 // max(alphabet - 1, 1))
    if kt > 4:
        t[3] = t[1]           # duplicated train rows: tie on best/second
    idx, dist, _ = oracle_int.knn2(q, t)
    cidx, cdist = oracle_cv2.knn2(q, t)
    assert np.array_equal(idx, cidx)
    assert np.array_equal(dist.view(np.uint32), cdist.view(np.uint32))
    for ratio, mutual in [(0.75, True), (0.75, False), (0.0, True), (0.9, True)]:
        a = oracle_int.match(q, t, ratio, mutual)
        b = oracle_cv2.match(q, t, ratio, mutual)
        for x, y in zip(a, b):
            assert np.array_equal(x, y), (ratio, mutual)


def test_oracle_empty():
    q = np.zeros((0, 128), np.uint8)
    t = np.zeros((4, 128), np.uint8)
    assert oracle_int.match(q, t)[0].size == 0
    assert oracle_int.match(t, q)[0].size == 0
    idx, dist, _ = oracle_int.knn2(t, q)
    assert (idx == -1).all() and np.isinf(dist).all()
