"""CPU: the float-path oracle (numpy restatement of lightglue.py:223-254, 417-426) against fixtures produced
by the reference's own filter_matches (tests/golden/make_golden_float.py)."""
import os

import numpy as np
import pytest

from oracle import oracle_float


def bf16_bits_to_f32(x):
    return (x.astype(np.uint16).astype(np.uint32) << 16).view(np.float32)


@pytest.mark.parametrize("dim", [128, 256])
def test_oracle_float_matches_reference_fixture(golden_dir, dim):
    g = np.load(os.path.join(golden_dir, f"float_mutual_{dim}.npz"))
    d0, d1 = bf16_bits_to_f32(g["d0_bf16"]), bf16_bits_to_f32(g["d1_bf16"])
    assert d0.shape[1] == dim
    sim = oracle_float.similarity(d0, d1)
    for name, thr in (("none", None), ("thr", float(g["threshold"]))):
        m0, m1, _ = oracle_float.filter_matches(sim, thr)
        # numpy's and torch's fp32 GEMMs may round differently: compare through the tolerance the float path states
        q, t = oracle_float.match_list(m0)
        agree, gap = oracle_float.agreement(sim, q, t, g[f"q_{name}"], g[f"t_{name}"])
        assert agree >= 0.995 and gap <= 1e-5, (name, agree, gap)
        assert len(g[f"q_{name}"]) > 100
    # exact duplicate column 5 / 6: the lowest index wins the column arg-max of its rows
    assert not np.any(g["t_none"] == 6) or np.any(g["t_none"] == 5)


def test_oracle_float_small_cases():
    d0 = np.eye(4, 128, dtype=np.float32)
    d1 = np.eye(6, 128, dtype=np.float32)[[2, 0, 1, 5, 4, 3]]
    q, t, s = oracle_float.match(d0, d1)
    assert np.array_equal(q, np.arange(4)) and np.array_equal(t, [1, 2, 0, 5]) and np.allclose(s, 1)
    q, t, s = oracle_float.match(d0, d1, threshold=1.5)
    assert q.size == 0
    q, t, s = oracle_float.match(np.zeros((0, 128), np.float32), d1)
    assert q.size == 0


@pytest.mark.parametrize("dim", [128, 256])
def test_oracle_assignment_head_matches_reference_fixture(golden_dir, dim):
    """numpy restatement of sigmoid_log_double_softmax + filter_matches (lightglue.py:196-254) against the
    outputs of the reference's own functions (tests/golden/make_golden_assign.py)."""
    g = np.load(os.path.join(golden_dir, f"assign_{dim}.npz"))
    d0, d1 = bf16_bits_to_f32(g["d0_bf16"]), bf16_bits_to_f32(g["d1_bf16"])
    scores = oracle_float.log_double_softmax(oracle_float.similarity(d0, d1), g["z0"], g["z1"])
    assert np.allclose(scores[7], g["scores_row7"], atol=2e-4)
    for name, th in (("none", None), ("thr", float(g["threshold"]))):
        m0, m1, ms0, ms1 = oracle_float.filter_log_matches(scores, th)
        # fp32 (reference) vs fp64 (restatement) may flip near-ties and rows sitting on the threshold
        assert np.mean(m0 == g[f"m0_{name}"]) >= 0.995 and np.mean(m1 == g[f"m1_{name}"]) >= 0.995
        same = m0 == g[f"m0_{name}"]
        assert np.allclose(ms0[same], g[f"ms0_{name}"][same], atol=1e-4)
        assert np.allclose(ms1, g[f"ms1_{name}"], atol=1e-4) or np.mean(np.isclose(ms1, g[f"ms1_{name}"], atol=1e-4)) >= 0.995
    assert (g["m0_thr"] > -1).sum() > 100 and (g["m0_none"] > -1).sum() > (g["m0_thr"] > -1).sum()
