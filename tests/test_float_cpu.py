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
