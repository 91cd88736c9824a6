import importlib

import numpy as np
import pytest


def test_exhaustive_and_window_counts(pkg):
    assert len(pkg.exhaustive_pairs(64)) == 2016
    assert len(pkg.exhaustive_pairs(257)) == 32896
    w = pkg.window_pairs(2000, 32)
    assert len(w) == 2000 * 32 - 32 * 33 // 2 == 63472
    assert (w[:, 1] > w[:, 0]).all() and (w[:, 1] - w[:, 0] <= 32).all()


def test_blocked_order_is_a_permutation(pkg):
    p = pkg.exhaustive_pairs(40)
    b = pkg.blocked_order(p, 16)
    assert sorted(map(tuple, b.tolist())) == sorted(map(tuple, p.tolist()))


def test_partition_covers_everything_once(pkg):
    p = pkg.blocked_order(pkg.exhaustive_pairs(37))
    k = np.random.default_rng(0).integers(100, 5000, size=37)
    for ws in (1, 2, 3, 8):
        seen = []
        costs = []
        for r in range(ws):
            pr, idx = pkg.partition(p, ws, r, k)
            assert np.array_equal(p[idx], pr)
            seen.append(idx)
            costs.append((k[pr[:, 0]].astype(np.int64) * k[pr[:, 1]]).sum())
        allidx = np.concatenate(seen)
        assert np.array_equal(np.sort(allidx), np.arange(len(p)))
        assert np.array_equal(allidx, np.arange(len(p)))          # contiguous, in rank order
        if ws > 1:
            assert max(costs) <= 1.25 * (sum(costs) / ws)


def test_synthetic_bank_has_planted_matches(pkg):
    from oracle import oracle_int
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    desc, kpi = syn.make_bank_tensor(3, 600, seed=1, device="cpu")
    assert desc.shape == (3, 768, 128) and (desc[:, 600:] == 0).all()
    q, t, d = oracle_int.match(desc[0, :600].numpy(), desc[1, :600].numpy())
    assert len(q) > 20
    desc2, _ = syn.make_bank_tensor(3, 600, seed=1, device="cpu")
    assert (desc == desc2).all()


def test_window_shard_covers_the_schedule(pkg):
    for n, w in ((2000, 32), (40, 8), (5, 32), (33, 32)):
        full = pkg.window_pairs(n, w)
        for ws in (1, 2, 3, 8):
            got, counts = [], []
            for r in range(ws):
                lo, hi, halo, lp, gp = pkg.window_shard(n, w, ws, r)
                assert halo <= n and (len(gp) == 0 or (gp.max() < halo and gp.min() >= lo))
                assert np.array_equal(lp + lo, gp)
                assert len(lp) == 0 or lp.max() < halo - lo
                got.append(gp)
                counts.append(len(gp))
            assert np.array_equal(np.concatenate(got), full)
            if n == 2000:
                assert max(counts) - min(counts) <= 2 * w          # balanced on pair counts
                lo, hi, halo, _, _ = pkg.window_shard(n, w, 8, 3)
                assert halo - lo <= 2000 // 8 + w + 2                 # ~282 images per rank instead of 2,000


def test_partition_by_owner_is_balanced_and_local_first(pkg):
    sched = importlib.import_module("3d_reconstruction_b200.scheduler")
    for n_img, ws in ((257, 8), (257, 4), (64, 8), (12, 2), (5, 2), (9, 4)):
        pairs = sched.blocked_order(sched.exhaustive_pairs(n_img))
        n_loc = -(-n_img // ws)
        seen = np.zeros(len(pairs), int)
        sizes = []
        for r in range(ws):
            p, idx, n_local = sched.partition_by_owner(pairs, ws, r, n_loc)
            assert np.array_equal(p, pairs[idx])
            seen[idx] += 1
            sizes.append(len(p))
            lo = r * n_loc
            assert ((p[:n_local] >= lo) & (p[:n_local] < lo + n_loc)).all()        # leading pairs: own images only
            assert not (((p[n_local:] >= lo) & (p[n_local:] < lo + n_loc)).all(axis=1)).any()
            assert np.all(np.diff(idx[:n_local]) > 0) and np.all(np.diff(idx[n_local:]) > 0)   # list order inside both parts
        assert (seen == 1).all()                                                     # every pair exactly once
        assert max(sizes) - min(sizes) <= max(1, ws), (n_img, ws, sizes)
    p, idx, n_local = sched.partition_by_owner(pairs, 1, 0, 9)
    assert len(p) == len(pairs) and n_local == len(pairs)


def test_partition_by_owner_weighted_targets(pkg):
    """weights = measured relative throughput of the ranks: the pair counts follow them, every pair is still owned
    exactly once, local pairs still come first, and equal weights reproduce the unweighted partition."""
    sched = importlib.import_module("3d_reconstruction_b200.scheduler")
    n_img, ws = 257, 8
    n_loc = -(-n_img // ws)
    pairs = pkg.blocked_order(pkg.exhaustive_pairs(n_img))
    w = [1.0, 1.02, 0.98, 1.0, 1.0, 1.0, 0.96, 1.04]
    tgt = sched.rank_targets(len(pairs), ws, w)
    assert int(tgt.sum()) == len(pairs) and tgt[7] > tgt[0] > tgt[6]
    seen = []
    for r in range(ws):
        p, idx, n_local = sched.partition_by_owner(pairs, ws, r, n_loc, weights=w)
        assert len(p) == int(tgt[r])
        assert np.all(p[:n_local] // n_loc == r) and np.array_equal(pairs[idx], p)
        seen.append(idx)
        p0, idx0, nl0 = sched.partition_by_owner(pairs, ws, r, n_loc)
        p1, idx1, nl1 = sched.partition_by_owner(pairs, ws, r, n_loc, weights=[1.0] * ws)
        assert np.array_equal(idx0, idx1) and nl0 == nl1
    assert np.array_equal(np.sort(np.concatenate(seen)), np.arange(len(pairs)))
    with pytest.raises(ValueError):
        sched.rank_targets(10, 2, [1.0, 0.0])
