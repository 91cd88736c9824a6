"""The match-list digest (matcher.digest_matches) against a numpy restatement, on CPU tensors: it is what
bench.py and the multi-rank GPU test use to prove that N ranks produce the lists one process produces."""
import importlib

import numpy as np
import torch

K = [0x9E3779B97F4A7C15, 0xC2B2AE3D27D4EB4F, 0x165667B19E3779F9, 0x27D4EB2F165667C5, 0xBF58476D1CE4E5B9, 0x94D049BB133111EB]


def _np_digest(q, t, d, offsets, pair_ids):
    k = [np.uint64(c).astype(np.int64) for c in K]
    cnt = np.diff(offsets)
    slot = np.repeat(np.arange(len(cnt)), cnt)
    with np.errstate(over="ignore"):
        x = (pair_ids[slot].astype(np.int64) * k[0] + q.astype(np.int64) * k[1] + t.astype(np.int64) * k[2] +
             d.astype(np.float32).view(np.int32).astype(np.int64) * k[3])
        x = (x ^ (x >> 30)) * k[4]
        x = (x ^ (x >> 27)) * k[5]
        x = x ^ (x >> 31)
        per = np.zeros(len(cnt), np.int64)
        np.add.at(per, slot, x)
        return per, int(per.sum().astype(np.uint64))


def test_digest_matches_numpy_and_is_shard_additive():
    m = importlib.import_module("3d_reconstruction_b200.matcher")
    rng = np.random.default_rng(0)
    cnt = rng.integers(0, 40, size=23)
    cnt[3] = 0
    off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int64)
    s = int(off[-1])
    q = rng.integers(0, 16384, size=s).astype(np.int32)
    t = rng.integers(0, 16384, size=s).astype(np.int32)
    d = rng.random(s).astype(np.float32) * 500
    fm = m.FlatMatches(torch.from_numpy(q), torch.from_numpy(t), torch.from_numpy(d), torch.from_numpy(off))
    per, tot = fm.digest()
    eper, etot = _np_digest(q, t, d, off, np.arange(23))
    assert np.array_equal(per.numpy(), eper) and tot == etot
    # two "ranks" holding slices of the list with their global pair ids: per-pair values equal, totals add mod 2^64
    cut = 9
    a = m.FlatMatches(fm.query_idx[:off[cut]], fm.train_idx[:off[cut]], fm.dist[:off[cut]], fm.offsets[:cut + 1])
    b = m.FlatMatches(fm.query_idx[off[cut]:], fm.train_idx[off[cut]:], fm.dist[off[cut]:], fm.offsets[cut:] - off[cut])
    pa, ta = a.digest(np.arange(cut))
    pb, tb = b.digest(np.arange(cut, 23))
    assert np.array_equal(torch.cat([pa, pb]).numpy(), eper)
    assert (ta + tb) % (1 << 64) == etot
    # any change of a record changes the digest; the order of the records inside a pair does not matter for the total
    d2 = d.copy()
    d2[5] = np.nextafter(d2[5], np.float32(1e9))
    assert _np_digest(q, t, d2, off, np.arange(23))[1] != etot
    assert m.FlatMatches(fm.query_idx[:0], fm.train_idx[:0], fm.dist[:0], torch.zeros(1, dtype=torch.int64)).digest()[1] == 0
