"""CPU restatement of the reference's mutual-arg-best rule on a raw similarity matrix -- TEST INFRASTRUCTURE ONLY.

Follows lightglue/lightglue.py of the reference:
  * similarity()       :223      sim = einsum('bmd,bnd->bmn', mdesc0, mdesc1)   (fp32)
  * filter_matches()   :235-254  row / column arg-max (:237), mutual check (:239-242), threshold on the
                                 winning score (:243-250), -1 for rejected rows (:252-253)
  * match_list()       :417-426  valid = m0 > -1; matches = stack([where(valid), m0[valid]]) -- ascending query index
Pinned against the reference itself: tests/golden/float_mutual_*.npz hold the outputs of the reference's
own filter_matches (imported from /root/reference/lightglue/lightglue.py by tests/golden/make_golden_float.py)
for seeded inputs, and tests/test_oracle.py checks this restatement against them.
The threshold here applies to the raw similarity (the reference exponentiates log-scores first, :243).
"""
from __future__ import annotations

import numpy as np


def similarity(d0: np.ndarray, d1: np.ndarray) -> np.ndarray:
    return np.asarray(d0, np.float32) @ np.asarray(d1, np.float32).T


def filter_matches(sim: np.ndarray, threshold=None):
    """-> (m0 int64 [M] (-1 = rejected), m1 int64 [N], mscores0 float32 [M]); first index wins ties (np.argmax)."""
    m, n = sim.shape
    if m == 0 or n == 0:
        return np.full(m, -1, np.int64), np.full(n, -1, np.int64), np.zeros(m, np.float32)
    a0 = sim.argmax(1)
    a1 = sim.argmax(0)
    mutual0 = np.arange(m) == a1[a0]
    mutual1 = np.arange(n) == a0[a1]
    best0 = sim[np.arange(m), a0]
    ms0 = np.where(mutual0, best0, 0).astype(np.float32)
    valid0 = mutual0 & (best0 > threshold) if threshold is not None else mutual0
    valid1 = mutual1 & valid0[a1]
    return np.where(valid0, a0, -1), np.where(valid1, a1, -1), ms0


def match_list(m0: np.ndarray):
    valid = m0 > -1
    return np.nonzero(valid)[0], m0[valid]


def match(d0, d1, threshold=None, mutual: bool = True):
    """(query idx, train idx, similarity) ascending in query index."""
    sim = similarity(d0, d1)
    if mutual:
        m0, _, _ = filter_matches(sim, threshold)
    else:
        a0 = sim.argmax(1) if sim.size else np.zeros(0, np.int64)
        best = sim[np.arange(len(a0)), a0] if sim.size else np.zeros(0, np.float32)
        m0 = np.where(best > threshold, a0, -1) if threshold is not None else a0
    q, t = match_list(np.asarray(m0))
    return q.astype(np.int32), t.astype(np.int32), sim[q, t].astype(np.float32)


def agreement(sim: np.ndarray, got_q, got_t, exp_q, exp_t):
    """Index agreement of two match lists + the largest similarity gap on rows where they differ
    (gap = how far the row's chosen similarity is from the row maximum): the float path's tolerance."""
    got = dict(zip(np.asarray(got_q).tolist(), np.asarray(got_t).tolist()))
    exp = dict(zip(np.asarray(exp_q).tolist(), np.asarray(exp_t).tolist()))
    rows = set(got) | set(exp)
    same = sum(1 for r in rows if got.get(r, -1) == exp.get(r, -1))
    gap = 0.0
    for r in rows:
        if got.get(r, -1) != exp.get(r, -1):
            for lst in (got, exp):
                if r in lst:
                    gap = max(gap, float(sim[r].max() - sim[r, lst[r]]))
    return same / max(len(rows), 1), gap


# ---------------------------------------------------------------------- the reference's assignment head
def log_sigmoid(z):
    z = np.asarray(z, np.float64)
    return np.minimum(z, 0) - np.log1p(np.exp(-np.abs(z)))


def log_softmax(x, axis):
    x = np.asarray(x, np.float64)
    m = x.max(axis=axis, keepdims=True)
    return x - (m + np.log(np.exp(x - m).sum(axis=axis, keepdims=True)))


def log_double_softmax(sim, z0, z1):
    """sigmoid_log_double_softmax, lightglue.py:196-208 -> the [M, N] core of the log assignment matrix
    (the extra dust-bin row / column at :206-207 is cut off again by filter_matches at :237). fp64."""
    sim = np.asarray(sim, np.float64)
    cert = log_sigmoid(np.asarray(z0).reshape(-1, 1)) + log_sigmoid(np.asarray(z1).reshape(1, -1))     # :200
    return log_softmax(sim, 1) + log_softmax(sim, 0) + cert                                              # :201-205


def filter_log_matches(scores, th=None):
    """filter_matches on a log assignment core, lightglue.py:235-254 -> (m0, m1, mscores0, mscores1)."""
    m, n = scores.shape
    if m == 0 or n == 0:
        return np.full(m, -1, np.int64), np.full(n, -1, np.int64), np.zeros(m, np.float32), np.zeros(n, np.float32)
    a0, a1 = scores.argmax(1), scores.argmax(0)                                  # :237-238
    mutual0 = np.arange(m) == a1[a0]                                             # :241
    mutual1 = np.arange(n) == a0[a1]
    max0_exp = np.exp(scores[np.arange(m), a0])                                  # :243
    ms0 = np.where(mutual0, max0_exp, 0)                                         # :245
    ms1 = np.where(mutual1, ms0[a1], 0)                                          # :246
    valid0 = mutual0 & (ms0 > th) if th is not None else mutual0                 # :247-250
    valid1 = mutual1 & valid0[a1]                                                # :251
    return (np.where(valid0, a0, -1), np.where(valid1, a1, -1), ms0.astype(np.float32), ms1.astype(np.float32))


def assign(mdesc0, mdesc1, z0, z1, th=None):
    """MatchAssignment.forward after the projections (:223-226) + filter_matches + match list (:417-426)."""
    scores = log_double_softmax(similarity(mdesc0, mdesc1), z0, z1)
    m0, m1, ms0, ms1 = filter_log_matches(scores, th)
    q, t = match_list(m0)
    return q.astype(np.int32), t.astype(np.int32), ms0[q], (m0, m1, ms0, ms1), scores


# ---------------------------------------------------------------------- timed CPU baseline of config 5 (bench.py)
def head_torch_cpu(mdesc0, mdesc1, z0=None, z1=None, th=0.1):
    """The reference's matcher head as it EXECUTES it -- torch fp32 ops on full M x N matrices -- on the host cores:
    sim = einsum (lightglue.py:223); with logits: log double soft-max incl. the dust-bin border (:196-208) and
    filter_matches (:235-254); without: filter_matches' arg-max / mutual / threshold steps on the raw similarity.
    Restated here (the reference package cannot be imported on the GPU box); bench.py times it beside the GPU
    float path. Inputs: float32 torch tensors [M, D], [N, D], logits [M], [N]. -> (m0 [M], mscores0 [M])."""
    import torch
    import torch.nn.functional as F
    sim = torch.einsum("md,nd->mn", mdesc0, mdesc1)
    if z0 is not None:
        m, n = sim.shape
        border = sim.new_zeros((m + 1, n + 1))
        border[:m, :n] = (F.log_softmax(sim, 1) + F.log_softmax(sim.t().contiguous(), 1).t()
                          + F.logsigmoid(z0)[:, None] + F.logsigmoid(z1)[None, :])
        border[:m, n] = F.logsigmoid(-z0)
        border[m, :n] = F.logsigmoid(-z1)
        core = border[:m, :n]
        best0, best1 = core.max(1), core.max(0)
        val0 = best0.values.exp()
    else:
        best0, best1 = sim.max(1), sim.max(0)
        val0 = best0.values
    a0, a1 = best0.indices, best1.indices
    mutual0 = torch.arange(a0.numel()) == a1[a0]
    ms0 = torch.where(mutual0, val0, val0.new_zeros(()))
    valid0 = mutual0 & (ms0 > th) if th is not None else mutual0
    return torch.where(valid0, a0, torch.full_like(a0, -1)), ms0
