"""GPU: BASELINE.json config 1 at its REAL size with REAL SIFT descriptors (tests/golden/sift_full.npz: every
descriptor of the reference's ystad_kloster/DSC_0001.JPG / DSC_0002.JPG, 56,041 x 52,421 rows, and the
nfeatures=16384 subsets) against cv2 LIVE on this box: the complete knnMatch(k=2) table (indices + float32
distance bits of every query row) and the ratio / cross-check lists, for every shipped pipeline variant.
Real SIFT rows carry the norm / near-tie structure the chunk-bound + refine proof of the default pipeline
leans on (csrc/match_refine.cuh), which synthetic rows only imitate."""
import os

import numpy as np
import pytest
import torch

from oracle import oracle_cv2

pytestmark = pytest.mark.gpu

VARIANTS = {"chunk256x64_refine": (1, 2, 5, 64), "chunk256x32_refine": (1, 2, 5, 32), "chunk_refine": (1, 2, 3, 64),
            "top1_verify": (1, 1, 3, 64), "top2_reject": (1, 0, 3, 64)}


def _bits(x):
    return np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def cases(golden_dir):
    g = np.load(os.path.join(golden_dir, "sift_full.npz"))
    q, t = g["q"], g["t"]
    q16, t16 = q[g["q_idx16k"]], t[g["t_idx16k"]]
    out = {
        "16k": (q16, t16),                      # cv2.SIFT_create(nfeatures=16384): 16,385 x 16,384
        "ragged_8k_16k": (q16[:8192], t16),     # its 8,192 strongest keypoints against all of the other image
        "full": (q, t),                         # nfeatures=0: 56,041 x 52,421
    }
    ref = {}
    for name, (a, b) in out.items():
        # two cv2 calls per case (forward k=2, reverse: its first column is knnMatch k=1); the three lists are
        # derived from them exactly as oracle_cv2.match does (Python-float compare of the float32 distances)
        idx, dist = oracle_cv2.knn2(a, b)
        nearest_q = oracle_cv2.knn2(b, a)[0][:, 0]
        rows = np.arange(len(a))
        lowe = dist[:, 0].astype(np.float64) < 0.75 * dist[:, 1].astype(np.float64)
        mut = nearest_q[idx[:, 0]] == rows

        def pick(mask):
            return rows[mask].astype(np.int32), idx[mask, 0], dist[mask, 0]
        ref[name] = {"knn": (idx, dist), "ratio_mutual": pick(lowe & mut), "ratio": pick(lowe), "mutual": pick(mut)}
    # ... and that derivation is oracle_cv2.match itself on the smallest case
    eq, et, ed = oracle_cv2.match(*out["ragged_8k_16k"], 0.75, True)
    gq, gt, gd = ref["ragged_8k_16k"]["ratio_mutual"]
    assert np.array_equal(eq, gq) and np.array_equal(et, gt) and np.array_equal(_bits(ed), _bits(gd))
    # the box's cv2 agrees with the container that made the fixture
    assert len(ref["full"]["ratio"][0]) == int(g["n_ratio_full"]) and len(ref["full"]["ratio_mutual"][0]) == int(g["n_mutual_full"])
    assert len(ref["16k"]["ratio"][0]) == int(g["n_ratio_16k"]) and len(ref["16k"]["ratio_mutual"][0]) == int(g["n_mutual_16k"])
    return out, ref


@pytest.mark.parametrize("variant", list(VARIANTS))
@pytest.mark.parametrize("case", ["16k", "ragged_8k_16k", "full"])
def test_real_sift_equals_cv2(pkg, cases, case, variant):
    data, ref = cases
    q, t = data[case]
    epi, pipeline, kernel, chunk = VARIANTS[variant]
    for name, val in (("epilogue", epi), ("pipeline", pipeline), ("kernel", kernel), ("chunk", chunk)):
        pkg.set_option(name, val)
    try:
        bank = pkg.DescriptorBank.from_arrays([q, t])
        idx, dist = pkg.knn2_pairs(bank, [[0, 1]])
        idx, dist = idx[0, :len(q)].cpu().numpy(), dist[0, :len(q)].cpu().numpy()
        eidx, edist = ref[case]["knn"]
        assert np.array_equal(idx, eidx), f"knn rows differ: {np.nonzero((idx != eidx).any(1))[0][:8]}"
        assert np.array_equal(_bits(dist), _bits(edist))
        for key, ratio, mutual in (("ratio_mutual", 0.75, True), ("ratio", 0.75, False), ("mutual", 0.0, True)):
            gq, gt, gd = pkg.match_pairs(bank, [[0, 1]], ratio=ratio, mutual=mutual).to_lists()[0]
            eq, et, ed = ref[case][key]
            assert np.array_equal(gq, eq) and np.array_equal(gt, et), (key, len(gq), len(eq))
            assert np.array_equal(_bits(gd), _bits(ed)), key
            assert len(gq) > 1000
        if variant == "chunk256x64_refine":
            st = pkg.match_stats(_last_workspace(pkg, bank))
            print(f"{case}: rows {st['rows']}, chunk-evaluated {st['rows_chunk_evaluated']}, fallback {st['rows_fallback']}, "
                  f"ratio-verify {st['rows_ratio_verify']}, matches {st['matches']}")
    finally:
        for name, val in (("epilogue", 1), ("pipeline", 2), ("kernel", 5), ("chunk", 64)):
            pkg.set_option(name, val)
        torch.cuda.synchronize()


def _last_workspace(pkg, bank):
    """Runs the default rule once more with a workspace we keep, so that its counters can be read."""
    ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(bank.k_stride, 1)), dtype=torch.uint8, device=bank.device)
    pkg.match_pairs(bank, [[0, 1]], ratio=0.75, mutual=True, workspace=ws)
    return ws


@pytest.mark.parametrize("knn_eval", [0, 1])
def test_knn2_chunk_pipeline_eval_variants(pkg, cases, golden_dir, knn_eval):
    """knnMatch(k=2) on the chunk pipeline with both evaluators of the two winning chunks (0 = tensor cores over
    gathered rows, the default; 1 = dp4a): the complete cv2 table on the real 16k / ragged cases, the tie fixture
    (equal distances -> lowest train index) and a ragged synthetic pair whose last chunk is partly padding."""
    data, ref = cases
    pkg.set_option("knn_eval", knn_eval)
    try:
        assert pkg.get_option("knn_eval") == knn_eval
        for case in ("16k", "ragged_8k_16k"):
            q, t = data[case]
            idx, dist = pkg.knn_match(q, t)
            eidx, edist = ref[case]["knn"]
            assert np.array_equal(idx, eidx), f"{case}: knn rows differ: {np.nonzero((idx != eidx).any(1))[0][:8]}"
            assert np.array_equal(_bits(dist), _bits(edist)), case
        g = np.load(os.path.join(golden_dir, "ties.npz"))
        idx, dist = pkg.knn_match(g["q"], g["t"])
        assert np.array_equal(idx, g["knn_idx"]) and np.array_equal(_bits(dist), _bits(g["knn_dist"]))
        rng = np.random.default_rng(11)
        base = rng.integers(0, 256, size=(40, 128))
        for kq, kt in ((700, 515), (300, 1), (129, 65), (64, 2000)):
            q = np.clip(base[rng.integers(0, 40, kq)] + rng.integers(-9, 10, (kq, 128)), 0, 255).astype(np.uint8)
            t = np.clip(base[rng.integers(0, 40, kt)] + rng.integers(-9, 10, (kt, 128)), 0, 255).astype(np.uint8)
            t[kt // 2] = t[0]                       # an exact duplicate column: stable order decides
            idx, dist = pkg.knn_match(q, t)
            eidx, edist = oracle_cv2.knn2(q, t)
            assert np.array_equal(idx, eidx), (kq, kt)
            assert np.array_equal(_bits(dist), _bits(edist)), (kq, kt)
    finally:
        pkg.set_option("knn_eval", 0)
        torch.cuda.synchronize()


@pytest.mark.parametrize("refine", [0, 1])
def test_refine_variants_equal_cv2(pkg, cases, refine):
    """Both forms of the refinement (0 = the fused kernel, the default; 1 = streaming classification + one warp per
    row to evaluate) on the real pairs, where a third of the rows needs its winning chunk evaluated, and on a
    bank built to send rows to the fallback sweep."""
    data, ref = cases
    pkg.set_option("refine", refine)
    try:
        assert pkg.get_option("refine") == refine
        for case in ("16k", "ragged_8k_16k", "full"):
            q, t = data[case]
            bank = pkg.DescriptorBank.from_arrays([q, t])
            for key, ratio, mutual in (("ratio_mutual", 0.75, True), ("ratio", 0.75, False), ("mutual", 0.0, True)):
                gq, gt, gd = pkg.match_pairs(bank, [[0, 1]], ratio=ratio, mutual=mutual).to_lists()[0]
                eq, et, ed = ref[case][key]
                assert np.array_equal(gq, eq) and np.array_equal(gt, et), (case, key, len(gq), len(eq))
                assert np.array_equal(_bits(gd), _bits(ed)), (case, key)
        rng = np.random.default_rng(3)
        arrays = [rng.integers(0, 256, size=(k, 128)).astype(np.uint8) for k in (3000, 2500, 1111)]   # uniform bytes: 10 % fall back
        bank = pkg.DescriptorBank.from_arrays(arrays)
        pairs = [[0, 1], [1, 2], [2, 0], [1, 0]]
        got = pkg.match_pairs(bank, pairs, ratio=0.9, mutual=True).to_lists()
        for (a, b), (gq, gt, gd) in zip(pairs, got):
            eq, et, ed = oracle_cv2.match(arrays[a], arrays[b], 0.9, True)
            assert np.array_equal(gq, eq) and np.array_equal(gt, et) and np.array_equal(_bits(gd), _bits(ed)), (a, b)
    finally:
        pkg.set_option("refine", 0)
        torch.cuda.synchronize()
