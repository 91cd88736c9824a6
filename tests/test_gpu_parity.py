"""GPU parity tests: the CUDA path, called through the C-ABI, against the oracle and the committed
cv2 fixtures. Bit-exact for indices; distances compared as float32 bit patterns."""
import importlib
import os

import numpy as np
import pytest
import torch

from oracle import oracle_int

pytestmark = pytest.mark.gpu


def _bits(x):
    return np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)


DEFAULT_KERNEL = 5
DEFAULT_PIPELINE = 2
DEFAULT_CHUNK = 64


@pytest.fixture(scope="module", params=[(1, 2, 3), (1, 2, 5, 32), (1, 2, 5, 64), (1, 1, 3), (1, 0, 3), (0, 0, 3)],
                ids=["chunk_refine", "chunk256x32_refine", "chunk256x64_refine", "top1_verify", "top2_reject", "top2_exact"])
def epi(pkg, request):
    """(epilogue, pipeline, kernel[, chunk]): the shipped variants -- the default chunk sweep on the norm-sorted
    bank + refine + exact fallback (N128 and N256 MMAs, 32- and 64-column chunks); the top-1 sweep + verify
    sweeps; the top-2 threshold-reject sweep with only the mutual verify; and the debug epilogue that sends
    every element through the insertion. (The cta_group::2 variants live in csrc/experimental/ and are only
    compiled with -DB200M_EXPERIMENTAL.)"""
    pkg.set_option("epilogue", request.param[0])
    pkg.set_option("pipeline", request.param[1])
    pkg.set_option("kernel", request.param[2])
    pkg.set_option("chunk", request.param[3] if len(request.param) > 3 else DEFAULT_CHUNK)
    yield request.param
    pkg.set_option("chunk", DEFAULT_CHUNK)
    pkg.set_option("epilogue", 1)
    pkg.set_option("pipeline", DEFAULT_PIPELINE)
    pkg.set_option("kernel", DEFAULT_KERNEL)


def test_bank_sort(pkg):
    """b200m_bank_sort_u8: rows ordered by (squared norm, row), padding last; perm / inv / chunk tables."""
    rng = np.random.default_rng(5)
    ks = [700, 3000, 1, 0, 2049]
    arrays = [rng.integers(0, 256, size=(k, 128)).astype(np.uint8) for k in ks]
    arrays[0][:300] = arrays[0][300:600]          # duplicate rows: equal norms keep their original order
    arrays[1][:, 64:] = 0
    bank = pkg.DescriptorBank.from_arrays(arrays)
    torch.cuda.synchronize()
    perm, inv, snorm, cmin, spread = [x.cpu().numpy() for x in bank.sort_views()]
    norms = bank.norms().cpu().numpy()
    desc = bank.desc.cpu().numpy()
    sdesc = bank.sorted_desc.cpu().numpy()
    for i, k in enumerate(ks):
        exp = np.argsort(norms[i], kind="stable")
        assert np.array_equal(perm[i], exp), i
        assert np.array_equal(inv[i][perm[i]], np.arange(bank.k_stride))
        assert np.array_equal(snorm[i], norms[i][exp])
        assert np.array_equal(sdesc[i], desc[i][exp])
        assert np.array_equal(cmin[i], snorm[i][::32])
        if k:
            assert (perm[i][:k] < k).all() and np.array_equal(np.sort(perm[i][:k]), np.arange(k))
            sp = max(int(snorm[i][min(c + 31, k - 1)] - snorm[i][c]) for c in range(0, k, 32))
            assert int(spread[i]) == sp, (i, int(spread[i]), sp)
        else:
            assert int(spread[i]) == 0


def test_pipelines_agree_on_fallback_heavy_data(pkg):
    """Small-alphabet descriptors (many exact ties, distances comparable to the norm spread of a chunk)
    push most rows of the chunk pipeline through its exact fallback; results must not change."""
    rng = np.random.default_rng(21)
    q = rng.integers(0, 3, size=(900, 128)).astype(np.uint8)
    t = rng.integers(0, 3, size=(1200, 128)).astype(np.uint8)
    t[100:400] = q[200:500]                        # planted exact matches
    t[500:520] = q[200:220]                        # ... some of them twice (ratio test must reject those)
    bank = pkg.DescriptorBank.from_arrays([q, t])
    res = {}
    for pl in (2, 1):
        pkg.set_option("pipeline", pl)
        try:
            res[pl] = [pkg.match_pairs(bank, [[0, 1], [1, 0]], ratio=r, mutual=m).to_lists() for r, m in
                       [(0.75, True), (0.75, False), (0.0, True), (0.9, True)]]
        finally:
            pkg.set_option("pipeline", DEFAULT_PIPELINE)
    n = 0
    for a, b in zip(res[2], res[1]):
        for x, y in zip(a, b):
            assert all(np.array_equal(u, v) for u, v in zip(x, y))
            n += len(x[0])
    assert n > 500
    for (a, b), (gq, gt, gd) in zip([(q, t), (t, q)], res[2][0]):
        eq, et, ed = oracle_int.match(a, b, 0.75, True)
        assert np.array_equal(gq, eq) and np.array_equal(gt, et) and np.array_equal(_bits(gd), _bits(ed))


def test_debug_dot_tile(pkg):
    """TMA + swizzle + smem descriptors + tcgen05.mma.kind::i8 + tcgen05.ld, without any epilogue."""
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, size=(128, 128)).astype(np.uint8)
    b = rng.integers(0, 256, size=(256, 128)).astype(np.uint8)
    a[0] = 255
    b[0] = 255
    ad, bd = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    out = torch.empty((128, 256), dtype=torch.int32, device="cuda")
    lib = pkg._lib
    lib.check(lib.lib().b200m_debug_dot_tile_u8(ad.data_ptr(), bd.data_ptr(), out.data_ptr(),
                                               torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    ref = a.astype(np.int64) @ b.astype(np.int64).T
    got = out.cpu().numpy().astype(np.int64)
    assert np.array_equal(got, ref), f"mismatching elements: {(got != ref).sum()} of {ref.size}"


def _check_pair(pkg, q, t):
    idx, dist = pkg.knn_match(q, t)
    eidx, edist, _ = oracle_int.knn2(q, t)
    assert np.array_equal(idx, eidx), f"knn idx mismatch rows: {np.nonzero((idx != eidx).any(1))[0][:10]}"
    assert np.array_equal(_bits(dist), _bits(edist))
    for ratio, mutual in [(0.75, True), (0.75, False), (0.0, True)]:
        got = pkg.match(q, t, ratio=ratio, mutual=mutual)
        exp = oracle_int.match(q, t, ratio, mutual)
        assert np.array_equal(got[0], exp[0]) and np.array_equal(got[1], exp[1]), (ratio, mutual)
        assert np.array_equal(_bits(got[2]), _bits(exp[2])), (ratio, mutual)


@pytest.mark.parametrize("name", ["sift_pair.npz", "ties.npz"])
def test_golden_fixture(pkg, epi, golden_dir, name):
    """BASELINE.json config 1 (real SIFT of DSC_0001/DSC_0002) and the tie fixture vs stored cv2 output."""
    g = np.load(os.path.join(golden_dir, name))
    idx, dist = pkg.knn_match(g["q"], g["t"])
    assert np.array_equal(idx, g["knn_idx"])
    assert np.array_equal(_bits(dist), _bits(g["knn_dist"]))
    ratio = 0.75 if name == "sift_pair.npz" else 0.0
    q, t, d = pkg.match(g["q"], g["t"], ratio=ratio, mutual=True)
    assert np.array_equal(q, g["m_q"]) and np.array_equal(t, g["m_t"])
    assert np.array_equal(_bits(d), _bits(g["m_d"]))
    assert len(q) > 0


@pytest.mark.parametrize("kq,kt", [(1, 1), (1, 2), (2, 1), (5, 300), (128, 256), (129, 257), (300, 5),
                                    (1000, 777), (640, 2048)])
def test_ragged_shapes_random(pkg, epi, kq, kt):
    rng = np.random.default_rng(kq * 7919 + kt)
    q = rng.integers(0, 256, size=(kq, 128)).astype(np.uint8)
    t = rng.integers(0, 256, size=(kt, 128)).astype(np.uint8)
    _check_pair(pkg, q, t)


def test_adversarial_ties_and_extremes(pkg, epi):
    rng = np.random.default_rng(11)
    base = rng.integers(0, 2, size=(16, 128)).astype(np.uint8) * 255
    q = base[rng.integers(0, 16, size=700)]
    t = base[rng.integers(0, 16, size=900)]
    q[10] = 0
    t[20:40] = 0          # all-zero rows are valid competitors (not padding)
    q[11] = 255
    t[41] = 255           # max magnitude: d^2 up to 128*255^2
    _check_pair(pkg, q, t)
    # identical sets: every row has distance 0 to itself, duplicates tie
    _check_pair(pkg, t, t)


def test_empty_inputs(pkg):
    q = np.zeros((0, 128), np.uint8)
    t = np.random.default_rng(0).integers(0, 256, size=(10, 128)).astype(np.uint8)
    assert pkg.match(q, t)[0].size == 0
    assert pkg.match(t, q)[0].size == 0
    idx, dist = pkg.knn_match(t, q)
    assert (idx == -1).all() and np.isinf(dist).all()


def test_synthetic_bank_batched(pkg, epi):
    """Several images, ragged K, exhaustive pairs in one call: planted matches must survive."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    ks = [1500, 1024, 999, 1300, 257]
    desc, kpi = syn.make_bank_tensor(len(ks), 1500, seed=3, device="cpu", k_per_image=ks)
    bank = pkg.DescriptorBank.from_packed(desc.cuda(), kpi)
    pairs = pkg.exhaustive_pairs(len(ks))
    pairs = np.concatenate([pairs, pairs[:, ::-1]], 0)      # both directions
    res = pkg.match_pairs(bank, pairs).to_lists()
    total = 0
    for (a, b), (q, t, d) in zip(pairs, res):
        eq, et, ed = oracle_int.match(desc[a, :ks[a]].numpy(), desc[b, :ks[b]].numpy(), 0.75, True)
        assert np.array_equal(q, eq) and np.array_equal(t, et), (a, b)
        assert np.array_equal(_bits(d), _bits(ed)), (a, b)
        total += len(q)
    assert total > 100
    # small internal batches must give identical results
    pkg.set_option("batch", 3)
    try:
        res2 = pkg.match_pairs(bank, pairs).to_lists()
    finally:
        pkg.set_option("batch", 0)
    for x, y in zip(res, res2):
        assert all(np.array_equal(u, v) for u, v in zip(x, y))


def test_cap_overflow_is_reported(pkg):
    rng = np.random.default_rng(2)
    t = rng.integers(0, 256, size=(300, 128)).astype(np.uint8)
    bank = pkg.DescriptorBank.from_arrays([t, t])
    res = pkg.match_pairs(bank, [[0, 1]], ratio=0.0, mutual=True, cap=16)
    assert int(res.count[0]) == 300
    with pytest.raises(OverflowError):
        res.to_lists()


def test_full_size_properties(pkg):
    """BASELINE config sizes (K=8192 and 16384): size-independent properties instead of a CPU oracle:
    self-match is the identity with distance 0; knn best of (i,i) is the row itself; results are
    deterministic across runs; mutual matches are one-to-one."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    for k in (8192, 16384):
        desc, kpi = syn.make_bank_tensor(3, k, seed=9, device="cuda", dup_frac=0.0)
        bank = pkg.DescriptorBank.from_packed(desc, kpi)
        r = pkg.match_pairs(bank, [[0, 0], [0, 1], [1, 0]], ratio=0.0, mutual=True)
        lists = r.to_lists()
        q, t, d = lists[0]
        # distinct rows (dup_frac=0) => self-match is exactly the identity with zero distance
        uniq = np.unique(desc[0].cpu().numpy(), axis=0).shape[0]
        if uniq == k:
            assert np.array_equal(q, np.arange(k)) and np.array_equal(t, q) and (d == 0).all()
        # (0,1) and (1,0) under pure mutual-NN (no ratio) are transposes of each other unless ties interfere
        q01, t01, d01 = lists[1]
        q10, t10, d10 = lists[2]
        a = set(zip(q01.tolist(), t01.tolist()))
        b = set(zip(t10.tolist(), q10.tolist()))
        assert len(a ^ b) <= max(2, len(a) // 1000)
        assert len(set(t01.tolist())) == len(t01)            # one-to-one
        r2 = pkg.match_pairs(bank, [[0, 0], [0, 1], [1, 0]], ratio=0.0, mutual=True).to_lists()
        for x, y in zip(lists, r2):
            assert all(np.array_equal(u, v) for u, v in zip(x, y))
        # sampled rows against the oracle at full size
        qd, td = desc[0].cpu().numpy(), desc[1].cpu().numpy()
        idx, dist = pkg.knn2_pairs(bank, [[0, 1]])
        rows = np.random.default_rng(k).choice(k, size=256, replace=False)
        eidx, edist, _ = oracle_int.knn2(qd[rows], td)
        assert np.array_equal(idx[0].cpu().numpy()[rows], eidx)
        assert np.array_equal(_bits(dist[0].cpu().numpy()[rows]), _bits(edist))


def test_lightglue_shaped_adapter(pkg, golden_dir):
    g = np.load(os.path.join(golden_dir, "sift_pair.npz"))
    m = pkg.B200Matcher().eval().to("cuda")
    feats0 = {"descriptors": torch.from_numpy(g["q"].astype(np.float32))[None], "keypoints": torch.zeros(1, len(g["q"]), 2),
              "image_size": torch.tensor([[1936., 1296.]])}
    feats1 = {"descriptors": torch.from_numpy(g["t"].astype(np.float32))[None], "keypoints": torch.zeros(1, len(g["t"]), 2),
              "image_size": torch.tensor([[1936., 1296.]])}
    out = m({"image0": feats0, "image1": feats1})
    matches = out["matches"][0]                       # what matching.py:124-128 reads
    idx0, idx1 = matches[..., 0].cpu().numpy(), matches[..., 1].cpu().numpy()
    assert np.array_equal(idx0, g["m_q"]) and np.array_equal(idx1, g["m_t"])
    m0 = out["matches0"][0].cpu().numpy()
    assert (m0[idx0] == idx1).all() and (m0 >= 0).sum() == len(idx0)


def test_lightglue_shaped_adapter_cuda_inputs(pkg, golden_dir):
    """matching.py:107-120 builds the feature tensors on the device: descriptors that already live on the GPU
    are verified and packed there (no host round trip) and give the same lists; non-integer values are refused."""
    g = np.load(os.path.join(golden_dir, "sift_pair.npz"))
    m = pkg.B200Matcher().eval().to("cuda")
    d0 = torch.from_numpy(g["q"].astype(np.float32))[None].cuda()
    d1 = torch.from_numpy(g["t"].astype(np.float32))[None].cuda()
    out = m({"image0": {"descriptors": d0}, "image1": {"descriptors": d1}})
    matches = out["matches"][0]
    assert np.array_equal(matches[..., 0].cpu().numpy(), g["m_q"]) and np.array_equal(matches[..., 1].cpu().numpy(), g["m_t"])
    # u8 tensors on the device take the same path
    bank = pkg.DescriptorBank.from_arrays([d0[0].to(torch.uint8), d1[0].to(torch.uint8)], device="cuda")
    q, t, _ = pkg.match_pairs(bank, [[0, 1]]).to_lists()[0]
    assert np.array_equal(q, g["m_q"]) and np.array_equal(t, g["m_t"])
    with pytest.raises(ValueError):
        m({"image0": {"descriptors": d0 + 0.25}, "image1": {"descriptors": d1}})
    with pytest.raises(ValueError):
        m({"image0": {"descriptors": d0 - 300.0}, "image1": {"descriptors": d1}})


def test_randomised_campaign(pkg):
    """tools/stress_parity.py, a slice of it: ragged banks from four generators (uniform bytes, SIFT-like with
    re-observed rows, tiny alphabets, constant rows), every ordered pair, random ratio / mutual settings; lists and
    float32 distance bits equal the integer oracle's."""
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "stress_parity", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "stress_parity.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.main(rounds=14, seed=123)


def test_rowmerge_handoff_stress_one_tile_items(pkg):
    """100,000 work items of a single column tile each (K = 256 per image): the 16 epilogue warps hand their row
    states over through the double-buffered `rowmerge` shared-memory slots (parity flips every item) as fast as
    the kernels can go -- a race there would show as a wrong or missing row. compute-sanitizer is closed on the
    GPU pool, so: every pipeline (chunk sweep, exact top-1, exact top-2) must produce the same per-pair digests
    twice in a row, and 400 sampled pairs must equal the integer oracle."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    n_img, k = 448, 256
    ks = np.full(n_img, k)
    ks[::7] = 255
    ks[3::11] = 129
    ks[5::13] = 128
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=17, device="cuda", k_per_image=ks.tolist(), p_shared=0.5)
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    rng = np.random.default_rng(0)
    a = rng.integers(0, n_img, size=100000)
    pairs = np.stack([a, (a + rng.integers(1, 4, size=a.size)) % n_img], 1).astype(np.int32)   # neighbours share rows
    pairs_dev = torch.from_numpy(pairs).cuda()
    digests = {}
    out = None
    try:
        for pl in (2, 1, 0):
            pkg.set_option("pipeline", pl)
            for rep in range(2):
                out = pkg.match_pairs(bank, pairs_dev, cap=256, out=out)
                out.check_overflow()
                per, tot = out.flatten().digest()
                digests.setdefault(pl, []).append((per.clone(), tot))
    finally:
        pkg.set_option("pipeline", DEFAULT_PIPELINE)
    ref = digests[2][0][0]
    for pl, runs in digests.items():
        for per, _ in runs:
            bad = torch.nonzero(per != ref).flatten()
            assert bad.numel() == 0, f"pipeline {pl}: {bad.numel()} pairs differ, first {pairs[bad[:5].cpu().numpy()].tolist()}"
    flat = pkg.match_pairs(bank, pairs_dev, cap=256).flatten().cpu()
    assert int(flat.offsets[-1]) > 100000
    d = desc.cpu().numpy()
    for p in rng.choice(len(pairs), size=400, replace=False).tolist():
        i, j = map(int, pairs[p])
        eq, et, ed = oracle_int.match(d[i, :ks[i]], d[j, :ks[j]], 0.75, True)
        q, t, dd = flat.pair(p)
        assert np.array_equal(q.numpy(), eq) and np.array_equal(t.numpy(), et), p
        assert np.array_equal(_bits(dd.numpy()), _bits(ed)), p
