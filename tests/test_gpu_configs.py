"""GPU: BASELINE.json configs 2 and 4 at their full sizes (parity-test cases, not bench lines):
size-independent properties + sampled pairs against the oracle; config 4 also checks that the
halo-sharded bank (what 8 ranks would hold) gives the same lists as the whole bank."""
import importlib
import time

import numpy as np
import pytest
import torch

from oracle import oracle_cv2, oracle_int

pytestmark = pytest.mark.gpu


def _check_sampled_cv2(desc, k, pairs, lists, sample):
    """GPU == cv2 itself (BFMatcher knnMatch k=2 + ratio + cross-check on this box) on sampled pairs:
    indices and the float32 distance bits."""
    for p in sample:
        a, b = map(int, pairs[p])
        eq, et, ed = oracle_cv2.match(desc[a, :k].cpu().numpy().astype(np.float32), desc[b, :k].cpu().numpy().astype(np.float32), 0.75, True)
        q, t, d = lists[p]
        assert np.array_equal(q, eq) and np.array_equal(t, et), (a, b)
        assert np.array_equal(np.asarray(d, np.float32).view(np.uint32), np.asarray(ed, np.float32).view(np.uint32)), (a, b)


def _check_sampled(desc, k, pairs, lists, sample):
    for p in sample:
        a, b = map(int, pairs[p])
        eq, et, ed = oracle_int.match(desc[a, :k].cpu().numpy(), desc[b, :k].cpu().numpy(), 0.75, True)
        q, t, d = lists[p]
        assert np.array_equal(q, eq) and np.array_equal(t, et), (a, b)
        assert np.array_equal(d.view(np.uint32), ed.view(np.uint32)), (a, b)


def test_config2_64_images_8192_exhaustive(pkg):
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    n_img, k = 64, 8192
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=1234, device="cuda")
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    pairs = pkg.blocked_order(pkg.exhaustive_pairs(n_img))
    assert len(pairs) == 2016
    out = pkg.match_pairs(bank, pairs, cap=4096)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = pkg.match_pairs(bank, pairs, cap=4096, out=out)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"config 2: {len(pairs) / dt:.0f} pairs/s ({2.0 * k * k * 128 * len(pairs) / dt / 1e12:.0f} TOP/s)")
    lists = out.to_lists()
    n = sum(len(q) for q, _, _ in lists)
    assert n > 10000
    for q, t, d in lists:
        assert len(set(t.tolist())) == len(t) and np.all(np.diff(q) > 0)      # one-to-one, ascending
    neighbours = [p for p, (a, b) in enumerate(pairs) if b - a == 1][:2]
    _check_sampled(desc, k, pairs, lists, neighbours + [len(pairs) // 2])
    sample = np.random.default_rng(2).choice(len(pairs), size=8, replace=False).tolist()
    _check_sampled_cv2(desc, k, pairs, lists, sample)


def test_config3_257_images_16384_exhaustive(pkg):
    """The bench workload itself (BASELINE.json config 3) once, with its properties and eight random pairs
    (plus two neighbouring ones, which share planted matches) against cv2 on the box."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    n_img, k = 257, 16384
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=1234, device="cuda")
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    pairs = pkg.blocked_order(pkg.exhaustive_pairs(n_img))
    assert len(pairs) == 32896
    out = pkg.match_pairs(bank, pairs, cap=4096)
    out.check_overflow()
    cnt = out.count.cpu().numpy()
    assert cnt.sum() == 1126467                      # what bench.py reports as matches_rank0 for this seed
    sample = np.random.default_rng(3).choice(len(pairs), size=8, replace=False).tolist()
    sample += [p for p, (a, b) in enumerate(pairs) if b - a == 1][:2]
    q_all, t_all, d_all = out.query_idx, out.train_idx, out.dist
    lists = {p: (q_all[p, :cnt[p]].cpu().numpy().astype(np.int64), t_all[p, :cnt[p]].cpu().numpy().astype(np.int64),
                 d_all[p, :cnt[p]].cpu().numpy()) for p in sample}
    for p in sample:
        q, t, _ = lists[p]
        assert len(set(t.tolist())) == len(t) and np.all(np.diff(q) > 0)
    _check_sampled_cv2(desc, k, pairs, lists, sample)


def test_config3_every_list_default_pipeline_equals_exact_pipeline(pkg):
    """The WHOLE bench workload (all 32,896 pairs of BASELINE.json config 3) through the default pipeline (chunk
    bounds on the norm-sorted bank + refine + fallback, whose shortcut proof depends on the data) and through the
    independent exact pipeline 1 (threshold-reject top-1 sweep + verify sweeps on the original bank): the 64-bit
    digest of every single pair's list (query rows, train rows, float32 distance bits) must agree."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    n_img, k = 257, 16384
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=1234, device="cuda")
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    pairs = pkg.blocked_order(pkg.exhaustive_pairs(n_img))
    per = {}
    out = None
    for pl in (2, 1):
        pkg.set_option("pipeline", pl)
        try:
            out = pkg.match_pairs(bank, pairs, cap=4096, out=out)
            out.check_overflow()
            flat = out.flatten()
            per[pl], total = flat.digest()
            cnt = (flat.offsets[1:] - flat.offsets[:-1]).clone()
            print(f"pipeline {pl}: {int(flat.offsets[-1])} matches, digest {total:016x}")
            if pl == 2:
                cnt2 = cnt
        finally:
            pkg.set_option("pipeline", 2)
    assert int(cnt2.sum()) == 1126467
    assert torch.equal(cnt, cnt2)
    bad = torch.nonzero(per[1] != per[2]).flatten()
    assert bad.numel() == 0, f"{bad.numel()} pairs differ between the pipelines, first: {pairs[bad[:5].cpu().numpy()].tolist()}"


def test_config4_window_schedule_and_halo_shards(pkg):
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    n_img, k, window = 2000, 8192, 32
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=77, device="cuda")
    pairs = pkg.window_pairs(n_img, window)
    assert len(pairs) == 63472
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    cap = 3072
    out = pkg.match_pairs(bank, pairs, cap=cap)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = pkg.match_pairs(bank, pairs, cap=cap, out=out)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"config 4: {len(pairs) / dt:.0f} pairs/s on one GPU ({2.0 * k * k * 128 * len(pairs) / dt / 1e12:.0f} TOP/s)")
    out.check_overflow()
    cnt = out.count.cpu().numpy()
    assert cnt.sum() > 100000
    q_all, t_all, d_all = out.query_idx.cpu().numpy(), out.train_idx.cpu().numpy(), out.dist.cpu().numpy()
    del bank
    # what ranks 0, 3 and 7 of an 8-GPU run hold (own images + 32-image halo) must reproduce their slice
    first = 0
    offs = {}
    for r in range(8):
        lo, hi, halo, lp, gp = pkg.window_shard(n_img, window, 8, r)
        offs[r] = (first, lo, halo, lp)
        first += len(gp)
    for r in (0, 3, 7):
        p0, lo, halo, lp = offs[r]
        sbank = pkg.DescriptorBank.from_packed(desc[lo:halo].clone(), kpi[lo:halo].clone())
        so = pkg.match_pairs(sbank, lp, cap=cap)
        sc = so.count.cpu().numpy()
        assert np.array_equal(sc, cnt[p0:p0 + len(lp)])
        sq, st_, sd = so.query_idx.cpu().numpy(), so.train_idx.cpu().numpy(), so.dist.cpu().numpy()
        for i in range(0, len(lp), 97):
            c = sc[i]
            assert np.array_equal(sq[i, :c], q_all[p0 + i, :c]) and np.array_equal(st_[i, :c], t_all[p0 + i, :c])
            assert np.array_equal(sd[i, :c].view(np.uint32), d_all[p0 + i, :c].view(np.uint32))
        del sbank, so
    sample = [0, 31337] + np.random.default_rng(4).choice(len(pairs), size=6, replace=False).tolist()
    lists = {p: (q_all[p, :cnt[p]].astype(np.int64), t_all[p, :cnt[p]].astype(np.int64), d_all[p, :cnt[p]]) for p in sample}
    _check_sampled(desc, k, pairs, lists, [0, 31337])
    _check_sampled_cv2(desc, k, pairs, lists, sample)


def test_config1_native_size_ragged(pkg):
    """BASELINE.json config 1 at cv2.SIFT's default nfeatures=0 sizes (56,041 x 52,421 keypoints for
    DSC_0001/DSC_0002): k_stride 56,064 exercises the global passes of the norm sort and the ragged last
    blocks / tiles. The real images are not on the GPU box, so SIFT-like synthetic rows stand in; rows are
    sampled against the oracle, the mutual list is checked for the properties cv2's cross-check guarantees."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    ks = [56041, 52421]
    desc, kpi = syn.make_bank_tensor(2, max(ks), seed=5, device="cuda", k_per_image=ks)
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    perm, inv, snorm, cmin, spread = bank.sort_views()
    norms = bank.norms()
    for i, k in enumerate(ks):
        assert torch.equal(snorm[i], torch.sort(norms[i], stable=True).values)
        assert torch.equal(perm[i].long(), torch.sort(norms[i], stable=True).indices)
    res = {}
    for pl in (2, 1):
        pkg.set_option("pipeline", pl)
        try:
            res[pl] = pkg.match_pairs(bank, [[0, 1], [1, 0]], ratio=0.75, mutual=True).to_lists()
        finally:
            pkg.set_option("pipeline", 2)
    for a, b in zip(res[2], res[1]):
        assert all(np.array_equal(u, v) for u, v in zip(a, b))
    q, t, d = res[2][0]
    assert len(q) > 5000 and len(set(t.tolist())) == len(t) and np.all(np.diff(q) > 0)
    # sampled matched and unmatched query rows against the oracle's knn (ratio + forward direction)
    qd, td = desc[0, :ks[0]].cpu().numpy(), desc[1, :ks[1]].cpu().numpy()
    rng = np.random.default_rng(0)
    rows = np.unique(np.concatenate([rng.choice(q, 128, replace=False), rng.choice(ks[0], 128, replace=False)]))
    eidx, edist, _ = oracle_int.knn2(qd[rows], td)
    matched = dict(zip(q.tolist(), zip(t.tolist(), d.tolist())))
    for r, (i1, _), (d1, d2) in zip(rows.tolist(), eidx.tolist(), edist.tolist()):
        passes = float(d1) < 0.75 * float(d2)
        if r in matched:
            assert passes and matched[r][0] == i1 and np.float32(matched[r][1]) == np.float32(d1)
        elif passes:
            # rejected by the cross-check only: some other query must be at least as near to i1
            dq = ((qd.astype(np.int64) - td[i1].astype(np.int64)) ** 2).sum(1)
            best = int(np.argmin(dq))
            assert best != r
