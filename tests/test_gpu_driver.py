"""GPU: the matching.py-shaped driver end to end (batched GPU matching of all view-graph edges + host
walk) against the restated reference loop driven by the cv2 oracle matcher, and the reference loop
itself driven by the product's LightGlue-shaped callable (the literal drop-in at matching.py:122)."""
import importlib
import os

import numpy as np
import pytest

from oracle import oracle_driver
from tests.scene import make_scene
from tests.stage_fixture import assert_same_stage, load_stage_ref

pytestmark = pytest.mark.gpu


def _same(a, b):
    pa, ma = a
    pb, mb = b
    assert [tuple(map(int, p)) for p in pa] == [tuple(map(int, p)) for p in pb]
    assert len(ma) == len(mb)
    for x, y in zip(ma, mb):
        for u, v in zip(x, y):
            assert np.array_equal(np.asarray(u), np.asarray(v))


def test_driver_equals_reference_loop(pkg, tmp_path):
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    desc, pts, size, k, codebook = make_scene()
    connection, start = drv.bow_view_graph(desc, k, codebook, cos_thr=0.0)
    ref = oracle_driver.reference_loop(desc, pts, size, connection, start, oracle_driver.cv2_matcher(), min_matches=60)
    assert len(ref[0]) >= 3
    # 1. batched driver
    got = drv.match_view_graph(desc, pts, connection, start, gates=drv.Gates(min_matches=60))
    _same(ref, got)
    # 2. the product matcher as the callable of the reference's own loop (matching.py:122 unchanged)
    via_callable = oracle_driver.reference_loop(desc, pts, size, connection, start, pkg.B200Matcher(), min_matches=60)
    _same(ref, via_callable)
    # 2b. geometric verification batched on the GPU instead of cv2 per edge: same accepted edges, same rows
    got_gpu = drv.match_view_graph(desc, pts, connection, start, gates=drv.Gates(min_matches=60), geometry="gpu")
    _same(ref, got_gpu)
    # 3. stage files for sfm.py
    drv.save_matches(*got, output_dir=str(tmp_path))
    all_matches = np.load(os.path.join(str(tmp_path), "all_matches.npy"), allow_pickle=True)
    assert len(all_matches) == len(ref[1]) and len(all_matches[0]) == 3


def test_vq_words_equal_scipy(pkg):
    """b200m_vq_f32 against scipy.cluster.vq.vq (what matching.py:27 calls): identical words, hence an
    identical view graph; float (DISK-like) and integer-valued (SIFT-like) descriptors, 200 words as bow.py:7."""
    from scipy.cluster.vq import kmeans, vq
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    rng = np.random.default_rng(3)
    imgs_f = [rng.standard_normal((int(n), 128)).astype(np.float32) for n in (2048, 1500, 1, 777)]
    imgs_i = [np.abs(rng.standard_normal((int(n), 128)) * 60).clip(0, 255).round().astype(np.float32) for n in (900, 2048, 33)]
    for imgs in (imgs_f, imgs_i):
        np.random.seed(0)
        codebook, _ = kmeans(np.vstack(imgs).astype("float"), 200, 1)
        got = drv.vq_words(imgs, codebook)
        total = same = 0
        for g, d in zip(got, imgs):
            exp, _ = vq(d.astype("float"), codebook)
            assert g.shape == exp.shape
            same += int((g == exp).sum())
            total += len(exp)
        assert same == total, (same, total)
    desc, pts, size, k, codebook = make_scene()
    c0, s0 = oracle_driver.view_graph(desc, k, codebook, cos_thr=0.2)
    c1, s1 = drv.bow_view_graph(desc, k, codebook, cos_thr=0.2, device="cuda")
    assert s0 == s1 and [list(map(int, a)) for a in c0] == [list(map(int, a)) for a in c1]


def test_driver_on_real_photographs(pkg, golden_dir):
    """Real SIFT features of six of the reference's own photographs (tests/golden/make_golden_scene.py): the BoW
    view graph with GPU visual words equals the scipy one, and the batched driver reproduces the restated
    reference loop (cv2 matcher per visited edge) row for row -- with cv2 geometry per edge and with the batched
    GPU geometry."""
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    g = np.load(os.path.join(golden_dir, "ystad_scene.npz"))
    n = 6
    desc = np.array([None] * n, dtype=object)
    pts = np.array([None] * n, dtype=object)
    size = np.array([None] * n, dtype=object)
    for i in range(n):
        desc[i], pts[i], size[i] = g[f"desc{i}"].astype(np.float32), g[f"pts{i}"], g[f"size{i}"]
    k, codebook = int(g["k"]), g["codebook"]
    conn_ref, start_ref = oracle_driver.view_graph(desc, k, codebook, cos_thr=0.0)
    conn, start = drv.bow_view_graph(desc, k, codebook, cos_thr=0.0, device="cuda")
    assert start == start_ref and [list(map(int, c)) for c in conn] == [list(map(int, c)) for c in conn_ref]
    ref = oracle_driver.reference_loop(desc, pts, size, conn, start, oracle_driver.cv2_matcher(), min_matches=60)
    assert len(ref[0]) >= 3 and min(len(m[0]) for m in ref[1]) >= 60
    got = drv.match_view_graph(desc, pts, conn, start, gates=drv.Gates(min_matches=60))
    _same(ref, got)
    # Batched GPU geometry on the same edges. These photographs include near-pure rotations (DSC_0001/0002: 437
    # RANSAC inliers, none of which cv2 places in front of both cameras within its depth limit), where the
    # cheirality count is ill-conditioned in any implementation -- so compare the well-conditioned edges only:
    # every edge the reference loop accepted has a cv2 pose-inlier count the GPU count matches within 15 %.
    from oracle import oracle_cv2, oracle_geometry
    for (a, b), row in zip(ref[0], ref[1]):
        k0, k1 = pts[a][row[0]], pts[b][row[1]]
        cv = oracle_geometry.verify_pair(k0, k1, drv.FOCAL_LENGTH)
        gp = pkg.essential_ransac([k0], [k1], drv.FOCAL_LENGTH, iters=2048, seed=1)
        assert cv is not None and int(cv[4].sum()) > 10
        assert abs(int(gp.pose_inliers[0]) - int(cv[4].sum())) <= 0.15 * int(cv[4].sum()) + 5, (a, b, gp.pose_inliers, cv[4].sum())
        assert int(gp.inliers[0]) >= 0.9 * int(cv[1].sum())
    got_gpu = drv.match_view_graph(desc, pts, conn, start, gates=drv.Gates(min_matches=60), geometry="gpu")
    assert len(got_gpu[0]) >= 3 and all(len(r) == 3 and len(r[0]) == len(r[1]) == len(r[2]) for r in got_gpu[1])


def test_driver_equals_the_reference_script_itself(pkg, tmp_path):
    """The product stage (GPU visual words -> view graph -> every directed edge matched in one batched GPU call ->
    host walk with the reference's default gates) against the files the UNMODIFIED /root/reference/matching.py
    wrote for the same inputs (tests/golden/stage_ref.npz, made by tests/golden/make_golden_stage.py); also the
    product matcher as the script's own `matcher` callable (matching.py:20,122)."""
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    s = load_stage_ref()
    conn, start = drv.bow_view_graph(s["desc"], s["k"], s["codebook"], device="cuda")
    assert start == s["start"] and [list(map(int, c)) if c else [] for c in conn] == s["connection"]
    pairs, matches = drv.match_view_graph(s["desc"], s["pts"], conn, start)
    assert_same_stage(s["img_pairs"], s["all_matches"], pairs, matches)
    via = oracle_driver.reference_loop(s["desc"], s["pts"], s["size"], conn, start, pkg.B200Matcher())
    assert_same_stage(s["img_pairs"], s["all_matches"], *via)
    # stage files in -> stage files out (run_matching_stage), in the reference's on-disk layout
    import joblib
    out = str(tmp_path)
    np.save(os.path.join(out, "all_descriptors.npy"), s["desc"])
    np.save(os.path.join(out, "all_points.npy"), s["pts"])
    np.save(os.path.join(out, "img_size.npy"), s["size"])
    joblib.dump((s["k"], s["codebook"]), os.path.join(out, "bow_codebook.plk"), compress=3)
    drv.run_matching_stage(out)
    img_pairs = np.load(os.path.join(out, "img_pairs.npy"), allow_pickle=True)
    all_matches = np.load(os.path.join(out, "all_matches.npy"), allow_pickle=True)
    assert_same_stage(s["img_pairs"], s["all_matches"], img_pairs, [list(r) for r in all_matches])


def test_driver_float_descriptors_dispatch(pkg):
    """The reference's own stage file holds FLOAT descriptors (DISK, feature_extraction.py:10,28,50): with
    MatchConfig(metric="ip_bf16") the driver packs a bf16 bank and runs the mutual arg-max rule with the score
    threshold; the walk equals the restated reference loop driven by the float oracle matcher (inputs are
    bf16-representable, so the two agree row for row). Unsupported settings raise instead of being ignored."""
    import torch

    from oracle import oracle_float
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    rng = np.random.default_rng(5)
    n_img, n_world, k_img = 6, 2600, 1400
    world = rng.standard_normal((n_world, 128)).astype(np.float32)
    world /= np.linalg.norm(world, axis=1, keepdims=True)
    xyz = rng.uniform([-4, -3, 8], [4, 3, 14], size=(n_world, 3))
    desc = np.array([None] * n_img, dtype=object)
    pts = np.array([None] * n_img, dtype=object)
    size = np.array([None] * n_img, dtype=object)
    for i in range(n_img):
        ids = rng.permutation(n_world)[:k_img]
        d = world[ids] + 0.04 * rng.standard_normal((k_img, 128)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        desc[i] = torch.from_numpy(d).to(torch.bfloat16).float().numpy()          # DISK-shaped: float32 [K, 128], unit norm
        cam = xyz[ids] + np.array([-(i - n_img / 2) * 0.5, 0.0, 0.0])
        pts[i] = (drv.FOCAL_LENGTH * cam[:, :2] / cam[:, 2:3] + rng.normal(0, 0.3, size=(k_img, 2))).astype(np.float32)
        size[i] = np.array([1296.0, 1936.0], dtype=np.float32)
    connection = [[j for j in range(n_img) if j != i] for i in range(n_img)]
    thr = 0.6

    def float_oracle_matcher(data):
        q, t, sc = oracle_float.match(data["image0"]["descriptors"][0].numpy(), data["image1"]["descriptors"][0].numpy(), thr, True)
        return {"matches": [torch.from_numpy(np.stack([q, t], -1).astype(np.int64)).reshape(-1, 2)], "scores": [torch.from_numpy(sc)]}

    ref = oracle_driver.reference_loop(desc, pts, size, connection, 0, float_oracle_matcher, min_matches=200)
    assert len(ref[0]) >= 3
    conf = pkg.MatchConfig(metric="ip_bf16", score_threshold=thr, mutual=True)
    got = drv.match_view_graph(desc, pts, connection, 0, conf=conf, gates=drv.Gates(min_matches=200))
    _same(ref, got)
    # the u8 path refuses float descriptors loudly, unknown settings raise
    with pytest.raises(ValueError):
        drv.match_view_graph(desc, pts, connection, 0, conf=pkg.MatchConfig())
    with pytest.raises(NotImplementedError):
        drv.match_view_graph(desc, pts, connection, 0, conf=pkg.MatchConfig(metric="hamming"))
    with pytest.raises(NotImplementedError):
        drv.match_view_graph(desc, pts, connection, 0, conf=pkg.MatchConfig(k=3))
    with pytest.raises(ValueError):
        drv.match_view_graph(desc, pts, connection, 0, conf=pkg.MatchConfig(score_threshold=0.5))
