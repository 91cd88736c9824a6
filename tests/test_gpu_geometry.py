"""GPU: batched essential-matrix RANSAC + pose recovery (matching.py:133-142) against cv2 and ground truth.

RANSAC is randomised on both sides, so parity is statistical and the tolerances are stated per test:
the minimal solver must reproduce cv2's own five-point solutions; on synthetic two-view scenes the inlier
sets must agree with the planted inliers (recall >= 0.93, precision >= 0.97), be no worse than cv2's on them,
overlap cv2's set (Jaccard >= 0.75: two good models differ on matches near the 1 px threshold), and the pose
must be within 0.5 deg of the truth.
"""
import numpy as np
import pytest

from oracle import oracle_geometry as og

pytestmark = pytest.mark.gpu
FOCAL = 2378.98305085


def _align(E):
    E = E / np.linalg.norm(E)
    return E


def _same_up_to_sign(A, B):
    return abs(float((_align(A) * _align(B)).sum()))


def _rot_angle_deg(Ra, Rb):
    c = (np.trace(Ra.T @ Rb) - 1) / 2
    return float(np.degrees(np.arccos(np.clip(c, -1, 1))))


@pytest.mark.parametrize("seed", [0, 1, 2, 3, 4, 5])
def test_five_point_solver_matches_cv2(pkg, seed):
    geo = __import__("importlib").import_module("3d_reconstruction_b200.geometry")
    k0, k1, _, _, _, E_true = og.two_view_scene(n=5, outliers=0.0, noise_px=0.0, seed=seed)
    x0, x1 = k0 / FOCAL, k1 / FOCAL
    mine = geo.five_point(x0, x1)
    assert 1 <= len(mine) <= 10
    # every candidate satisfies the five epipolar constraints and the essential-matrix constraints
    for E in mine:
        h0 = np.concatenate([x0, np.ones((5, 1))], 1)
        h1 = np.concatenate([x1, np.ones((5, 1))], 1)
        assert np.abs(np.einsum("ni,ij,nj->n", h1, E, h0)).max() < 1e-9
        assert abs(np.linalg.det(E)) < 1e-9
        assert np.abs(2 * E @ E.T @ E - np.trace(E @ E.T) * E).max() < 1e-8
    # the true matrix is among them, and so is every solution cv2's solver returns
    assert max(_same_up_to_sign(E, E_true) for E in mine) > 1 - 1e-8
    ref = og.five_point_solutions(x0, x1)
    assert len(ref) >= 1
    for Er in ref:
        assert max(_same_up_to_sign(E, Er) for E in mine) > 1 - 1e-6
    assert len(mine) >= len(ref)


@pytest.mark.parametrize("n,outliers,seed", [(800, 0.3, 0), (2000, 0.5, 1), (300, 0.1, 2), (60, 0.3, 3)])
def test_ransac_against_truth_and_cv2(pkg, n, outliers, seed):
    geo = __import__("importlib").import_module("3d_reconstruction_b200.geometry")
    k0, k1, good, R, t, E_true = og.two_view_scene(n=n, outliers=outliers, seed=seed)
    res = geo.essential_ransac([k0], [k1], FOCAL, threshold=1.0, iters=1024, seed=7)
    mask, pose_mask = res.masks(0)
    assert res.inliers[0] == mask.sum() and res.pose_inliers[0] == pose_mask.sum()
    # planted inliers: recall and precision
    assert (mask & good).sum() >= 0.93 * good.sum()
    assert (mask & good).sum() >= 0.97 * mask.sum()
    assert _same_up_to_sign(res.E[0], E_true) > 0.999
    assert _rot_angle_deg(res.R[0], R) < 0.5
    assert float(np.degrees(np.arccos(np.clip(res.t[0] @ t, -1, 1)))) < 2.0
    assert abs(np.linalg.det(res.R[0]) - 1) < 1e-9 and abs(np.linalg.norm(res.t[0]) - 1) < 1e-9
    assert pose_mask.sum() >= 0.98 * mask.sum() and not (pose_mask & ~mask).any()
    # cv2 (the reference's calls) on the same data
    E, cmask, cR, ct, cpose = og.verify_pair(k0, k1, FOCAL)
    inter, union = (mask & cmask).sum(), (mask | cmask).sum()
    assert inter / union >= 0.75, (inter, union)       # two good models differ on the matches near the 1 px threshold
    assert (mask & good).sum() >= (cmask & good).sum() - 0.03 * good.sum()       # not worse than cv2 on the planted inliers
    # cv2 stops as soon as prob = 0.999 is reached and often keeps a slightly worse model: one-sided
    assert int(mask.sum()) >= 0.95 * int(cmask.sum()) - 3 and int(mask.sum()) <= good.sum() * 1.03 + 3
    assert _rot_angle_deg(res.R[0], cR) < 0.7
    assert int(pose_mask.sum()) >= 0.95 * int(cpose.sum()) - 3


def test_ransac_batch_ragged_degenerate_and_deterministic(pkg):
    geo = __import__("importlib").import_module("3d_reconstruction_b200.geometry")
    scenes = [og.two_view_scene(n=500, outliers=0.4, seed=10), og.two_view_scene(n=4, outliers=0.0, seed=11),
              og.two_view_scene(n=1200, outliers=0.2, seed=12, yaw=-0.2, baseline=0.6)]
    rng = np.random.default_rng(0)
    noise0, noise1 = rng.uniform(-600, 600, (400, 2)), rng.uniform(-600, 600, (400, 2))
    p0 = [s[0] for s in scenes] + [noise0, np.zeros((0, 2))]
    p1 = [s[1] for s in scenes] + [noise1, np.zeros((0, 2))]
    a = geo.essential_ransac(p0, p1, FOCAL, iters=1024, seed=3)
    b = geo.essential_ransac(p0, p1, FOCAL, iters=1024, seed=3)
    assert np.array_equal(a.mask, b.mask) and np.array_equal(a.E, b.E) and np.array_equal(a.pose_mask, b.pose_mask)
    for i in (0, 2):
        mask, _ = a.masks(i)
        good = scenes[i][2]
        assert (mask & good).sum() >= 0.93 * good.sum() and (mask & good).sum() >= 0.97 * mask.sum()
        assert _rot_angle_deg(a.R[i], scenes[i][3]) < 0.5
        # the same pair alone gives the same answer as inside the batch? (seeded per pair index, so compare with truth only)
    assert a.inliers[1] == 0 and a.pose_inliers[1] == 0 and not a.E[1].any()          # fewer than five matches
    assert a.inliers[3] < 40                                                           # pure noise: no consensus
    assert a.inliers[4] == 0
    c = geo.essential_ransac(p0, p1, FOCAL, iters=1024, seed=4)
    m0, _ = a.masks(0)
    m1, _ = c.masks(0)
    assert (m0 & m1).sum() / (m0 | m1).sum() > 0.85                                     # another seed: same consensus set
