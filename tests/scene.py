"""A small synthetic SfM scene for the driver tests: 3-D points seen by cameras on an arc, SIFT-like u8
descriptors per 3-D point re-observed with integer noise, keypoints in the reference's centre-origin /
y-up convention (feature_extraction.py:34), plus a BoW codebook built the way bow.py does."""
import numpy as np

FOCAL = 2378.98305085
SIZE = (1296.0, 1936.0)


def make_scene(n_img=7, n_world=2500, k_img=900, seed=0, codebook_k=800):
    rng = np.random.default_rng(seed)
    world = rng.uniform([-4, -3, 8], [4, 3, 14], size=(n_world, 3))
    base = np.abs(rng.standard_normal((n_world, 128)))
    base = np.clip(np.rint(base / np.linalg.norm(base, axis=1, keepdims=True) * 512), 0, 255)
    descs, points, sizes = [], [], []
    for i in range(n_img):
        ang = (i - n_img / 2) * 0.06
        R = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
        t = np.array([-(i - n_img / 2) * 0.5, 0.0, 0.0])
        cam = world @ R.T + t
        uv = FOCAL * cam[:, :2] / cam[:, 2:3]
        vis = (np.abs(uv[:, 0]) < SIZE[0] / 2) & (np.abs(uv[:, 1]) < SIZE[1] / 2)
        ids = np.nonzero(vis)[0]
        ids = rng.permutation(ids)[:k_img]
        d = np.clip(base[ids] + rng.integers(-5, 6, size=(len(ids), 128)), 0, 255).astype(np.float32)
        n_fresh = k_img // 5
        fresh = np.abs(rng.standard_normal((n_fresh, 128)))
        fresh = np.clip(np.rint(fresh / np.linalg.norm(fresh, axis=1, keepdims=True) * 512), 0, 255).astype(np.float32)
        pts = np.concatenate([uv[ids] + rng.normal(0, 0.3, size=(len(ids), 2)),
                              rng.uniform([-SIZE[0] / 2, -SIZE[1] / 2], [SIZE[0] / 2, SIZE[1] / 2], size=(n_fresh, 2))])
        dd = np.concatenate([d, fresh])
        order = rng.permutation(len(dd))
        descs.append(dd[order])
        points.append(pts[order])
        sizes.append(np.array(SIZE))
    all_desc = np.array([None] * n_img, dtype=object)
    all_pts = np.array([None] * n_img, dtype=object)
    img_size = np.array([None] * n_img, dtype=object)
    for i in range(n_img):
        all_desc[i], all_pts[i], img_size[i] = descs[i], points[i], sizes[i]
    # bow.py:13-23: stack everything, scipy kmeans (k, 1 iteration)
    from scipy.cluster.vq import kmeans
    stack = np.vstack([d.astype("float") for d in descs])
    np.random.seed(seed)
    codebook, _ = kmeans(stack, codebook_k, 1)
    return all_desc, all_pts, img_size, len(codebook), codebook
