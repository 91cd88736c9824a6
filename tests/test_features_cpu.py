"""CPU: features.extract_features / build_codebook write the stage files in the layout the reference's
feature_extraction.py:34-53 and bow.py:23-25 produce, and driver.load_features / load_codebook read them back."""
import importlib
import os

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")


def _texture(seed, w=320, h=240):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 255, (h // 8, w // 8, 3), dtype=np.uint8)
    img = cv2.resize(img, (w, h), interpolation=cv2.INTER_CUBIC)
    for _ in range(40):
        c = (int(rng.integers(0, w)), int(rng.integers(0, h)))
        cv2.circle(img, c, int(rng.integers(3, 15)), tuple(int(x) for x in rng.integers(0, 255, 3)), -1)
    return img


def test_stage_files_have_the_reference_layout(pkg, tmp_path):
    feat = importlib.import_module("3d_reconstruction_b200.features")
    drv = importlib.import_module("3d_reconstruction_b200.driver")
    img_dir, out_dir = str(tmp_path / "imgs"), str(tmp_path / "output")
    os.makedirs(img_dir)
    imgs = {"b.png": _texture(1), "a.PNG": _texture(2, 300, 200), "c.tif": _texture(3)}
    for name, img in imgs.items():
        cv2.imwrite(os.path.join(img_dir, name), img)
    open(os.path.join(img_dir, "notes.txt"), "w").write("not an image")
    os.makedirs(os.path.join(img_dir, "sub.png"))                      # a directory is not a file, whatever its name
    names = feat.extract_features(img_dir, out_dir, max_keypoints=150)
    assert names == ["a.PNG", "b.png", "c.tif"] == open(os.path.join(out_dir, "img_list.txt")).read().split()
    desc, pts, size = drv.load_features(out_dir)
    colours = np.load(os.path.join(out_dir, "all_colors.npy"), allow_pickle=True)
    assert desc.dtype == object and len(desc) == len(pts) == len(size) == len(colours) == 3
    for i, name in enumerate(names):
        img = cv2.imread(os.path.join(img_dir, name))
        h, w = img.shape[:2]
        k = len(desc[i])
        assert 20 < k <= 150 and desc[i].shape == (k, 128) and desc[i].dtype == np.uint8
        assert pts[i].shape == (k, 2) and colours[i].shape == (k, 3)
        assert list(size[i]) == [w, h]
        # centre origin, y up (feature_extraction.py:34) -> back to pixels as matching.py:109-110 does
        x, y = pts[i][:, 0] + w / 2, -pts[i][:, 1] + h / 2
        assert x.min() >= 0 and x.max() < w and y.min() >= 0 and y.max() < h
        assert np.array_equal(colours[i], img[y.astype(int), x.astype(int), :])          # :32
        kp, d = cv2.SIFT_create(nfeatures=150).detectAndCompute(cv2.cvtColor(img, cv2.COLOR_BGR2GRAY), None)
        assert np.array_equal(desc[i], d[:150].astype(np.uint8))
        assert np.allclose(np.stack([x, y], 1), np.array([q.pt for q in kp[:150]]))
    k, codebook = feat.build_codebook(out_dir, k=12, iters=1, seed=0)
    k2, cb2 = drv.load_codebook(out_dir)
    assert k == k2 == 12 and cb2.shape[1] == 128 and cb2.dtype == np.float64 and np.array_equal(codebook, cb2)
    # the files feed the view-graph builder as they are
    conn, start = drv.bow_view_graph(desc, k, cb2, cos_thr=0.0)
    assert len(conn) == 3 and 0 <= start < 3


def test_image_without_keypoints_gives_empty_arrays(pkg):
    feat = importlib.import_module("3d_reconstruction_b200.features")
    d, p, s, c = feat.sift_features(np.zeros((64, 48, 3), np.uint8), 100)
    assert d.shape == (0, 128) and p.shape == (0, 2) and c.shape == (0, 3) and list(s) == [48, 64]
