"""Generates tests/golden/ystad_scene.npz: real SIFT features of the reference's own photographs
(/root/reference/ystad_kloster/DSC_0001..0006.JPG) in the layout feature_extraction.py:34-50 writes
(descriptors [K,128], centre-origin / y-up keypoints, image size), plus a BoW codebook built the way
bow.py:7-25 does (scipy kmeans, k = 200, one iteration) -- run in the build container, where
/root/reference exists:

    python tests/golden/make_golden_scene.py

The matching stage (matching.py) then runs on real data in the tests: the batched GPU driver against the
restated reference loop with the cv2 matcher. Extractor: cv2.SIFT_create(nfeatures=2500) -- the extractor
BASELINE.json names for this path (the reference ships DISK, which needs downloaded weights).
"""
import os

import cv2
import numpy as np
from scipy.cluster.vq import kmeans

HERE = os.path.dirname(os.path.abspath(__file__))
IMG_DIR = "/root/reference/ystad_kloster"
N_IMG, N_FEAT, K_WORDS = 6, 2500, 200


def main():
    names = sorted(f for f in os.listdir(IMG_DIR) if f.split(".")[-1] in ("JPG", "jpg"))[:N_IMG]
    sift = cv2.SIFT_create(nfeatures=N_FEAT)
    descs, pts, sizes = [], [], []
    for n in names:
        img = cv2.imread(os.path.join(IMG_DIR, n))                       # feature_extraction.py:25
        kp, d = sift.detectAndCompute(cv2.cvtColor(img, cv2.COLOR_BGR2GRAY), None)
        d = d[:N_FEAT]
        xy = np.array([k.pt for k in kp[:N_FEAT]], dtype=np.float64)
        size = np.array([img.shape[1], img.shape[0]], dtype=np.float64)  # (w, h) like feats['image_size']
        pts.append(np.stack([xy[:, 0] - size[0] / 2, -(xy[:, 1] - size[1] / 2)], -1).astype(np.float32))   # :34
        assert np.array_equal(d, np.rint(d)) and d.min() >= 0 and d.max() <= 255
        descs.append(d.astype(np.uint8))
        sizes.append(size)
    np.random.seed(0)
    stack = np.vstack([d.astype("float") for d in descs])
    codebook, _ = kmeans(stack[::4], K_WORDS, 1)                          # bow.py:23 (subsampled to keep it quick)
    out = {"names": np.array(names), "codebook": codebook.astype(np.float64), "k": np.int64(len(codebook)),
           "cv2_version": cv2.__version__}
    for i in range(N_IMG):
        out[f"desc{i}"] = descs[i]
        out[f"pts{i}"] = pts[i]
        out[f"size{i}"] = sizes[i]
    np.savez_compressed(os.path.join(HERE, "ystad_scene.npz"), **out)
    print([len(d) for d in descs], "codebook", codebook.shape, os.path.getsize(os.path.join(HERE, "ystad_scene.npz")))


if __name__ == "__main__":
    main()
