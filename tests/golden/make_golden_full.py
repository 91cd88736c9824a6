"""Generates tests/golden/sift_full.npz: the REAL-SIZE form of BASELINE.json config 1 -- every SIFT
descriptor cv2 finds in the reference's own ystad_kloster/DSC_0001.JPG and DSC_0002.JPG
(cv2.SIFT_create(nfeatures=0): 56,041 and 52,421 rows), plus the row lists of the nfeatures=16384 runs
(cv2 keeps the strongest keypoints, so those descriptor sets are row subsets of the full ones, in a
different order). Run in the build container, where /root/reference exists:

    python tests/golden/make_golden_full.py

Descriptors only (13.9 MB of u8 before compression): expected outputs are produced by cv2 LIVE on the GPU
box in tests/test_gpu_real_sift.py (same image => same cv2 4.13.0), the sanity counts recorded here are a
cross-check that the box's cv2 agrees with this container's.
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle_cv2  # noqa: E402

REF = "/root/reference/ystad_kloster"


def main():
    out = {"cv2_version": cv2.__version__}
    full, sub = [], []
    for tag, name in (("q", "DSC_0001.JPG"), ("t", "DSC_0002.JPG")):
        img = cv2.imread(os.path.join(REF, name), cv2.IMREAD_GRAYSCALE)
        kp0, d0 = cv2.SIFT_create(nfeatures=0).detectAndCompute(img, None)
        kp1, d1 = cv2.SIFT_create(nfeatures=16384).detectAndCompute(img, None)
        assert np.array_equal(d0, np.rint(d0)) and d0.min() >= 0 and d0.max() <= 255
        where = {(k.pt, k.size, k.angle, k.octave): i for i, k in enumerate(kp0)}
        idx = np.array([where[(k.pt, k.size, k.angle, k.octave)] for k in kp1], dtype=np.int32)
        assert np.array_equal(d0[idx], d1)
        out[tag] = d0.astype(np.uint8)
        out[tag + "_idx16k"] = idx
        full.append(out[tag])
        sub.append(out[tag][idx])
    # sanity counts (cv2 here); the tests recompute everything with cv2 on the box
    for name, (q, t) in (("full", full), ("16k", sub)):
        rq, _, _ = oracle_cv2.match(q, t, 0.75, False)
        mq, _, _ = oracle_cv2.match(q, t, 0.75, True)
        out[f"n_ratio_{name}"] = np.int64(len(rq))
        out[f"n_mutual_{name}"] = np.int64(len(mq))
        print(name, q.shape, t.shape, "ratio survivors", len(rq), "ratio + mutual", len(mq))
    path = os.path.join(HERE, "sift_full.npz")
    np.savez_compressed(path, **out)
    print(os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
