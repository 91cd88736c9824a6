"""Generates the committed fixtures from the reference's own sample images (run in the build
container, where /root/reference exists; the GPU box only reads the resulting .npz files).

    python tests/golden/make_golden.py

sift_pair.npz : cv2.SIFT_create(nfeatures=N) descriptors of ystad_kloster/DSC_0001.JPG and
                DSC_0002.JPG (BASELINE.json config 1) and what cv2 4.13.0 returns for them:
                BFMatcher(NORM_L2).knnMatch(k=2) indices + float32 distances, and the match list
                after Lowe ratio 0.75 + cross-check.
ties.npz      : an adversarial small-alphabet case with duplicated rows (tie-break pins), same outputs.
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle_cv2  # noqa: E402

REF = "/root/reference/ystad_kloster"
N = 3072


def main():
    sift = cv2.SIFT_create(nfeatures=N)
    descs = []
    for name in ("DSC_0001.JPG", "DSC_0002.JPG"):
        img = cv2.imread(os.path.join(REF, name), cv2.IMREAD_GRAYSCALE)
        _, d = sift.detectAndCompute(img, None)
        assert d.dtype == np.float32 and np.array_equal(d, np.rint(d)) and d.min() >= 0 and d.max() <= 255
        descs.append(d.astype(np.uint8))
    q, t = descs
    idx, dist = oracle_cv2.knn2(q, t)
    mq, mt, md = oracle_cv2.match(q, t, 0.75, True)
    rq, rt, rd = oracle_cv2.match(q, t, 0.75, False)
    np.savez_compressed(os.path.join(HERE, "sift_pair.npz"), q=q, t=t, knn_idx=idx, knn_dist=dist,
                        m_q=mq, m_t=mt, m_d=md, r_q=rq, r_t=rt, r_d=rd, cv2_version=cv2.__version__)
    print("sift_pair", q.shape, t.shape, "ratio survivors", len(rq), "mutual", len(mq))

    rng = np.random.default_rng(7)
    base = rng.integers(0, 3, size=(40, 128)).astype(np.uint8) * 40
    q = base[rng.integers(0, 40, size=300)]
    t = base[rng.integers(0, 40, size=520)]
    t[100:110] = 0
    q[5] = 255
    t[7] = 255
    idx, dist = oracle_cv2.knn2(q, t)
    mq, mt, md = oracle_cv2.match(q, t, 0.0, True)
    np.savez_compressed(os.path.join(HERE, "ties.npz"), q=q, t=t, knn_idx=idx, knn_dist=dist,
                        m_q=mq, m_t=mt, m_d=md, cv2_version=cv2.__version__)
    print("ties", q.shape, t.shape, "mutual(no ratio)", len(mq))


if __name__ == "__main__":
    main()
