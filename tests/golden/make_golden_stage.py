"""Generates tests/golden/stage_ref.npz by running the reference's OWN script -- /root/reference/matching.py,
unmodified, under runpy -- on real SIFT features of its own photographs. Run in the build container
(where /root/reference exists):

    python tests/golden/make_golden_stage.py

matching.py imports only `LightGlue` and `rbd` from its matcher package (matching.py:4-5; the package itself
needs kornia + downloaded weights, neither available offline) and reads its inputs from ./output/
(matching.py:13-16). So the script is run from a scratch directory with
  * output/all_descriptors.npy, all_points.npy, img_size.npy in feature_extraction.py:34-50's layout
    (cv2.SIFT_create(nfeatures=N_FEAT) of the first N_IMG ystad_kloster images, the extractor BASELINE.json
    names), output/bow_codebook.plk built the way bow.py:7-25 does (scipy kmeans, k = 200, 1 iteration);
  * sys.modules['lightglue'] / ['lightglue.utils'] stubbed: LightGlue(features=...) returns a matcher with the
    LightGlue dict contract (lightglue.py:326-341, 442-450) whose rule is the cv2 one BASELINE.json names
    (oracle_cv2.match), rbd is the four lines of lightglue/utils.py:64-67 restated.
Everything else -- BoW view graph, BFS order, gates (> 8, > 10, >= 500, >= 30 %), cv2 essential-matrix check,
3-D point ids, the files it saves -- is the reference's own code executing. Its outputs
(output/img_pairs.npy, output/all_matches.npy) and its inputs are committed as the fixture that pins
oracle/oracle_driver.py (CPU test) and 3d_reconstruction_b200/driver.py (GPU test).
"""
import os
import runpy
import sys
import tempfile
import types

import cv2
import joblib
import numpy as np
import torch
from scipy.cluster.vq import kmeans

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle_cv2  # noqa: E402

REF = "/root/reference"
IMG_DIR = os.path.join(REF, "ystad_kloster")
N_IMG, N_FEAT, K_WORDS = 12, 4096, 200


def extract():
    names = sorted(f for f in os.listdir(IMG_DIR) if f.split(".")[-1] in ("JPG", "jpg"))[:N_IMG]
    feats = np.array([None] * N_IMG, dtype=object)
    points = np.array([None] * N_IMG, dtype=object)
    sizes = np.array([None] * N_IMG, dtype=object)
    for i, n in enumerate(names):
        img = cv2.imread(os.path.join(IMG_DIR, n))                                  # feature_extraction.py:25
        kp, d = cv2.SIFT_create(nfeatures=N_FEAT).detectAndCompute(cv2.cvtColor(img, cv2.COLOR_BGR2GRAY), None)
        kp, d = kp[:N_FEAT], d[:N_FEAT]
        assert len(kp) == N_FEAT and np.array_equal(d, np.rint(d)) and d.min() >= 0 and d.max() <= 255
        xy = np.array([k.pt for k in kp], dtype=np.float32)
        size = np.array([img.shape[1], img.shape[0]], dtype=np.float32)             # feats['image_size'] is (w, h)
        points[i] = np.array([np.array([pt[0] - size[0] / 2, -(pt[1] - size[1] / 2)]) for pt in xy])   # :34
        feats[i] = d                                                                 # float32, integer valued
        sizes[i] = size
    return names, feats, points, sizes


def stub_modules():
    class LightGlue:                                                                 # matching.py:20
        def __init__(self, features=None, **conf):
            pass

        def eval(self):
            return self

        def to(self, device):
            return self

        def __call__(self, data):                                                    # matching.py:122
            q = data["image0"]["descriptors"][0].cpu().numpy()
            t = data["image1"]["descriptors"][0].cpu().numpy()
            qi, ti, di = oracle_cv2.match(q, t, 0.75, True)
            m = torch.from_numpy(np.stack([qi, ti], -1).astype(np.int64)).reshape(-1, 2)
            m0 = torch.full((1, q.shape[0]), -1, dtype=torch.long)
            m1 = torch.full((1, t.shape[0]), -1, dtype=torch.long)
            m0[0, m[:, 0]] = m[:, 1]
            m1[0, m[:, 1]] = m[:, 0]
            return {"matches0": m0, "matches1": m1, "matching_scores0": torch.zeros(1, q.shape[0]),
                    "matching_scores1": torch.zeros(1, t.shape[0]), "stop": 0, "matches": [m],
                    "scores": [torch.from_numpy(di)]}

    def rbd(data: dict) -> dict:                                                     # lightglue/utils.py:64-67
        return {k: v[0] if isinstance(v, (torch.Tensor, np.ndarray, list)) else v for k, v in data.items()}

    pkg = types.ModuleType("lightglue")
    pkg.LightGlue = LightGlue
    utils = types.ModuleType("lightglue.utils")
    utils.rbd = rbd
    pkg.utils = utils
    sys.modules["lightglue"] = pkg
    sys.modules["lightglue.utils"] = utils


def main():
    names, feats, points, sizes = extract()
    np.random.seed(0)
    stack = feats[0].astype("float")                                                 # bow.py:13-17
    for i in range(1, N_IMG):
        stack = np.vstack((feats[i].astype("float"), stack))
    codebook, _ = kmeans(stack, K_WORDS, 1)                                          # bow.py:23
    stub_modules()
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.makedirs(os.path.join(tmp, "output"))
        np.save(os.path.join(tmp, "output", "all_descriptors.npy"), feats)
        np.save(os.path.join(tmp, "output", "all_points.npy"), points)
        np.save(os.path.join(tmp, "output", "img_size.npy"), sizes)
        joblib.dump((len(codebook), codebook), os.path.join(tmp, "output", "bow_codebook.plk"), compress=3)   # bow.py:25
        os.chdir(tmp)
        try:
            cv2.setRNGSeed(0)
            g = runpy.run_path(os.path.join(REF, "matching.py"), run_name="__main__")
        finally:
            os.chdir(cwd)
        img_pairs = np.load(os.path.join(tmp, "output", "img_pairs.npy"))
        all_matches = np.load(os.path.join(tmp, "output", "all_matches.npy"), allow_pickle=True)
    out = {"names": np.array(names), "codebook": codebook.astype(np.float64), "k": np.int64(len(codebook)),
           "cv2_version": cv2.__version__, "img_pairs": np.asarray(img_pairs, dtype=np.int64).reshape(-1, 2),
           "start": np.int64(g["start"]), "n_pairs": np.int64(len(all_matches))}
    out["connection"] = np.array([np.array(c if c else [], dtype=np.int64) for c in g["connection"]], dtype=object)
    for i in range(N_IMG):
        out[f"desc{i}"] = feats[i].astype(np.uint8)
        out[f"pts{i}"] = points[i].astype(np.float32)
        out[f"size{i}"] = sizes[i].astype(np.float32)
    for p, row in enumerate(all_matches):
        out[f"idx0_{p}"] = np.asarray(row[0], dtype=np.int64)
        out[f"idx1_{p}"] = np.asarray(row[1], dtype=np.int64)
        out[f"ids_{p}"] = np.asarray(row[2], dtype=np.int64)
    path = os.path.join(HERE, "stage_ref.npz")
    np.savez_compressed(path, **out)
    print("start", g["start"], "pairs", out["img_pairs"].tolist(), [len(r[0]) for r in all_matches], os.path.getsize(path))


if __name__ == "__main__":
    main()
