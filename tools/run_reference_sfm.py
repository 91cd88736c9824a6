"""Drop-in check of the stage files: runs the UNMODIFIED reference sfm.py (/root/reference/sfm.py, build container
only) on img_pairs.npy / all_matches.npy written by this repo's save_matches.

The matching stage itself needs a B200; here its result comes from the restated reference loop with the cv2
matcher (oracle/oracle_driver.py), which tests/test_gpu_driver.py::test_driver_on_real_photographs shows to be
row-for-row what the GPU driver produces on the same fixture (tests/golden/ystad_scene.npz: six of the
reference's photographs). Everything is written under gpurun_out/sfm_run/ (scratch)."""
import importlib
import os
import subprocess
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle_driver  # noqa: E402

drv = importlib.import_module("3d_reconstruction_b200.driver")


def main():
    g = np.load(os.path.join(ROOT, "tests", "golden", "ystad_scene.npz"))
    n = 6
    run = os.path.join(ROOT, "gpurun_out", "sfm_run")
    out = os.path.join(run, "output")
    os.makedirs(out, exist_ok=True)
    desc = np.array([None] * n, dtype=object)
    pts = np.array([None] * n, dtype=object)
    size = np.array([None] * n, dtype=object)
    cols = np.array([None] * n, dtype=object)
    names = [str(x) for x in g["names"]]
    for i in range(n):
        desc[i], pts[i], size[i] = g[f"desc{i}"].astype(np.float32), g[f"pts{i}"], g[f"size{i}"]
        img = cv2.imread(os.path.join("/root/reference/ystad_kloster", names[i]))
        px = pts[i][:, 0] + size[i][0] / 2
        py = -pts[i][:, 1] + size[i][1] / 2
        cols[i] = [img[int(y), int(x), :] for x, y in zip(px, py)]          # feature_extraction.py:32
    conn, start = oracle_driver.view_graph(desc, int(g["k"]), g["codebook"], cos_thr=0.0)
    pairs, matches = oracle_driver.reference_loop(desc, pts, size, conn, start, oracle_driver.cv2_matcher(), min_matches=60)
    drv.save_matches(pairs, matches, out)                                    # this repo's writer
    np.save(os.path.join(out, "all_points.npy"), pts)
    np.save(os.path.join(out, "all_colors.npy"), cols)
    np.save(os.path.join(out, "img_size.npy"), size)
    with open(os.path.join(out, "img_list.txt"), "wt") as f:
        f.write("\n".join(names) + "\n")
    r = subprocess.run([sys.executable, "/root/reference/sfm.py"], cwd=run, capture_output=True, text=True)
    print(r.stdout[-400:], r.stderr[-600:])
    assert r.returncode == 0, "reference sfm.py failed on these stage files"
    cams = np.load(os.path.join(out, "cameras_extrinsic.npy"), allow_pickle=True)
    p3d = np.load(os.path.join(out, "points_3d.npy"))
    print("reference sfm.py ran unchanged:", len(pairs), "pairs ->", len(cams), "cameras,", len(p3d), "3-D points,",
          os.path.getsize(os.path.join(out, "result.ply")), "bytes of PLY")


if __name__ == "__main__":
    main()
