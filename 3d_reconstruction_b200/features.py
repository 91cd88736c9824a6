"""Producer of the stage files the matching stage reads (host side, cv2; not part of the GPU hot path).

The reference's extractor script (feature_extraction.py) runs DISK, whose weights have to be downloaded; the path
BASELINE.json names works on OpenCV SIFT descriptors. This module writes SIFT features in exactly the on-disk
layout feature_extraction.py:14-53 produces, so that `driver.run_matching_stage` -- and after it the reference's own
sfm.py -- can be run on a directory of photographs without copying test code:

    output/img_list.txt          one image file name per line                         (:14, :24)
    output/all_descriptors.npy   object array, one [K_i, 128] array per image         (:38, :50)
    output/all_points.npy        object array, [K_i, 2]: x - w/2, -(y - h/2)           (:34, :37, :41)
    output/img_size.npy          object array, (w, h) per image                       (:36, :44)
    output/all_colors.npy        object array, BGR pixel under every keypoint          (:32, :39, :47)
    output/bow_codebook.plk      (k, codebook) of bow.py:23-25 (build_codebook)

Descriptors are stored as uint8 (OpenCV's SIFT values are integers in [0, 255]); the reference code only ever calls
`.astype("float")` on them (bow.py:12, matching.py:27), so either dtype reads the same.
"""
from __future__ import annotations

import os

import numpy as np

IMAGE_SUFFIXES = ("JPG", "jpg", "PNG", "png", "RAW", "raw", "TIF", "tif")     # feature_extraction.py:23


def list_images(img_dir: str):
    """Sorted regular files of img_dir with one of the suffixes the reference accepts (feature_extraction.py:13, :23)."""
    names = sorted(n for n in os.listdir(img_dir) if os.path.isfile(os.path.join(img_dir, n)))
    return [n for n in names if n.split(".")[-1] in IMAGE_SUFFIXES]


def sift_features(img_bgr: np.ndarray, max_keypoints: int = 2048, sift=None):
    """One image -> (descriptors uint8 [K, 128], points float64 [K, 2] centre-origin / y-up, size (w, h), colours
    uint8 [K, 3]). K <= max_keypoints: OpenCV may return a few more than `nfeatures` (it keeps every orientation of a
    kept keypoint), the surplus is cut so that the reference's keypoint budget (feature_extraction.py:10) holds."""
    import cv2
    if sift is None:
        sift = cv2.SIFT_create(nfeatures=max_keypoints)
    gray = cv2.cvtColor(img_bgr, cv2.COLOR_BGR2GRAY) if img_bgr.ndim == 3 else img_bgr
    kps, desc = sift.detectAndCompute(gray, None)
    h, w = gray.shape[:2]
    size = np.array([w, h], dtype=np.float64)
    if desc is None or not len(kps):
        return np.zeros((0, 128), np.uint8), np.zeros((0, 2), np.float64), size, np.zeros((0, 3), np.uint8)
    kps, desc = kps[:max_keypoints], desc[:max_keypoints]
    if not (np.array_equal(desc, np.rint(desc)) and desc.min() >= 0 and desc.max() <= 255):
        raise ValueError("SIFT descriptors are expected to be integers in [0, 255] (OpenCV's convention)")
    xy = np.array([kp.pt for kp in kps], dtype=np.float64)
    pts = np.stack([xy[:, 0] - w / 2, -(xy[:, 1] - h / 2)], axis=1)
    colour_src = img_bgr if img_bgr.ndim == 3 else np.repeat(gray[:, :, None], 3, axis=2)
    colours = colour_src[xy[:, 1].astype(np.int64), xy[:, 0].astype(np.int64), :]      # pixel under the keypoint, truncated
    return desc.astype(np.uint8), pts, size, colours


def extract_features(img_dir: str, output_dir: str = "output", max_keypoints: int = 2048, limit: int | None = None):
    """feature_extraction.py as a function, with SIFT as the extractor: every image of img_dir -> the five stage files
    under output_dir. limit: only the first `limit` images (a quick run on part of a data set). -> list of file names."""
    import cv2
    names = list_images(img_dir)
    if limit is not None:
        names = names[:limit]
    n = len(names)
    descs, points, sizes, colours = (np.array([None] * n, dtype=object) for _ in range(4))
    sift = cv2.SIFT_create(nfeatures=max_keypoints)
    for i, name in enumerate(names):
        img = cv2.imread(os.path.join(img_dir, name))
        if img is None:
            raise IOError(f"cannot read image {os.path.join(img_dir, name)}")
        descs[i], points[i], sizes[i], colours[i] = sift_features(img, max_keypoints, sift)
    os.makedirs(output_dir, exist_ok=True)
    with open(os.path.join(output_dir, "img_list.txt"), "wt") as f:
        f.writelines(name + "\n" for name in names)
    np.save(os.path.join(output_dir, "all_points.npy"), points)
    np.save(os.path.join(output_dir, "img_size.npy"), sizes)
    np.save(os.path.join(output_dir, "all_colors.npy"), colours)
    np.save(os.path.join(output_dir, "all_descriptors.npy"), descs)
    return names


def build_codebook(output_dir: str = "output", k: int = 200, iters: int = 1, seed: int | None = None):
    """bow.py as a function: scipy k-means (k = 200, one iteration, bow.py:7-8, :23) over the stacked descriptors of
    all_descriptors.npy, dumped as (k, codebook) to bow_codebook.plk (:25). seed: numpy seed for a repeatable run
    (scipy draws the initial centres from numpy's global generator; the reference leaves it unseeded)."""
    import joblib
    from scipy.cluster.vq import kmeans
    all_desc = np.load(os.path.join(output_dir, "all_descriptors.npy"), allow_pickle=True)
    stack = np.vstack([np.asarray(d).astype("float") for d in all_desc if len(d)])
    if seed is not None:
        np.random.seed(seed)
    codebook, _ = kmeans(stack, k, iters)
    joblib.dump((k, codebook), os.path.join(output_dir, "bow_codebook.plk"), compress=3)
    return k, codebook
