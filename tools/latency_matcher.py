"""Per-call latency of the drop-in callable (matching.py:122 shape): one pair per call, CUDA float tensors in."""
import importlib
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_reconstruction_b200")
syn = importlib.import_module("3d_reconstruction_b200.synthetic")


def run(k, reps=20, where="cuda"):
    desc, _ = syn.make_bank_tensor(2, k if k % 256 == 0 else (k // 256 + 1) * 256, seed=7, device="cuda")
    d0 = desc[0, :k].float()[None].to(where)
    d1 = desc[1, :k].float()[None].to(where)
    kp = torch.zeros((1, k, 2), device=where)
    m = pkg.B200Matcher()
    data = {"image0": {"descriptors": d0, "keypoints": kp, "image_size": torch.tensor([[1936., 1296.]])},
            "image1": {"descriptors": d1, "keypoints": kp, "image_size": torch.tensor([[1936., 1296.]])}}
    out = m(data)
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        out = m(data)
        n = out["matches"][0].shape[0]
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    ts.sort()
    print(json.dumps({"k": k, "inputs": where, "median_ms": round(1e3 * ts[len(ts) // 2], 3), "min_ms": round(1e3 * ts[0], 3),
                      "matches": int(n)}), flush=True)


if __name__ == "__main__":
    for where in ("cuda", "cpu"):
        for k in (2048, 8192, 16384):
            run(k, where=where)
