"""Small end-to-end case for compute-sanitizer (development aid): every kernel of the default pipeline,
the fallback path (tie-heavy data) and the float path, a few hundred keypoints each."""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_reconstruction_b200")
syn = importlib.import_module("3d_reconstruction_b200.synthetic")

desc, kpi = syn.make_bank_tensor(3, 700, seed=5, device="cpu", k_per_image=[700, 515, 300])
bank = pkg.DescriptorBank.from_packed(desc.cuda(), kpi)
n = int(pkg.match_pairs(bank, [[0, 1], [1, 2], [2, 0]]).count.sum())
rng = np.random.default_rng(1)
q = rng.integers(0, 3, size=(300, 128)).astype(np.uint8)
t = rng.integers(0, 3, size=(400, 128)).astype(np.uint8)
t[:100] = q[:100]
m = len(pkg.match(q, t)[0])
idx, dist = pkg.knn_match(q, t)
d0 = torch.nn.functional.normalize(torch.randn(300, 256), dim=1)
f = len(pkg.match_float(d0, d0[torch.randperm(300)] + 0.01 * torch.randn(300, 256), score_threshold=0.5)[0])
torch.cuda.synchronize()
print("sanitize case done:", n, m, idx.shape, f)
