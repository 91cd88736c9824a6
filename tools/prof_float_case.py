"""One small float workload for ncu: python tools/prof_float_case.py <dim> <n_pairs> [head]"""
import importlib, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_reconstruction_b200")
dim, n_pairs = int(sys.argv[1]), int(sys.argv[2])
head = len(sys.argv) > 3
n_img, k = 64, 2048
g = torch.Generator(device="cuda").manual_seed(dim)
unit = torch.nn.functional.normalize(torch.randn((n_img, k, dim), generator=g, device="cuda"), dim=2)
kpi = torch.full((n_img,), k, dtype=torch.int32)
rng = np.random.default_rng(0)
pairs = torch.from_numpy(np.stack([rng.integers(0, n_img, n_pairs), rng.integers(0, n_img, n_pairs)], 1).astype(np.int32)).cuda()
if head:
    bank = pkg.FloatBank((4.5 * unit).to(torch.bfloat16).contiguous(), kpi)
    z = torch.randn((n_img, k), generator=g, device="cuda")
    for _ in range(2):
        out = pkg.assign_pairs(bank, z, pairs, filter_threshold=0.1, cap=k)
else:
    bank = pkg.FloatBank(unit.to(torch.bfloat16).contiguous(), kpi)
    for _ in range(2):
        out = pkg.match_pairs_float(bank, pairs, score_threshold=0.2, mutual=True, cap=k)
torch.cuda.synchronize()
print("pairs", n_pairs, "matches", int(out.count.sum()))
