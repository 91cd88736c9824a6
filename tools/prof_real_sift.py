"""One real-SIFT workload for ncu (tests/golden/sift_full.npz, both directions): python tools/prof_real_sift.py [reps]"""
import importlib, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3d_reconstruction_b200")
g = np.load(os.path.join(ROOT, "tests", "golden", "sift_full.npz"))
bank = pkg.DescriptorBank.from_arrays([g["q"], g["t"]])
pairs = torch.tensor([[0, 1], [1, 0]] * 8, dtype=torch.int32, device="cuda")
ks = int(bank.k_stride)
ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(ks, len(pairs))), dtype=torch.uint8, device="cuda")
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    out = pkg.match_pairs(bank, pairs, cap=ks, workspace=ws)
torch.cuda.synchronize()
print("matches per pair", out.count[:2].tolist())
