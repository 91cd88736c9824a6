"""One small workload for ncu: python tools/prof_case.py <n_img> <k> <mode> [reps]"""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_reconstruction_b200")
syn = importlib.import_module("3d_reconstruction_b200.synthetic")
n_img, k, mode = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
desc, kpi = syn.make_bank_tensor(n_img, k, seed=1234, device="cuda")
bank = pkg.DescriptorBank.from_packed(desc, kpi)
pairs = torch.from_numpy(pkg.blocked_order(pkg.exhaustive_pairs(n_img))).cuda()
pkg.set_option("epilogue", mode)
out = None
for _ in range(reps):
    out = pkg.match_pairs(bank, pairs, cap=k, out=out)
torch.cuda.synchronize()
print("pairs", len(pairs), "matches", int(out.count.sum()))
