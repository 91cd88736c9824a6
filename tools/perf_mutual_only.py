"""Mutual-only matching (ratio test off) on BASELINE config 3 shapes, refine = 0 (fused) against refine = 1 (streaming
classification + one warp per row): python tools/perf_mutual_only.py [n_img] [k]"""
import importlib, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_reconstruction_b200")
syn = importlib.import_module("3d_reconstruction_b200.synthetic")
n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 33
k = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
desc, kpi = syn.make_bank_tensor(n_img, k, seed=1234, device="cuda")
bank = pkg.DescriptorBank.from_packed(desc, kpi)
und = pkg.blocked_order(pkg.exhaustive_pairs(n_img))[:256]
pairs = torch.from_numpy(np.ascontiguousarray(np.concatenate([und, und[:, ::-1]], 0), dtype=np.int32)).cuda()
ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(k, len(pairs))), dtype=torch.uint8, device="cuda")
ref = None
for refine in (0, 1, 0, 1):
    pkg.set_option("refine", refine)
    out = pkg.match_pairs(bank, pairs, ratio=0.0, mutual=True, cap=k, workspace=ws)
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); pkg.match_pairs(bank, pairs, ratio=0.0, mutual=True, cap=k, out=out, workspace=ws); e1.record()
        torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    pkg.set_option("profile", 1); pkg.profile_read()
    pkg.match_pairs(bank, pairs, ratio=0.0, mutual=True, cap=k, out=out, workspace=ws); torch.cuda.synchronize()
    pr = pkg.profile_read(); pkg.set_option("profile", 0)
    flat = out.flatten()
    sig = (int(flat.offsets[-1]), int(flat.query_idx.sum()), int(flat.train_idx.sum()), int(flat.dist.view(torch.int32).to(torch.int64).sum()))
    if ref is None:
        ref = sig
    print(f"refine={refine}: {np.median(ts):.2f} ms for {len(pairs)} pairs = {len(pairs) / np.median(ts) * 1e3:.0f} pairs/s; "
          f"refine_ms {pr['refine_ms']:.2f} first {pr['first_sweep_ms']:.2f} verify {pr['verify_sweep_ms']:.2f}; same lists: {sig == ref}")
pkg.set_option("refine", 0)
