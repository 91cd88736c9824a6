"""Quick device-side timing of the sweep on a synthetic bank (development aid, not the bench)."""
import importlib
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_reconstruction_b200")
syn = importlib.import_module("3d_reconstruction_b200.synthetic")


def run(n_img, k, modes=((1, 2, 5, 64), (1, 2, 5, 32), (2, 2, 5, 32)), reps=3, max_pairs=None, grid=0):
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=1234, device="cuda")
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    pairs = pkg.blocked_order(pkg.exhaustive_pairs(n_img))
    if max_pairs:
        pairs = pairs[:max_pairs]
    P = len(pairs)
    pairs_t = torch.from_numpy(pairs).cuda()
    cap = k
    out = None
    for mode in modes:
        pkg.set_option("epilogue", mode[0])
        pkg.set_option("pipeline", mode[1])
        pkg.set_option("kernel", mode[2])
        pkg.set_option("chunk", mode[3] if len(mode) > 3 else 64)
        pkg.set_option("grid", grid)
        out = pkg.match_pairs(bank, pairs_t, cap=cap, out=out)
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            pkg.match_pairs(bank, pairs_t, cap=cap, out=out)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) / 1e3)
        t = min(ts)
        ops = 2.0 * k * k * 128 * P
        print(json.dumps({"n_img": n_img, "k": k, "pairs": P, "mode": mode, "grid": grid, "sec": round(t, 5),
                          "pairs_per_s": round(P / t, 1), "TOPS": round(ops / t / 1e12, 1),
                          "frac_of_4500": round(ops / t / 4.5e15, 3),
                          "matches": int(out.count.sum().item())}), flush=True)
    pkg.set_option("epilogue", 1)
    pkg.set_option("pipeline", 2)
    pkg.set_option("kernel", 5)
    pkg.set_option("grid", 0)
    pkg.set_option("chunk", 64)


if __name__ == "__main__":
    run(16, 8192)
    run(24, 16384)
