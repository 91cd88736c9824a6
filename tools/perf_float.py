"""BASELINE.json config 5 timing (development aid): random pairs over a bank of 2,048-keypoint float images."""
import importlib
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_reconstruction_b200")


def run(dim, n_img=128, k=2048, n_pairs=4096, reps=5):
    g = torch.Generator(device="cuda").manual_seed(dim)
    desc = torch.nn.functional.normalize(torch.randn((n_img, k, dim), generator=g, device="cuda"), dim=2).to(torch.bfloat16)
    bank = pkg.FloatBank(desc.contiguous(), torch.full((n_img,), k, dtype=torch.int32))
    rng = np.random.default_rng(0)
    pairs = np.stack([rng.integers(0, n_img, n_pairs), rng.integers(0, n_img, n_pairs)], 1).astype(np.int32)
    pt = torch.from_numpy(pairs).cuda()
    out = pkg.match_pairs_float(bank, pt, score_threshold=0.2, mutual=True, cap=k)
    ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes_bf16(k, n_pairs)), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        pkg.match_pairs_float(bank, pt, score_threshold=0.2, mutual=True, cap=k, out=out, workspace=ws)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / 1e3)
    t = min(ts)
    flops = 2.0 * k * k * dim * n_pairs          # algorithmic (one direction); the kernel sweeps both for the mutual check
    print(json.dumps({"dim": dim, "k": k, "pairs": n_pairs, "sec": round(t, 5), "pairs_per_s": round(n_pairs / t, 1),
                      "algorithmic_TFLOPs": round(flops / t / 1e12, 1), "executed_TFLOPs": round(2 * flops / t / 1e12, 1),
                      "bank_MB": desc.numel() * 2 / 1e6}), flush=True)


def run_assign(dim, n_img=128, k=2048, n_pairs=4096, reps=5):
    """The fused assignment head (log double soft-max + filter_matches): four sweeps per pair."""
    g = torch.Generator(device="cuda").manual_seed(dim)
    desc = (4.5 * torch.nn.functional.normalize(torch.randn((n_img, k, dim), generator=g, device="cuda"), dim=2)).to(torch.bfloat16)
    bank = pkg.FloatBank(desc.contiguous(), torch.full((n_img,), k, dtype=torch.int32))
    z = torch.randn((n_img, k), generator=g, device="cuda")
    rng = np.random.default_rng(0)
    pairs = np.stack([rng.integers(0, n_img, n_pairs), rng.integers(0, n_img, n_pairs)], 1).astype(np.int32)
    pt = torch.from_numpy(pairs).cuda()
    out = pkg.assign_pairs(bank, z, pt, filter_threshold=0.1, cap=k)
    ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes_assign(k, n_pairs)), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    pkg.set_option("profile", 1)
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        pkg.assign_pairs(bank, z, pt, filter_threshold=0.1, cap=k, out=out, workspace=ws)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / 1e3)
    prof = pkg.profile_read()
    pkg.set_option("profile", 0)
    t = min(ts)
    flops = 2.0 * k * k * dim * n_pairs
    print(json.dumps({"head": "log double soft-max", "dim": dim, "k": k, "pairs": n_pairs, "sec": round(t, 5),
                      "pairs_per_s": round(n_pairs / t, 1), "algorithmic_TFLOPs": round(flops / t / 1e12, 1),
                      "executed_TFLOPs": round(4 * flops / t / 1e12, 1), "profile": str(prof)}), flush=True)


if __name__ == "__main__":
    for mode in (0, 1):
        pkg.set_option("float_sweep", mode)
        print(f"float_sweep = {mode} ({'single sweep, fused column maximum' if mode == 0 else 'one sweep per direction'})", flush=True)
        run(256)
        run(128)
        run_assign(256)
        run_assign(128)
