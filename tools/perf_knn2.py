"""knnMatch(k=2) throughput on BASELINE config 3 shapes (development aid): the chunk pipeline (b200m_knn2_sorted_u8)
against the exact top-2 sweep (b200m_knn2_u8), with the per-phase device times of the library profiler."""
import importlib
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_reconstruction_b200")
syn = importlib.import_module("3d_reconstruction_b200.synthetic")


def run(n_img=33, k=16384, reps=3):
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=1234, device="cuda")
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    pairs = torch.from_numpy(pkg.blocked_order(pkg.exhaustive_pairs(n_img))[:512]).cuda()
    P = len(pairs)
    ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(k, P)), dtype=torch.uint8, device="cuda")
    res = {}
    for name, pipeline in (("chunk_pipeline", 2), ("exact_top2_sweep", 1)):
        pkg.set_option("pipeline", pipeline)
        idx, dist = pkg.knn2_pairs(bank, pairs, workspace=ws)
        torch.cuda.synchronize()
        pkg.set_option("profile", 1)
        pkg.profile_read()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            idx, dist = pkg.knn2_pairs(bank, pairs, workspace=ws)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) / 1e3)
        prof = pkg.profile_read()
        pkg.set_option("profile", 0)
        t = min(ts)
        res[name] = (idx.clone(), dist.clone())
        st = pkg.match_stats(ws) if pipeline == 2 else {}
        print(json.dumps({"knn2": name, "k": k, "pairs": P, "sec": round(t, 5), "pairs_per_s": round(P / t, 1),
                          "TOPS": round(2.0 * k * k * 128 * P / t / 1e12, 1),
                          "per_rep_ms": {kk: round(v / reps, 3) for kk, v in prof.items() if kk.endswith("_ms")},
                          "stats": st}), flush=True)
    pkg.set_option("pipeline", 2)
    a, b = res["chunk_pipeline"], res["exact_top2_sweep"]
    print("identical:", bool(torch.equal(a[0], b[0]) and torch.equal(a[1].view(torch.int32), b[1].view(torch.int32))))


if __name__ == "__main__":
    run()
