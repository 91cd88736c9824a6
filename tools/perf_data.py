"""How much the default pipeline's throughput depends on the data (development aid).

The chunk-bound pipeline settles a row cheaply when the bounds separate; rows they cannot settle are evaluated
exactly (refine), sent to the exact fallback sweep or to the ratio verify sweep. This tool reports, for several
kinds of banks, pairs/s + algorithmic TOP/s of pipeline 2 (default) and pipeline 1 (exact top-1 sweep + verify),
the row counters of pipeline 2, and that both pipelines return the same lists:
  synthetic   the bench generator (BASELINE config 3 shapes)
  real_sift   every SIFT descriptor of the reference's DSC_0001 / DSC_0002 (tests/golden/sift_full.npz), both directions
  uniform     uniform random bytes (wide norm spread inside a chunk)
  near_dup    SIFT-like rows, half of them re-observed from a small world: many near ties
  alphabet    rows drawn from 8 prototypes: exact duplicates everywhere (d1 == d2: the bounds reject every row)"""
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3d_reconstruction_b200")
syn = importlib.import_module("3d_reconstruction_b200.synthetic")
sys.path.insert(0, os.path.join(ROOT, "tools"))
from stress_parity import gen  # noqa: E402


def banks():
    desc, kpi = syn.make_bank_tensor(24, 16384, seed=1234, device="cuda")
    yield "synthetic", pkg.DescriptorBank.from_packed(desc, kpi), pkg.blocked_order(pkg.exhaustive_pairs(24))
    g = np.load(os.path.join(ROOT, "tests", "golden", "sift_full.npz"))
    yield "real_sift", pkg.DescriptorBank.from_arrays([g["q"], g["t"]]), np.array([[0, 1], [1, 0]] * 8, np.int32)
    rng = np.random.default_rng(5)
    world = np.abs(rng.standard_normal((600, 128)))
    world = np.clip(np.rint(world / np.linalg.norm(world, axis=1, keepdims=True) * 512), 0, 255)
    for name, kind in (("uniform", 0), ("near_dup", 1), ("alphabet", 2)):
        arrays = [gen(rng, 8192, kind, world) for _ in range(12)]
        yield name, pkg.DescriptorBank.from_arrays(arrays), pkg.exhaustive_pairs(12)


def run():
    for name, bank, pairs in banks():
        k = bank.k_host().astype(np.int64)
        ops = float(sum(2.0 * k[a] * k[b] * 128 for a, b in np.asarray(pairs).tolist()))
        pd = torch.from_numpy(np.ascontiguousarray(pairs, dtype=np.int32)).cuda()
        P = len(pairs)
        cap = int(bank.k_stride)
        ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(bank.k_stride, P)), dtype=torch.uint8, device="cuda")
        out = {}
        line = {"bank": name, "images": int(bank.n_img), "k_stride": int(bank.k_stride), "pairs": P}
        for pl in (2, 1):
            pkg.set_option("pipeline", pl)
            try:
                res = pkg.match_pairs(bank, pd, cap=cap, workspace=ws)
                torch.cuda.synchronize()
                ts = []
                for _ in range(3):
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    res = pkg.match_pairs(bank, pd, cap=cap, workspace=ws)
                    e1.record()
                    torch.cuda.synchronize()
                    ts.append(e0.elapsed_time(e1) / 1e3)
                t = min(ts)
                pkg.set_option("profile", 1)
                pkg.profile_read()
                res = pkg.match_pairs(bank, pd, cap=cap, workspace=ws)
                torch.cuda.synchronize()
                prof = pkg.profile_read()
                pkg.set_option("profile", 0)
                out[pl] = res.flatten().digest()[1]
                line[f"pipeline{pl}"] = {"pairs_per_s": round(P / t, 1), "algorithmic_TOPs": round(ops / t / 1e12, 1), "ms": round(t * 1e3, 3),
                                         "kernel_ms": {kk: round(v, 3) for kk, v in prof.items() if kk.endswith("_ms")}}
                if pl == 2:
                    st = pkg.match_stats(ws)
                    rows = max(1, st["rows"])
                    line["rows"] = {kk: st[kk] for kk in ("rows", "rows_chunk_evaluated", "rows_fallback", "rows_ratio_verify", "matches")}
                    line["rows_pct"] = {kk: round(100.0 * st[kk] / rows, 3) for kk in ("rows_chunk_evaluated", "rows_fallback", "rows_ratio_verify")}
            finally:
                pkg.set_option("pipeline", 2)
        line["pipelines_agree"] = bool(out[1] == out[2])
        print(json.dumps(line), flush=True)
        del ws, bank
        torch.cuda.empty_cache()


if __name__ == "__main__":
    run()
