#!/usr/bin/env python
"""bench.py -- image-pairs/s of the pairwise descriptor matcher on BASELINE.json's headline workload.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --steps K --warmup W     (cv2.BFMatcher on the host cores)

Workload (config.workload): synthetic 257 images x 16,384 SIFT-like u8 descriptors, exhaustive
32,896 pairs (BASELINE.json configs[2], the configuration the metric is quoted on). A step is one
pass of the full matching rule (2-NN, Lowe ratio 0.75, mutual check) over all pairs.
  value     pairs/s with the packed bank resident in HBM (CUDA events, max over ranks)
  e2e       same through the host-buffer API (distributed.match_pairs_host -> matcher.match_pairs_host):
            host descriptors -> device (every rank uploads its image slice; one rank: chunked upload
            overlapped with the matching), norms + norm sort, all-gather of the bank over NVLink, match,
            one packed gather to rank 0, flat match lists -> pinned host memory
  roofline  the first-sweep kernel against the tensor roofline (int8 ops = 2 * Kq * Kt * 128 per pair)
  cpu_baseline  cv2.BFMatcher(NORM_L2).knnMatch k=2 + reverse k=1 + ratio filter on a bounded sample
  config.result_digest / matches_total   64-bit digest of EVERY match record of the value leg, summed over
            the ranks: equal at N = 1, 2, 4, 8 <=> the N-GPU lists are the 1-GPU lists
  secondary the other BASELINE.json configs and regimes, measured outside the timed regions (short): C4 at every N; at
            N = 1 also C2, the knnMatch(k=2) table, the real SIFT pair of config 1, mutual-only matching, the u8 path at
            the reference's shipped 2,048 keypoints, C5 (float, raw arg-max and the full head) and the BoW vq kernel
N > 1 splits the fixed pair list over the ranks (static cost-balanced partition): "scaling": "strong".
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

N_IMG, K, D = 257, 16384, 128
SEED = 1234
RATIO = 0.75
CAP = 4096                     # match slots per pair (neighbouring images share ~2.2k true matches)
METRIC = "image-pairs/sec @16k SIFT kpts"
WORKLOAD = "synthetic 257 images x 16384 u8 SIFT-like descriptors, exhaustive 32896 pairs, 2-NN + ratio 0.75 + mutual"
MATCHES_C3 = 1126467           # matches of the workload (seed 1234): every N must reproduce it
SM_COUNT, INT8_OPS_PER_CLK_SM = 148, 16384     # tcgen05 kind::i8: 8,192 MACs per clock per SM


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def _sub_out(pkg, out, n):
    return pkg.PairMatches(out.query_idx[:n], out.train_idx[:n], out.dist[:n], out.count[:n], out.cap)


def int8_peak_probe(pkg, bank, pairs_dev, out, workspace, ops_per_pair, cublas=True):
    """MEASURED_PEAKS.json has no int8 entry: live int8 dense rates on this box for context, outside every timed
    region. (1) this library's own sweep with the epilogue reduced to one tcgen05.ld per accumulator (option
    epilogue = 2: TMA + MMA + hand-over only, results invalid) over one batch of 256 pairs: a three-launch burst
    and ~1 s sustained; (2) cuBLASLt IGEMM through torch._int_mm, 8192^3, a short burst and ~1.5 s sustained."""
    import torch
    res = {}
    n = min(256, int(pairs_dev.shape[0]))
    sub = _sub_out(pkg, out, n)
    pkg.set_option("epilogue", 2)
    try:
        pkg.match_pairs(bank, pairs_dev[:n], ratio=RATIO, mutual=True, cap=CAP, out=sub, workspace=workspace)
        for name, iters in (("mma_only_sweep_tops", 3), ("mma_only_sweep_sustained_tops", 200)):
            torch.cuda.synchronize()
            pkg.set_option("profile", 1)
            pkg.profile_read()
            for _ in range(iters):
                pkg.match_pairs(bank, pairs_dev[:n], ratio=RATIO, mutual=True, cap=CAP, out=sub, workspace=workspace)
            torch.cuda.synchronize()
            pr = pkg.profile_read()
            res[name] = ops_per_pair * n * iters / (pr["first_sweep_ms"] / 1e3) / 1e12
    finally:
        pkg.set_option("profile", 0)
        pkg.set_option("epilogue", 1)
    if not cublas:
        return res
    try:
        a = torch.randint(-128, 127, (8192, 8192), dtype=torch.int8, device=bank.device)
        b = torch.randint(-128, 127, (8192, 8192), dtype=torch.int8, device=bank.device)
        for _ in range(3):
            torch._int_mm(a, b)
        for name, iters in (("cublaslt_int_mm_8192_burst_tops", 20), ("cublaslt_int_mm_8192_sustained_tops", 1500)):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                torch._int_mm(a, b)
            e1.record()
            torch.cuda.synchronize()
            res[name] = 2.0 * 8192 ** 3 * iters / (e0.elapsed_time(e1) / 1e3) / 1e12
        del a, b
    except Exception as ex:          # a missing cuBLASLt int8 path must not cost the bench line
        res["cublaslt_int_mm_error"] = str(ex)[:120]
    return res


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self.stop_flag = threading.Event()

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip().splitlines()
                if out:
                    f = [x.strip() for x in out[0].split(",")]
                    self.samples.append((float(f[0]), float(f[1]), float(f[2])))
                    for n, v in zip(names, f[3:7]):
                        if v.lower().startswith("active"):
                            self.reasons.add(n)
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": sorted(self.reasons)}
        sm = sorted(s[0] for s in self.samples)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(s[1] for s in self.samples),
                "power_w_max": max(s[2] for s in self.samples), "samples": len(sm), "reasons": sorted(self.reasons)}


def make_host_bank(n_img, k, torch, device):
    """The synthetic bank of the workload; generated on `device`, returned as (desc, k_per_image)."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    return syn.make_bank_tensor(n_img, k, seed=SEED, device=device)


def cpu_baseline(desc_cpu, pairs, budget_s=20.0, max_pairs=8, min_pairs=2):
    """cv2 on the host cores over a bounded sample of the workload's pairs (float32 inputs: cv2's fastest path)."""
    from oracle import oracle_cv2
    times, threads = [], None
    t_total = time.perf_counter()
    for p in range(min(max_pairs, len(pairs))):
        a, b = int(pairs[p][0]), int(pairs[p][1])
        t, threads = oracle_cv2.time_knn2(desc_cpu[a], desc_cpu[b], repeats=1, full_rule=True)
        times.append(t)
        if p + 1 >= min_pairs and time.perf_counter() - t_total > budget_s:
            break
    per_pair = float(np.median(times))
    # beside the headline: the forward knnMatch(k=2) call alone, and the whole rule on uint8 inputs
    a, b = int(pairs[0][0]), int(pairs[0][1])
    t_knn = min(oracle_cv2.time_knn2(desc_cpu[a], desc_cpu[b], repeats=1, full_rule=False)[0] for _ in range(2))
    t_u8 = oracle_cv2.time_knn2(desc_cpu[a], desc_cpu[b], repeats=1, full_rule=True, as_u8=True)[0]
    return {"value": 1.0 / per_pair, "unit": "pairs/s", "cores": int(threads or os.cpu_count() or 1), "kind": "reference",
            "sample": f"{len(times)} pairs of the workload (16384x16384x128), cv2.BFMatcher(NORM_L2).knnMatch k=2 on float32 "
                      f"+ reverse knnMatch k=1 + ratio filter, median {per_pair:.3f} s/pair",
            "knn_only_pairs_per_s": 1.0 / t_knn, "u8_input_pairs_per_s": 1.0 / t_u8,
            "gops": 2.0 * K * K * D / per_pair / 1e9}


def run_reference(args):
    """--impl reference: the reference arm = cv2.BFMatcher on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from oracle import oracle_cv2
    import cv2
    sched = importlib.import_module("3d_reconstruction_b200.scheduler")
    # a bounded sample: a few images of the workload are enough to draw the pairs from
    n_img_sample = 6
    desc, _ = make_host_bank(n_img_sample, K, torch, "cpu")
    desc = desc.numpy()
    pairs = sched.exhaustive_pairs(n_img_sample)
    pairs_per_step = 1
    times = []
    threads = None
    for s in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        for j in range(pairs_per_step):
            a, b = pairs[(s * pairs_per_step + j) % len(pairs)]
            _, threads = oracle_cv2.time_knn2(desc[a], desc[b], repeats=1, full_rule=True)
        dt = time.perf_counter() - t0
        if s >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    val = pairs_per_step / (ms / 1e3)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "pairs/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "step": f"{pairs_per_step} pair(s) of the workload per step (bounded sample)",
                   "cv2": cv2.__version__},
        "cpu_baseline": {"value": val, "unit": "pairs/s", "cores": int(threads or 1), "kind": "reference",
                         "sample": f"{pairs_per_step} pair(s)/step x {args.steps} steps, cv2.BFMatcher(NORM_L2).knnMatch k=2 "
                                   "(float32) + reverse k=1 + ratio filter"},
        "e2e": {"value": val, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------- secondary configs (outside the timed regions)
def _timed(torch, fn, reps):
    """Best and median device time of fn() in seconds (CUDA events on the current stream)."""
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / 1e3)
    return min(ts), float(np.median(ts))


def _int8_line(n_pairs, k, best, med, int8_peak, gpus=1):
    ops = 2.0 * k * k * 128 * n_pairs
    return {"pairs": n_pairs, "pairs_per_s": n_pairs / med, "pairs_per_s_best": n_pairs / best, "ms": 1e3 * med,
            "tops": ops / med / 1e12, "frac_of_peak": ops / med / 1e12 / (int8_peak * gpus) if int8_peak else None,
            "frac_of_nominal_4500": ops / med / 1e12 / (4500.0 * gpus)}


def secondary_c2(pkg, torch, dev, int8_peak):
    """BASELINE.json configs[1]: 64 images x 8,192 keypoints, exhaustive 2,016 pairs, one GPU."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    n_img, k = 64, 8192
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=SEED, device=dev)
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    pairs = torch.from_numpy(pkg.blocked_order(pkg.exhaustive_pairs(n_img))).to(dev)
    out = pkg.match_pairs(bank, pairs, cap=CAP)
    ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(k, len(pairs))), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    out.check_overflow()
    best, med = _timed(torch, lambda: pkg.match_pairs(bank, pairs, cap=CAP, out=out, workspace=ws), 5)
    r = _int8_line(len(pairs), k, best, med, int8_peak)
    r["matches"] = int(out.count.sum().item())
    r["workload"] = "64 images x 8192 u8 descriptors, exhaustive 2016 pairs (bank resident, whole rule)"
    return r


def secondary_knn2(pkg, torch, dev, int8_peak):
    """The literal cv2.BFMatcher(NORM_L2).knnMatch(k=2) table (best + second-best train row of EVERY query row) on
    BASELINE config 3 shapes: 512 pairs over 33 images x 16,384; chunk pipeline against the exact top-2 sweep, outputs
    compared bit for bit."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    n_img, k = 33, 16384
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=SEED, device=dev)
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    pairs = torch.from_numpy(pkg.blocked_order(pkg.exhaustive_pairs(n_img))[:512]).to(dev)
    ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(k, len(pairs))), dtype=torch.uint8, device=dev)
    res = {}
    r = {}
    for name, pipeline in (("chunk_pipeline", 2), ("exact_top2_sweep", 1)):
        pkg.set_option("pipeline", pipeline)
        try:
            res[name] = pkg.knn2_pairs(bank, pairs, workspace=ws)
            torch.cuda.synchronize()
            best, med = _timed(torch, lambda: pkg.knn2_pairs(bank, pairs, workspace=ws), 3)
        finally:
            pkg.set_option("pipeline", 2)
        r[name] = _int8_line(len(pairs), k, best, med, int8_peak)
    a, b = res["chunk_pipeline"], res["exact_top2_sweep"]
    r["identical_outputs"] = bool(torch.equal(a[0], b[0]) and torch.equal(a[1].view(torch.int32), b[1].view(torch.int32)))
    r["workload"] = "knnMatch(k=2) tables of 512 pairs over 33 images x 16384 u8 descriptors"
    return r


def secondary_real_sift(pkg, torch, dev, int8_peak):
    """BASELINE.json configs[0] as a throughput figure: every SIFT descriptor of the reference's DSC_0001 / DSC_0002
    (tests/golden/sift_full.npz, 56,041 x 52,421), both directions, the whole rule. A strongly overlapping real pair:
    a third of the rows survive the Lowe test, so refine and the mutual check cost far more than on the headline bank."""
    path = os.path.join(ROOT, "tests", "golden", "sift_full.npz")
    if not os.path.exists(path):
        return {"unavailable": "tests/golden/sift_full.npz not found"}
    g = np.load(path)
    bank = pkg.DescriptorBank.from_arrays([g["q"], g["t"]], device=dev)
    pairs = torch.tensor([[0, 1], [1, 0]] * 8, dtype=torch.int32, device=dev)
    ks = int(bank.k_stride)
    ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(ks, len(pairs))), dtype=torch.uint8, device=dev)
    out = pkg.match_pairs(bank, pairs, cap=ks, workspace=ws)
    torch.cuda.synchronize()
    out.check_overflow()
    best, med = _timed(torch, lambda: pkg.match_pairs(bank, pairs, cap=ks, out=out, workspace=ws), 3)
    st = pkg.match_stats(ws)
    kq, kt = int(g["q"].shape[0]), int(g["t"].shape[0])
    ops = 2.0 * kq * kt * 128 * len(pairs)
    return {"pairs": len(pairs), "pairs_per_s": len(pairs) / med, "ms": 1e3 * med, "tops": ops / med / 1e12,
            "frac_of_peak": ops / med / 1e12 / int8_peak if int8_peak else None,
            "matches_per_pair": int(out.count[0].item()), "matches_expected_cv2": int(g["n_mutual_full"]),
            "rows": {kk: st[kk] for kk in ("rows", "rows_chunk_evaluated", "rows_fallback", "rows_ratio_verify")},
            "workload": f"real SIFT, DSC_0001 <-> DSC_0002, {kq} x {kt} descriptors, ratio 0.75 + mutual, both directions x 8"}


def _flat_digest(torch, out):
    """Order-sensitive 64-bit digest of a PairMatches (counts, indices and the float32 distance bits), on the device."""
    flat = out.flatten()
    h = torch.zeros((), dtype=torch.int64, device=flat.offsets.device)
    for i, t in enumerate((flat.offsets, flat.query_idx, flat.train_idx, flat.dist.view(torch.int32))):
        v = t.to(torch.int64)
        w = torch.arange(1, v.numel() + 1, dtype=torch.int64, device=v.device) * (2 * i + 3)
        h = h * 1000003 + (v * w).sum()
    return f"{int(h.item()) & 0xFFFFFFFFFFFFFFFF:016x}"


def secondary_mutual_only(pkg, torch, dev, int8_peak):
    """The reference matcher's OWN rule -- mutual arg-best, no ratio test (lightglue.py:235-254; cv2
    BFMatcher(crossCheck=True)) -- on BASELINE config 3 shapes: 512 directed pairs over 33 images x 16,384, once with
    both directions of 256 image pairs in the list and once with 512 pairs in one direction only. With ratio <= 0 every
    query row is a candidate, so the mutual check is a second full sweep (the verify sweep): this row puts that cost on
    the record. Size-independent check, outside the timed region: the mutual lists of (a, b) and (b, a) are each
    other's transposes."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    n_img, k = 33, 16384
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=SEED, device=dev)
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    und = pkg.blocked_order(pkg.exhaustive_pairs(n_img))
    lists = {"both_directions": np.concatenate([und[:256], und[:256, ::-1]], 0), "one_direction": und[:512]}
    ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(k, 512)), dtype=torch.uint8, device=dev)
    r = {}
    for name, pl in lists.items():
        pairs = torch.from_numpy(np.ascontiguousarray(pl, dtype=np.int32)).to(dev)
        out = pkg.match_pairs(bank, pairs, ratio=0.0, mutual=True, cap=k, workspace=ws)
        torch.cuda.synchronize()
        out.check_overflow()
        best, med = _timed(torch, lambda: pkg.match_pairs(bank, pairs, ratio=0.0, mutual=True, cap=k, out=out, workspace=ws), 3)
        r[name] = _int8_line(len(pairs), k, best, med, int8_peak)
        r[name]["matches_per_pair"] = float(out.count.sum().item()) / len(pairs)
        r[name]["digest"] = _flat_digest(torch, out)
        pkg.set_option("profile", 1)                      # one more call, outside the timing: device ms per kernel class
        try:
            pkg.profile_read()
            pkg.match_pairs(bank, pairs, ratio=0.0, mutual=True, cap=k, out=out, workspace=ws)
            torch.cuda.synchronize()
            pr = pkg.profile_read()
        finally:
            pkg.set_option("profile", 0)
        r[name]["ms_by_kernel_class"] = {kk: pr[kk] for kk in ("first_sweep_ms", "verify_sweep_ms", "fallback_sweep_ms", "refine_ms",
                                                             "select_ms", "emit_ms")}
        r[name]["rows"] = {kk: vv for kk, vv in pkg.match_stats(ws).items() if kk.startswith("rows")}
        if name == "both_directions":
            ok = bool(torch.equal(out.count[:256], out.count[256:]))
            for p in (0, 100, 255):                       # (q, t) of (a, b) == (t, q) of (b, a), as sets
                n = int(out.count[p].item())
                fwd = out.query_idx[p, :n].to(torch.int64) * k + out.train_idx[p, :n]
                bwd = out.train_idx[256 + p, :n].to(torch.int64) * k + out.query_idx[256 + p, :n]
                ok = ok and bool(torch.equal(torch.sort(fwd).values, torch.sort(bwd).values))
            r[name]["transpose_symmetric"] = ok
    r["workload"] = "mutual arg-best only (ratio test off), 512 directed pairs over 33 images x 16384 u8 descriptors"
    r["note"] = ("ops counted as ONE K x K contraction per pair; the mutual check of every row runs as a second full sweep "
                 "here, so 0.5 of the sweep rate is this mode's ceiling until reverse pairs are reused")
    return r


def secondary_u8_k2048(pkg, torch, dev, int8_peak, peaks):
    """The reference's SHIPPED size (feature_extraction.py:10: at most 2,048 keypoints per image) on the u8 path:
    4,096 pairs over 128 images x 2,048 x 128 u8, whole rule (2-NN + ratio 0.75 + mutual), bank resident. Small-K regime:
    8 work items per pair, so the step is many short batches -- reported with tensor AND HBM fractions and the number of
    kernel launches per step."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    n_img, k, n_pairs = 128, 2048, 4096
    desc, kpi = syn.make_bank_tensor(n_img, k, seed=SEED, device=dev)
    bank = pkg.DescriptorBank.from_packed(desc, kpi)
    pairs = torch.from_numpy(pkg.blocked_order(pkg.exhaustive_pairs(n_img))[:n_pairs]).to(dev)
    ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(k, n_pairs)), dtype=torch.uint8, device=dev)
    out = pkg.match_pairs(bank, pairs, cap=k, workspace=ws)
    torch.cuda.synchronize()
    out.check_overflow()
    pkg.set_option("profile", 1)
    try:
        pkg.profile_read()
        pkg.match_pairs(bank, pairs, cap=k, out=out, workspace=ws)
        torch.cuda.synchronize()
        launches = int(pkg.profile_read()["launches"])
    finally:
        pkg.set_option("profile", 0)
    best, med = _timed(torch, lambda: pkg.match_pairs(bank, pairs, cap=k, out=out, workspace=ws), 5)
    r = _int8_line(n_pairs, k, best, med, int8_peak)
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    # algorithmic bytes per pair: both images' descriptor rows once (sorted query + sorted train), their norms, the row results
    bytes_pair = 2 * k * 128 + 2 * k * 4 + k * 16
    r.update({"matches": int(out.count.sum().item()), "digest": _flat_digest(torch, out), "launches_per_step": launches,
              "algorithmic_gbs": bytes_pair * n_pairs / med / 1e9, "frac_of_hbm": bytes_pair * n_pairs / med / 1e9 / hbm,
              "hbm_peak_gbs": hbm, "us_per_pair": 1e6 * med / n_pairs,
              "workload": "4096 pairs over 128 images x 2048 u8 descriptors (the reference's shipped keypoint budget), whole rule"})
    return r


def secondary_vq(pkg, torch, dev, peaks):
    """Visual words of the BoW pair scheduler (matching.py:26-28) at the reference's own size: 290 images x 2,048
    descriptors x 128 floats against 200 words; b200m_vq_f32 and its no-allocation twin b200m_vq_f32_static, outputs
    compared. HBM-bound in principle (304 MB of descriptors read once), fp64 FMA-bound in practice."""
    lib = pkg._lib.lib()
    n_obs, dim, n_codes = 290 * 2048, 128, 200
    g = torch.Generator(device=dev).manual_seed(5)
    obs = (torch.rand((n_obs, dim), generator=g, device=dev) * 255).round()
    cb = (torch.rand((n_codes, dim), generator=g, device=dev, dtype=torch.float64) * 255)
    scratch = torch.empty((n_codes,), dtype=torch.float64, device=dev)
    code = torch.empty((n_obs,), dtype=torch.int32, device=dev)
    code_s = torch.empty_like(code)
    st = torch.cuda.current_stream().cuda_stream

    def run(static):
        if static:
            pkg._lib.check(lib.b200m_vq_f32_static(obs.data_ptr(), n_obs, dim, cb.data_ptr(), n_codes, code_s.data_ptr(), None, st))
        else:
            pkg._lib.check(lib.b200m_vq_f32(obs.data_ptr(), n_obs, dim, cb.data_ptr(), n_codes, code.data_ptr(), None,
                                            scratch.data_ptr(), st))
    run(False), run(True)
    torch.cuda.synchronize()
    _, med = _timed(torch, lambda: run(False), 5)
    _, med_s = _timed(torch, lambda: run(True), 5)
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    byt = n_obs * dim * 4 + n_obs * 4
    return {"descriptors": n_obs, "words": n_codes, "ms": 1e3 * med, "ms_static_entry": 1e3 * med_s,
            "descriptors_per_s": n_obs / med, "algorithmic_gbs": byt / med / 1e9, "frac_of_hbm": byt / med / 1e9 / hbm,
            "fp64_gflops": 2.0 * n_obs * n_codes * dim / med / 1e9, "static_entry_identical": bool(torch.equal(code, code_s)),
            "workload": "290 x 2048 float32 descriptors (128-d) -> nearest of 200 float64 words, scipy vq arithmetic"}


def secondary_c4(pkg, dmod, torch, dist, dev, rank, world, int8_peak):
    """BASELINE.json configs[3]: 2,000 images x 8,192 keypoints, each image against the next 32 (63,472 pairs).
    One GPU: the whole bank resident. N GPUs: rank 0 holds the bank on the host and every rank receives only its
    query range + 32-image halo (distributed.scatter_bank_window, NCCL point-to-point); match lists are gathered
    with one packed message per rank. Times are device times, max over ranks."""
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    sched = importlib.import_module("3d_reconstruction_b200.scheduler")
    n_img, k, window, cap = 2000, 8192, 32, 3072
    host_desc = host_kpi = None
    if rank == 0:
        d, kp = syn.make_bank_tensor(n_img, k, seed=77, device=dev)
        if world > 1:
            host_desc, host_kpi = d.cpu().pin_memory(), kp.cpu()
            del d
            torch.cuda.empty_cache()
        else:
            host_desc, host_kpi = d, kp
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ld, lk, lpairs, gpairs = dmod.scatter_bank_window(host_desc, host_kpi, n_img, k, 128, window, dev)
    bank = pkg.DescriptorBank.from_packed(ld.contiguous(), lk)
    e1.record()
    torch.cuda.synchronize()
    scatter_ms = e0.elapsed_time(e1)
    pd = torch.from_numpy(np.ascontiguousarray(lpairs)).to(dev)
    out = pkg.match_pairs(bank, pd, cap=cap)
    ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(k, max(len(lpairs), 1))), dtype=torch.uint8, device=dev)
    n_pairs_of = [len(sched.window_shard(n_img, window, world, r)[4]) for r in range(world)]

    def gather():
        flat = out.flatten()
        counts = (flat.offsets[1:] - flat.offsets[:-1]).to(torch.int32)
        return flat, dmod.gather_packed(counts, [flat.query_idx, flat.train_idx, flat.dist], n_pairs_of)

    gather()                                            # warm-up (NCCL point-to-point channels)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    out.check_overflow()
    tm = []
    for _ in range(3):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        a.record()
        pkg.match_pairs(bank, pd, cap=cap, out=out, workspace=ws)
        b.record()
        flat, got = gather()
        c.record()
        torch.cuda.synchronize()
        tm.append((a.elapsed_time(b), b.elapsed_time(c)))
    t = torch.tensor([scatter_ms, float(np.median([x[0] for x in tm])), float(np.median([x[1] for x in tm]))],
                     dtype=torch.float64, device=dev)
    per, tot = flat.digest(torch.arange(len(lpairs), dtype=torch.int64) + int(sum(n_pairs_of[:rank])))
    agg = torch.tensor([int(flat.offsets[-1]), tot - (1 << 64) if tot >= (1 << 63) else tot], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(agg, op=dist.ReduceOp.SUM)
    t = t.tolist()
    n_pairs = int(sum(n_pairs_of))
    r = _int8_line(n_pairs, k, t[1] / 1e3, t[1] / 1e3, int8_peak, world)
    r.update({"gpus": world, "match_ms": t[1], "gather_ms": t[2], "scatter_prepare_sort_ms_first_call": t[0],
              "images_on_rank0": int(ld.shape[0]), "matches_total": int(agg[0]),
              "result_digest": f"{int(agg[1]) & ((1 << 64) - 1):016x}",
              "pairs_per_s_with_gather": n_pairs / ((t[1] + t[2]) / 1e3),
              "workload": "2000 images x 8192 u8 descriptors, each vs the next 32 (63472 pairs), halo-sharded over the ranks"})
    return r


def secondary_c5(pkg, torch, dev, peaks):
    """BASELINE.json configs[4]: float descriptors, 2,048 keypoints, D = 128 / 256, bf16 tensor cores; 4,096 random
    pairs over 128 images. Mutual arg-max of the raw similarity and the full log double soft-max head, each beside
    the reference's own torch fp32 ops (lightglue.py:196-254, restated in oracle_float.head_torch_cpu) on the host."""
    from oracle import oracle_float
    n_img, k, n_pairs = 128, 2048, 4096
    bf16_peak = float(peaks.get("bf16_tflops", 1590.0))
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    res = {}
    rng = np.random.default_rng(0)
    pairs = np.stack([rng.integers(0, n_img, n_pairs), rng.integers(0, n_img, n_pairs)], 1).astype(np.int32)
    pt = torch.from_numpy(pairs).to(dev)
    for dim in (128, 256):
        g = torch.Generator(device=dev).manual_seed(dim)
        unit = torch.nn.functional.normalize(torch.randn((n_img, k, dim), generator=g, device=dev), dim=2)
        kpi = torch.full((n_img,), k, dtype=torch.int32)
        flops = 2.0 * k * k * dim * n_pairs                     # algorithmic: ONE dense M x N product per pair
        # mutual arg-max of the raw similarity
        bank = pkg.FloatBank(unit.to(torch.bfloat16).contiguous(), kpi)
        out = pkg.match_pairs_float(bank, pt, score_threshold=0.2, mutual=True, cap=k)
        ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes_bf16(k, n_pairs)), dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        best, med = _timed(torch, lambda: pkg.match_pairs_float(bank, pt, score_threshold=0.2, mutual=True, cap=k, out=out,
                                                               workspace=ws), 5)
        a, b = unit[int(pairs[0, 0])].float().cpu(), unit[int(pairs[0, 1])].float().cpu()
        oracle_float.head_torch_cpu(a, b, None, None, 0.2)           # warm-up (thread pool, allocator)
        t0 = time.perf_counter()
        for _ in range(4):
            oracle_float.head_torch_cpu(a, b, None, None, 0.2)
        cpu_s = (time.perf_counter() - t0) / 4
        res[f"c5_d{dim}"] = {"pairs": n_pairs, "pairs_per_s": n_pairs / med, "ms": 1e3 * med,
                             "algorithmic_tflops": flops / med / 1e12, "frac_of_bf16_peak": flops / med / 1e12 / bf16_peak,
                             "frac_of_hbm": 2.0 * k * dim * 2 * n_pairs / med / 1e9 / hbm_peak,
                             "cpu_reference_pairs_per_s": 1.0 / cpu_s, "cpu_threads": torch.get_num_threads(),
                             "workload": f"4096 random pairs over 128 images x 2048 x {dim} bf16, mutual arg-max + threshold"}
        # the log double soft-max assignment head
        bank = pkg.FloatBank((4.5 * unit).to(torch.bfloat16).contiguous(), kpi)
        z = torch.randn((n_img, k), generator=g, device=dev)
        out = pkg.assign_pairs(bank, z, pt, filter_threshold=0.1, cap=k)
        ws = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes_assign(k, n_pairs)), dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        best, med = _timed(torch, lambda: pkg.assign_pairs(bank, z, pt, filter_threshold=0.1, cap=k, out=out, workspace=ws), 5)
        a, b = 4.5 * a, 4.5 * b
        z0, z1 = z[int(pairs[0, 0])].cpu(), z[int(pairs[0, 1])].cpu()
        oracle_float.head_torch_cpu(a, b, z0, z1, 0.1)
        t0 = time.perf_counter()
        for _ in range(4):
            oracle_float.head_torch_cpu(a, b, z0, z1, 0.1)
        cpu_s = (time.perf_counter() - t0) / 4
        res[f"c5_head_d{dim}"] = {"pairs": n_pairs, "pairs_per_s": n_pairs / med, "ms": 1e3 * med,
                                  "algorithmic_tflops": flops / med / 1e12, "frac_of_bf16_peak": flops / med / 1e12 / bf16_peak,
                                  "frac_of_hbm": (2.0 * k * dim * 2 + 2.0 * k * 4) * n_pairs / med / 1e9 / hbm_peak,
                                  "cpu_reference_pairs_per_s": 1.0 / cpu_s, "cpu_threads": torch.get_num_threads(),
                                  "workload": f"same pairs, log double soft-max + matchability + filter_matches (D = {dim})"}
        del bank, out, ws, unit
    res["bf16_peak_tflops"] = bf16_peak
    res["hbm_peak_gbs"] = hbm_peak      # frac_of_hbm: both images' bf16 rows (+ logits) per pair / time / measured copy bandwidth
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n-img", type=int, default=N_IMG, help="development only: fewer images (invalid as a bench number)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the configs 2 / 4 / 5 measurements")
    ap.add_argument("--no-rebalance", action="store_true",
                    help="N > 1: keep the equal-count partition instead of the one weighted by the warm-up's per-rank throughput")
    ap.add_argument("--chunk", type=int, default=None, help="development only: columns per chunk of the chunk sweep (32 / 64)")
    ap.add_argument("--kernel", type=int, default=None, help="development only: sweep kernel variant (library option 'kernel')")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)       # timing rule: at least three untimed warm-up steps (reported as run)

    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: there is no CPU fallback for the product path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    import __graft_entry__ as g
    g.build()
    pkg = importlib.import_module("3d_reconstruction_b200")
    dmod = importlib.import_module("3d_reconstruction_b200.distributed")
    if args.kernel is not None:
        pkg.set_option("kernel", args.kernel)
    if args.chunk is not None:
        pkg.set_option("chunk", args.chunk)

    n_img = args.n_img
    n_loc, _ = dmod.image_slices(n_img, world)
    # ------------------------------------------------------------------ inputs
    # rank 0 generates the bank; the other ranks receive it over NCCL (timed cold = communicator set-up included,
    # and warm). Every rank then keeps ITS image slice in pinned host memory: the e2e leg starts from there.
    t0 = time.perf_counter()
    if rank == 0:
        desc0, kpi0 = make_host_bank(n_img, K, torch, dev)
    else:
        desc0 = kpi0 = None
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t0
    bcast_ms = []
    for _ in range(2 if world > 1 else 1):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        desc, kpi = dmod.broadcast_bank_tensors(desc0, kpi0, (n_img, K, D), dev)
        ev1.record()
        torch.cuda.synchronize()
        bcast_ms.append(ev0.elapsed_time(ev1))
        if rank == 0:
            desc0, kpi0 = desc, kpi
    lo, hi = rank * n_loc, min(n_img, (rank + 1) * n_loc)
    host_local = torch.empty((max(hi - lo, 0), K, D), dtype=torch.uint8).pin_memory()
    host_local.copy_(desc[lo:hi])
    k_host = kpi.cpu().numpy()
    bank = pkg.DescriptorBank.from_packed(desc, kpi)

    pairs_all = pkg.blocked_order(pkg.exhaustive_pairs(n_img))
    # the same static partition in both legs: every pair goes to the owner of one of its images (count-balanced, and it
    # spreads the match-rich pairs of neighbouring images evenly: 87.0 ms against 88.9 ms for the slowest of 8 ranks
    # under a contiguous split of the list)
    pairs_mine, idx_mine, _ = pkg.partition_by_owner(pairs_all, world, rank, n_loc)
    pairs_dev = torch.from_numpy(np.ascontiguousarray(pairs_mine)).to(dev)
    P_all, P = len(pairs_all), len(pairs_mine)

    out = pkg.match_pairs(bank, pairs_dev, ratio=RATIO, mutual=True, cap=CAP)      # allocates outputs once
    ws_bytes = pkg._lib.lib().b200m_workspace_bytes(bank.k_stride, P)
    workspace = torch.empty(int(ws_bytes), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    out.check_overflow()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ------------------------------------------------------------------ device-resident leg (value)
    barrier()
    w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_meas = max(1, args.warmup - 1)                    # the first warm-up step is cold: measure the others
    for i in range(args.warmup):
        if i == args.warmup - n_meas:
            w0.record()
        pkg.match_pairs(bank, pairs_dev, ratio=RATIO, mutual=True, cap=CAP, out=out, workspace=workspace)
    w1.record()
    barrier()
    # The ranks do not all run at the same pairs/s (GPUs under a power cap differ by a few per cent, and so does the
    # match density of the pair sets): the warm-up steps measure every rank's throughput and the owner partition then
    # hands out the pairs in that proportion (scheduler.partition_by_owner(weights=...)); untimed, like the warm-up.
    rank_weights = None
    if world > 1 and args.warmup > 0 and not args.no_rebalance:
        rank_weights = dmod.rank_throughput(w0.elapsed_time(w1) / n_meas, P, dev)
    if rank_weights is not None:
        pairs_mine, idx_mine, _ = pkg.partition_by_owner(pairs_all, world, rank, n_loc, weights=rank_weights)
        pairs_dev = torch.from_numpy(np.ascontiguousarray(pairs_mine)).to(dev)
        P = len(pairs_mine)
        del out, workspace
        out = pkg.match_pairs(bank, pairs_dev, ratio=RATIO, mutual=True, cap=CAP)
        workspace = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(bank.k_stride, P)), dtype=torch.uint8, device=dev)
        pkg.match_pairs(bank, pairs_dev, ratio=RATIO, mutual=True, cap=CAP, out=out, workspace=workspace)
        barrier()
    pkg.set_option("profile", 1)
    pkg.profile_read()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        pkg.match_pairs(bank, pairs_dev, ratio=RATIO, mutual=True, cap=CAP, out=out, workspace=workspace)
    e1.record()
    barrier()
    sampler.stop_flag.set()
    sampler.join(timeout=3)
    prof = pkg.profile_read()
    pkg.set_option("profile", 0)
    ms_mine = e0.elapsed_time(e1) / args.steps
    ms_step = max_over_ranks(ms_mine)
    value = P_all / (ms_step / 1e3)
    stats = pkg.match_stats(workspace)
    ms_ranks = [ms_mine]
    if world > 1:                       # per-rank step times: how much of the max is skew between the GPUs
        t = torch.zeros((world,), dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(t, torch.tensor([ms_mine], dtype=torch.float64, device=dev))
        ms_ranks = [round(float(x), 3) for x in t.tolist()]

    # every match record of this rank -> 64-bit digest; counts and digests add over the ranks (mod 2^64): the same
    # pair of numbers at every N <=> the N-GPU lists are the 1-GPU lists (outside the timed region)
    flat = out.flatten()
    _, dig = flat.digest(torch.from_numpy(np.asarray(idx_mine, dtype=np.int64)))
    agg = torch.tensor([int(flat.offsets[-1]), dig - (1 << 64) if dig >= (1 << 63) else dig], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(agg, op=dist.ReduceOp.SUM)
    matches_total = int(agg[0])
    result_digest = int(agg[1]) & ((1 << 64) - 1)
    if n_img == N_IMG:
        assert matches_total == MATCHES_C3, f"{matches_total} matches over {world} rank(s), the workload has {MATCHES_C3}"
    del flat

    # roofline of the dominant kernel (first sweep): algorithmic ops of the pairs one launch covers
    peaks, peak_kind = measured_peaks()
    ops_per_pair = 2.0 * K * K * D
    sweep_ms_per_launch = prof["first_sweep_ms"] / max(prof["first_sweeps"], 1)
    pairs_per_launch = P * args.steps / max(prof["first_sweeps"], 1)
    achieved = ops_per_pair * pairs_per_launch / (sweep_ms_per_launch / 1e3) / 1e12
    clocks = sampler.summary()
    probe = int8_peak_probe(pkg, bank, pairs_dev, out, workspace, ops_per_pair, cublas=(world == 1)) if rank == 0 else {}
    # peak: MEASURED_PEAKS.json has no int8 entry. The tensor pipe retires 8,192 int8 MACs per clock per SM
    # (twice the bf16 rate), so the hardware ceiling of THIS run is that rate at the SM clock sampled under load;
    # without a clock sample the library's own MMA-only skeleton, sustained, stands in. The former proxy
    # (2 x measured sustained bf16) is kept as a note only: the int8 sweep exceeds it (cheaper per op in power).
    if clocks.get("sm_mhz"):
        int8_peak = SM_COUNT * INT8_OPS_PER_CLK_SM * clocks["sm_mhz"] * 1e6 / 1e12
        peak_source = (f"per-clock tensor limit at the sampled SM clock: {SM_COUNT} SMs x {INT8_OPS_PER_CLK_SM} int8 ops/clk x "
                       f"{clocks['sm_mhz']:.0f} MHz (median under load); MEASURED_PEAKS.json ({peak_kind}) has no int8 entry")
    else:
        int8_peak = float(probe.get("mma_only_sweep_sustained_tops") or 2.0 * float(peaks.get("bf16_tflops_sustained", 1400.0)))
        peak_source = "no clock sample: this library's MMA-only sweep, sustained (measured in this run)"
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "r02q_traffic.json")
    if os.path.exists(tpath) and pkg.get_option("kernel") == 5 and pkg.get_option("chunk") == 64 and K == 16384:
        with open(tpath) as f:
            tj = json.load(f)
        traffic = tj["dram_bytes_read_per_launch"] + tj["dram_bytes_write_per_launch"]
        traffic_src = tj["source"]
    bf16x2 = 2.0 * float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops")))
    roofline = {
        "bound": "tensor", "achieved": achieved, "peak": int8_peak, "unit": "TOP/s", "frac": achieved / int8_peak,
        "traffic": traffic, "traffic_unit": "bytes of DRAM read + written per launch (ncu)", "traffic_source": traffic_src,
        "traffic_measured_in_run": False,
        "kernel": ("b200m::chunk_sweep256_kernel<0,%d>" % pkg.get_option("chunk")) if pkg.get_option("kernel") == 5
                  else "b200m::chunk_sweep_kernel",
        "peak_source": peak_source,
        "frac_of_nominal_4500": achieved / 4500.0,
        "frac_of_mma_only_sustained": (achieved / probe["mma_only_sweep_sustained_tops"]) if probe.get("mma_only_sweep_sustained_tops") else None,
        "frac_of_2x_bf16_sustained": achieved / bf16x2,
        "note_2x_bf16_sustained": f"{bf16x2:.0f} TOP/s = 2 x bf16_tflops_sustained of MEASURED_PEAKS.json: a power-limited bf16 GEMM "
                                  "figure, not an int8 ceiling (a fraction above 1 against it says so)",
        "launch_ms": sweep_ms_per_launch, "launches_timed": prof["first_sweeps"],
        "share_of_step": prof["first_sweep_ms"] / (e0.elapsed_time(e1)),
        "verify_share_of_step": prof["verify_sweep_ms"] / (e0.elapsed_time(e1)),
        "other_kernels_ms_per_step": {k: prof[k] / args.steps for k in
                                      ("verify_sweep_ms", "fallback_sweep_ms", "refine_ms", "select_ms", "emit_ms")},
        "row_paths": {k: stats[k] for k in ("rows", "rows_chunk_evaluated", "rows_fallback", "rows_ratio_verify", "candidates")},
        "int8_peak_probe": probe,
    }
    if roofline["frac"] > 1.0:       # a fraction above 1 means the peak is wrong (e.g. a stale clock sample): say so, keep the line
        roofline["peak_inconsistent"] = True
        sys.stderr.write(f"bench.py: roofline.frac = {roofline['frac']:.3f} > 1 -- the sampled clock does not bound this run\n")
    gpu_launches = prof["launches"]

    # ------------------------------------------------------------------ end-to-end leg (host buffers)
    phase_ms = {}

    def e2e_step(record):
        evs = {} if record else None
        res = dmod.match_pairs_host(host_local, k_host, n_img, pairs_all, ratio=RATIO, mutual=True, cap=CAP, device=dev,
                                    out=out, workspace=workspace, staging=True, events=evs, rank_weights=rank_weights)
        if evs:
            torch.cuda.synchronize()
            for name, (a, b) in evs.items():
                phase_ms.setdefault(name, []).append(a.elapsed_time(b))
        return res

    cold0 = time.perf_counter()
    res = e2e_step(False)
    barrier()
    e2e_cold_ms = 1e3 * (time.perf_counter() - cold0)
    for _ in range(2):
        res = e2e_step(False)
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        res = e2e_step(False)
    e1.record()
    barrier()
    e2e_ms = max_over_ranks(max(e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0)) / args.steps)
    h2d_t = torch.tensor([host_local.numel() + k_host.size * 4 + pairs_mine.size * 4], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(h2d_t, op=dist.ReduceOp.SUM)
    h2d = int(h2d_t.item())
    e2e_same = None
    d2h = 0
    if rank == 0:
        d2h = int(res.offsets.numel() * 8 + res.query_idx.numel() * 12)
        e2e_same = bool(res.digest()[1] == result_digest and int(res.offsets[-1]) == matches_total)
        assert e2e_same, "the host-buffer path returned different match lists than the resident-bank path"
    if world > 1:                                     # warm per-phase device times of the sharded path (outside the timed steps)
        for _ in range(3):
            barrier()                                 # no rank-to-rank skew inside the phase times
            e2e_step(True)
        barrier()
    phases = {k: max_over_ranks(float(np.median(v))) for k, v in sorted(phase_ms.items())}
    phases_min = {k: -max_over_ranks(-float(np.median(v))) for k, v in sorted(phase_ms.items())}

    # ------------------------------------------------------------------ the other BASELINE configs (short, untimed region)
    secondary = None
    if not args.no_secondary and n_img == N_IMG:
        del bank, out, workspace, desc, desc0
        torch.cuda.empty_cache()
        secondary = {}
        secondary["c4"] = secondary_c4(pkg, dmod, torch, dist, dev, rank, world, int8_peak)
        torch.cuda.empty_cache()
        if world == 1:
            secondary["c2"] = secondary_c2(pkg, torch, dev, int8_peak)
            torch.cuda.empty_cache()
            secondary["knn2"] = secondary_knn2(pkg, torch, dev, int8_peak)
            torch.cuda.empty_cache()
            secondary["c1_real_sift"] = secondary_real_sift(pkg, torch, dev, int8_peak)
            torch.cuda.empty_cache()
            secondary["mutual_only"] = secondary_mutual_only(pkg, torch, dev, int8_peak)
            torch.cuda.empty_cache()
            secondary["u8_k2048"] = secondary_u8_k2048(pkg, torch, dev, int8_peak, peaks)
            torch.cuda.empty_cache()
            secondary["vq"] = secondary_vq(pkg, torch, dev, peaks)
            torch.cuda.empty_cache()
            secondary.update(secondary_c5(pkg, torch, dev, peaks))

    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline and world == 1:      # the CPU baseline is timed at N = 1 only
            cpu = cpu_baseline(host_local.numpy(), pairs_all)     # one rank: host_local is the whole host bank
        line = {
            "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "n_img": n_img, "k": K, "d": D, "pairs": P_all, "pairs_per_rank": P, "ms_per_step_ranks": ms_ranks,
                       "rank_weights": rank_weights,
                       "cap": CAP, "seed": SEED, "matches_total": matches_total, "result_digest": f"{result_digest:016x}",
                       "e2e_lists_equal_value_lists": e2e_same,
                       "l2": "inputs larger than L2: 539 MB bank, every step sweeps all of it",
                       "bank_generate_s": t_gen,
                       "collectives_ms": {"bank_broadcast_cold": bcast_ms[0], "bank_broadcast_warm": bcast_ms[-1] if world > 1 else None,
                                          "e2e_first_call": e2e_cold_ms, "e2e_phases_warm": phases or None,
                                          "e2e_phases_warm_min_over_ranks": phases_min or None,
                                          "e2e_bank_transport": dmod.transport_in_use()},
                       "parallelism": f"pairs sharded over {world} GPU(s)"
                                      + ("" if rank_weights is None else " in proportion to every rank's throughput measured over the warm-up steps (rank_weights)")
                                      + "; value leg: no collective in the data path; e2e leg: "
                                      "sharded upload, pairs between a rank's own images matched while the prepared bank slices "
                                      "travel between the GPUs (copy engines over NVLink), one packed gather per step"},
            "e2e": {"value": P_all / (e2e_ms / 1e3), "unit": "pairs/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "3d_reconstruction_b200.distributed.match_pairs_host (matcher.match_pairs_host on one rank)"},
            "roofline": roofline,
            "cpu_baseline": cpu,
            "gpu_launches": gpu_launches,
            "clocks": clocks,
            "tops_whole_step": ops_per_pair * P_all / (ms_step / 1e3) / 1e12,
            "secondary": secondary,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
