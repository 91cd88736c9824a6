#!/usr/bin/env python
"""bench.py -- image-pairs/s of the pairwise descriptor matcher on BASELINE.json's headline workload.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --steps K --warmup W     (cv2.BFMatcher on the host cores)

Workload (config.workload): synthetic 257 images x 16,384 SIFT-like u8 descriptors, exhaustive
32,896 pairs (BASELINE.json configs[2], the configuration the metric is quoted on). A step is one
pass of the full matching rule (2-NN, Lowe ratio 0.75, mutual check) over all pairs.
  value     pairs/s with the packed bank resident in HBM (CUDA events, max over ranks)
  e2e       same through the host-buffer API: bank + pair list host->device, match lists device->host
  roofline  the first-sweep kernel against the tensor roofline (int8 ops = 2 * Kq * Kt * 128 per pair)
  cpu_baseline  cv2.BFMatcher(NORM_L2).knnMatch k=2 + reverse k=1 + ratio filter on a bounded sample
N > 1 splits the fixed pair list over the ranks (static cost-balanced partition, bank broadcast over
NCCL once, match lists gathered to rank 0 in the e2e leg): "scaling": "strong".
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


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def int8_peak_probe(pkg, bank, pairs_dev, out, workspace, ops_per_pair):
    """MEASURED_PEAKS.json has no int8 entry: two live int8 dense rates on this box for context, outside every
    timed region. (1) cuBLASLt IGEMM through torch._int_mm, 8192^3, a short burst and ~1.5 s sustained;
    (2) this library's own sweep with the epilogue reduced to one tcgen05.ld per accumulator (option
    epilogue = 2: TMA + MMA + hand-over only, results invalid) over one batch of 256 pairs."""
    import torch
    res = {}
    n = min(256, int(pairs_dev.shape[0]))
    pkg.set_option("epilogue", 2)
    try:
        pkg.match_pairs(bank, pairs_dev[:n], ratio=RATIO, mutual=True, cap=CAP, out=out, workspace=workspace)
        pkg.set_option("profile", 1)
        pkg.profile_read()
        for _ in range(3):
            pkg.match_pairs(bank, pairs_dev[:n], ratio=RATIO, mutual=True, cap=CAP, out=out, workspace=workspace)
        torch.cuda.synchronize()
        pr = pkg.profile_read()
        res["mma_only_sweep_tops"] = ops_per_pair * n * 3 / (pr["first_sweep_ms"] / 1e3) / 1e12
    finally:
        pkg.set_option("profile", 0)
        pkg.set_option("epilogue", 1)
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
    return {"value": 1.0 / per_pair, "unit": "pairs/s", "cores": int(threads or os.cpu_count() or 1), "kind": "reference",
            "sample": f"{len(times)} pairs of the workload (16384x16384x128), cv2.BFMatcher(NORM_L2).knnMatch k=2 on float32 "
                      f"+ reverse knnMatch k=1 + ratio filter, median {per_pair:.3f} s/pair"}


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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n-img", type=int, default=N_IMG, help="development only: fewer images (invalid as a bench number)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    # ------------------------------------------------------------------ inputs
    # rank 0 generates the bank and keeps a pinned host copy (the e2e leg starts from it); the
    # other ranks receive it over NCCL (NVLink) -- the one broadcast of the path.
    t0 = time.perf_counter()
    if rank == 0:
        desc0, kpi0 = make_host_bank(n_img, K, torch, dev)
        host_desc = torch.empty(desc0.shape, dtype=torch.uint8).pin_memory()
        host_desc.copy_(desc0)
        host_kpi = kpi0.cpu()
    else:
        desc0 = kpi0 = None
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    desc, kpi = dmod.broadcast_bank_tensors(desc0, kpi0, (n_img, K, D), dev)
    ev1.record()
    torch.cuda.synchronize()
    bcast_ms = ev0.elapsed_time(ev1)
    bank = pkg.DescriptorBank.from_packed(desc, kpi)

    pairs_all = pkg.blocked_order(pkg.exhaustive_pairs(n_img))
    k_host = np.full(n_img, K)
    pairs_mine, _ = dmod.my_pairs(pairs_all, k_host)
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
    for _ in range(args.warmup):
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
    ms_step = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    value = P_all / (ms_step / 1e3)
    n_matches = int(out.count.sum().item())

    # roofline of the dominant kernel (first sweep): algorithmic ops of the pairs one launch covers
    peaks, peak_kind = measured_peaks()
    ops_per_pair = 2.0 * K * K * D
    sweep_ms_per_launch = prof["first_sweep_ms"] / max(prof["first_sweeps"], 1)
    pairs_per_launch = P * args.steps / max(prof["first_sweeps"], 1)
    achieved = ops_per_pair * pairs_per_launch / (sweep_ms_per_launch / 1e3) / 1e12
    int8_peak = 2.0 * float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops")))
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "r01e_traffic.json")
    if os.path.exists(tpath) and pkg.get_option("kernel") == 5 and pkg.get_option("chunk") == 64 and K == 16384:
        with open(tpath) as f:
            tj = json.load(f)
        traffic = tj["dram_bytes_read_per_launch"] + tj["dram_bytes_write_per_launch"]
        traffic_src = tj["source"]
    roofline = {
        "bound": "tensor", "achieved": achieved, "peak": int8_peak, "unit": "TFLOP/s", "frac": achieved / int8_peak,
        "traffic": traffic, "traffic_unit": "bytes of DRAM read + written per launch (ncu)", "traffic_source": traffic_src,
        "kernel": ("b200m::chunk_sweep256_kernel<0,%d>" % pkg.get_option("chunk")) if pkg.get_option("kernel") == 5
                  else "b200m::chunk_sweep_kernel",
        "peak_source": f"2 x bf16_tflops_sustained of MEASURED_PEAKS.json ({peak_kind}): int8 tcgen05 rate is twice the bf16 rate; "
                       "nominal dense int8 is 4500",
        "frac_of_nominal_4500": achieved / 4500.0,
        "launch_ms": sweep_ms_per_launch, "launches_timed": prof["first_sweeps"],
        "share_of_step": prof["first_sweep_ms"] / (e0.elapsed_time(e1)),
        "verify_share_of_step": prof["verify_sweep_ms"] / (e0.elapsed_time(e1)),
        "other_kernels_ms_per_step": {k: prof[k] / args.steps for k in
                                      ("verify_sweep_ms", "fallback_sweep_ms", "refine_ms", "select_ms", "emit_ms")},
    }
    gpu_launches = prof["launches"]
    if world == 1:
        roofline["int8_peak_probe"] = int8_peak_probe(pkg, bank, pairs_dev, out, workspace, ops_per_pair)

    # ------------------------------------------------------------------ end-to-end leg (host buffers)
    # every step: rank 0 host bank -> device (pinned), NCCL broadcast, local match, flat lists gathered
    # to rank 0 and copied to the host.
    def e2e_step():
        if rank == 0:
            d = host_desc.to(dev, non_blocking=True)
            kp = host_kpi.to(dev, non_blocking=True)
        else:
            d = kp = None
        d, kp = dmod.broadcast_bank_tensors(d, kp, (n_img, K, D), dev)
        b = pkg.DescriptorBank.from_packed(d, kp)
        pd = torch.from_numpy(np.ascontiguousarray(pairs_mine)).to(dev, non_blocking=True)
        r = pkg.match_pairs(b, pd, ratio=RATIO, mutual=True, cap=CAP, out=out, workspace=workspace)
        flat = r.flatten()
        got = dmod.gather_flat(flat.offsets, [flat.query_idx, flat.train_idx, flat.dist])
        if rank == 0:
            off, (fq, ft, fd) = got
            return off.cpu(), fq.cpu(), ft.cpu(), fd.cpu()
        return None

    for _ in range(max(1, min(args.warmup, 2))):
        res = e2e_step()
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        res = e2e_step()
    e1.record()
    barrier()
    e2e_ms = max_over_ranks(max(e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0)) / args.steps)
    h2d = int(host_desc.numel() + host_kpi.numel() * 4 + pairs_mine.size * 4) if rank == 0 else 0
    d2h = int(res[0].numel() * 8 + res[1].numel() * 12) if rank == 0 else 0

    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline and world == 1:      # the CPU baseline is timed at N = 1 only
            cpu = cpu_baseline(host_desc.numpy(), pairs_all)
        line = {
            "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "n_img": n_img, "k": K, "d": D, "pairs": P_all, "pairs_per_rank": P,
                       "cap": CAP, "seed": SEED, "matches_rank0": n_matches,
                       "l2": "inputs larger than L2: 539 MB bank, every step sweeps all of it",
                       "bank_broadcast_ms": bcast_ms, "bank_generate_s": t_gen,
                       "parallelism": f"pairs sharded over {world} GPU(s), no collective in the data path"},
            "e2e": {"value": P_all / (e2e_ms / 1e3), "unit": "pairs/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "roofline": roofline,
            "cpu_baseline": cpu,
            "gpu_launches": gpu_launches,
            "clocks": sampler.summary(),
            "tops_whole_step": ops_per_pair * P_all / (ms_step / 1e3) / 1e12,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
