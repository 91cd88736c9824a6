"""BASELINE.json config 4 over N GPUs (torchrun): 2,000 images x 8,192 keypoints, each image against the next
32; rank 0 holds the bank on the host, every rank receives only its query range + 32-image halo
(distributed.scatter_bank_window, point-to-point over NCCL), matches its pairs, flat lists are gathered to
rank 0, which checks sampled pairs against the integer oracle. Prints one JSON line (development aid).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/run_config4_sharded.py [n_img]
"""
import importlib
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3d_reconstruction_b200")
dmod = importlib.import_module("3d_reconstruction_b200.distributed")
syn = importlib.import_module("3d_reconstruction_b200.synthetic")


def main():
    n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
    k, window, cap = 8192, 32, 3072
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    host_desc = host_kpi = None
    if rank == 0:
        d, kp = syn.make_bank_tensor(n_img, k, seed=77, device=dev)
        host_desc, host_kpi = d.cpu().pin_memory(), kp.cpu()
        del d
        torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ld, lk, lpairs, gpairs = dmod.scatter_bank_window(host_desc, host_kpi, n_img, k, 128, window, dev)
    bank = pkg.DescriptorBank.from_packed(ld.contiguous(), lk)
    torch.cuda.synchronize()
    t_scatter = time.perf_counter() - t0
    res = pkg.match_pairs(bank, lpairs, cap=cap)          # warm-up pass (allocations, NCCL connections)
    flat = res.flatten()
    if world > 1:
        dmod.gather_flat(flat.offsets, [flat.query_idx, flat.train_idx, flat.dist])
        dist.barrier()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    res = pkg.match_pairs(bank, lpairs, cap=cap, out=res)
    res.check_overflow()
    flat = res.flatten()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    got = dmod.gather_flat(flat.offsets, [flat.query_idx, flat.train_idx, flat.dist]) if world > 1 else (flat.offsets, [flat.query_idx, flat.train_idx, flat.dist])
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    times = torch.tensor([t_scatter, t2 - t1, t3 - t2], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    if rank == 0:
        off, (fq, ft, fd) = got
        off, fq, ft, fd = off.cpu().numpy(), fq.cpu().numpy(), ft.cpu().numpy(), fd.cpu().numpy()
        pairs = pkg.window_pairs(n_img, window)
        assert len(off) == len(pairs) + 1
        from oracle import oracle_int
        rng = np.random.default_rng(0)
        for p in rng.choice(len(pairs), size=6, replace=False).tolist() + [len(pairs) - 1]:
            a, b = map(int, pairs[p])
            eq, et, ed = oracle_int.match(host_desc[a].numpy(), host_desc[b].numpy(), 0.75, True)
            s, e = off[p], off[p + 1]
            assert np.array_equal(fq[s:e], eq) and np.array_equal(ft[s:e], et)
            assert np.array_equal(fd[s:e].view(np.uint32), np.asarray(ed, np.float32).view(np.uint32))
        tt = times.cpu().numpy()
        print(json.dumps({"config": 4, "n_img": n_img, "pairs": int(len(pairs)), "gpus": world, "images_on_rank0": int(ld.shape[0]),
                          "scatter_and_sort_s_first_call": round(float(tt[0]), 3), "match_s": round(float(tt[1]), 3),
                          "gather_s": round(float(tt[2]), 3), "pairs_per_s_match": round(len(pairs) / float(tt[1]), 1),
                          "matches": int(off[-1]), "sampled_pairs_equal_oracle": True}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
