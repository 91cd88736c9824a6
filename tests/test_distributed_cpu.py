"""CPU, gloo, world_size 2: the N>1 host logic (static partition + variable-length gather) gives the
same flat result as one rank. The compute itself is stubbed by the oracle here (no GPU in this test)."""
import importlib
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle_int


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg = importlib.import_module("3d_reconstruction_b200")
    dmod = importlib.import_module("3d_reconstruction_b200.distributed")
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    ks = [300, 200, 260, 120, 280]
    if rank == 0:
        desc, kpi = syn.make_bank_tensor(len(ks), 300, seed=4, device="cpu", k_per_image=ks)
    else:
        desc = kpi = None
    desc, kpi = dmod.broadcast_bank_tensors(desc, kpi, (len(ks), 512, 128), "cpu")
    pairs = pkg.blocked_order(pkg.exhaustive_pairs(len(ks)), 2)
    mine, idx = dmod.my_pairs(pairs, np.asarray(ks))
    offs, q, t, d = [0], [], [], []
    for a, b in mine:
        mq, mt, md = oracle_int.match(desc[a, :ks[a]].numpy(), desc[b, :ks[b]].numpy())
        q.append(mq); t.append(mt); d.append(md)
        offs.append(offs[-1] + len(mq))
    cat = lambda xs, dt: torch.from_numpy(np.concatenate(xs).astype(dt)) if xs else torch.zeros(0, dtype=torch.from_numpy(np.zeros(0, dt)).dtype)
    got = dmod.gather_flat(torch.tensor(offs, dtype=torch.int64), [cat(q, np.int32), cat(t, np.int32), cat(d, np.float32)])
    if rank == 0:
        goff, (gq, gt, gd) = got
        torch.save({"off": goff, "q": gq, "t": gt, "d": gd, "pairs": torch.from_numpy(pairs), "desc": desc, "ks": ks}, out_path)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_partition_and_gather(tmp_path):
    out = str(tmp_path / "res.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = torch.load(out)
    pairs, ks, desc = r["pairs"].numpy(), r["ks"], r["desc"]
    assert len(r["off"]) == len(pairs) + 1
    total = 0
    for p, (a, b) in enumerate(pairs):
        eq, et, ed = oracle_int.match(desc[a, :ks[a]].numpy(), desc[b, :ks[b]].numpy())
        lo, hi = int(r["off"][p]), int(r["off"][p + 1])
        assert np.array_equal(r["q"][lo:hi].numpy(), eq) and np.array_equal(r["t"][lo:hi].numpy(), et)
        assert np.array_equal(r["d"][lo:hi].numpy().view(np.uint32), ed.view(np.uint32))
        total += hi - lo
    assert total == len(r["q"]) and total > 0


def _worker_window(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dmod = importlib.import_module("3d_reconstruction_b200.distributed")
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    n_img, window, k = 9, 3, 200
    if rank == 0:
        desc, kpi = syn.make_bank_tensor(n_img, k, seed=8, device="cpu")
    else:
        desc = kpi = None
    d, kp, lp, gp = dmod.scatter_bank_window(desc, kpi, n_img, 256, 128, window, "cpu")
    torch.save({"d": d, "k": kp, "lp": torch.from_numpy(lp), "gp": torch.from_numpy(gp)}, os.path.join(out_dir, f"w{rank}.pt"))
    if rank == 0:
        torch.save({"desc": desc, "kpi": kpi}, os.path.join(out_dir, "full.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_window_shard_with_halo(tmp_path, pkg):
    """config-4 style sharding: every rank holds only its images + halo, and its local pair list addresses
    exactly the same descriptors as the global schedule does in the full bank."""
    mp.spawn(_worker_window, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    full = torch.load(str(tmp_path / "full.pt"))
    gps = []
    for r in range(2):
        w = torch.load(str(tmp_path / f"w{r}.pt"))
        assert w["d"].shape[0] < full["desc"].shape[0]                      # a shard, not the whole bank
        for (la, lb), (ga, gb) in zip(w["lp"].tolist(), w["gp"].tolist()):
            assert torch.equal(w["d"][la], full["desc"][ga]) and torch.equal(w["d"][lb], full["desc"][gb])
            assert int(w["k"][la]) == int(full["kpi"][ga])
        gps.append(w["gp"].numpy())
    assert np.array_equal(np.concatenate(gps), pkg.window_pairs(9, 3))
