"""GPU, NCCL, 2 ranks (skipped on a one-GPU box): the N-GPU result is BITWISE the 1-GPU result (SURVEY 8e;
the reference's contract is one ordered row per pair, matching.py:174,188-189). Two processes run the sharded
host-buffer path (every rank uploads and prepares its image slice, in-place all-gather of the bank over NVLink,
static pair partition, one packed gather) and the gathered flat lists are compared, array for array, with what
one process computes for the whole pair list."""
import importlib
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _inputs():
    syn = importlib.import_module("3d_reconstruction_b200.synthetic")
    ks = [1500, 1024, 999, 1300, 257, 1536, 0, 1100, 640, 1290, 1400, 777]
    desc, kpi = syn.make_bank_tensor(len(ks), 1500, seed=3, device="cpu", k_per_image=ks)
    sched = importlib.import_module("3d_reconstruction_b200.scheduler")
    pairs = sched.blocked_order(sched.exhaustive_pairs(len(ks)), 4)
    pairs = np.concatenate([pairs, pairs[::3, ::-1]], 0)            # some pairs in the other direction too
    return desc, kpi, pairs


def _worker(rank, world, port, out_path):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    dmod = importlib.import_module("3d_reconstruction_b200.distributed")
    desc, kpi, pairs = _inputs()
    n_img = desc.shape[0]
    n_loc, _ = dmod.image_slices(n_img, world)
    local = desc[rank * n_loc: min(n_img, (rank + 1) * n_loc)].clone().pin_memory()
    res = {}
    for transport in ("auto", "nccl", "auto"):                       # copy engines over symmetric memory (twice: cached buffers), NCCL
        got = dmod.match_pairs_host(local, kpi.numpy(), n_img, pairs, cap=1536, device=dev, transport=transport, block=2)
        if rank == 0:
            res[transport] = {"off": got.offsets.clone(), "q": got.query_idx.clone(), "t": got.train_idx.clone(),
                              "d": got.dist.clone()}
        else:
            assert got is None
    if rank == 0:
        res["transport"] = dmod.transport_in_use()
        torch.save(res, out_path)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_ranks_equal_one_rank_bitwise(pkg, tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "gathered.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    res = torch.load(out)
    desc, kpi, pairs = _inputs()
    bank = pkg.DescriptorBank.from_packed(desc.cuda(), kpi)
    ref = pkg.match_pairs(bank, pairs, cap=1536).flatten().cpu()
    print("bank transport of the 'auto' calls:", res["transport"])
    for transport in ("auto", "nccl"):
        r = res[transport]
        assert torch.equal(r["off"], ref.offsets), transport
        assert torch.equal(r["q"], ref.query_idx) and torch.equal(r["t"], ref.train_idx), transport
        assert torch.equal(r["d"].view(torch.int32), ref.dist.view(torch.int32)), transport
    assert int(ref.offsets[-1]) > 1000


def _worker_small_cap(rank, world, port, out_path):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    dmod = importlib.import_module("3d_reconstruction_b200.distributed")
    desc, kpi, pairs = _inputs()
    n_img = desc.shape[0]
    n_loc, _ = dmod.image_slices(n_img, world)
    local = desc[rank * n_loc: min(n_img, (rank + 1) * n_loc)].clone().pin_memory()
    # only rank 1 can overflow: rank 0's images are emptied, so every pair rank 0 may own has an empty side, while the
    # pairs between two of rank 1's images are always rank 1's (partition_by_owner)
    kp = kpi.numpy().copy()
    kp[:n_loc] = 0
    seen = []
    for cap in (4, 1536):                      # too small for rank 1's pairs, then a call that must still work
        try:
            got = dmod.match_pairs_host(local, kp, n_img, pairs, cap=cap, device=dev, transport="nccl", block=2)
            seen.append("ok" if (got is None) == (rank != 0) else "bad")
        except OverflowError:
            seen.append("OverflowError")
    torch.save(seen, f"{out_path}.{rank}")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_overflow_on_one_rank_raises_on_both(pkg, tmp_path):
    """cap too small for pairs that only ONE rank owns: both ranks raise OverflowError (instead of one raising and the
    other hanging in the gather), and the next call on the same process group works."""
    import torch.multiprocessing as mp
    out = str(tmp_path / "seen")
    mp.spawn(_worker_small_cap, args=(2, _free_port(), out), nprocs=2, join=True)
    a, b = torch.load(out + ".0"), torch.load(out + ".1")
    assert a == b == ["OverflowError", "ok"]
