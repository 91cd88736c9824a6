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


class _FakeBank:
    """CPU stand-in with the table layout of DescriptorBank (bank.table_sections): the all-gather of image slices
    is host logic + collectives and runs under gloo; the kernels that fill a slice need a GPU."""

    def __init__(self, n_img, ks):
        bank_mod = importlib.import_module("3d_reconstruction_b200.bank")
        self.n_img, self.k_stride = n_img, ks
        R = n_img * ks
        self.desc = torch.zeros((n_img, ks, 128), dtype=torch.uint8)
        self.sqnorm = torch.zeros((R + R // 32,), dtype=torch.int32)
        self.sorted_desc = torch.zeros_like(self.desc)
        self.sort_table = torch.zeros((3 * R + R // 32 + 2 * ((n_img + 3) & ~3),), dtype=torch.int32)
        self.table_sections = lambda: bank_mod.DescriptorBank.table_sections(self)


def _fill(bank, lo, hi, seed):
    """Deterministic stand-in for prepare_range: every section of the images [lo, hi) gets values that depend on
    (section, image, offset) only."""
    for s, (t, per) in enumerate(bank.table_sections()):
        for img in range(lo, hi):
            g = torch.Generator().manual_seed(seed * 1000 + s * 100 + img)
            t[img * per:(img + 1) * per] = torch.randint(0, 200, (per,), generator=g).to(t.dtype)


def _worker_sharded(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dmod = importlib.import_module("3d_reconstruction_b200.distributed")
    sched = importlib.import_module("3d_reconstruction_b200.scheduler")
    n_img = 5
    n_loc, n_pad = dmod.image_slices(n_img, world)
    bank = _FakeBank(n_pad, 256)
    _fill(bank, rank * n_loc, (rank + 1) * n_loc, seed=7)
    dmod.allgather_bank(bank, n_loc)
    torch.save([t.clone() for t, _ in bank.table_sections()], os.path.join(out_dir, f"bank{rank}.pt"))
    # packed gather of variable-length lists: rank r's pairs produce (global pair index % 5) + 1 matches each
    pairs = sched.exhaustive_pairs(n_img)
    mine, idx = sched.partition(pairs, world, rank)
    n_pairs_of = [len(sched.partition(pairs, world, r)[1]) for r in range(world)]
    counts = torch.tensor([(int(i) % 5) + 1 for i in idx], dtype=torch.int32)
    q = torch.cat([torch.arange(c, dtype=torch.int32) + 100 * int(i) for c, i in zip(counts.tolist(), idx)])
    d = q.to(torch.float32) * 0.5
    got = dmod.gather_packed(counts, [q, q + 1, d], n_pairs_of)
    if rank == 0:
        off, (gq, gt, gd) = got
        torch.save({"off": off, "q": gq, "t": gt, "d": gd}, os.path.join(out_dir, "gathered.pt"))
    else:
        assert got is None
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_bank_allgather_and_packed_gather(tmp_path, pkg):
    """The sharded host-buffer path (distributed.match_pairs_host) minus the kernels: every rank fills the table
    sections of ITS image slice, the in-place all-gather leaves the whole bank on every rank; the packed gather
    returns exactly the single-process flat lists."""
    mp.spawn(_worker_sharded, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    n_img = 5
    whole = _FakeBank(6, 256)
    _fill(whole, 0, 6, seed=7)
    exp = [t for t, _ in whole.table_sections()]
    for r in range(2):
        got = torch.load(str(tmp_path / f"bank{r}.pt"))
        assert len(got) == len(exp) == 10
        for g, e in zip(got, exp):
            assert torch.equal(g, e)
    res = torch.load(str(tmp_path / "gathered.pt"))
    n_pairs = n_img * (n_img - 1) // 2
    counts = [(i % 5) + 1 for i in range(n_pairs)]
    assert res["off"].tolist() == np.concatenate([[0], np.cumsum(counts)]).tolist()
    eq = torch.cat([torch.arange(c, dtype=torch.int32) + 100 * i for i, c in enumerate(counts)])
    assert torch.equal(res["q"], eq) and torch.equal(res["t"], eq + 1) and torch.equal(res["d"], eq.float() * 0.5)


def test_upload_segments_follow_the_pair_order(pkg):
    m = importlib.import_module("3d_reconstruction_b200.matcher")
    pairs = pkg.blocked_order(pkg.exhaustive_pairs(70), 16)
    segs = m.upload_segments(pairs, 70, 16)
    assert [s[0] for s in segs] == [0, 1, 2, 3, 4]                       # blocks become necessary one after the other
    assert segs[0][1] == 0 and segs[-1][2] == len(pairs)
    for (need, a, b), nxt in zip(segs, segs[1:] + [None]):
        assert b > a and pairs[a:b].max() // 16 <= need
        if nxt:
            assert nxt[1] == b
    row_major = pkg.exhaustive_pairs(70)                                   # (0, 69) comes early: one big late segment
    segs = m.upload_segments(row_major, 70, 16)
    assert segs[-1][0] == 4 and segs[-1][2] == len(row_major) and sum(b - a for _, a, b in segs) == len(row_major)
    assert m.upload_segments(np.zeros((0, 2), np.int64), 70, 16) == []


def test_upload_blocks_ramp_and_segments_on_boundaries(pkg):
    m = importlib.import_module("3d_reconstruction_b200.matcher")
    assert m.upload_blocks(33, 12, ramp=True) == [0, 2, 6, 14, 26, 33]
    assert m.upload_blocks(33, 12) == [0, 12, 24, 33]
    assert m.upload_blocks(1, 12, ramp=True) == [0, 1] and m.upload_blocks(0, 12, ramp=True) == [0]
    bounds = m.upload_blocks(33, 12, ramp=True)
    pairs = pkg.exhaustive_pairs(33)
    need = np.searchsorted(np.asarray(bounds[1:]), pairs.max(axis=1), side="right")
    pairs = pairs[np.argsort(need, kind="stable")]                         # what distributed.match_pairs_host does
    segs = m.upload_segments(pairs, 33, bounds)
    assert [s[0] for s in segs] == [0, 1, 2, 3, 4]
    assert (segs[0][1], segs[0][2]) == (0, 1)                              # the pair (0, 1) can start after two images
    for need_b, a, b in segs:
        assert pairs[a:b].max() < bounds[need_b + 1] and pairs[a:b].max() >= bounds[need_b]
    assert sum(b - a for _, a, b in segs) == len(pairs)


def _throughput_worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dmod = importlib.import_module("3d_reconstruction_b200.distributed")
    sched = importlib.import_module("3d_reconstruction_b200.scheduler")
    w = dmod.rank_throughput(100.0 if rank == 0 else 110.0, 1000, "cpu")      # rank 1 is 10 % slower
    pairs = sched.exhaustive_pairs(40)
    p, idx, n_local = sched.partition_by_owner(pairs, world, rank, 20, weights=w)
    torch.save({"w": w, "idx": idx}, f"{out_path}.{rank}")
    dist.barrier()
    dist.destroy_process_group()


def test_rank_throughput_weights_agree_on_all_ranks(tmp_path):
    """distributed.rank_throughput: every rank ends up with the same weights (they feed a partition that must be
    identical everywhere), the slower rank gets fewer pairs, nothing is lost."""
    out = str(tmp_path / "w")
    mp.spawn(_throughput_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    a, b = torch.load(out + ".0", weights_only=False), torch.load(out + ".1", weights_only=False)
    assert a["w"] == b["w"] and a["w"][0] > a["w"][1]
    assert abs(a["w"][0] / a["w"][1] - 1.1) < 0.01
    assert len(a["idx"]) > len(b["idx"])
    assert np.array_equal(np.sort(np.concatenate([a["idx"], b["idx"]])), np.arange(40 * 39 // 2))


def _count_check_worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dmod = importlib.import_module("3d_reconstruction_b200.distributed")
    seen = []
    # (counts of rank 0, counts of rank 1, cap): only rank 1 overflows / only rank 0 holds a bad pair / an empty slice / fine
    cases = [([3, 5, 0], [2, 9], 8), ([4, -1], [1, 1, 1], 8), ([], [8, 8], 8), ([1, 2], [], 1)]
    for c0, c1, cap in cases:
        mine = torch.tensor(c0 if rank == 0 else c1, dtype=torch.int32)
        try:
            dmod.check_counts_everywhere(mine, cap)
            seen.append("ok")
        except (OverflowError, ValueError) as ex:
            seen.append(type(ex).__name__)
        # the collective that follows in match_pairs_host must still line up on both ranks after an error
        t = torch.tensor([rank + 1])
        dist.all_reduce(t)
        assert int(t) == 3
    agreed = [dmod.agree_all(ok, "cpu") for ok in (True, rank == 0, False)]
    torch.save({"seen": seen, "agreed": agreed}, f"{out_path}.{rank}")
    dist.barrier()
    dist.destroy_process_group()


def test_count_errors_are_raised_on_every_rank(tmp_path):
    """distributed.check_counts_everywhere / agree_all: an overflow (count > cap) or a bad pair (count = -1) on ONE
    rank raises the same exception on BOTH, so no rank walks into gather_packed alone; a transport flag that only one
    rank sets is false everywhere."""
    out = str(tmp_path / "c")
    mp.spawn(_count_check_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    a, b = torch.load(out + ".0", weights_only=False), torch.load(out + ".1", weights_only=False)
    assert a["seen"] == b["seen"] == ["OverflowError", "ValueError", "ok", "OverflowError"]
    assert a["agreed"] == b["agreed"] == [True, False, False]
