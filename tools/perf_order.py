"""Does the pair ORDER / segmentation of the sharded host path cost sweep time? (development aid)
Times match_pairs on one GPU for rank 0's share of config 3 under (a) the contiguous partition of the value leg,
(b) the owner partition in one call, (c) the owner partition in the segments the sharded path launches."""
import importlib
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_reconstruction_b200")
syn = importlib.import_module("3d_reconstruction_b200.synthetic")
sch = importlib.import_module("3d_reconstruction_b200.scheduler")
M = importlib.import_module("3d_reconstruction_b200.matcher")

ws = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n_img, k = 257, 16384
desc, kpi = syn.make_bank_tensor(n_img, k, seed=1234, device="cuda")
bank = pkg.DescriptorBank.from_packed(desc, kpi)
pairs = sch.blocked_order(sch.exhaustive_pairs(n_img))
n_loc = -(-n_img // ws)
cases = {}
cases["contiguous"] = [sch.partition(pairs, ws, 0, np.full(n_img, k))[0]]
own, idx, n_local = sch.partition_by_owner(pairs, ws, 0, n_loc)
cases["owner_one_call"] = [own]
segs = M.upload_segments(own[:n_local], n_loc, 12)
cases["owner_segments"] = [own[a:b] for _, a, b in segs] + [own[n_local:]]
cases["owner_local_then_rest"] = [own[:n_local], own[n_local:]]
P = len(own)
out = M._new_out(P, 4096, bank.device)
wsb = torch.empty(int(pkg._lib.lib().b200m_workspace_bytes(k, P)), dtype=torch.uint8, device="cuda")
for name, lists in cases.items():
    devs = [torch.from_numpy(np.ascontiguousarray(l)).cuda() for l in lists]
    ts = []
    for rep in range(4):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        at = 0
        for d in devs:
            pkg.match_pairs(bank, d, cap=4096, out=M._out_rows(out, at, at + len(d)), workspace=wsb)
            at += len(d)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    print(json.dumps({"case": name, "pairs": int(sum(len(l) for l in lists)), "calls": len(lists), "ms": [round(t, 2) for t in ts]}), flush=True)
