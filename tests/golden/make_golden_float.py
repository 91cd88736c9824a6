"""Generates tests/golden/float_mutual_{128,256}.npz with the REFERENCE's own filter_matches
(lightglue/lightglue.py:235-254, imported by path: lightglue/__init__.py needs kornia) -- run in the
build container, where /root/reference exists:

    python tests/golden/make_golden_float.py

Inputs are bf16-representable fp32 descriptors (stored as their upper 16 bits) with planted
correspondences; outputs are what the reference computes in fp32: sim = einsum('bmd,bnd->bmn')
(:223), filter_matches(scores, th) on the similarity padded to [1, M+1, N+1] (:235-254) with
th = exp(threshold) (the reference thresholds exp(score), :243), and the match list (:417-426).
"""
import importlib.util
import math
import os
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/lightglue/lightglue.py"


def load_reference():
    warnings.simplefilter("ignore")
    spec = importlib.util.spec_from_file_location("ref_lightglue", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def make_inputs(m, n, dim, seed):
    g = torch.Generator().manual_seed(seed)
    world = torch.nn.functional.normalize(torch.randn(max(m, n), dim, generator=g), dim=1)
    d0 = world[torch.randperm(max(m, n), generator=g)[:m]] + 0.05 * torch.randn(m, dim, generator=g)
    d1 = world[torch.randperm(max(m, n), generator=g)[:n]] + 0.05 * torch.randn(n, dim, generator=g)
    d0 = torch.nn.functional.normalize(d0, dim=1).to(torch.bfloat16)
    d1 = torch.nn.functional.normalize(d1, dim=1).to(torch.bfloat16)
    d1[5] = d1[6]            # an exact duplicate column: ties
    return d0, d1


def main():
    ref = load_reference()
    for dim, m, n, thr in ((128, 420, 390, 0.6), (256, 400, 380, 0.6)):
        d0, d1 = make_inputs(m, n, dim, 100 + dim)
        f0, f1 = d0.float()[None], d1.float()[None]
        sim = torch.einsum("bmd,bnd->bmn", f0, f1)                       # lightglue.py:223
        scores = sim.new_full((1, m + 1, n + 1), 0)
        scores[:, :m, :n] = sim
        out = {}
        for name, th in (("none", None), ("thr", math.exp(thr))):
            m0, m1, ms0, ms1 = ref.filter_matches(scores, th)              # lightglue.py:235-254
            valid = m0[0] > -1                                             # lightglue.py:419-425
            q = torch.where(valid)[0]
            out[f"m0_{name}"] = m0[0].numpy()
            out[f"m1_{name}"] = m1[0].numpy()
            out[f"q_{name}"] = q.numpy()
            out[f"t_{name}"] = m0[0][valid].numpy()
        np.savez_compressed(os.path.join(HERE, f"float_mutual_{dim}.npz"),
                            d0_bf16=d0.view(torch.int16).numpy(), d1_bf16=d1.view(torch.int16).numpy(),
                            threshold=np.float32(thr), torch_version=torch.__version__, **out)
        print(dim, "matches (no threshold)", len(out["q_none"]), "with threshold", len(out["q_thr"]))


if __name__ == "__main__":
    main()
