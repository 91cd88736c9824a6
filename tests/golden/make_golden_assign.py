"""Generates tests/golden/assign_{128,256}.npz with the REFERENCE's own assignment head:
sigmoid_log_double_softmax (lightglue/lightglue.py:196-208) and filter_matches (:235-254), imported by
path from /root/reference (lightglue/__init__.py needs kornia) -- run in the build container:

    python tests/golden/make_golden_assign.py

Inputs: projected descriptors mdesc0 / mdesc1 (what MatchAssignment.forward holds after :220-222),
bf16-representable and stored as their upper 16 bits, with planted correspondences scaled so that a true
match has similarity ~ 20 (a peaked soft-max, like a trained head), and matchability logits z0 / z1.
Outputs: m0, m1, mscores0, mscores1 of filter_matches(scores, 0.1) and of filter_matches(scores, None).
"""
import os

import numpy as np
import torch

from make_golden_float import HERE, load_reference


def make_inputs(m, n, dim, seed, scale=4.5):
    g = torch.Generator().manual_seed(seed)
    w = int(1.4 * max(m, n))          # a share of the points has no partner in the other image
    world = torch.nn.functional.normalize(torch.randn(w, dim, generator=g), dim=1)
    d0 = world[torch.randperm(w, generator=g)[:m]] + 0.04 * torch.randn(m, dim, generator=g)
    d1 = world[torch.randperm(w, generator=g)[:n]] + 0.04 * torch.randn(n, dim, generator=g)
    d0 = (scale * torch.nn.functional.normalize(d0, dim=1)).to(torch.bfloat16)
    d1 = (scale * torch.nn.functional.normalize(d1, dim=1)).to(torch.bfloat16)
    z0 = 1.0 + 1.5 * torch.randn(m, generator=g)
    z1 = 1.0 + 1.5 * torch.randn(n, generator=g)
    return d0, d1, z0, z1


def main():
    ref = load_reference()
    for dim, m, n in ((128, 420, 390), (256, 400, 380)):
        d0, d1, z0, z1 = make_inputs(m, n, dim, 300 + dim)
        sim = torch.einsum("bmd,bnd->bmn", d0.float()[None], d1.float()[None])        # lightglue.py:223
        scores = ref.sigmoid_log_double_softmax(sim, z0[None, :, None], z1[None, :, None])   # :226
        out = {}
        for name, th in (("none", None), ("thr", 0.1)):
            m0, m1, ms0, ms1 = ref.filter_matches(scores, th)                             # :235-254
            out[f"m0_{name}"] = m0[0].numpy()
            out[f"m1_{name}"] = m1[0].numpy()
            out[f"ms0_{name}"] = ms0[0].numpy()
            out[f"ms1_{name}"] = ms1[0].numpy()
        np.savez_compressed(os.path.join(HERE, f"assign_{dim}.npz"),
                            d0_bf16=d0.view(torch.int16).numpy(), d1_bf16=d1.view(torch.int16).numpy(),
                            z0=z0.numpy(), z1=z1.numpy(), threshold=np.float32(0.1),
                            scores_row7=scores[0, 7, :-1].numpy(), torch_version=torch.__version__, **out)
        print(dim, "mutual", int((out["m0_none"] > -1).sum()), "above 0.1", int((out["m0_thr"] > -1).sum()))


if __name__ == "__main__":
    main()
