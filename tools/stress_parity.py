"""Randomised parity campaign of the default pipeline against the integer oracle (development aid):
ragged banks, mixed generators (uniform bytes, SIFT-like rows with planted near-duplicates, tiny alphabets
that tie everywhere, constant rows), both pair directions, ratio on/off, mutual on/off."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3d_reconstruction_b200")
from oracle import oracle_int  # noqa: E402


def gen(rng, k, kind, world):
    if kind == 0:
        return rng.integers(0, 256, size=(k, 128)).astype(np.uint8)
    if kind == 1:       # SIFT-like rows, a share re-observed from a common world with small noise
        a = np.abs(rng.standard_normal((k, 128)))
        a = np.clip(np.rint(a / np.linalg.norm(a, axis=1, keepdims=True) * 512), 0, 255)
        n = k // 2
        pick = rng.integers(0, len(world), size=n)
        a[:n] = np.clip(world[pick] + rng.integers(-4, 5, size=(n, 128)), 0, 255)
        return a[rng.permutation(k)].astype(np.uint8)
    if kind == 2:       # tiny alphabet: ties everywhere
        base = rng.integers(0, 2, size=(8, 128)).astype(np.uint8) * int(rng.integers(1, 256))
        return base[rng.integers(0, 8, size=k)]
    a = np.full((k, 128), int(rng.integers(0, 256)), np.uint8)      # constant rows + a few odd ones
    a[rng.integers(0, k, size=max(1, k // 10))] = rng.integers(0, 256, size=128).astype(np.uint8)
    return a


def main(rounds=40, seed=0):
    rng = np.random.default_rng(seed)
    checked = 0
    for r in range(rounds):
        n_img = int(rng.integers(2, 6))
        world = np.abs(rng.standard_normal((600, 128)))
        world = np.clip(np.rint(world / np.linalg.norm(world, axis=1, keepdims=True) * 512), 0, 255)
        kind = int(rng.integers(0, 4))
        sizes = [int(rng.choice([1, 2, 3, 31, 32, 33, 63, 64, 65, 200, 255, 256, 257, 511, 900, 1500, 2100])) for _ in range(n_img)]
        descs = [gen(rng, k, kind, world) for k in sizes]
        bank = pkg.DescriptorBank.from_arrays(descs, device="cuda")
        pairs = [(a, b) for a in range(n_img) for b in range(n_img)]
        ratio = float(rng.choice([0.75, 0.6, 0.9, 0.0]))
        mutual = bool(rng.integers(0, 2))
        res = pkg.match_pairs(bank, pairs, ratio=ratio, mutual=mutual)
        res.check_overflow()
        kidx, kdist = pkg.knn2_pairs(bank, pairs)           # the knnMatch(k=2) face on the same bank (chunk pipeline + tensor-core evaluator)
        kidx, kdist = kidx.cpu().numpy(), kdist.cpu().numpy()
        for pi, (a, b) in enumerate(pairs):
            eidx, edist, _ = oracle_int.knn2(descs[a], descs[b])
            n = len(descs[a])
            if not (np.array_equal(kidx[pi, :n], eidx) and np.array_equal(kdist[pi, :n].view(np.uint32), np.asarray(edist, np.float32).view(np.uint32))):
                bad = np.nonzero((kidx[pi, :n] != eidx).any(1))[0][:5]
                print("KNN MISMATCH round", r, "kind", kind, "pair", (a, b), "sizes", sizes[a], sizes[b], "rows", bad, flush=True)
                sys.exit(1)
        for (a, b), (q, t, d) in zip(pairs, res.to_lists()):
            eq, et, ed = oracle_int.match(descs[a], descs[b], ratio, mutual)
            ok = np.array_equal(q, eq) and np.array_equal(t, et) and np.array_equal(
                np.asarray(d, np.float32).view(np.uint32), np.asarray(ed, np.float32).view(np.uint32))
            if not ok:
                print("MISMATCH round", r, "kind", kind, "pair", (a, b), "sizes", sizes[a], sizes[b], "ratio", ratio, "mutual", mutual,
                      "got", len(q), "expected", len(eq), flush=True)
                sys.exit(1)
            checked += 1
    print("stress parity OK:", checked, "pairs over", rounds, "banks (match lists and knnMatch tables)", flush=True)


if __name__ == "__main__":
    # usage: stress_parity.py [rounds] [seed] [pipeline] [kernel] [chunk]
    if len(sys.argv) > 3:
        pkg.set_option("pipeline", int(sys.argv[3]))
    if len(sys.argv) > 4:
        pkg.set_option("kernel", int(sys.argv[4]))
    if len(sys.argv) > 5:
        pkg.set_option("chunk", int(sys.argv[5]))
    main(rounds=int(sys.argv[1]) if len(sys.argv) > 1 else 40, seed=int(sys.argv[2]) if len(sys.argv) > 2 else 0)
