"""GPU: the C-ABI used from plain C (examples/b200m_demo.c: no Python, no torch in the process) -- the binding a
maintainer's own host code would make around one image pair (matching.py:107-128). The demo is compiled here with
gcc against include/b200m.h and the in-tree library, run as a subprocess, and its match list and knnMatch(k=2)
table are compared bit for bit with the oracle."""
import os
import struct
import subprocess

import numpy as np
import pytest

from oracle import oracle_int

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_demo(out_dir):
    exe = os.path.join(str(out_dir), "b200m_demo")
    lib = os.path.join(ROOT, "3d_reconstruction_b200", "lib")
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = ["gcc", "-O2", "-std=c11", "-Wall", "-Werror", os.path.join(ROOT, "examples", "b200m_demo.c"),
           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(cuda, "include"), "-L" + lib, "-lb200m",
           "-L" + os.path.join(cuda, "lib64"), "-lcudart", "-Wl,-rpath," + lib, "-Wl,-rpath," + os.path.join(cuda, "lib64"),
           "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_demo_compiles_as_plain_c(pkg, tmp_path):
    """Not a GPU test: include/b200m.h is valid C11 and every entry point the demo uses links from the library."""
    assert os.path.exists(build_demo(tmp_path))


def _bits(x):
    return np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)


@pytest.mark.gpu
@pytest.mark.parametrize("kq,kt,ratio,mutual", [(700, 515, 0.75, 1), (3000, 4100, 0.75, 1), (257, 64, 0.0, 1), (900, 900, 0.8, 0)])
def test_demo_equals_oracle(pkg, tmp_path, kq, kt, ratio, mutual):
    exe = build_demo(tmp_path)
    rng = np.random.default_rng(kq * 7 + kt)
    world = np.abs(rng.standard_normal((max(kq, kt), 128)))
    world = np.clip(np.rint(world / np.linalg.norm(world, axis=1, keepdims=True) * 512), 0, 255)
    q = np.clip(world[rng.permutation(len(world))[:kq]] + rng.integers(-5, 6, (kq, 128)), 0, 255).astype(np.uint8)
    t = np.clip(world[rng.permutation(len(world))[:kt]] + rng.integers(-5, 6, (kt, 128)), 0, 255).astype(np.uint8)
    t[kt // 3] = t[0]                                   # a duplicate column: the stable order decides
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(fin, "wb") as f:
        f.write(struct.pack("<ii", kq, kt))
        f.write(q.tobytes())
        f.write(t.tobytes())
    r = subprocess.run([exe, fin, fout, str(ratio), str(mutual)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = open(fout, "rb").read()
    count = struct.unpack_from("<i", raw, 0)[0]
    rec = np.frombuffer(raw, dtype=np.dtype([("q", "<i4"), ("t", "<i4"), ("d", "<f4")]), count=count, offset=4)
    knn = np.frombuffer(raw, dtype=np.dtype([("i0", "<i4"), ("i1", "<i4"), ("d0", "<f4"), ("d1", "<f4")]), count=kq, offset=4 + 12 * count)
    eq, et, ed = oracle_int.match(q, t, ratio, bool(mutual))
    assert count == len(eq) and count > 10
    assert np.array_equal(rec["q"], eq) and np.array_equal(rec["t"], et) and np.array_equal(_bits(rec["d"]), _bits(ed))
    eidx, edist, _ = oracle_int.knn2(q, t)
    assert np.array_equal(np.stack([knn["i0"], knn["i1"]], 1), eidx)
    assert np.array_equal(_bits(np.stack([knn["d0"], knn["d1"]], 1)), _bits(edist))
