"""b200m_vq_f32_static: visual-word assignment for hosts that cannot allocate device memory.

Who needs it: a plugin that only ever sees the device buffers its engine hands it (the engine owns the pool), a
binding that links libb200m.so but not the CUDA runtime, a service that must not cudaMalloc once it is up. Such a
caller cannot supply the float64 [n_codes] scratch b200m_vq_f32 asks for, so this variant keeps the squared codebook
norms in a 32 KiB array built into the library's device image (B200M_VQ_STATIC_MAX_CODES = 4096 rows).

The file name sorts after every other test file on purpose: pytest -x stops at the first failure, and a failure
here should not hide the parity tests. The GPU test has to pass both in a fresh process and after the rest of the
suite has used the library (the address of the built-in array is looked up once per device, on the first call).
"""
import ctypes as C
import re
import os

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _max_codes():
    src = open(os.path.join(ROOT, "include", "b200m.h")).read()
    return int(re.search(r"#define\s+B200M_VQ_STATIC_MAX_CODES\s+(\d+)", src).group(1))


def test_static_vq_refuses_codebooks_beyond_its_array(pkg):
    """The size check comes before anything touches a device, so it can be exercised anywhere: one row more than
    the built-in array holds is B200M_E_BAD_ARG, and the message names the limit and the entry point to use
    instead. Nothing is dereferenced (the pointers are dummies)."""
    lib = pkg._lib.lib()
    limit = _max_codes()
    assert limit == 4096
    dummy = C.c_void_p(256)
    rc = lib.b200m_vq_f32_static(dummy, 1, 128, dummy, limit + 1, dummy, None, None)
    assert rc == -1
    msg = lib.b200m_last_error_string().decode()
    assert str(limit) in msg and "b200m_vq_f32" in msg and str(limit + 1) in msg
    # n_obs == 0 is a no-op, as for b200m_vq_f32
    assert lib.b200m_vq_f32_static(None, 0, 128, None, 200, None, None, None) == 0


def _vq_both(pkg, obs_np, codebook_np):
    """Runs b200m_vq_f32 (scratch from torch) and b200m_vq_f32_static (no scratch argument) on the same device
    inputs through the C-ABI and returns (codes, dists) of each. Inputs and outputs are allocated by the test,
    as a caller of either entry point would; the static call itself is given nothing else."""
    lib = pkg._lib.lib()
    dev = torch.device("cuda:0")
    obs = torch.from_numpy(np.ascontiguousarray(obs_np, dtype=np.float32)).to(dev)
    cb = torch.from_numpy(np.ascontiguousarray(codebook_np, dtype=np.float64)).to(dev)
    n_obs, dim = obs.shape
    n_codes = cb.shape[0]
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream().cuda_stream
        code_a = torch.full((n_obs,), -7, dtype=torch.int32, device=dev)
        dist_a = torch.full((n_obs,), -7.0, dtype=torch.float64, device=dev)
        scratch = torch.empty((n_codes,), dtype=torch.float64, device=dev)
        pkg._lib.check(lib.b200m_vq_f32(obs.data_ptr(), n_obs, dim, cb.data_ptr(), n_codes, code_a.data_ptr(),
                                        dist_a.data_ptr(), scratch.data_ptr(), st))
        code_b = torch.full((n_obs,), -9, dtype=torch.int32, device=dev)
        dist_b = torch.full((n_obs,), -9.0, dtype=torch.float64, device=dev)
        pkg._lib.check(lib.b200m_vq_f32_static(obs.data_ptr(), n_obs, dim, cb.data_ptr(), n_codes, code_b.data_ptr(),
                                               dist_b.data_ptr(), st))
        torch.cuda.synchronize(dev)
        return code_a.cpu().numpy(), dist_a.cpu().numpy(), code_b.cpu().numpy(), dist_b.cpu().numpy()


@pytest.mark.gpu
def test_static_vq_equals_vq_on_two_calls_in_a_row(pkg):
    """Two calls back to back through the C-ABI, each checked against b200m_vq_f32 on the same inputs. Both entry
    points run the same two kernels in the same order, so the bar is the one the vq tests use for integer work --
    identical words -- and here also bit-identical float64 distances.

    The second call uses a SMALLER codebook, of a different descriptor length, than the first: the built-in array
    still holds the first call's norms beyond the new n_codes, and a correct call must neither need them cleared nor
    read them. The first call also covers the full 4096-row limit; a third repeats the first codebook to show the
    array is reusable after a different call overwrote part of it."""
    from scipy.cluster.vq import vq
    rng = np.random.default_rng(20260611)
    limit = _max_codes()
    # call 1: the largest codebook the built-in array takes, SIFT-like integer-valued descriptors, ragged n_obs
    obs1 = np.abs(rng.standard_normal((5000 + 13, 128)) * 60).clip(0, 255).round().astype(np.float32)
    cb1 = np.abs(rng.standard_normal((limit, 128)) * 60).clip(0, 255).astype(np.float64)
    # call 2: 200 words as bow.py:7, float descriptors of another length
    obs2 = rng.standard_normal((2048 + 1, 256)).astype(np.float32)
    cb2 = rng.standard_normal((200, 256))
    first = None
    for k, (obs, cb) in enumerate(((obs1, cb1), (obs2, cb2), (obs1, cb1))):
        code_a, dist_a, code_b, dist_b = _vq_both(pkg, obs, cb)
        assert (code_b >= 0).all() and (code_b < cb.shape[0]).all()
        assert np.array_equal(code_a, code_b), f"call {k}: words differ in {(code_a != code_b).sum()} rows"
        assert np.array_equal(dist_a.view(np.int64), dist_b.view(np.int64)), f"call {k}: distances differ"
        if k == 1:   # small enough to check against what the reference calls (matching.py:27), like the vq tests do
            exp, _ = vq(obs.astype("float"), cb)
            assert np.array_equal(code_b, exp)
        if k == 0:
            first = (code_b.copy(), dist_b.copy())
        if k == 2:
            assert np.array_equal(first[0], code_b) and np.array_equal(first[1].view(np.int64), dist_b.view(np.int64))
