"""CPU: the C-ABI library loads, exports every symbol include/b200m.h declares, and fails loudly
(no CPU fallback) when there is no sm_100 device."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "b200m.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200m_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(pkg):
    syms = _header_symbols()
    assert sorted(pkg.SYMBOLS) == syms
    dll = ctypes.CDLL(pkg.LIB_PATH)
    for s in syms:
        assert hasattr(dll, s), s
    assert dll.b200m_version() == 120


def test_options_roundtrip(pkg):
    pkg.set_option("batch", 7)
    assert pkg.get_option("batch") == 7
    pkg.set_option("batch", 0)
    with pytest.raises(pkg.B200MError):
        pkg.set_option("nonsense", 1)


def test_options_are_read_afresh_by_every_call(pkg):
    """Entry points work on a per-call copy of the knobs (one consistent view from the workspace check to the last
    batch); the copy must not outlive the call: a set_option shows in the very next size query, both ways."""
    lib = pkg._lib.lib()
    base = lib.b200m_workspace_bytes(256, 100)
    pkg.set_option("batch", 7)
    small = lib.b200m_workspace_bytes(256, 100)
    pkg.set_option("batch", 0)
    assert small < base and lib.b200m_workspace_bytes(256, 100) == base
    # the head's column partial sums (quadratic in k_stride) are only part of the size when the single-sweep
    # log-sum-exp can run ("float_sweep" = 0)
    one = lib.b200m_workspace_bytes_assign(2048, 64)
    pkg.set_option("float_sweep", 1)
    two = lib.b200m_workspace_bytes_assign(2048, 64)
    pkg.set_option("float_sweep", 0)
    assert one - two == 64 * (2048 // 256) * 2048 * 4 and lib.b200m_workspace_bytes_assign(2048, 64) == one


def test_workspace_bytes_saturates(pkg):
    lib = pkg._lib.lib()
    small = lib.b200m_workspace_bytes(256, 1)
    assert small >= 256 * 164
    big1 = lib.b200m_workspace_bytes(16384, 1024)
    big2 = lib.b200m_workspace_bytes(16384, 32896)
    assert big1 == big2 and big1 < (2 << 30)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(pkg):
    lib = pkg._lib.lib()
    assert lib.b200m_device_check() == pkg._lib.E_NO_DEVICE
    buf = np.zeros(256 * 128, np.uint8)
    k = np.array([1], np.int32)
    n = np.zeros(256, np.int32)
    rc = lib.b200m_bank_prepare_u8(buf.ctypes.data, 1, 256, k.ctypes.data, n.ctypes.data, None)
    assert rc == pkg._lib.E_NO_DEVICE
    with pytest.raises(RuntimeError):
        pkg.DescriptorBank.from_arrays([np.zeros((4, 128), np.uint8)], device="cpu")
    with pytest.raises((RuntimeError, AssertionError)):
        pkg.match(np.zeros((4, 128), np.uint8), np.zeros((4, 128), np.uint8))


def test_bad_arguments(pkg):
    lib = pkg._lib.lib()
    assert lib.b200m_bank_prepare_u8(None, 1, 256, None, None, None) == -1
    buf = np.zeros(8, np.uint8)
    assert lib.b200m_bank_prepare_u8(buf.ctypes.data, 1, 100, buf.ctypes.data, buf.ctypes.data, None) == -1
    assert b"256" in lib.b200m_last_error_string()


def test_u8_packing_validation(pkg):
    ok = pkg.to_u8_descriptors(np.arange(256, dtype=np.float32).repeat(64).reshape(128, 128) % 256)
    assert ok.dtype == np.uint8
    with pytest.raises(ValueError):
        pkg.to_u8_descriptors(np.full((2, 128), 0.5, np.float32))
    with pytest.raises(ValueError):
        pkg.to_u8_descriptors(np.full((2, 128), 256.0, np.float32))
    with pytest.raises(ValueError):
        pkg.to_u8_descriptors(np.zeros((2, 64), np.uint8))
