"""ctypes binding of the C-ABI in include/b200m.h (lib/libb200m.so, built in-tree by build.py).

The product path has no CPU fallback: if the shared library is missing, loading raises; if the
machine has no sm_100 GPU, every compute entry point returns B200M_E_NO_DEVICE and we raise.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200m.so")

# every symbol include/b200m.h declares (tests check the library exports all of them)
SYMBOLS = (
    "b200m_version",
    "b200m_last_error_string",
    "b200m_device_check",
    "b200m_sm_count",
    "b200m_norm_table_elems",
    "b200m_bank_prepare_u8",
    "b200m_sort_table_elems",
    "b200m_bank_sort_u8",
    "b200m_bank_prepare_range_u8",
    "b200m_bank_sort_range_u8",
    "b200m_workspace_bytes",
    "b200m_knn2_u8",
    "b200m_knn2_sorted_u8",
    "b200m_match_pairs_u8",
    "b200m_workspace_bytes_bf16",
    "b200m_match_pairs_bf16",
    "b200m_workspace_bytes_assign",
    "b200m_workspace_bytes_ransac",
    "b200m_essential_ransac_f32",
    "b200m_five_point_f64",
    "b200m_assign_pairs_bf16",
    "b200m_vq_f32",
    "b200m_vq_f32_static",
    "b200m_compact_matches",
    "b200m_match_stats_read",
    "b200m_profile_read",
    "b200m_profile_detail",
    "b200m_set_option",
    "b200m_get_option",
    "b200m_debug_dot_tile_u8",
)

E_NO_DEVICE = -4


class B200MError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b200m error {code}: {msg}")
        self.code = code


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build the CUDA extension first (python -c 'import __graft_entry__ as g; g.build()'). "
            "There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, sz, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t, C.c_double
    L.b200m_version.restype = C.c_int
    L.b200m_last_error_string.restype = C.c_char_p
    L.b200m_device_check.restype = C.c_int
    L.b200m_sm_count.restype = C.c_int
    L.b200m_norm_table_elems.restype = i64
    L.b200m_norm_table_elems.argtypes = [i32, i64]
    L.b200m_bank_prepare_u8.restype = C.c_int
    L.b200m_bank_prepare_u8.argtypes = [vp, i32, i64, vp, vp, vp]
    L.b200m_sort_table_elems.restype = i64
    L.b200m_sort_table_elems.argtypes = [i32, i64]
    L.b200m_bank_sort_u8.restype = C.c_int
    L.b200m_bank_sort_u8.argtypes = [vp, vp, vp, i32, i64, vp, vp, vp]
    L.b200m_bank_prepare_range_u8.restype = C.c_int
    L.b200m_bank_prepare_range_u8.argtypes = [vp, i32, i64, vp, vp, i32, i32, vp]
    L.b200m_bank_sort_range_u8.restype = C.c_int
    L.b200m_bank_sort_range_u8.argtypes = [vp, vp, vp, i32, i64, vp, vp, i32, i32, vp]
    L.b200m_workspace_bytes.restype = sz
    L.b200m_workspace_bytes.argtypes = [i64, i64]
    L.b200m_knn2_u8.restype = C.c_int
    L.b200m_knn2_u8.argtypes = [vp, vp, vp, i32, i64, vp, i64, vp, vp, vp, sz, vp]
    L.b200m_knn2_sorted_u8.restype = C.c_int
    L.b200m_knn2_sorted_u8.argtypes = [vp, vp, vp, vp, vp, i32, i64, vp, i64, vp, vp, vp, sz, vp]
    L.b200m_match_pairs_u8.restype = C.c_int
    L.b200m_match_pairs_u8.argtypes = [vp, vp, vp, vp, vp, i32, i64, vp, i64, dbl, C.c_int, vp, vp, vp, vp, i64, vp, sz, vp]
    L.b200m_workspace_bytes_bf16.restype = sz
    L.b200m_workspace_bytes_bf16.argtypes = [i64, i64]
    L.b200m_workspace_bytes_assign.restype = sz
    L.b200m_workspace_bytes_assign.argtypes = [i64, i64]
    L.b200m_assign_pairs_bf16.restype = C.c_int
    L.b200m_assign_pairs_bf16.argtypes = [vp, vp, vp, i32, i64, i32, vp, i64, C.c_float, C.c_int, vp, vp, vp, vp, i64,
                                          vp, sz, vp]
    L.b200m_workspace_bytes_ransac.restype = sz
    L.b200m_workspace_bytes_ransac.argtypes = [i64, i32]
    L.b200m_essential_ransac_f32.restype = C.c_int
    L.b200m_essential_ransac_f32.argtypes = [vp, vp, i64, i32, C.c_float, C.c_uint64, C.c_float, vp, vp, vp, vp, vp, vp, vp,
                                             vp, sz, vp]
    L.b200m_five_point_f64.restype = C.c_int
    L.b200m_five_point_f64.argtypes = [vp, vp, vp, vp, vp]
    L.b200m_match_pairs_bf16.restype = C.c_int
    L.b200m_match_pairs_bf16.argtypes = [vp, vp, i32, i64, i32, vp, i64, C.c_float, C.c_int, C.c_int, vp, vp, vp, vp, i64,
                                         vp, sz, vp]
    L.b200m_vq_f32.restype = C.c_int
    L.b200m_vq_f32.argtypes = [vp, i64, i32, vp, i32, vp, vp, vp, vp]
    L.b200m_vq_f32_static.restype = C.c_int
    L.b200m_vq_f32_static.argtypes = [vp, i64, i32, vp, i32, vp, vp, vp]
    L.b200m_compact_matches.restype = C.c_int
    L.b200m_compact_matches.argtypes = [vp, vp, vp, vp, vp, i64, i64, vp, vp, vp, vp]
    L.b200m_profile_read.restype = C.c_int
    L.b200m_profile_read.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(i64), C.POINTER(i64)]
    L.b200m_match_stats_read.restype = C.c_int
    L.b200m_match_stats_read.argtypes = [vp, C.POINTER(i64), C.c_int, vp]
    L.b200m_profile_detail.restype = C.c_int
    L.b200m_profile_detail.argtypes = [C.POINTER(C.c_double), C.c_int]
    L.b200m_set_option.restype = C.c_int
    L.b200m_set_option.argtypes = [C.c_char_p, i64]
    L.b200m_get_option.restype = i64
    L.b200m_get_option.argtypes = [C.c_char_p]
    L.b200m_debug_dot_tile_u8.restype = C.c_int
    L.b200m_debug_dot_tile_u8.argtypes = [vp, vp, vp, vp]
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        raise B200MError(rc, lib().b200m_last_error_string().decode("utf-8", "replace"))


def set_option(name: str, value: int) -> None:
    check(lib().b200m_set_option(name.encode(), int(value)))


def get_option(name: str) -> int:
    return int(lib().b200m_get_option(name.encode()))


STAT_FIELDS = ("rows", "rows_chunk_evaluated", "rows_fallback", "candidates", "rows_ratio_verify", "matches")


def match_stats(workspace) -> dict:
    """Counters of the last match_pairs call that used `workspace` (a torch uint8 CUDA tensor); blocks."""
    import torch
    out = (C.c_int64 * 6)()
    with torch.cuda.device(workspace.device):
        check(lib().b200m_match_stats_read(workspace.data_ptr(), out, 6, torch.cuda.current_stream().cuda_stream))
    return dict(zip(STAT_FIELDS, (int(x) for x in out)))


def profile_read():
    """-> dict(first_sweep_ms, verify_sweep_ms, first_sweeps, launches, fallback_sweep_ms, refine_ms, select_ms,
    emit_ms); resets the counters."""
    det = (C.c_double * 6)()
    check(lib().b200m_profile_detail(det, 6))
    a, b = C.c_double(0), C.c_double(0)
    n, l = C.c_int64(0), C.c_int64(0)
    check(lib().b200m_profile_read(C.byref(a), C.byref(b), C.byref(n), C.byref(l)))
    return {"first_sweep_ms": a.value, "verify_sweep_ms": b.value, "first_sweeps": n.value, "launches": l.value,
            "fallback_sweep_ms": det[2], "refine_ms": det[3], "select_ms": det[4], "emit_ms": det[5]}
