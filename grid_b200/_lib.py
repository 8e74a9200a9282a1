"""ctypes binding of libgridb200.so (the C ABI declared in include/gridb200.h).

There is no fallback: if the shared library is missing or a call fails, the caller gets an exception.
"""

from __future__ import annotations

import ctypes as C
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "csrc" / "libgridb200.so"

_i64, _i32, _u64, _f64, _vp, _sz = C.c_int64, C.c_int32, C.c_uint64, C.c_double, C.c_void_p, C.c_size_t

# name -> (restype, argtypes); must list every symbol include/gridb200.h declares (tests/test_abi.py checks)
SIGNATURES = {
    "gridb200_last_error": (C.c_char_p, []),
    "gridb200_version": (C.c_char_p, []),
    "gridb200_launch_count": (_u64, []),
    "gridb200_col_raw_stats": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp]),
    "gridb200_col_row_raw_stats": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp]),
    "gridb200_row_stats_fused_ws_bytes": (_sz, [_i64]),
    "gridb200_row_stats_fused": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gridb200_depth_filter": (C.c_int, [_vp, _vp, _i64, _f64, _f64, _vp, _vp]),
    "gridb200_row_stats": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp]),
    "gridb200_row_scales": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _vp, _vp]),
    "gridb200_col_norm_stats_ws_bytes": (_sz, [_i64, _i64]),
    "gridb200_col_norm_stats": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gridb200_col_finalize": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    "gridb200_select_ws_bytes": (_sz, [_i64]),
    "gridb200_select": (C.c_int, [_vp, _i64, _f64, _vp, _vp, _vp, _sz, _vp]),
    "gridb200_select_static": (C.c_int, [_vp, _i64, _f64, _vp, _vp, _vp]),
    "gridb200_knn_col_filter": (C.c_int, [_vp, _vp, _i64, _f64, _f64, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gridb200_transform": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _i64, _vp, _i64,
                                     _vp, _vp, _vp]),
    "gridb200_knn_prepare": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _i32, _vp, _i64, _vp, _vp]),
    "gridb200_gram": (C.c_int, [_vp, _i64, _i64, _vp, _i64, _i64, _i64, _i32, _vp, _i64, C.c_int, C.c_int, _vp]),
    "gridb200_gram_set_max_pairs": (C.c_int, [C.c_int]),
    "gridb200_topk_init": (C.c_int, [_vp, _vp, _i64, _i32, _vp]),
    "gridb200_topk_merge": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i32, _vp, _vp, _vp]),
    "gridb200_topk_merge_t": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i32, _vp, _vp, _vp]),
    "gridb200_topk_merge_lists": (C.c_int, [_vp, _vp, _i64, _i32, _i32, _vp, _vp, _vp]),
    "gridb200_gram_topk_ws_bytes": (_sz, []),
    "gridb200_gram_topk": (C.c_int, [_vp, _i64, _i64, _vp, _i64, _i64, _i64, _i32, _vp, _vp, _i64, _i64, _vp, _i64, _vp, _i64,
                                     _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gridb200_topk_merge_cand": (C.c_int, [_vp, _vp, _vp, _i64, _i32, _i32, _vp, _vp, _vp]),
    "gridb200_split_tf32": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _vp, _vp]),
    "gridb200_gram_f32": (C.c_int, [_vp, _vp, _i64, _i64, _vp, _vp, _i64, _i64, _i64, _vp, _i64, _vp]),
    "gridb200_topk_merge_f32": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i32, _vp, _vp, _vp]),
    "gridb200_dipcn": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp, _i64, _i32, _i64, _i32, _i32, _vp, _i64, _vp, _vp]),
    "gridb200_synth_depth": (C.c_int, [_u64, _i64, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _i64, _vp]),
    "gridb200_ingest_ws_bytes": (_sz, [_i64]),
    "gridb200_ingest_text": (C.c_int, [_vp, _i64, C.c_char_p, _i32, _i64, _i64, _vp, _i64, _i64, _i64, _vp, _vp, _vp,
                                       _sz, _vp]),
    "gridb200_depth_pack_u16": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _i64, _vp, _vp]),
    "gridb200_depth_unpack_u16": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _i64, _i64, _vp]),
    "gridb200_depth_packed_row_words": (_i64, [_i64, _i32]),
    "gridb200_depth_pack_bits": (C.c_int, [_vp, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _i64, _i64, _vp, _vp, _i64, _vp, _vp]),
    "gridb200_depth_unpack_bits": (C.c_int, [_vp, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _i64, _vp, _vp, _i64, _i64, _vp]),
    "gridb200_depth_unpack_bits_u16": (C.c_int, [_vp, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _i64, _vp]),
    "gridb200_col_row_raw_stats_u16": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp]),
    "gridb200_row_stats_fused_u16": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gridb200_row_means_exact_u16": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gridb200_col_norm_stats_u16": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _vp,
                                              _sz, _vp]),
    "gridb200_transform_u16": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _i64, _vp, _i64,
                                         _vp, _vp, _vp]),
    "gridb200_hi_phase": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _vp]),
    "gridb200_hi_level_pass": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _vp]),
    "gridb200_hi_impute": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _f64, _vp, _vp]),
    "gridb200_row_means_exact": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gridb200_format_z_ws_bytes": (_sz, [_i64, _i64]),
    "gridb200_format_z_measure": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _sz, _vp]),
    "gridb200_format_z_write": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp, _i64, _vp]),
    "gridb200_parse_z_ws_bytes": (_sz, [_i64, _i64]),
    "gridb200_parse_z_rows": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _i64, _vp, _i64, _vp, _vp, _sz, _vp]),
    "gridb200_format_nbrs_ws_bytes": (_sz, [_i64]),
    "gridb200_format_nbrs_measure": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _i32, _i64, _i64, _i64, _vp, _vp, _sz, _vp]),
    "gridb200_format_nbrs_write": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _i32, _i64, _i64, _i64, _vp, _vp, _i64, _vp]),
    "gridb200_hash_tokens": (C.c_int, [_vp, _vp, _i64, _vp, _vp]),
    "gridb200_parse_nbrs_rows": (C.c_int, [_vp, _vp, _i64, _i64, _vp, _vp, _vp, _vp, _i64, _i32, _vp, _i64, _vp, _vp, _vp,
                                           _sz, _vp]),
    "gridb200_region_coverage": (C.c_int, [_vp, _i64, _vp, _i32, C.c_char_p, _i32, _i64, _i64, _vp, _vp]),
    "gridb200_count_reads_ws_bytes": (_sz, [_i32, _i32]),
    "gridb200_count_reads": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _vp, _i32, _vp, _i32, _vp, _i32, _vp, _vp, _i32,
                                       _vp, _i32, _i32, _i32, _vp, _vp, _sz, _vp]),
}


class GridB200Error(RuntimeError):
    """A libgridb200 call returned a non-zero status."""

    def __init__(self, fn: str, code: int, msg: str):
        super().__init__(f"{fn} failed ({code}): {msg}")
        self.code = code


_lib = None


def load() -> C.CDLL:
    """Load the library once; raises if it has not been built (python grid_b200/csrc/build.py)."""
    global _lib
    if _lib is None:
        import os

        path = Path(os.environ.get("GRIDB200_LIB", LIB_PATH))  # A/B runs of two builds of the SAME library
        if not path.exists():
            raise ImportError(f"{path} is missing: build it with `python grid_b200/csrc/build.py` "
                              "(there is no CPU fallback)")
        lib = C.CDLL(str(path))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def call(name: str, *args):
    """Call an int-status entry point and raise GridB200Error on failure."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise GridB200Error(name, rc, lib.gridb200_last_error().decode())


def ptr(t) -> int | None:
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()
