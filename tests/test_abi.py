"""The C-ABI library loads without a GPU and exports every symbol include/gridb200.h declares."""
import ctypes
import re
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parents[1]


def declared():
    text = (REPO / "include" / "gridb200.h").read_text()
    return re.findall(r"GRIDB200_API\s+[\w\s\*]+?\b(gridb200_\w+)\s*\(", text)


@pytest.fixture(scope="module")
def lib_path():
    from grid_b200.csrc import build

    return build.build()


def test_header_declares_the_path():
    names = declared()
    assert len(names) == len(set(names)) >= 20
    for must in ("gridb200_col_raw_stats", "gridb200_row_stats", "gridb200_col_norm_stats", "gridb200_select",
                 "gridb200_transform", "gridb200_gram", "gridb200_topk_merge", "gridb200_dipcn", "gridb200_ingest_text"):
        assert must in names


def test_library_exports_every_declared_symbol(lib_path):
    lib = ctypes.CDLL(str(lib_path))
    for name in declared():
        assert hasattr(lib, name), f"{name} declared in gridb200.h but not exported"


def test_python_binding_covers_the_header(lib_path):
    from grid_b200 import _lib

    assert set(_lib.SIGNATURES) == set(declared())
    lib = _lib.load()
    assert lib.gridb200_version().startswith(b"gridb200")
    assert lib.gridb200_launch_count() == 0
    assert lib.gridb200_select_ws_bytes(1000) > 1000 * 8
    assert lib.gridb200_col_norm_stats_ws_bytes(100, 43) > 0        # few bins: row slabs need partial buffers
    # many bins: row slabs only even out the last wave (6 slabs of 544 rows at config 2), none for a short matrix
    assert lib.gridb200_col_norm_stats_ws_bytes(3202, 2_900_000) == 6 * 2_900_000 * 16
    assert lib.gridb200_col_norm_stats_ws_bytes(200, 2_900_000) == 0


def test_bad_arguments_are_reported_not_crashed(lib_path):
    from grid_b200 import _lib

    with pytest.raises(_lib.GridB200Error, match="null"):
        _lib.call("gridb200_col_raw_stats", None, 4, 4, 4, None, None, None)
    with pytest.raises(_lib.GridB200Error, match="multiple of 4"):
        _lib.call("gridb200_col_raw_stats", 4096, 4, 5, 5, 4096, 4096, None)
    with pytest.raises(_lib.GridB200Error, match="kdim"):
        _lib.call("gridb200_gram", 4096, 4, 4, 4096, 4, 4, 70, 200, 4096, 4, 0, 1, None)
    with pytest.raises(_lib.GridB200Error, match="vmax"):
        _lib.call("gridb200_gram", 4096, 4, 4, 4096, 4, 4, 64, 5000, 4096, 4, 0, 1, None)
