"""The distance formatting of the neighbours file (grid_b200/csrc/nbr_text.cuh, compiled here with g++) against Python's
own "%.2f": the device writer and grid_b200.io.save_neighbors must print the same bytes."""
import ctypes as C
import math
import subprocess
from pathlib import Path

import numpy as np
import pytest

REPO = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    out = tmp_path_factory.mktemp("step5") / "libstep5_host.so"
    subprocess.run(["g++", "-O1", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", str(out),
                    str(REPO / "tests" / "csrc" / "step5_host.cpp")], check=True)
    lib = C.CDLL(str(out))
    lib.nt_dist_hundredths.argtypes, lib.nt_dist_hundredths.restype = [C.c_longlong, C.c_double], C.c_ulonglong
    lib.nt_fixed2_hundredths.argtypes, lib.nt_fixed2_hundredths.restype = [C.c_double], C.c_ulonglong
    lib.nt_put_fixed2.argtypes, lib.nt_put_fixed2.restype = [C.c_ulonglong, C.c_char_p], C.c_int
    return lib


def text_of(lib, q: int) -> str:
    buf = C.create_string_buffer(40)
    n = lib.nt_put_fixed2(q, buf)
    assert n > 0
    return buf.raw[:n].decode()


def python_dist(d2: int, r_use: int) -> str:
    s = math.sqrt(d2 / 1e4)
    return f"{(s * s) / (2 * (r_use or 1)):.2f}"


def test_fixed2_equals_python_formatting(lib):
    rng = np.random.default_rng(1)
    vals = list(rng.random(20000) * 10.0 ** rng.integers(-4, 9, 20000))
    vals += [0.0, 0.005, 0.015, 0.025, 0.125, 0.375, 1.005, 2.675, 1e-9, 0.994999999999, 0.995, 1234567.895, 2.0 ** 43]
    vals += [k / 8 for k in range(0, 4000)]          # exactly representable .125/.375/.625/.875: true ties -> half-even
    vals += [np.nextafter(k + 0.005, 0) for k in range(50)] + [np.nextafter(k + 0.005, 9e9) for k in range(50)]
    for v in vals:
        v = float(v)
        assert text_of(lib, lib.nt_fixed2_hundredths(v)) == f"{v:.2f}", v


def test_distance_field_equals_the_host_writer(lib):
    rng = np.random.default_rng(2)
    for r_use in (0, 1, 7, 38, 1000, 2_609_999):
        d2 = rng.integers(0, 400 * 400 * max(r_use, 1) + 5, 6000).tolist()
        d2 += [(2 * m + 1) * 100 * max(r_use, 1) for m in range(300)]   # exact %.2f ties: the sqrt/square noise decides
        d2 += [0, 1, 2, 9_999, 10_000, (1 << 52) + 12345]
        for v in d2:
            assert text_of(lib, lib.nt_dist_hundredths(v, float(2 * (r_use or 1)))) == python_dist(v, r_use), (v, r_use)
