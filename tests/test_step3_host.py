"""Step-3 producers on the CPU: the oracle against the reference-generated goldens and the reference's own known
answers; the shipped line grammar / read predicate (grid_b200/csrc/region_line.cuh, compiled here with g++ through
tests/csrc/step3_host.cpp) fuzzed against the oracle; the host-side decision and region-table logic."""
import ctypes as C
import gzip
import json
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "oracle"))

import step3_oracle as so  # noqa: E402
from gen_golden_cases import STEP3_COVERAGE, STEP3_READS  # noqa: E402

from grid_b200 import synth  # noqa: E402
from grid_b200.utils import count_reads as gcr  # noqa: E402
from grid_b200.utils import mosdepth as gmd  # noqa: E402

GOLD = json.loads((REPO / "tests" / "golden" / "step3.json").read_text())


# ---- oracle pinned on the unmodified reference's outputs -------------------------------------------------------------
@pytest.mark.parametrize("case", sorted(STEP3_COVERAGE))
def test_coverage_oracle_matches_reference_golden(case, tmp_path):
    paths = synth.write_region_files(tmp_path, **STEP3_COVERAGE[case])
    for name, chrom, lo, hi in synth.REGION_QUERIES:
        assert [so.compute_region_coverage(p, chrom, lo, hi) for p in paths] == GOLD["coverage"][case][name], name


@pytest.mark.parametrize("case", sorted(STEP3_READS))
def test_read_count_oracle_matches_reference_golden(case):
    kw = STEP3_READS[case]
    cols, off = synth.read_columns(kw["seed"], kw["ns"], kw["mean_reads"])
    got = so.count_reads_regions(cols, off, synth.READ_REGIONS, kw["flags"], kw["min_mapq"])
    assert got.tolist() == GOLD["reads"][case]


def _bed(tmp_path, lines):
    f = tmp_path / "regions.bed.gz"
    with gzip.open(f, "wt") as fh:
        for line in lines:
            fh.write(line + "\n")
    return f


# the reference's known answers, test/test_mosdepth.py:23-56
REFERENCE_KNOWN_ANSWERS = [
    (["chr6\t1000\t2000\t30.0"], ("chr6", 1000, 2000), 3000),
    (["chr6\t0\t2000\t40.0"], ("chr6", 1000, 2000), 4000),
    (["chr6\t1000\t2000\t20.0", "chr6\t2000\t3000\t40.0"], ("chr6", 1000, 3000), 3000),
    (["chr1\t1000\t2000\t30.0"], ("chr6", 1000, 2000), 0),
    (["chr6\t5000\t6000\t30.0"], ("chr6", 1000, 2000), 0),
    ([], ("chr6", 1000, 2000), 0),
]


@pytest.mark.parametrize("lines,query,want", REFERENCE_KNOWN_ANSWERS)
def test_coverage_oracle_on_the_reference_known_answers(lines, query, want, tmp_path):
    assert so.compute_region_coverage(_bed(tmp_path, lines), *query) == want
    assert so.compute_region_coverage_text("".join(ln + "\n" for ln in lines), *query) == want


# ---- the shipped grammar, compiled for the host ----------------------------------------------------------------------
@pytest.fixture(scope="module")
def hostlib(tmp_path_factory):
    out = tmp_path_factory.mktemp("step3") / "libstep3_host.so"
    subprocess.run(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-o", str(out), str(REPO / "tests" / "csrc" / "step3_host.cpp")],
                   check=True)
    lib = C.CDLL(str(out))
    lib.rl_region_coverage.argtypes = [C.c_char_p, C.c_longlong, C.c_char_p, C.c_int, C.c_longlong, C.c_longlong, C.c_void_p]
    lib.rl_count_reads.argtypes = [C.c_void_p] * 5 + [C.c_longlong, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                                      C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    lib.rl_bucket_check.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int]
    return lib


def grammar(lib, text: bytes, chrom: str, start: int, end: int):
    out = (C.c_ulonglong * 4)()
    lib.rl_region_coverage(text, len(text), chrom.encode(), len(chrom.encode()), start, end, out)
    return list(out)


def decided(lib, text: bytes, chrom, start, end):
    """What region_coverage_batch returns for this text given the device integers (replay on irregular / near-tie)."""
    weighted, covered, lines, irregular = grammar(lib, text, chrom, start, end)
    res = None if irregular else gmd.decide_coverage(weighted, covered, lines)
    return ("replay", None) if res is None else ("exact", res)


def test_grammar_on_reference_known_answers(hostlib):
    for lines, query, want in REFERENCE_KNOWN_ANSWERS:
        text = "".join(ln + "\n" for ln in lines).encode()
        assert decided(hostlib, text, *query) == ("exact", want)


@pytest.mark.parametrize("case", sorted(STEP3_COVERAGE))
def test_grammar_matches_golden_on_cohort_files(hostlib, case, tmp_path):
    paths = synth.write_region_files(tmp_path, **STEP3_COVERAGE[case])
    for name, chrom, lo, hi in synth.REGION_QUERIES:
        got = []
        for p in paths:
            kind, v = decided(hostlib, gzip.open(p, "rb").read(), chrom, lo, hi)
            got.append(v if kind == "exact" else so.compute_region_coverage(p, chrom, lo, hi))
        assert got == GOLD["coverage"][case][name], name


IRREGULAR_LINES = [
    b"", b" chr6\t1\t2\t3.00", b"\tchr6\t1\t2\t3.00", b"chr6\t1\t2", b"chr6\t1\t2\t", b"chr6\t1\t2\t\t", b"chr6\t+1\t2\t3.00",
    b"chr6\t1\t2\t3.000", b"chr6\t1\t2\t1e1", b"chr6\t1\t2\tnan", b"chr6\t1\t2\t-3.00", b"chr6\t1_0\t20\t3.00", b"chr6\t1\t2\t.5",
    b"chr6\t1\t2\t3.0 x", b"chr6\t1\t2\t3.0 \tx", b"chr6\t12345678901\t2\t3.00", b"chr6\t1\t2\t41943.04", b"chr6\t1\t2\t1234567890.0",
    b"\x1cchr6\t1\t2\t3.00", b"\xc2\xa0chr6\t1\t2\t3.00", b"chr6\t 1\t2\t3.00", b"chr6\t1.0\t2\t3.00",
]
REGULAR_LINES = [
    b"chr6\t1\t2\t3", b"chr6\t1\t2\t3.", b"chr6\t1\t2\t3.5", b"chr6\t1\t2\t3.50\t", b"chr6\t1\t2\t3.50 ", b"chr6\t1\t2\t3.50\tx\ty",
    b"chr6\t001\t2\t03.50", b"chr6 \t1\t2\t3.50", b"c\t1\t2\t0.00", b"chr6\t1\t2\t41943.03", b"chr6\t9999999999\t2\t3.50 \x0b\x0c",
]


def test_every_line_outside_the_grammar_is_reported(hostlib):
    """A line the device does not understand must send the file to the host replay, on any chromosome."""
    for ln in IRREGULAR_LINES:
        text = b"chr5\t0\t10\t1.00\n" + ln + b"\nchr6\t0\t10\t2.00\n"
        assert grammar(hostlib, text, "chr6", 0, 10)[3] == 1, ln
    for ln in REGULAR_LINES:
        text = b"chr5\t0\t10\t1.00\n" + ln + b"\n"
        assert grammar(hostlib, text, "chr6", 0, 10)[3] == 0, ln
        # and on regular lines the decision agrees with Python (or defers)
        kind, v = decided(hostlib, text, "chr6", 0, 5)
        assert kind == "replay" or v == so.compute_region_coverage_text(text.decode(), "chr6", 0, 5), ln


def test_grammar_fuzz_against_python_semantics(hostlib):
    """Random files built from regular and hostile fragments: whenever the device integers are trusted the result
    equals the reference arithmetic; whenever Python raises, the grammar has reported the file."""
    rng = np.random.default_rng(20251104)
    chroms = ["chr6", "chr60", "chr", "6", "chr6 "]
    exact = replay = 0
    for _ in range(1500):
        lines = []
        for _ in range(int(rng.integers(0, 12))):
            u = rng.random()
            if u < 0.06:
                lines.append(IRREGULAR_LINES[int(rng.integers(len(IRREGULAR_LINES)))])
                continue
            s = int(rng.integers(0, 50)) * 100
            e = s + int(rng.choice([100, 250, 1000, 0, -100]))
            dp = f"{int(rng.integers(0, 6000)) / 100:.{int(rng.integers(0, 3))}f}"
            tail = str(rng.choice(["", "", "", " ", "\t", "\tq", " \x0c"]))
            lines.append(f"{chroms[int(rng.integers(len(chroms)))]}\t{s}\t{e}\t{dp}{tail}".encode())
        text = b"".join(ln + b"\n" for ln in lines)
        lo = int(rng.integers(0, 3000))
        hi = lo + int(rng.integers(-10, 3000))
        try:
            want = so.compute_region_coverage_text(text.decode(), "chr6", lo, hi)
        except Exception:
            want = "raises"
        weighted, covered, nlines, irregular = grammar(hostlib, text, "chr6", lo, hi)
        if want == "raises":
            assert irregular > 0, text
            continue
        kind, v = decided(hostlib, text, "chr6", lo, hi)
        if kind == "exact":
            assert v == want, (text, lo, hi)
            exact += 1
        else:
            replay += 1
    assert exact > 1000 and replay < 400


def test_exact_half_ties_and_near_ties_are_deferred():
    assert gmd.decide_coverage(4001000, 2000, 2) is None            # exactly .5: the reference's fp64 noise decides
    assert gmd.decide_coverage(4001001, 2000, 2) == 2001
    assert gmd.decide_coverage(4000999, 2000, 2) == 2000
    assert gmd.decide_coverage(0, 0, 0) == 0
    assert gmd.decide_coverage(0, 500, 1) == 0
    # a chromosome-wide region: 3e6 lines, the tie window widens but is still decided by integers
    lines, covered = 3_000_000, 3_000_000_000
    weighted = 3000 * covered + covered // 2 + 1                      # one unit above the tie: inside the noise band
    assert gmd.decide_coverage(weighted, covered, lines) is None
    assert gmd.decide_coverage(3000 * covered + covered // 2 + 30_000_000, covered, lines) == 3001
    assert gmd.decide_coverage(weighted, 1 << 40, lines) is None      # beyond the documented range


def test_decision_never_contradicts_the_float_evaluation():
    """Random small tables: when decide_coverage answers, Python's float evaluation gives the same integer."""
    rng = np.random.default_rng(7)
    answered = 0
    for _ in range(4000):
        n = int(rng.integers(1, 40))
        h = rng.integers(0, 9000, n)
        ov = rng.integers(1, 1001, n)
        region_cov = 0.0
        for a, b in zip(h.tolist(), ov.tolist()):
            region_cov += float(f"{a // 100}.{a % 100:02d}") * b
        want = int(round(100 * (region_cov / int(ov.sum()))))
        got = gmd.decide_coverage(int((h * ov).sum()), int(ov.sum()), n)
        if got is not None:
            assert got == want
            answered += 1
    assert answered > 3900


# ---- read counting ---------------------------------------------------------------------------------------------------
def host_counts(lib, cols, off, regions, flags, min_mapq, use_directory=True):
    bounds, lo, hi = gcr.region_table(regions)
    bits = gcr.flag_bitmap(flags)
    directory = gcr.bucket_directory(bounds) if use_directory else None
    first, shift, per_tid, n_tid = directory if directory else (None, 0, 0, 0)
    out = np.zeros((len(off) - 1, len(regions)), dtype=np.int64)
    for s in range(len(off) - 1):
        sl = [np.ascontiguousarray(c[off[s]:off[s + 1]]) for c in cols]
        lib.rl_count_reads(*(c.ctypes.data for c in sl), len(sl[0]), bits.ctypes.data, int(min_mapq), bounds.ctypes.data,
                           len(bounds), lo.ctypes.data, hi.ctypes.data, len(regions),
                           first.ctypes.data if first is not None else None, shift, per_tid, n_tid, out[s].ctypes.data)
    return out


@pytest.mark.parametrize("case", sorted(STEP3_READS))
def test_read_predicate_and_region_table_match_golden(hostlib, case):
    kw = STEP3_READS[case]
    cols, off = synth.read_columns(kw["seed"], kw["ns"], kw["mean_reads"])
    assert host_counts(hostlib, cols, off, synth.READ_REGIONS, kw["flags"], kw["min_mapq"]).tolist() == GOLD["reads"][case]


def test_region_table_handles_overlap_nesting_and_empty_regions(hostlib):
    rng = np.random.default_rng(11)
    cols, off = synth.read_columns(77, 4, 2500)
    lo0 = synth.LPA[1] - 6000
    regions = []
    for _ in range(60):
        a = lo0 + int(rng.integers(0, 55_000))
        regions.append((int(rng.choice([5, 6, 6, 6, 7])), a, a + int(rng.integers(-50, 20_000))))
    regions += [(6, -100, synth.LPA[1]), (6, 0, (1 << 31) + 5), (-1, 0, 10), (6, synth.LPA[2], synth.LPA[2])]
    want = so.count_reads_regions(cols, off, regions, [83, 99, 147, 163], 1)
    assert host_counts(hostlib, cols, off, regions, [83, 99, 147, 163], 1).tolist() == want.tolist()
    assert want.sum() > 0


def test_bucket_directory_equals_the_binary_search(hostlib, monkeypatch):
    rng = np.random.default_rng(5)
    for nb, max_tid, max_pos, cap in [(1, 0, 1000, 1 << 17), (2, 3, 5000, 1 << 17), (1468, 24, 250_000_000, 1 << 17),
                                      (4000, 2, 3_000_000, 64), (300, 40, (1 << 31) - 1, 1 << 12)]:
        monkeypatch.setattr(gcr, "MAX_BUCKETS", cap)
        keys = (rng.integers(0, max_tid + 1, nb).astype(np.int64) << 32) | rng.integers(0, max_pos + 1, nb).astype(np.int64)
        bounds = np.unique(keys)
        first, shift, per_tid, n_tid = gcr.bucket_directory(bounds)
        assert len(first) == n_tid * per_tid + 1 <= max(cap, n_tid * 2) + 1 and first[-1] <= len(bounds)
        tid = np.concatenate([rng.integers(0, max_tid + 3, 6000), bounds >> 32, bounds >> 32, bounds >> 32]).astype(np.int32)
        bp = (bounds & 0xFFFFFFFF)
        pos = np.concatenate([rng.integers(0, min(max_pos * 2 + 10, (1 << 31) - 1), 6000), bp, np.maximum(bp - 1, 0),
                              np.minimum(bp + 1, (1 << 31) - 1)]).astype(np.int32)
        assert hostlib.rl_bucket_check(tid.ctypes.data, pos.ctypes.data, len(tid), bounds.ctypes.data, len(bounds),
                                       first.ctypes.data, shift, per_tid, n_tid) == 1, (nb, max_tid)


def test_counts_with_and_without_the_directory_agree(hostlib):
    kw = STEP3_READS["mapq1"]
    cols, off = synth.read_columns(kw["seed"], kw["ns"], kw["mean_reads"])
    a = host_counts(hostlib, cols, off, synth.READ_REGIONS, kw["flags"], 1, use_directory=False)
    assert a.tolist() == GOLD["reads"]["mapq1"]
    many = list(range(60, 180))
    c = host_counts(hostlib, cols, off, synth.READ_REGIONS, many, 1)
    assert c.tolist() == so.count_reads_regions(cols, off, synth.READ_REGIONS, many, 1).tolist() and c.sum() > 0


def test_flag_bitmap_is_an_exact_value_test():
    bits = gcr.flag_bitmap([83, 147, None, "99", 70000, -1, True])
    on = [f for f in range(65536) if (bits[f >> 5] >> (f & 31)) & 1]
    assert on == [83, 147]
