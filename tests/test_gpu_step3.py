"""Step-3 producers on the GPU: region coverage and read counting through the C ABI against the
reference-generated goldens, the oracle, and size-independent properties."""
import gzip
import json
import sys
from pathlib import Path

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO / "oracle"))

import step3_oracle as so  # noqa: E402
from gen_golden_cases import STEP3_COVERAGE, STEP3_READS  # noqa: E402

from grid_b200 import synth  # noqa: E402

GOLD = json.loads((REPO / "tests" / "golden" / "step3.json").read_text())


@pytest.fixture(scope="module")
def mods():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200.utils import count_reads, mosdepth

    return mosdepth, count_reads


@pytest.mark.parametrize("case", sorted(STEP3_COVERAGE))
def test_region_coverage_equals_the_reference(case, tmp_path, mods):
    md, _ = mods
    paths = synth.write_region_files(tmp_path, **STEP3_COVERAGE[case])
    for name, chrom, lo, hi in synth.REGION_QUERIES:
        assert md.region_coverage_batch(paths, chrom, lo, hi, threads=3) == GOLD["coverage"][case][name], name
    # one file per call, tiny batches and one big batch all agree
    name, chrom, lo, hi = synth.REGION_QUERIES[0]
    assert [md.compute_region_coverage(p, chrom, lo, hi) for p in paths] == GOLD["coverage"][case][name]
    assert md.region_coverage_batch(paths, chrom, lo, hi, batch_bytes=1) == GOLD["coverage"][case][name]


def _bed(path, lines, newline_at_end=True):
    text = "\n".join(lines) + ("\n" if lines and newline_at_end else "")
    with gzip.open(path, "wt", newline="") as fh:
        fh.write(text)
    return path


def test_reference_known_answers(tmp_path, mods):
    """test/test_mosdepth.py:23-56."""
    md, _ = mods
    cases = [(["chr6\t1000\t2000\t30.0"], ("chr6", 1000, 2000), 3000),
             (["chr6\t0\t2000\t40.0"], ("chr6", 1000, 2000), 4000),
             (["chr6\t1000\t2000\t20.0", "chr6\t2000\t3000\t40.0"], ("chr6", 1000, 3000), 3000),
             (["chr1\t1000\t2000\t30.0"], ("chr6", 1000, 2000), 0),
             (["chr6\t5000\t6000\t30.0"], ("chr6", 1000, 2000), 0),
             ([], ("chr6", 1000, 2000), 0)]
    for k, (lines, q, want) in enumerate(cases):
        assert md.compute_region_coverage(_bed(tmp_path / f"k{k}.bed.gz", lines), *q) == want
    # all of them in one batch (an empty file in the middle of the buffer), same query
    paths = [_bed(tmp_path / f"b{k}.bed.gz", lines) for k, (lines, _, _) in enumerate(cases)]
    want = [so.compute_region_coverage(p, "chr6", 1000, 2000) for p in paths]
    assert md.region_coverage_batch(paths, "chr6", 1000, 2000) == want


def test_irregular_files_behave_like_the_reference(tmp_path, mods):
    """Lines outside the device grammar: the value Python's int()/float() rules give, or the reference's exception."""
    md, _ = mods
    good = "chr6\t1000\t2000\t20.00"
    files = {
        "exp": [good, "chr6\t2000\t3000\t4e1"],                # float() accepts it -> 3000
        "plus": [good, "chr6\t+2000\t3000\t40.00"],            # int() accepts it
        "blank_line": [good, "", good],                        # ValueError: not enough values to unpack
        "short": [good, "chr6\t2000\t3000"],
        "word": [good, "chr5\t2000\t3000\tabc"],               # other chromosome, still converted first -> ValueError
        "crlf": [good + "\r", "chr6\t2000\t3000\t40.00\r"],
        "no_final_newline": [good, "chr6\t2000\t3000\t40.00"],
        "three_decimals": [good, "chr6\t2000\t3000\t40.004"],
        "half_tie": ["chr6\t1000\t2000\t20.00", "chr6\t2000\t3000\t20.01"],  # exact .5: replayed in fp64
    }
    paths, want = [], []
    for name, lines in files.items():
        p = _bed(tmp_path / f"{name}.bed.gz", lines, newline_at_end=name != "no_final_newline")
        paths.append(p)
        try:
            want.append(so.compute_region_coverage(p, "chr6", 1000, 3000))
        except Exception as e:
            want.append(type(e))
    got = md.region_coverage_batch(paths, "chr6", 1000, 3000)
    got = [type(g) if isinstance(g, Exception) else g for g in got]
    assert got == want
    assert ValueError in want and 3000 in want
    with pytest.raises(ValueError):
        md.compute_region_coverage(paths[list(files).index("blank_line")], "chr6", 1000, 3000)
    missing = md.region_coverage_batch([tmp_path / "nope.bed.gz"], "chr6", 0, 10)[0]
    assert isinstance(missing, FileNotFoundError)


def test_collect_region_coverage_writes_the_reference_format(tmp_path, mods):
    md, _ = mods
    kw = STEP3_COVERAGE["by1000"]
    synth.write_region_files(tmp_path / "work", **kw)
    ids = synth.sample_ids(kw["n"])
    (tmp_path / "samples.txt").write_text("\n".join(ids + ["S9999999"]) + "\n")
    cfg = {"samples_file": str(tmp_path / "samples.txt"), "chrom": "chr6", "start_bp": synth.LPA[1], "end_bp": synth.LPA[2],
           "output_dir": str(tmp_path / "out"), "threads": 2,
           "mosdepth": {"output_file_prefix": "cov", "work_dir": str(tmp_path / "work")}}
    md.collect_region_coverage(cfg, None)
    lines = (tmp_path / "out" / "cov.tsv").read_text().splitlines()
    assert lines[0] == f"Sample\tchr6:{synth.LPA[1]}-{synth.LPA[2]}"
    assert lines[1:] == [f"{s}\t{v}" for s, v in zip(ids, GOLD["coverage"]["by1000"]["lpa"])]


def test_region_coverage_large_batch_against_exact_integers(mods):
    """Full-size property (size-independent): H and covered are linear over disjoint regions; checked on a 64 MB batch."""
    md, _ = mods
    from grid_b200._lib import call, ptr

    rng = np.random.default_rng(5)
    nfiles, nl = 16, 120_000
    texts, per_file = [], []
    for f in range(nfiles):
        h = rng.integers(0, 9000, nl)
        rows = [f"chr6\t{1000 * j}\t{1000 * (j + 1)}\t{a // 100}.{a % 100:02d}\n" for j, a in enumerate(h.tolist())]
        texts.append("".join(rows).encode())
        per_file.append(h)
    off = np.concatenate([[0], np.cumsum([len(t) for t in texts])]).astype(np.int64)
    text = torch.frombuffer(bytearray(b"".join(texts)), dtype=torch.uint8).cuda()
    off_d = torch.from_numpy(off).cuda()

    def run(lo, hi):
        out = torch.empty((nfiles, 4), dtype=torch.int64, device="cuda")
        call("gridb200_region_coverage", ptr(text), int(off[-1]), ptr(off_d), nfiles, b"chr6", 4, lo, hi, ptr(out),
             torch.cuda.current_stream().cuda_stream)
        return out.cpu().numpy()

    a, b, c = 12_345_678, 60_000_500, 119_999_999
    whole, left, right = run(a, c), run(a, b), run(b, c)
    assert (whole[:, 3] == 0).all()
    assert (whole[:, :2] == left[:, :2] + right[:, :2]).all()
    assert (whole[:, 1] == c - a).all()
    for f in (0, nfiles - 1):
        j = np.arange(nl)
        ov = np.clip(np.minimum(c, 1000 * (j + 1)) - np.maximum(a, 1000 * j), 0, None)
        assert whole[f, 0] == int((per_file[f] * ov).sum()) and whole[f, 2] == int((ov > 0).sum())


# ---- read counting ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", sorted(STEP3_READS))
def test_read_counts_equal_the_reference(case, mods):
    _, cr = mods
    kw = STEP3_READS[case]
    cols, off = synth.read_columns(kw["seed"], kw["ns"], kw["mean_reads"])
    got = cr.count_reads_columns(cols, off, synth.READ_REGIONS, kw["flags"], kw["min_mapq"]).cpu().numpy()
    assert got.tolist() == GOLD["reads"][case]
    # the reference's one-region call shape, through an injected decoder
    s = 1
    sl = tuple(c[off[s]:off[s + 1]] for c in cols)
    one = cr.count_reads_in_region("x.cram", "ref.fa", "chr6", synth.LPA[1], synth.LPA[2], kw["flags"], kw["min_mapq"],
                                   reader=lambda *a: (sl, 6))
    assert one == GOLD["reads"][case][s][0]


def test_read_counts_many_regions_and_ragged_samples(mods):
    """734 random overlapping regions (more intervals than the shared-memory histogram holds in one variant), samples
    from empty to 300k reads, block boundaries that straddle samples."""
    _, cr = mods
    rng = np.random.default_rng(3)
    cols, off = synth.read_columns(99, 23, 60_000)
    lo0 = synth.LPA[1] - 6000
    for g in (734, 2500):
        regions = []
        for _ in range(g):
            a = lo0 + int(rng.integers(0, 55_000))
            regions.append((int(rng.choice([5, 6, 6, 6, 7])), a, a + int(rng.integers(-20, 9_000))))
        got = cr.count_reads_columns(cols, off, regions, [83, 99, 147, 163], 1).cpu().numpy()
        plain = cr.count_reads_columns(cols, off, regions, [83, 99, 147, 163], 1, use_directory=False).cpu().numpy()
        assert (got == plain).all()          # bucket directory vs plain binary search over the boundary table
        many = list(range(60, 180))
        assert (cr.count_reads_columns(cols, off, regions[:50], many, 1).cpu().numpy()
                == so.count_reads_regions(cols, off, regions[:50], many, 1)).all()
        sub = rng.choice(g, 40, replace=False)
        want = so.count_reads_regions(cols, off, [regions[k] for k in sub], [83, 99, 147, 163], 1)
        assert (got[:, sub] == want).all()
        assert got.sum() > 0


def test_read_counts_full_size_properties(mods):
    """2e7 reads on the device: counts over a partition of a region add up to the region's count; a batch equals its
    samples counted one by one; nothing passes an empty flag set."""
    _, cr = mods
    n, ns = 20_000_000, 7
    rng = np.random.default_rng(1)
    flag = synth.READ_FLAG_PALETTE[rng.integers(0, 16, n)]
    mapq = rng.integers(0, 61, n).astype(np.uint8)
    tid = rng.integers(5, 8, n).astype(np.int32)
    ntid = np.where(rng.random(n) < 0.9, tid, tid + 1).astype(np.int32)
    pos = rng.integers(0, 1_000_000, n).astype(np.int32)
    cols = tuple(torch.from_numpy(c).cuda() for c in (flag, mapq, tid, ntid, pos))
    off = np.array([0, 1, 2049, 2049, 5_000_000, 5_000_007, 19_999_999, n], dtype=np.int64)
    cuts = [0, 1, 1000, 250_000, 250_001, 999_999, 1_000_000]
    parts = [(6, a, b) for a, b in zip(cuts[:-1], cuts[1:])]
    regions = [(6, 0, 1_000_000)] + parts
    flags = [83, 99, 147, 163]
    got = cr.count_reads_columns(cols, off, regions, flags, 10).cpu().numpy()
    assert (got[:, 0] == got[:, 1:].sum(axis=1)).all()
    ok = np.isin(flag, flags) & (mapq >= 10) & (tid == ntid) & (tid == 6)
    want = [int(ok[off[s]:off[s + 1]].sum()) for s in range(ns)]
    assert got[:, 0].tolist() == want and sum(want) > 1_000_000
    s = 3
    one = cr.count_reads_columns(tuple(c[off[s]:off[s + 1]].clone() for c in cols), [0, int(off[s + 1] - off[s])],
                                 regions, flags, 10).cpu().numpy()
    assert (one[0] == got[s]).all()
    assert int(cr.count_reads_columns(cols, off, regions, [], 0).sum()) == 0


def test_count_reads_step_writes_the_reference_format(tmp_path, mods):
    _, cr = mods
    kw = STEP3_READS["mapq1"]
    cols, off = synth.read_columns(kw["seed"], kw["ns"], kw["mean_reads"])
    ids = synth.sample_ids(kw["ns"])
    (tmp_path / "crams").mkdir()
    for s in ids:
        (tmp_path / "crams" / f"{s}.cram").write_bytes(b"")
    (tmp_path / "samples.txt").write_text("\n".join(ids + ["S7777777"]) + "\n")
    cfg = {"directory_loc": str(tmp_path / "crams"), "samples_file": str(tmp_path / "samples.txt"), "file_type": "cram",
           "chrom": "chr6", "start_bp": synth.LPA[1], "end_bp": synth.LPA[2], "min_mapq": 1, "output_dir": str(tmp_path / "out"),
           "count_reads": {"flags": kw["flags"], "output_file_prefix": "read_counts"}, "reference_genome": "ref.fa"}

    def reader(path, ref, chrom, start, end):
        s = ids.index(Path(path).stem)
        if s == 4:
            raise OSError("truncated file")
        return tuple(c[off[s]:off[s + 1]] for c in cols), 6

    cr.count_reads(cfg, None, reader=reader)
    lines = (tmp_path / "out" / "read_counts.tsv").read_text().splitlines()
    assert lines[0] == f"Sample\tchr6:{synth.LPA[1]}-{synth.LPA[2]}"
    want = [f"{s}\t{'Error' if k == 4 else GOLD['reads']['mapq1'][k][0]}" for k, s in enumerate(ids)]
    assert lines[1:] == want
