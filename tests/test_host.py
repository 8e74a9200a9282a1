"""Host-side logic (formats, line filters, drop-in installation) -- CPU only."""
import gzip
import sys
import types

import numpy as np
import pytest

from grid_b200 import io as gio
from grid_b200 import synth
from oracle import grid_oracle as O


def test_synth_is_deterministic_and_sane():
    a = synth.depth_hundredths(5, 40, 300)
    assert (a == synth.depth_hundredths(5, 40, 300)).all()
    assert (a[10:20, 100:200] == synth.depth_hundredths(5, 10, 100, row0=10, col0=100)).all()  # counter-based
    assert 0.0003 < (a == 0).mean() < 0.003
    m = a[a > 0] / 100
    assert 15 < m.mean() < 60
    assert synth.read_counts(5, 10, [42599, 1000]).shape == (10, 2)


def test_parse_regions_matches_oracle_filters(tmp_path):
    lines = ["chr6\t1000\t2000\t25.50", "chr6\t2000\t3000\t0.00", "chr60\t1000\t2000\t7.25", "chr5\t1000\t2000\t9.00",
             "chr6\t9000\t10000\t31.10", "chr6\t3000\t4000\t12.34", "short\tline", "chr6\t3000\t4000\t13.37", ""]
    p = tmp_path / "S.regions.bed.gz"
    with gzip.open(p, "wt") as f:
        f.write("\n".join(lines))
    for chrom, lo, hi, mask in (("chr6", 1000, 4000, {}), ("6", 1000, 4000, {"chr6": {3}}), (None, None, None, {}),
                                ("chr6", None, None, {"chr60": {1, 2}})):
        want = O.scan_regions(p, chrom, lo, hi, mask)
        s, e, h = gio.parse_regions_text(gzip.open(p, "rb").read(), chrom, lo, hi, mask)
        assert list(zip(s.tolist(), e.tolist(), (h / 100).tolist())) == want


def test_step4_file_roundtrip_and_codes(tmp_path):
    zq = np.array([[0, -1, 250, gio.ZQ_NA, gio.ZQ_NEG_ZERO], [-200, 199, 1, -100, 32766]], dtype=np.int16)
    p = tmp_path / "n.tsv.gz"
    gio.write_normalized_output(p, ["A", "B"], zq, np.array([1.0, 2.0, 0.9994, np.nan, 1.5]),
                                np.array([0.5, 1000.5, np.nan, 1.0, 2.0]), np.array([30.004, 28.005]))
    txt = gzip.open(p, "rt").read().splitlines()
    assert txt[0] == "2\t5\t1.000\t2.000\t0.999\tNA\t1.500"
    assert txt[1] == "2\t5\t0.500\t1000.500\tNA\t1.000\t2.000"
    assert txt[2] == "A\t30.00\t0.00\t-0.01\t2.50\tNA\t-0.00"
    assert txt[3] == "B\t28.00\t-2.00\t1.99\t0.01\t-1.00\t327.66"
    ids, ratios, back, scales = gio.read_normalized_data(p)
    assert ids == ["A", "B"] and scales.tolist() == [30.0, 28.0]
    want = zq.copy()
    want[0, 4] = 0
    assert (back == want).all()
    # the oracle's reader sees the same numbers
    oid, oratio, oz, _ = O.read_normalized(p)
    np.testing.assert_array_equal(np.isnan(oratio), np.isnan(ratios))
    np.testing.assert_array_equal(np.where(np.isnan(oz), gio.ZQ_NA, np.rint(oz * 100)), want)


def test_neighbor_file_matches_oracle_writer(tmp_path):
    ids = ["S1", "S2", "S3"]
    scales = np.array([30.0, 28.004, 31.999])
    idx = np.array([[1, 2], [0, 2], [0, -1]], dtype=np.int32)
    d2 = np.array([[1234500, 99999999], [1234500, 5], [7, 0]], dtype=np.int64)
    gio.save_neighbors(tmp_path / "a.gz", ids, scales, idx, d2, 38)
    O.save_neighbors(tmp_path / "b.gz", ids, dict(zip(ids, scales)), [[1, 2], [0, 2], [0]],
                     [[123.45, 9999.9999], [123.45, 0.0005], [0.0007]], 38)
    assert gzip.open(tmp_path / "a.gz", "rt").read() == gzip.open(tmp_path / "b.gz", "rt").read()
    nid, nsc, lists = gio.load_neighbors(tmp_path / "a.gz")
    assert nid == ids and lists[2] == [("S1", 30.0)]


def test_one_directory_listing_finds_what_the_per_sample_glob_finds(tmp_path):
    names = ["S1_LPA.regions.bed.gz", "S10_LPA.regions.bed.gz", "xS2y.regions.bed.gz", "S3.regions.bed.gz.csi", "S4regions.bed.gz",
             "S5.regions.bed.gz", "S5_b.regions.bed.gz", ".hidden_S6.regions.bed.gz", "S7.mosdepth.global.dist.txt", "regions.bed.gz",
             "S[8].regions.bed.gz", "S9.other.gz"]
    for n in names:
        (tmp_path / n).write_bytes(b"")
    ids = ["S1", "S10", "S2", "S3", "S4", "S5", "S6", "S7", "S[8]", "S?", "S9", "regions", "nobody", "S"]
    want = [gio.find_bed_gz_for_individual(i, tmp_path) for i in ids]          # the reference's lookup, one glob per id
    assert gio.find_bed_gz_for_individuals(ids, tmp_path) == want
    assert want[0].name in ("S1_LPA.regions.bed.gz", "S10_LPA.regions.bed.gz")   # (the substring alias the reference has)
    assert want[ids.index("nobody")] == tmp_path / "nobody.regions.bed.gz"


def test_one_directory_listing_random_names(tmp_path):
    """Seeded random cohorts of overlapping ids (prefixes of each other, with and without region suffixes)."""
    rng = np.random.default_rng(3)
    alphabet = ["S", "1", "2", "0", "_", "a", "."]
    ids = sorted({"".join(rng.choice(alphabet, size=int(rng.integers(1, 6)))) for _ in range(60)} - {".", ".."})
    for sid in ids[::2]:
        suffix = str(rng.choice(["", "_LPA", ".x", "regions"]))
        (tmp_path / f"{sid}{suffix}.regions.bed.gz").write_bytes(b"")
    want = [gio.find_bed_gz_for_individual(i, tmp_path) for i in ids]
    assert gio.find_bed_gz_for_individuals(ids, tmp_path) == want
    assert sum(p.exists() for p in want) >= len(ids) // 2


def test_install_rebinds_reference_steps(monkeypatch):
    pkg = types.ModuleType("grid")
    utils = types.ModuleType("grid.utils")
    mods = {"grid": pkg, "grid.utils": utils}
    for name, fn in (("normalize_mosdepth", "normalize_mosdepth"), ("find_neighbors", "find_neighbors"),
                     ("compute_dipcn", "compute_diploid_genotypes"), ("hi_inference", "hi_inference")):
        m = types.ModuleType(f"grid.utils.{name}")
        setattr(m, fn, lambda config, console: "reference")
        setattr(utils, name, m)
        mods[f"grid.utils.{name}"] = m
    md = types.ModuleType("grid.utils.mosdepth")           # helper-level target; grid.utils.count_reads is absent here
    md.compute_region_coverage = lambda *a: "reference"     # (it needs pysam), which install() must tolerate
    utils.mosdepth = md
    mods["grid.utils.mosdepth"] = md
    for k, v in mods.items():
        monkeypatch.setitem(sys.modules, k, v)
    monkeypatch.setitem(sys.modules, "grid.utils.count_reads", None)   # import raises ImportError
    import grid_b200

    grid_b200.install()
    from grid_b200.utils import mosdepth as our_md

    assert md.compute_region_coverage is our_md.compute_region_coverage
    from grid_b200.utils import compute_dipcn, find_neighbors, hi_inference, normalize_mosdepth

    assert mods["grid.utils.hi_inference"].hi_inference is hi_inference.hi_inference
    assert mods["grid.utils.normalize_mosdepth"].normalize_mosdepth is normalize_mosdepth.normalize_mosdepth
    assert mods["grid.utils.find_neighbors"].find_neighbors is find_neighbors.find_neighbors
    assert mods["grid.utils.compute_dipcn"].compute_diploid_genotypes is compute_dipcn.compute_diploid_genotypes


def test_config_errors_are_logged_not_raised(capsys):
    from grid_b200.utils.find_neighbors import find_neighbors
    from grid_b200.utils.normalize_mosdepth import normalize_mosdepth

    assert normalize_mosdepth({}, None) is None and "Config error" in capsys.readouterr().out
    assert find_neighbors({"mosdepth": {}}, None) is None and "Config error" in capsys.readouterr().out


def test_product_never_imports_the_oracle():
    import pathlib

    for f in pathlib.Path(gio.__file__).parent.rglob("*.py"):
        assert "oracle" not in f.read_text().replace("the oracle", ""), f


def test_gzip_members_with_size_field_interoperate(tmp_path):
    """Files written by grid_b200.textio are ordinary gzip (any reader inflates them) whose members can also be
    located without inflating; files from other writers fall back to streaming."""
    import gzip
    import random

    from grid_b200 import textio as T

    rnd = random.Random(3)
    chunks = [("".join(rnd.choice("0123456789.\t-") for _ in range(rnd.randrange(1, 50000))) + "\n").encode() for _ in range(9)]
    raw = b"".join(T._deflate_member([c[: len(c) // 2], c[len(c) // 2:]], level) for c, level in zip(chunks, [1, 3, 6, 9, None, 0, 2, 4, 5]))
    p = tmp_path / "m.gz"
    p.write_bytes(raw)
    assert gzip.open(p, "rb").read() == b"".join(chunks)
    spans = T._split_members(raw)
    assert len(spans) == 9 and spans[0][0] == 0 and spans[-1][1] == len(raw)
    with T._open_text(p, 3) as f:
        assert isinstance(f, T._MemberStream)
        assert f.readline() == chunks[0]
        rest = b""
        while True:
            blk = f.read(70000)
            if not blk:
                break
            rest += blk
        assert rest == b"".join(chunks[1:])
    # a corrupted member is caught by its CRC
    bad = bytearray(raw)
    bad[spans[4][0] + T._HDR_LEN + 5] ^= 0x55
    (tmp_path / "bad.gz").write_bytes(bytes(bad))
    with pytest.raises(Exception):
        with T._open_text(tmp_path / "bad.gz", 2) as f:
            while f.read(1 << 20):
                pass
    # somebody else's gzip: streaming reader
    q = tmp_path / "plain.gz"
    q.write_bytes(gzip.compress(b"a\tb\n"))
    with T._open_text(q, 2) as f:
        assert not isinstance(f, T._MemberStream) and f.readline() == b"a\tb\n"
    assert T._fixed3_line([1.0, float("nan"), 2.0005, -0.0004]) == "1.000\tNA\t2.001\t-0.000"


def test_handoff_entries_are_tied_to_the_file(tmp_path):
    import os
    import time

    from grid_b200 import handoff as H

    H.clear()
    p = tmp_path / "a.gz"
    p.write_bytes(b"one")
    H.put(p, x=1)
    assert H.take(p) == {"x": 1} and H.take(p) is None          # consumed
    H.put(p, x=2)
    time.sleep(0.01)
    p.write_bytes(b"two")                                        # same size, new mtime
    os.utime(p, ns=(time.time_ns(), time.time_ns()))
    assert H.take(p) is None
    H.put(p, x=3)
    p.unlink()
    assert H.take(p) is None
    q = tmp_path / "b.gz"
    q.write_bytes(b"q")
    (tmp_path / "link.gz").symlink_to(q)
    H.put(tmp_path / "link.gz", x=4)
    assert H.take(q) == {"x": 4}                                 # keyed by real path
    assert H.enabled({}) and H.enabled({"gpu": None}) and not H.enabled({"gpu": {"device_handoff": False}})


def test_multi_gpu_host_helpers():
    """Pure host logic of the multi-GPU paths: bin slices (contiguous, quantum-aligned, covering), the device list of the
    additive `gpu` config key, the sharding rule, and the Gram tile count bench.py's roofline uses."""
    torch = pytest.importorskip("torch")
    from grid_b200 import engine, multigpu

    for b, world, q in ((2_900_000, 8, 64), (1000, 3, 4), (63, 4, 64), (1, 2, 4)):
        sl = multigpu.bin_slices(b, world, q)
        assert len(sl) == world and sl[0][0] == 0 and sl[-1][1] == b
        assert all(a <= z for a, z in sl) and all(sl[r][1] == sl[r + 1][0] for r in range(world - 1))
        assert all(a % q == 0 for a, _ in sl if a < b)
    assert multigpu.configured_devices({}) is None
    if torch.cuda.device_count() >= 1:
        assert multigpu.configured_devices({"gpu": {"devices": [0]}}) is None   # one device = the single-GPU path
    else:
        with pytest.raises(ValueError):                                            # an index beyond the visible devices
            multigpu.configured_devices({"gpu": {"devices": [0]}})
    with pytest.raises(ValueError):
        multigpu.configured_devices({"gpu": {"devices": "some"}})
    assert multigpu.knn_sharding({}, 3202) == "bins" and multigpu.knn_sharding({}, 50_000) == "rows"
    assert multigpu.knn_sharding({"gpu": {"sharding": "rows"}}, 100) == "rows"
    with pytest.raises(ValueError):
        multigpu.knn_sharding({"gpu": {"sharding": "columns"}}, 100)
    # 3202 rows = 13 tile rows of 256 (the last 130 high): upper triangle incl. the diagonal, clipped to the matrix
    sizes = [256] * 12 + [130]
    want = sum(sizes[i] * sizes[j] for i in range(13) for j in range(i, 13))
    assert engine.gram_tile_elems(3202, 3202, True) == want == 5_528_068
    assert engine.gram_tile_elems(3202, 1000, False) == 3_202_000
    # the host format's byte count: codes + 12 bytes per escape + the two tables of a relative format
    pk = engine.PackedDepth(torch.zeros((4, 44), dtype=torch.int32), 4, 100, torch.tensor([7, 9]), torch.tensor([1.5, 700.0]),
                            11, torch.zeros(4, dtype=torch.int32), torch.zeros(100, dtype=torch.int32))
    assert pk.relative and pk.nbytes == 4 * 44 * 4 + 2 * 12 + (4 + 100) * 4 and pk.n_big == 1
    assert not engine.PackedDepth(torch.zeros((4, 104), dtype=torch.int16), 4, 100, torch.zeros(0, dtype=torch.int64),
                                  torch.zeros(0), 16).relative


def test_bench_cli_and_numa_binding_degrade_gracefully(monkeypatch):
    """bench.py's argument surface the docs name, and the NUMA binding helper on a box without NVML / GPUs: it reports, it
    does not raise, and it leaves the affinity alone."""
    import os

    import bench

    monkeypatch.setattr(sys, "argv", ["bench.py", "--gpus", "1", "--host-format", "abs", "--device-format", "f32", "--timeline",
                                      "--ring-bins", "2900000", "--ring-regen"])
    a = bench.parse_args()
    assert a.host_format == "abs" and a.device_format == "f32" and a.timeline and a.ring_regen and a.ring_bins == 2_900_000
    before = os.sched_getaffinity(0)
    out = bench.bind_to_gpu_numa_node(0)
    assert out is None or "error" in out or out["cpus"] >= 1
    if out is None or "error" in out:
        assert os.sched_getaffinity(0) == before
    os.sched_setaffinity(0, before)
