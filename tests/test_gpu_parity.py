"""Parity of the CUDA path (through the C ABI) against the oracle and the reference's golden outputs.

Bars (north_star): neighbour index sets and integer counts bit-exact except inside exact distance ties;
normalised depths (the %.2f / %.3f text fields) string-equal; dipCN within 1e-5 relative.
"""
import gzip
from pathlib import Path

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from grid_b200 import synth  # noqa: E402
from oracle import grid_oracle as O  # noqa: E402

from helpers import assert_dipcn_files_equal, assert_neighbor_files_equal, assert_norm_files_equal  # noqa: E402
from test_oracle_golden import CASES  # noqa: E402

DIPCN_REL = 1e-5


@pytest.fixture(scope="module")
def eng():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200 import engine

    return engine


def _run_steps(cfg):
    from grid_b200.utils.compute_dipcn import compute_diploid_genotypes
    from grid_b200.utils.find_neighbors import find_neighbors
    from grid_b200.utils.normalize_mosdepth import normalize_mosdepth

    normalize_mosdepth(cfg, None)
    find_neighbors(cfg, None)
    compute_diploid_genotypes(cfg, None)
    od = Path(cfg["output_dir"])
    return od / "norm.tsv.gz", od / "nbrs.zMax2.0.tsv.gz", od / "dipcn.tsv"


@pytest.mark.parametrize("handoff", [True, False])
@pytest.mark.parametrize("case", sorted(CASES))
def test_file_chain_matches_reference_golden(case, handoff, tmp_path, golden, eng, monkeypatch):
    """BASELINE config 1 (N=100, LPA window) and the masked/truncated edge case, through the three step functions --
    once with the in-process device hand-off between the steps, once with every step parsing the previous file."""
    from grid_b200 import handoff as H, io as gio, textio

    cfg = CASES[case](tmp_path)
    cfg["gpu"] = {"device_handoff": handoff}
    H.clear()
    if handoff:  # prove the files are NOT parsed again: both readers are out of order
        def boom(*a, **k):
            raise AssertionError("a hand-off file was parsed although this process had just written it")
        monkeypatch.setattr(textio, "read_normalized_data", boom)
        monkeypatch.setattr("grid_b200.utils.compute_dipcn.load_neighbors", boom)
    norm, nbrs, dip = _run_steps(cfg)
    assert_norm_files_equal(norm, golden / case / "norm.tsv")
    assert_neighbor_files_equal(nbrs, golden / case / "nbrs.zMax2.0.tsv")
    assert_dipcn_files_equal(dip, golden / case / "dipcn.tsv", rel=DIPCN_REL)


def test_steps_interchange_with_the_oracle(tmp_path, eng):
    """Any step may be run by either implementation: oracle step 4 -> CUDA step 5 -> oracle step 6, and back."""
    cfg = synth.write_cohort_files(tmp_path, 99, 40, 120)
    from grid_b200.utils.compute_dipcn import compute_diploid_genotypes
    from grid_b200.utils.find_neighbors import find_neighbors

    O.normalize_mosdepth(cfg)
    find_neighbors(cfg, None)
    mine = gzip.open(Path(cfg["output_dir"]) / "nbrs.zMax2.0.tsv.gz", "rt").read()
    (tmp_path / "mine.tsv").write_text(mine)
    O.find_neighbors(cfg)
    assert_neighbor_files_equal(tmp_path / "mine.tsv", Path(cfg["output_dir"]) / "nbrs.zMax2.0.tsv.gz")
    want = O.compute_diploid_genotypes(cfg).read_text()
    compute_diploid_genotypes(cfg, None)
    (tmp_path / "want.tsv").write_text(want)
    assert_dipcn_files_equal(Path(cfg["output_dir"]) / "dipcn.tsv", tmp_path / "want.tsv", rel=DIPCN_REL)


def test_synth_device_equals_numpy(eng):
    for (n, b, r0, c0) in ((37, 1001, 0, 0), (8, 4096, 1000, 123456), (3, 5, 7, 9)):
        dev = eng.synth_depth(11, n, b, r0, c0).cpu().numpy()[:, :b]
        np.testing.assert_array_equal(dev, synth.depth_matrix(11, n, b, r0, c0))


def _array_case(seed, n, b, holes=0.0, **kw):
    """(fp32 matrix for the device, the same depths as the decimals the reference would parse, counts)."""
    h = synth.depth_hundredths(seed, n, b)
    if holes:
        rng = np.random.default_rng(seed)
        h[rng.random(h.shape) < holes] = 0
    counts = synth.read_counts(seed, n, [42599, 1500, 900000])
    return (h / 100.0).astype(np.float32), h / 100.0, counts


def _random_cases(count):
    """Seeded sweep over shapes and knobs (incl. zmax 6.0 -> |v| up to 600 > 511: the integer SIMT Gram takes over)."""
    rng = np.random.default_rng(2025)
    out = []
    for t in range(count):
        out.append((100 + t, int(rng.integers(20, 300)), int(rng.integers(64, 6000)), float(rng.choice([0.0, 0.02, 0.3])),
                    dict(zmax=float(rng.choice([1.0, 2.0, 3.0, 5.0, 6.0])), top_frac=float(rng.choice([0.05, 0.1, 0.3])),
                         num_neighbors=int(rng.choice([3, 50])), n_nbr=int(rng.choice([2, 30])),
                         min_depth=float(rng.choice([20, 28])), max_depth=float(rng.choice([100, 45])))))
    return out


@pytest.mark.parametrize("seed,n,b,holes,params", [
    (1, 64, 300, 0.01, dict(num_neighbors=20, n_nbr=8)),
    (2, 257, 2100, 0.0, dict(num_neighbors=500, n_nbr=300)),          # k > N-1
    (3, 130, 5000, 0.05, dict(min_depth=30, max_depth=40, top_frac=0.5, num_neighbors=40, n_nbr=10, zmax=1.5)),
    (4, 33, 1030, 0.0, dict(top_frac=0.0, frac_r=0.5, sigma2_max=0.9, num_neighbors=5, n_nbr=5)),
] + _random_cases(12))
def test_array_chain_matches_oracle(seed, n, b, holes, params, eng):
    depth, decimals, counts = _array_case(seed, n, b, holes)
    want = O.chain_arrays(decimals, counts, **params)
    p = eng.Params(**params)
    ld = (b + 3) // 4 * 4
    dd = np.zeros((n, ld), dtype=np.float32)
    dd[:, :b] = depth
    tdepth = torch.from_numpy(dd).cuda()
    res = eng.normalize(tdepth, b, p, want_zq=True, want_zh=True)
    # step 4 intermediates
    np.testing.assert_array_equal(np.nonzero(res.valid.cpu().numpy())[0], want["valid"])
    v = want["valid"]
    np.testing.assert_allclose(res.raw_mean.cpu().numpy(), want["raw_means"], rtol=1e-14)
    np.testing.assert_array_equal(res.scale_q.cpu().numpy(), want["scale_q"])
    np.testing.assert_allclose(res.mu.cpu().numpy()[v], want["mu"], rtol=1e-13)
    np.testing.assert_allclose(res.s2.cpu().numpy()[v], want["s2"], rtol=1e-11, atol=1e-18)
    np.testing.assert_allclose(res.ratio.cpu().numpy()[v], want["ratio"], rtol=1e-11, equal_nan=True)
    sel = np.nonzero(res.sel_pos.cpu().numpy() >= 0)[0]
    np.testing.assert_array_equal(sel, v[want["sel"]])
    assert res.rw == len(want["sel"]) and res.r_use == want["r_use"]
    np.testing.assert_array_equal(np.nonzero(res.keep.cpu().numpy())[0], v[want["sel"]][want["keep"]])
    zq = res.zq.cpu().numpy()[:, :res.rw].astype(np.int64)
    zq[zq == -32767] = 0
    np.testing.assert_array_equal(zq, want["zq"])                              # every %.2f field, exactly
    # tensor-core operand = clipped, NA->0, dropped bins zeroed; norms exact
    zh = eng.zh_rowmajor(res.zh).cpu().numpy()[:, :res.rw].astype(np.int64)
    np.testing.assert_array_equal(zh[:, want["keep"]], want["zc"])
    assert (np.delete(zh, want["keep"], axis=1) == 0).all()
    np.testing.assert_array_equal(res.norms.cpu().numpy(), (want["zc"] ** 2).sum(axis=1))
    # step 5: both Gram implementations
    for impl in (1, 0):
        idx, d2 = eng.knn(res.zh, res.kdim, res.norms, p.num_neighbors, p.zclip, impl=impl)
        np.testing.assert_array_equal(d2.cpu().numpy(), want["nbr_d2"])        # exact integers
        np.testing.assert_array_equal(idx.cpu().numpy(), want["nbr_idx"])      # (d2, index) order = oracle's
    # step 6
    cn = eng.dipcn(torch.from_numpy(counts).cuda(), torch.ones(n, dtype=torch.uint8).cuda(), res.scale_q, idx, p.n_nbr)
    np.testing.assert_allclose(cn.cpu().numpy(), want["dipcn"], rtol=DIPCN_REL)


def _kblocked(a: np.ndarray) -> torch.Tensor:
    """[n, k] integer matrix -> the k-blocked fp16 operand [k/64, n, 64] gridb200_gram reads."""
    n, k = a.shape
    return torch.from_numpy(np.ascontiguousarray(a.astype(np.float16).reshape(n, k // 64, 64).transpose(1, 0, 2))).cuda()


def test_gram_tensor_core_is_exact_at_extremes(eng):
    """All-(+-vmax) rows are the worst case for fp32 accumulation: the chunked drain must stay exact."""
    from grid_b200._lib import GridB200Error, call, ptr

    for vmax, n, k, impls in ((200, 300, 8192, (1, 0)), (511, 140, 4096, (1, 0)), (1, 129, 64, (1, 0)), (2048, 70, 512, (1,))):
        rng = np.random.default_rng(vmax)
        a = rng.choice([-vmax, vmax], size=(n, k)).astype(np.int64)
        a[0] = vmax
        a[1] = -vmax
        a[2, ::2] = 0
        t = _kblocked(a)
        want = a @ a.T
        for impl in impls:
            dot = torch.full((n, n + (n % 2)), -7, dtype=torch.int64, device="cuda")
            call("gridb200_gram", ptr(t), n, n, ptr(t), n, n, k, vmax, ptr(dot), dot.stride(0), 0, impl,
                 torch.cuda.current_stream().cuda_stream)
            np.testing.assert_array_equal(dot.cpu().numpy()[:, :n], want)
            call("gridb200_gram", ptr(t), n, n, ptr(t), n, n, k, vmax, ptr(dot), dot.stride(0), 1, impl,
                 torch.cuda.current_stream().cuda_stream)                    # accumulate
            np.testing.assert_array_equal(dot.cpu().numpy()[:, :n], 2 * want)
    with pytest.raises(GridB200Error, match="vmax"):                          # fp32 chunks cannot be exact there
        call("gridb200_gram", ptr(t), 70, 70, ptr(t), 70, 70, 512, 2048, ptr(dot), dot.stride(0), 0, 0,
             torch.cuda.current_stream().cuda_stream)


def test_gram_rectangular_blocks_and_long_k(eng):
    """Query block != reference block (the blocked / ring schedule), ragged edges, and a K long enough for many
    chunk drains, int32 overflow flushes and K splits."""
    from grid_b200._lib import call, ptr

    rng = np.random.default_rng(5)
    nq, nr, k = 200, 333, 64 * 700
    q = rng.integers(-200, 201, size=(nq, k)).astype(np.int64)
    r = rng.integers(-200, 201, size=(nr, k)).astype(np.int64)
    q[3] = 200
    r[5] = 200                                                               # 200*200*44800 = 1.8e9 > 2^30
    r[6] = -200
    tq, tr = _kblocked(q), _kblocked(r)
    dot = torch.empty((nq, nr + 1), dtype=torch.int64, device="cuda")
    call("gridb200_gram", ptr(tq), nq, nq, ptr(tr), nr, nr, k, 200, ptr(dot), dot.stride(0), 0, 0,
         torch.cuda.current_stream().cuda_stream)
    np.testing.assert_array_equal(dot.cpu().numpy()[:, :nr], q @ r.T)
    # a query block in the middle of the reference operand (what engine.knn passes for blocked queries)
    q0, nqb = 70, 150
    dot2 = torch.empty((nqb, nr + 1), dtype=torch.int64, device="cuda")
    call("gridb200_gram", tr.data_ptr() + q0 * 64 * 2, nqb, nr, ptr(tr), nr, nr, k, 200, ptr(dot2), dot2.stride(0), 0, 0,
         torch.cuda.current_stream().cuda_stream)
    np.testing.assert_array_equal(dot2.cpu().numpy()[:, :nr], r[q0:q0 + nqb] @ r.T)


def test_gram_symmetric_many_tiles(eng):
    """Same-operand call with several 256-row tiles: only the upper triangle is computed and then mirrored; the
    result must equal the general (two-operand) path bit for bit, with and without accumulation."""
    from grid_b200._lib import call, ptr

    rng = np.random.default_rng(11)
    n, k = 1100, 64 * 90
    a = rng.integers(-200, 201, size=(n, k)).astype(np.int64)
    t = _kblocked(a)
    want = a @ a.T
    dot = torch.full((n, n), 3, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    call("gridb200_gram", ptr(t), n, n, ptr(t), n, n, k, 200, ptr(dot), n, 0, 0, st)
    np.testing.assert_array_equal(dot.cpu().numpy(), want)
    call("gridb200_gram", ptr(t), n, n, ptr(t), n, n, k, 200, ptr(dot), n, 1, 0, st)
    np.testing.assert_array_equal(dot.cpu().numpy(), 2 * want)
    t2 = t.clone()                                                            # different pointer: general path
    call("gridb200_gram", ptr(t), n, n, ptr(t2), n, n, k, 200, ptr(dot), n, 0, 0, st)
    np.testing.assert_array_equal(dot.cpu().numpy(), want)


def test_knn_properties_at_scale(eng):
    """Size-independent checks on a cohort too large for the file oracle: symmetry of the exact distances,
    sortedness, self-exclusion, and agreement with brute-force int64 on sampled rows."""
    n, b = 1500, 40000
    depth = eng.synth_depth(21, n, b)
    p = eng.Params(num_neighbors=50, n_nbr=30)
    res = eng.normalize(depth, b, p, want_zq=False, want_zh=True)
    idx, d2 = eng.knn(res.zh, res.kdim, res.norms, 50, p.zclip)
    idx, d2 = idx.cpu().numpy(), d2.cpu().numpy()
    assert (np.diff(d2, axis=1) >= 0).all() and (idx != np.arange(n)[:, None]).all() and (idx >= 0).all()
    zh = eng.zh_rowmajor(res.zh).cpu().numpy().astype(np.int64)
    rows = np.random.default_rng(0).choice(n, 40, replace=False)
    full = ((zh[rows, None, :] - zh[None, :, :]) ** 2).sum(axis=2) if n * b * 40 < 2e8 else None
    for t, i in enumerate(rows):
        dd = ((zh[i][None, :] - zh) ** 2).sum(axis=1)
        dd[i] = np.iinfo(np.int64).max
        order = np.lexsort((np.arange(n), dd))[:50]
        np.testing.assert_array_equal(d2[i], dd[order])
        np.testing.assert_array_equal(idx[i], order)
    # d(i,j) == d(j,i) wherever both directions made the lists
    look = {(i, int(j)): int(d) for i in range(n) for j, d in zip(idx[i], d2[i])}
    assert all(look[(j, i)] == d for (i, j), d in look.items() if (j, i) in look)


def test_transform_saturation_and_negative_zero(eng):
    """An absurd outlier saturates the int16 code (counted, clamped) without disturbing any other cell; values in
    (-0.005, 0) keep their sign in the text code."""
    n, b = 48, 256
    h = synth.depth_hundredths(8, n, b)
    h[7, 100] = 4_000_000                       # 40000x coverage in one bin
    dec = h / 100.0
    want = O.chain_arrays(dec, synth.read_counts(8, n, [1000]), num_neighbors=5, n_nbr=5, max_depth=1e9)
    res = eng.normalize(torch.from_numpy(dec.astype(np.float32)).cuda(), b, eng.Params(max_depth=1e9), want_zq=True,
                        want_zh=True)
    zq = res.zq.cpu().numpy()[:, :res.rw].astype(np.int64)
    exp = want["zq"].copy()
    big = np.abs(exp) > 32766
    big &= exp != -32768
    assert big.sum() >= 1 and int(res.extra["sat"].item()) == int(big.sum())
    np.testing.assert_array_equal(zq[big], np.sign(exp[big]) * 32766)
    negzero = zq == -32767
    zq[negzero] = 0
    np.testing.assert_array_equal(zq[~big], exp[~big])
    z = O.normalize_matrix(np.where(dec[:, want["valid"]] > 0, dec[:, want["valid"]], np.nan))[0][:, want["sel"]]
    np.testing.assert_array_equal(negzero, (np.signbit(z) & (exp == 0)))


def test_errors_surface_as_exceptions(eng):
    from grid_b200._lib import GridB200Error

    depth = eng.synth_depth(1, 16, 64)
    with pytest.raises(IndexError):
        eng.normalize(depth, 64, eng.Params(top_frac=1.0))   # the reference's list index is out of range here too
    with pytest.raises(ValueError):
        eng.normalize(depth.double(), 64, eng.Params())
    with pytest.raises(GridB200Error):
        eng.knn(torch.zeros((1, 4, 64), dtype=torch.float16, device="cuda"), 64, torch.zeros(4, dtype=torch.int64, device="cuda"),
                2, 5000)


def _ingest(eng, text: bytes, chrom, lo, hi, mask: dict, bin0: int, b: int):
    """One gridb200_ingest_text call -> (row of hundredths as int64, stats)."""
    from grid_b200._lib import call, load, ptr

    want = (chrom if chrom.startswith("chr") else "chr" + chrom).encode() if chrom else b""
    t = torch.frombuffer(bytearray(text), dtype=torch.uint8).cuda() if text else torch.zeros(1, dtype=torch.uint8).cuda()
    row = torch.full((b,), -1.0, dtype=torch.float32, device="cuda")
    stats = torch.zeros(8, dtype=torch.int64, device="cuda")
    ws = torch.empty(load().gridb200_ingest_ws_bytes(b), dtype=torch.uint8, device="cuda")
    bits, nkb = None, 0
    kbs = mask.get(want.decode(), set()) if mask else set()
    if kbs:
        nkb = max(kbs) + 1
        arr = np.zeros((nkb + 31) // 32, dtype=np.uint32)
        for kb in kbs:
            arr[kb >> 5] |= np.uint32(1 << (kb & 31))
        bits = torch.from_numpy(arr).cuda()
    call("gridb200_ingest_text", ptr(t), len(text), want, len(want), lo if lo is not None else -1,
         hi if hi is not None else -1, ptr(bits), nkb, bin0, b, ptr(row), ptr(stats), ptr(ws), ws.numel(),
         torch.cuda.current_stream().cuda_stream)
    return np.rint(row.cpu().numpy().astype(np.float64) * 100).astype(np.int64), stats.cpu().numpy()


def test_device_ingest_matches_the_reference_line_filters(eng):
    """The device text parser against the host restatement of normalize_mosdepth.py:262-285 on hand-written lines
    (the reference's own known answers: test_normalize_extra.py:28-115) and on a generated file."""
    from grid_b200 import io as gio

    lines = ["chr6\t1000\t2000\t25.50", "chr6\t2000\t3000\t0.00", "chr60\t4000\t5000\t7.25", "chr5\t1000\t2000\t9.00",
             "chr6\t9000\t10000\t31.10", "chr6\t3000\t4000\t12.34", "short\tline", "chr6\t3000\t4000\t13.37", "",
             "chr6\t5000\t6000\t8", "chr6\t6000\t7000\t8.5\textra\tcols", "chr6\t7000\t7500\t41943.03\r",
             "chr6\t8000\t9000\t", "chr6\t10000\t11000"]
    text = "\n".join(lines).encode()                                           # no trailing newline
    b = 12
    for chrom, lo, hi, mask in (("chr6", 1000, 4000, {}), ("6", 1000, 4000, {"chr6": {3}}), (None, None, None, {}),
                                ("chr6", None, None, {"chr6": {1, 2, 9}}), ("chr6", 6500, 6500, {})):
        s, e, h = gio.parse_regions_text(text, chrom, lo, hi, mask)
        want = np.zeros(b, dtype=np.int64)
        want[s // 1000] = h                                                    # later duplicates overwrite
        got, st = _ingest(eng, text, chrom, lo, hi, mask, 0, b)
        np.testing.assert_array_equal(got, want)
        assert st[1] == 0 and st[0] == len(s)                                  # nothing irregular; kept lines agree
    # with chr6 as prefix the chr60 line is admitted (startswith) and reported as foreign
    assert _ingest(eng, text, "chr6", None, None, {}, 0, b)[1][4] == 1
    # lines outside the strict grammar are counted, never guessed
    for bad in ("chr6\t1000\t2000\t1e1", "chr6\t1000\t2000\t2.345", "chr6\t1000\t2500x\t3.00", "chr6\t1500\t2500\t3.00",
                "chr6\t1000\t3000\t3.00", "chr6\t99000\t100000\t3.00", "chr6\t1000\t2000\t-3.00", "chr6\t1000\t2000\t41943.04"):
        got, st = _ingest(eng, (bad + "\nchr6\t0\t1000\t1.00\n").encode(), "chr6", None, None, {}, 0, b)
        assert st[1] == 1 and got[0] == 100 and got[1:].sum() == 0, bad
    got, st = _ingest(eng, b"", "chr6", None, None, {}, 0, b)
    assert got.sum() == 0 and st.sum() == 0
    # a generated sample file, whole-chromosome, vs the host parser
    n_bins = 5000
    hund = synth.depth_hundredths(77, 1, n_bins)[0]
    body = "".join(f"chr1\t{1000 * j}\t{1000 * (j + 1)}\t{hund[j] // 100}.{hund[j] % 100:02d}\n" for j in range(n_bins))
    got, st = _ingest(eng, body.encode(), "chr1", None, None, {}, 0, n_bins)
    np.testing.assert_array_equal(got, hund)
    assert st[0] == (hund > 0).sum() and st[2] == n_bins


def test_device_and_host_packers_agree(tmp_path, eng):
    """pack_depth_matrix_device (GPU parse) and pack_depth_matrix (host parse) build the same matrix."""
    from grid_b200 import io as gio
    from grid_b200.utils import normalize_mosdepth as nm

    cfg = synth.write_cohort_files(tmp_path, 5, 12, 300)
    samples = [ln.strip() for ln in open(cfg["samples_file"]) if ln.strip()]
    wd = cfg["mosdepth"]["work_dir"]
    ind = gio.map_mosdepth_files_to_samples(wd, samples)
    for chrom, lo, hi in ((cfg.get("chrom"), cfg.get("start_bp"), cfg.get("end_bp")), (cfg.get("chrom"), None, None)):
        ids_h, bins_h, dep_h = nm.pack_depth_matrix(ind, wd, chrom, lo, hi, {}, 2, None)
        ids_d, bins_d, dep_d = nm.pack_depth_matrix_device(ind, wd, chrom, lo, hi, {}, 2, None)
        assert ids_h == ids_d
        dd = dep_d.cpu().numpy()
        col = {(int(s), int(e)): j for j, (s, e) in enumerate(bins_d)}
        seen = np.zeros(dd.shape[1], dtype=bool)
        for j, (s, e) in enumerate(bins_h):
            np.testing.assert_array_equal(dd[:, col[(int(s), int(e))]], dep_h[:, j])
            seen[col[(int(s), int(e))]] = True
        assert (dd[:, ~seen] == 0).all()                                       # grid bins nobody has stay absent


@pytest.mark.parametrize("bad_frac", [0.0, 0.02, 0.125, 0.2, 1.0])
def test_fused_row_stats_equal_the_full_pass(bad_frac, eng):
    """Per-sample sums over the valid bins derived from the all-bin totals (few failing bins: column fix-up; many:
    the full pass, chosen on the device) must equal gridb200_row_stats exactly, whatever the share of failing bins."""
    from grid_b200._lib import call, load, ptr

    n, b = 70, 4099
    h = synth.depth_hundredths(3, n, b)
    h[5] = 0                                                                   # a sample with no data at all
    h[:, 100:140] = 0                                                          # bins nobody has (not "bad": nothing to subtract)
    ld = (b + 3) // 4 * 4
    dd = np.zeros((n, ld), dtype=np.float32)
    dd[:, :b] = h / 100.0
    depth = torch.from_numpy(dd).cuda()
    rng = np.random.default_rng(int(bad_frac * 1000))
    valid_np = (rng.random(b) >= bad_frac).astype(np.uint8)
    valid = torch.from_numpy(valid_np).cuda()
    st = torch.cuda.current_stream().cuda_stream
    i64 = dict(dtype=torch.int64, device="cuda")
    i32 = dict(dtype=torch.int32, device="cuda")
    colsum, colcnt = torch.empty(b, **i64), torch.empty(b, **i32)
    rs_all, rc_all = torch.empty(n, **i64), torch.empty(n, **i32)
    call("gridb200_col_row_raw_stats", ptr(depth), n, b, ld, ptr(colsum), ptr(colcnt), ptr(rs_all), ptr(rc_all), st)
    np.testing.assert_array_equal(colsum.cpu().numpy(), h.sum(axis=0))
    np.testing.assert_array_equal(colcnt.cpu().numpy(), (h > 0).sum(axis=0))
    np.testing.assert_array_equal(rs_all.cpu().numpy(), h.sum(axis=1))
    np.testing.assert_array_equal(rc_all.cpu().numpy(), (h > 0).sum(axis=1))
    want_s, want_c = torch.empty(n, **i64), torch.empty(n, **i32)
    call("gridb200_row_stats", ptr(depth), n, b, ld, ptr(valid), ptr(want_s), ptr(want_c), st)
    got_s, got_c = torch.full((n,), -1, **i64), torch.full((n,), -1, **i32)
    ws = torch.empty(load().gridb200_row_stats_fused_ws_bytes(b), dtype=torch.uint8, device="cuda")
    call("gridb200_row_stats_fused", ptr(depth), n, b, ld, ptr(valid), ptr(colcnt), ptr(rs_all), ptr(rc_all), ptr(got_s),
         ptr(got_c), ptr(ws), ws.numel(), st)
    np.testing.assert_array_equal(got_s.cpu().numpy(), want_s.cpu().numpy())
    np.testing.assert_array_equal(got_c.cpu().numpy(), want_c.cpu().numpy())
    np.testing.assert_array_equal(want_s.cpu().numpy(), (h * valid_np[None, :]).sum(axis=1))
    mode = int(ws.view(torch.int32)[1].item())
    n_bad = int(((valid_np == 0) & ((h > 0).sum(axis=0) > 0)).sum())
    assert mode == (1 if n_bad * 8 <= b else 0)


@pytest.mark.parametrize("nn", [5, 200])
def test_dipcn_exon_rule_matches_the_reference_golden(nn, golden, eng):  # compute_dipcn_dir/compute_diploid_cn.py:8-63
    g = np.load(golden / "exon.npz")
    raw = g["raw"]
    cols = np.stack([raw[:, 0], raw[:, 1] + raw[:, 2], raw[:, 0] + raw[:, 1] + raw[:, 2], raw[:, 3]], axis=1)
    got = eng.dipcn(torch.from_numpy(cols.astype(np.int32)).cuda(), torch.from_numpy(g["has"].astype(np.uint8)).cuda(),
                    torch.from_numpy(g["scales"]).cuda(), torch.from_numpy(g["idx"]).cuda(), nn, rule="exon").cpu().numpy()
    want = g[f"out_n{nn}"]
    np.testing.assert_array_equal(np.isnan(got), np.isnan(want))
    np.testing.assert_allclose(got[~np.isnan(want)], want[~np.isnan(want)], rtol=1e-14)


def test_dipcn_batch_over_regions_both_rules(eng):
    """C5 shape in small: G = 734 regions at once; absent counts, zero counts, -1 padded lists, query offset."""
    rng = np.random.default_rng(8)
    n, g, k, q0, nq = 300, 734, 64, 37, 120
    counts = rng.integers(0, 3000, (n, g)).astype(np.int32)
    counts[rng.random((n, g)) < 0.05] = 0
    has = (rng.random(n) > 0.1).astype(np.uint8)
    scales = np.round(rng.uniform(0.5, 2.0, n), 2)
    idx = np.stack([rng.permutation(np.delete(np.arange(n), i))[:k] for i in range(n)]).astype(np.int32)
    idx[rng.random(idx.shape) < 0.03] = -1
    for rule, fn in (("pipeline", O.dipcn_arrays), ("exon", O.dipcn_arrays_exon)):
        want = fn(counts, has, scales, idx, 40)[q0:q0 + nq]
        got = eng.dipcn(torch.from_numpy(counts).cuda(), torch.from_numpy(has).cuda(), torch.from_numpy(scales).cuda(),
                        torch.from_numpy(idx[q0:q0 + nq]).cuda(), 40, q_gid0=q0, rule=rule).cpu().numpy()
        np.testing.assert_array_equal(np.isnan(got), np.isnan(want))
        np.testing.assert_allclose(got[~np.isnan(want)], want[~np.isnan(want)], rtol=1e-13)


# ---- step-4 hand-off text on the device (csrc/textio.cu) ------------------------------------------------------
def _codes(rng, n, rw):
    c = rng.integers(-450, 451, size=(n, rw)).astype(np.int16)
    special = rng.random((n, rw))
    c[special < 0.05] = -32768          # NA
    c[(special >= 0.05) & (special < 0.07)] = -32767   # -0.00
    c[(special >= 0.07) & (special < 0.09)] = rng.choice(np.array([32766, -32766, 999, 1000, -1000, 9999, 10000, -10000, 0, 5, -5], dtype=np.int16),
                                                         size=int(((special >= 0.07) & (special < 0.09)).sum()))
    return c


def _host_text(codes):
    from grid_b200 import io as gio

    tab = gio._z_table()
    return "".join("\t".join(tab[int(v) + 32768] for v in row) + "\n" for row in codes).encode()


@pytest.mark.parametrize("n,rw,ld", [(1, 1, 8), (3, 7, 8), (5, 2048, 2048), (4, 2049, 2056), (7, 6001, 6003), (300, 38, 40)])
def test_device_formatter_and_parser_match_the_host_writer(n, rw, ld, eng):
    from grid_b200._lib import call, load, ptr

    rng = np.random.default_rng(n * 1000 + rw)
    codes = _codes(rng, n, rw)
    zq = torch.full((n, ld), 12345, dtype=torch.int16, device="cuda")
    zq[:, :rw] = torch.from_numpy(codes).cuda()
    st = torch.cuda.current_stream().cuda_stream
    ws = torch.empty(load().gridb200_format_z_ws_bytes(n, rw), dtype=torch.uint8, device="cuda")
    row_off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    call("gridb200_format_z_measure", ptr(zq), n, rw, ld, ptr(row_off), ptr(ws), ws.numel(), st)
    off = row_off.cpu().numpy()
    want = _host_text(codes)
    assert int(off[-1]) == len(want)
    text = torch.zeros(int(off[-1]) + 64, dtype=torch.uint8, device="cuda")
    call("gridb200_format_z_write", ptr(zq), n, rw, ld, ptr(row_off), ptr(ws), text.data_ptr() + 3, int(off[-1]), st)  # odd phase
    got = bytes(text[3:3 + int(off[-1])].cpu().numpy())
    assert got == want
    assert int(text[3 + int(off[-1]):].sum()) == 0 and int(text[:3].sum()) == 0  # nothing written outside
    # ...and back: every line gets an "id\tscale\t" prefix of a different length, as in the file
    lines = want.split(b"\n")[:-1]
    full = b"".join(f"S{i:0{1 + i % 5}d}\t1.{i % 100:02d}\t".encode() + ln + b"\n" for i, ln in enumerate(lines))
    offs = np.concatenate([[0], np.cumsum([len(f"S{i:0{1 + i % 5}d}\t1.{i % 100:02d}\t") + len(ln) + 1 for i, ln in enumerate(lines)])])
    t_text = torch.frombuffer(bytearray(full), dtype=torch.uint8).cuda()
    t_off = torch.from_numpy(offs.astype(np.int64)).cuda()
    back = torch.full((n, ld), -32768, dtype=torch.int16, device="cuda")
    stats = torch.zeros(4, dtype=torch.int64, device="cuda")
    max_line = int(np.diff(offs).max())
    ws2 = torch.empty(load().gridb200_parse_z_ws_bytes(n, max_line), dtype=torch.uint8, device="cuda")
    call("gridb200_parse_z_rows", ptr(t_text), ptr(t_off), n, max_line, 2, rw, ptr(back), ld, ptr(stats), ptr(ws2), ws2.numel(), st)
    s = stats.cpu().numpy()
    assert (s[0], s[1], s[2]) == (n * rw, 0, 0)
    expect = codes.copy()
    expect[expect == -32767] = 0  # "-0.00" reads back as zero
    np.testing.assert_array_equal(back[:, :rw].cpu().numpy(), expect)


def test_device_parser_flags_what_it_does_not_understand(eng):
    from grid_b200._lib import call, load, ptr

    good = b"A\t1.00\t0.10\tNA\tnan\t-3.25\n"
    cases = {b"A\t1.00\t0.1\tNA\tnan\t-3.25\n": (1, 0), b"A\t1.00\t1e-1\tNA\tnan\t-3.25\n": (1, 0),
             b"A\t1.00\t0.10\tNA\tnan\n": (0, 1), b"A\t1.00\t0.10\tNA\tnan\t-3.25\t7.00\n": (0, 1),
             b"A\t1.00\t0.10\tNaN\tnan\t-3.25\n": (1, 0), b"A\t1.00\t500.00\tNA\tnan\t-3.25\n": (1, 0), good: (0, 0)}
    st = torch.cuda.current_stream().cuda_stream
    for text, (bad_tok, bad_row) in cases.items():
        t = torch.frombuffer(bytearray(text), dtype=torch.uint8).cuda()
        off = torch.tensor([0, len(text)], dtype=torch.int64, device="cuda")
        zq = torch.full((1, 8), -32768, dtype=torch.int16, device="cuda")
        stats = torch.zeros(4, dtype=torch.int64, device="cuda")
        ws = torch.empty(load().gridb200_parse_z_ws_bytes(1, len(text)), dtype=torch.uint8, device="cuda")
        call("gridb200_parse_z_rows", ptr(t), ptr(off), 1, len(text), 2, 4, ptr(zq), 8, ptr(stats), ptr(ws), ws.numel(), st)
        s = stats.cpu().numpy()
        assert (int(s[1]), int(s[2])) == (bad_tok, bad_row), text
    assert zq[0, :4].cpu().tolist() == [10, -32768, -32768, -325]


def test_device_text_file_round_trip_and_fallback(tmp_path, eng):
    from grid_b200 import io as gio, textio

    rng = np.random.default_rng(4)
    n, rw = 37, 4500
    codes = _codes(rng, n, rw)
    ids = synth.sample_ids(n)
    mu, ratio, raw = rng.uniform(0.5, 2, rw), rng.uniform(50, 300, rw), rng.uniform(20, 40, n)
    ratio[3] = np.nan
    zq = torch.from_numpy(codes).cuda()
    old_member = textio.MEMBER_BYTES
    textio.MEMBER_BYTES = 100_000  # several gzip members
    try:
        textio.write_normalized_output(tmp_path / "dev.tsv.gz", ids, zq, rw, mu, ratio, raw, threads=4)
    finally:
        textio.MEMBER_BYTES = old_member
    gio.write_normalized_output(tmp_path / "host.tsv.gz", ids, codes, mu, ratio, raw)
    assert gzip.open(tmp_path / "dev.tsv.gz", "rb").read() == gzip.open(tmp_path / "host.tsv.gz", "rb").read()
    ids2, ratios2, zq2, scales2 = textio.read_normalized_data(tmp_path / "host.tsv.gz")
    h_ids, h_ratios, h_zq, h_scales = gio.read_normalized_data(tmp_path / "host.tsv.gz")
    assert ids2 == h_ids
    np.testing.assert_array_equal(ratios2, h_ratios)
    np.testing.assert_array_equal(scales2, h_scales)
    np.testing.assert_array_equal(zq2[:, :rw].cpu().numpy(), h_zq)
    # a file the strict grammar rejects (one-decimal value) goes through the host reader and still arrives
    with gzip.open(tmp_path / "odd.tsv.gz", "wt") as f:
        f.write("2\t3\t1.000\t1.000\t1.000\n2\t3\t100.000\t100.000\tNA\nA\t30.00\t0.5\tNA\t-1.25\nB\t31.00\t0.25\t0.00\t2.00\n")
    ids3, _, zq3, _ = textio.read_normalized_data(tmp_path / "odd.tsv.gz")
    assert ids3 == ["A", "B"] and zq3[:, :3].cpu().tolist() == [[50, -32768, -125], [25, 0, 200]]


def test_packed_host_format_round_trips_bit_for_bit(eng):
    """uint16 hundredths + escape list -> the same fp32 matrix (incl. values that do not fit 16 bits, odd widths)."""
    n, b = 300, 4099
    depth = eng.synth_depth(11, n, b)
    ld = depth.stride(0)
    depth[5, 17] = 655.35      # first value that needs an escape
    depth[5, 18] = 655.34      # last that does not
    depth[299, 4098] = 12345.67
    depth[0, 0] = 41943.03
    depth[7, 9] = 50000.0      # beyond the exact-hundredths range: carried verbatim
    want = depth.clone()
    packed = eng.pack_depth(depth, b, chunk_rows=64)
    assert packed.esc_cell.numel() == 4 and packed.codes.dtype == torch.uint16
    assert sorted(packed.esc_cell.tolist()) == sorted([5 * b + 17, 299 * b + 4098, 0, 7 * b + 9])
    out = torch.full_like(depth, -1.0)
    eng.upload_packed(packed, out, chunk_rows=77)
    torch.cuda.synchronize()
    assert torch.equal(out[:, :b].view(torch.int32), want[:, :b].view(torch.int32))
    # and a second upload into the same buffers (staging reuse) after changing the source
    packed.codes[3, 4] = 1234
    eng.upload_packed(packed, out, chunk_rows=77)
    torch.cuda.synchronize()
    assert float(out[3, 4]) == np.float32(12.34)


@pytest.mark.parametrize("bits", [9, 13, 15, "auto"])
def test_bit_packed_host_format_round_trips_bit_for_bit(eng, bits):
    """Hundredths in 9..15 bits per cell (groups of 32 cells, little-endian bit stream) + escape list -> the same fp32
    matrix: ragged last group, values on both sides of the escape code, absent cells, odd chunking."""
    n, b = 130, 4099
    depth = eng.synth_depth(12, n, b)
    width = eng.choose_bits(depth, b) if bits == "auto" else bits
    if bits == "auto":
        assert width == 13                                   # a 30x cohort: depths below 81.91
    esc = ((1 << width) - 1) / 100.0
    depth[5, 17] = esc                  # first value that needs an escape
    depth[5, 18] = esc - 0.01           # last that does not
    depth[129, 4098] = 12345.67
    depth[0, 0] = 0.0                   # absent
    depth[7, 31], depth[7, 32] = 0.01, 0.02   # group boundary
    want = depth.clone()
    packed = eng.pack_depth(depth, b, chunk_rows=50, bits=bits)
    assert packed.bits == width and packed.codes.dtype == torch.int32
    assert packed.codes.shape[1] == ((b + 31) // 32 * width + 3) // 4 * 4
    n_esc = int((torch.round(want[:, :b].double() * 100) >= (1 << width) - 1).sum())
    assert packed.esc_cell.numel() == n_esc and {5 * b + 17, 129 * b + 4098} <= set(packed.esc_cell.tolist())
    out = torch.full_like(depth, -1.0)
    eng.upload_packed(packed, out, chunk_rows=37)
    torch.cuda.synchronize()
    assert torch.equal(out[:, :b].view(torch.int32), want[:, :b].view(torch.int32))


def test_handoff_is_dropped_when_the_file_changes(tmp_path, eng):
    """A step-4 file rewritten by somebody else (here: the oracle) between the steps must be parsed, not trusted."""
    from grid_b200 import handoff as H
    from grid_b200.utils.find_neighbors import find_neighbors
    from grid_b200.utils.normalize_mosdepth import normalize_mosdepth

    cfg = synth.write_cohort_files(tmp_path, 5, 30, 90)
    H.clear()
    normalize_mosdepth(cfg, None)
    norm = Path(cfg["output_dir"]) / "norm.tsv.gz"
    assert H._CACHE, "step 4 should have left a hand-off entry"
    # another cohort's step-4 output lands on the same path
    cfg2 = synth.write_cohort_files(tmp_path / "other", 6, 30, 90)
    O.normalize_mosdepth(cfg2)
    norm.write_bytes((Path(cfg2["output_dir"]) / "norm.tsv.gz").read_bytes())
    find_neighbors(cfg, None)
    got = gzip.open(Path(cfg["output_dir"]) / "nbrs.zMax2.0.tsv.gz", "rt").read()
    O.find_neighbors(cfg2)
    (tmp_path / "got.tsv").write_text(got)
    assert_neighbor_files_equal(tmp_path / "got.tsv", Path(cfg2["output_dir"]) / "nbrs.zMax2.0.tsv.gz")


def _tiny_step4(path, ids, scales, ratios, rows):
    """A step-4 file written by hand (what any writer, including the reference, may produce)."""
    n, rw = len(ids), len(ratios)
    fmt3 = lambda v: "NA" if v != v else f"{v:.3f}"  # noqa: E731
    with gzip.open(path, "wt") as f:
        f.write(f"{n}\t{rw}\t" + "\t".join("1.000" for _ in ratios) + "\n")
        f.write(f"{n}\t{rw}\t" + "\t".join(fmt3(v) for v in ratios) + "\n")
        for sid, sc, row in zip(ids, scales, rows):
            f.write(f"{sid}\t{sc:.2f}\t" + "\t".join("NA" if v != v else f"{v:.2f}" for v in row) + "\n")


@pytest.mark.parametrize("name,ids,ratios,rows,knobs", [
    ("squared_distance_3_4_5", ["A", "B"], [100.0, 100.0], [[0.0, 0.0], [3.0, 4.0]], {}),   # test_neighbors.py:69-74
    ("single_sample", ["A"], [100.0, 120.0], [[0.5, -0.5]], {}),
    ("k_beyond_n", ["A", "B", "C"], [100.0, 90.0, 80.0], [[0.1, 0.2, 0.3], [0.1, 0.2, 0.4], [1.5, -1.5, 0.0]], {"num_neighbors": 500}),
    ("na_cells_and_clip", ["A", "B", "C"], [100.0, 90.0], [[float("nan"), 5.0], [-7.25, float("nan")], [0.0, 1.99]], {}),
    ("all_ratios_nan_keeps_all", ["A", "B", "C"], [float("nan")] * 3, [[0.1, 0.2, 0.3], [0.3, 0.2, 0.1], [0.0, 0.0, 0.0]], {}),
    ("sigma2_max_drops_every_region", ["A", "B", "C"], [2000.0, 3000.0], [[0.1, 0.2], [0.3, 0.2], [0.0, 0.1]], {"sigma2_max": 1000.0}),
    ("frac_r_half", ["A", "B", "C", "D"], [10.0, 40.0, 30.0, 20.0], [[0.1, 1.0, 0.3, -0.4], [0.3, -1.0, 0.1, 0.4], [0.0, 0.5, 0.0, 0.0], [1.0, 1.0, 1.0, 1.0]], {"frac_r": 0.5}),
    ("exact_ties", ["A", "B", "C", "D"], [100.0], [[0.0], [1.0], [-1.0], [1.0]], {}),
])
def test_find_neighbors_edge_cases_match_the_oracle(name, ids, ratios, rows, knobs, tmp_path, eng):
    """Hand-written step-4 files through our step 5 and through the oracle's: the neighbour files must agree
    (helpers.assert_neighbor_files_equal allows reordering only inside exact distance ties)."""
    from grid_b200.utils.find_neighbors import find_neighbors

    od = tmp_path / "out"
    od.mkdir()
    cfg = {"output_dir": str(od), "output_file_type": "tsv",
           "mosdepth": {"normalize": {"output_file_prefix": "norm"},
                        "neighbors": {"output_file_prefix": "nbrs", "zmax": 2.0, "num_neighbors": 2, **knobs}}}
    _tiny_step4(od / "norm.tsv.gz", ids, [30.0 + i for i in range(len(ids))], ratios, rows)
    if name == "sigma2_max_drops_every_region":  # N x 0 matrix: scikit-learn refuses it, and so do we
        with pytest.raises(ValueError, match="0 feature"):
            O.find_neighbors(cfg)
        with pytest.raises(ValueError, match="0 feature"):
            find_neighbors(cfg, None)
        return
    find_neighbors(cfg, None)
    mine = od / "mine.tsv"
    mine.write_text(gzip.open(od / "nbrs.zMax2.0.tsv.gz", "rt").read())
    O.find_neighbors(cfg)
    assert_neighbor_files_equal(mine, od / "nbrs.zMax2.0.tsv.gz")
    if name == "squared_distance_3_4_5":  # clipped at zmax = 2: (0,0)-(2,2) -> 8 / (2 * 2 regions) = 2.00
        assert mine.read_text().splitlines()[0].split("\t")[4] == "2.00"


def _mutate_step4_inputs(cfg, name):
    """Config / input variations of step 4 that the goldens do not cover."""
    work = Path(cfg["mosdepth"]["work_dir"])
    n = cfg["mosdepth"]["normalize"]
    if name == "genome_wide":            # chrom = None: no prefix test, no window
        cfg.update(chrom=None, start_bp=None, end_bp=None)
    elif name == "chrom_without_prefix":  # "1" is compared as "chr1" (norm_chrom)
        cfg.update(chrom="1")
    elif name == "sub_window":           # closed-interval overlap test on both ends
        cfg.update(start_bp=20_500, end_bp=60_000)
    elif name == "tight_depth_filter":
        n.update(min_depth=33, max_depth=36)
    elif name == "top_frac_half":
        n.update(top_frac=0.5)
    elif name == "top_frac_zero":        # threshold = smallest ratio, strict '>' drops exactly the minimum
        n.update(top_frac=0.0)
    elif name == "empty_and_short_files":
        files = sorted(work.glob("*.regions.bed.gz"))
        with gzip.open(files[3], "wt") as f:   # a sample without any line: removed from the output
            f.write("")
        with gzip.open(files[5], "wt") as f:   # only bins that the depth filter rejects, plus malformed lines
            f.write("chr1\t1000\t2000\t0.01\nshort\tline\nchr1\t2000\t3000\t0.00\n")
    elif name == "duplicate_bin_last_wins":
        f0 = sorted(work.glob("*.regions.bed.gz"))[0]
        text = gzip.open(f0, "rt").read()
        with gzip.open(f0, "wt") as f:
            f.write(text + "chr1\t7000\t8000\t41.25\n")
    elif name == "mask_whole_stretch":
        Path(n["repeat_mask_file"]).write_text("chr1\t30000\t45999\n1\t70000\t70001\n")
    return cfg


@pytest.mark.parametrize("name", ["genome_wide", "chrom_without_prefix", "sub_window", "tight_depth_filter", "top_frac_half",
                                  "top_frac_zero", "empty_and_short_files", "duplicate_bin_last_wins", "mask_whole_stretch"])
@pytest.mark.parametrize("device_ingest", [True, False])
def test_normalize_mosdepth_variations_match_the_oracle(name, device_ingest, tmp_path, eng):
    from grid_b200.utils.normalize_mosdepth import normalize_mosdepth

    cfg = _mutate_step4_inputs(synth.write_cohort_files(tmp_path, 17, 36, 110), name)
    cfg["gpu"] = {"device_ingest": device_ingest, "device_handoff": False}
    normalize_mosdepth(cfg, None)
    out = Path(cfg["output_dir"]) / "norm.tsv.gz"
    mine = tmp_path / "mine.tsv"
    mine.write_text(gzip.open(out, "rt").read())
    out.unlink()
    assert O.normalize_mosdepth(cfg) is not None
    assert_norm_files_equal(mine, out)


@pytest.mark.parametrize("b", [6, 43, 130, 1000, 4100, 70001])
def test_sample_means_on_exact_two_decimal_ties_take_numpys_value(b, eng):
    """np.nanmean is a pairwise floating-point sum; when the exact mean is k + 0.005 its last ulp decides what "%.2f"
    prints (normalize_mosdepth.py:120, :548).  Rows built to sit exactly on such ties must print as numpy's do."""
    from grid_b200._lib import call, ptr

    rng = np.random.default_rng(b)
    n = 64
    h = rng.integers(1500, 6000, size=(n, b)).astype(np.int64)
    h[rng.random((n, b)) < 0.02] = 0                       # absent cells (NaN in the reference's matrix)
    valid = (rng.random(b) < 0.8).astype(np.uint8)
    valid[:2] = 1
    ties = 0
    for i in range(n):                                     # nudge one valid, present cell so that 2*sum == cnt (mod 2*cnt)
        cols = np.nonzero(valid & (h[i] > 0))[0]
        cnt = len(cols)
        if cnt % 2 or i % 4 == 3:                          # odd counts cannot tie; leave every fourth row alone
            continue
        s = int(h[i, cols].sum())
        delta = (cnt // 2 - s) % cnt
        h[i, cols[0]] += delta
        ties += 1
    assert ties > 10
    x = np.where(h > 0, h / 100.0, np.nan)
    # the reference's matrix is a fresh C-order array (normalize_mosdepth.py:404): row sums are numpy's PAIRWISE sums
    want_mean = np.nanmean(np.ascontiguousarray(x[:, valid.astype(bool)]), axis=1)
    want_q = np.array([float("%.2f" % v) for v in want_mean])
    ld = (b + 3) // 4 * 4
    depth = torch.zeros((n, ld), dtype=torch.float32, device="cuda")
    depth[:, :b] = torch.from_numpy((h / 100.0).astype(np.float32)).cuda()
    vmask = valid.astype(bool)
    rowsum = torch.from_numpy(np.where(vmask, h, 0).sum(axis=1)).cuda()
    rowcnt = torch.from_numpy(((h > 0) & vmask).sum(axis=1).astype(np.int32)).cuda()
    raw, q, f = (torch.empty(n, dtype=torch.float64, device="cuda") for _ in range(3))
    st = torch.cuda.current_stream().cuda_stream
    tvalid = torch.from_numpy(valid).cuda()
    call("gridb200_row_scales", ptr(rowsum), ptr(rowcnt), n, ptr(raw), ptr(q), ptr(f), st)
    before = q.cpu().numpy().copy()
    call("gridb200_row_means_exact", ptr(depth), n, b, ld, ptr(tvalid), ptr(rowsum), ptr(rowcnt), ptr(raw), ptr(q), st)
    np.testing.assert_array_equal(q.cpu().numpy(), want_q)
    # rows on a tie now carry numpy's double itself; and the exercise is not vacuous: some ties did print differently
    tie_rows = np.array([(2 * int(s)) % (2 * int(c)) == int(c) for s, c in zip(rowsum.cpu().tolist(), rowcnt.cpu().tolist())])
    np.testing.assert_array_equal(raw.cpu().numpy()[tie_rows], want_mean[tie_rows])
    if b >= 130:
        assert (before != want_q).any()
