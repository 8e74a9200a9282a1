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


@pytest.mark.parametrize("case", sorted(CASES))
def test_file_chain_matches_reference_golden(case, tmp_path, golden, eng):
    """BASELINE config 1 (N=100, LPA window) and the masked/truncated edge case, through the three step functions."""
    norm, nbrs, dip = _run_steps(CASES[case](tmp_path))
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


@pytest.mark.parametrize("seed,n,b,holes,params", [
    (1, 64, 300, 0.01, dict(num_neighbors=20, n_nbr=8)),
    (2, 257, 2100, 0.0, dict(num_neighbors=500, n_nbr=300)),          # k > N-1
    (3, 130, 5000, 0.05, dict(min_depth=30, max_depth=40, top_frac=0.5, num_neighbors=40, n_nbr=10, zmax=1.5)),
    (4, 33, 1030, 0.0, dict(top_frac=0.0, frac_r=0.5, sigma2_max=0.9, num_neighbors=5, n_nbr=5)),
])
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
