"""The depth matrix in 2 bytes per cell on the device (engine.Depth16, the *_u16 entry points): every step-4 result must be
BIT-IDENTICAL to the fp32 storage -- same integers go in -- including cells that need the escape table (depth >= 655.35)
in valid and in filtered-out bins, absent cells, per-sample means on exact %.2f ties, ragged shapes."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200 import engine

    return engine


def _same(a, b, name):
    if a is None or b is None:
        assert a is None and b is None, name
        return
    if a.dtype.is_floating_point:
        assert torch.equal(a.view(torch.int64 if a.dtype == torch.float64 else torch.int16),
                           b.view(torch.int64 if b.dtype == torch.float64 else torch.int16)), name   # bit patterns (NaN-safe)
    else:
        assert torch.equal(a, b), name


def _compare_normalize(eng, depth, b, p, want_escapes):
    d16 = eng.to_depth16(depth, b)
    assert d16.n_esc == want_escapes
    r32 = eng.normalize(depth, b, p, want_zq=True, want_zh=True)
    r16 = eng.normalize(d16, b, p, want_zq=True, want_zh=True)
    for name in ("colsum", "colcnt", "valid", "raw_mean", "scale_q", "row_factor", "mu", "s2", "ratio", "sel_pos", "keep",
                 "zq", "zh", "norms"):
        _same(getattr(r32, name), getattr(r16, name), name)
    assert (r32.rw, r32.r_use, r32.zscale) == (r16.rw, r16.r_use, r16.zscale)
    _same(r32.extra["rowsum"], r16.extra["rowsum"], "rowsum")
    return r32


@pytest.mark.parametrize("n,b", [(300, 5003), (37, 1100), (513, 2048)])
def test_u16_storage_gives_bit_identical_step4(eng, n, b):
    depth = eng.synth_depth(21 + n, n, b)
    rng = np.random.default_rng(n)
    cells = [(int(rng.integers(n)), int(rng.integers(b))) for _ in range(40)]
    for t, (i, j) in enumerate(cells):
        depth[i, j] = [655.35, 700.0, 1234.56, 41000.0, 655.36][t % 5]          # need the escape table
    depth[3, 5], depth[3, 6] = 655.34, 0.0                                       # last value that fits; an absent cell
    depth[:, 40:44] *= 8.0                                                        # bins that FAIL the depth filter (mean > 100)...
    depth[:, 40:44] = torch.round(depth[:, 40:44] * 100) / 100
    depth[7, 41], depth[8, 43] = 900.0, 2000.5                                   # ... with escapes inside them
    want = len({c for c in cells} | {(7, 41), (8, 43)})
    r = _compare_normalize(eng, depth, b, eng.Params(), want)
    assert int(r.valid[40:44].sum()) == 0 and r.rw > 0


def test_u16_storage_on_exact_tie_rows_and_few_bins(eng):
    """Few bins: row slabs with partial buffers, and per-sample means exactly on %.2f ties (numpy's pairwise sum replayed
    from the 2-byte cells, one of them an escape)."""
    rng = np.random.default_rng(4)
    n, b = 64, 130
    h = rng.integers(2500, 5000, size=(n, b)).astype(np.int64)
    h[rng.random((n, b)) < 0.02] = 0
    h[5, 9] = 70000                                                               # an escape inside a tie row
    for i in range(n):
        cols = np.nonzero(h[i] > 0)[0]
        cnt = len(cols)
        if cnt % 2 == 0 and i % 3 != 2:
            h[i, cols[0] if i != 5 else cols[1]] += (cnt // 2 - int(h[i, cols].sum())) % cnt
    ld = (b + 3) // 4 * 4
    depth = torch.zeros((n, ld), dtype=torch.float32, device="cuda")
    depth[:, :b] = torch.from_numpy((h / 100.0).astype(np.float32)).cuda()
    r = _compare_normalize(eng, depth, b, eng.Params(min_depth=1.0, max_depth=1e4), 1)
    x = np.where(h > 0, h / 100.0, np.nan)
    want_q = np.array([float("%.2f" % v) for v in np.nanmean(np.ascontiguousarray(x), axis=1)])
    np.testing.assert_array_equal(r.scale_q.cpu().numpy(), want_q)


def test_chain_and_host_upload_through_u16(eng):
    """run_chain on Depth16 = run_chain on fp32; the packed host formats (13 and 16 bits) upload straight into the 2-byte
    device matrix with the same codes and escape table as the device-side conversion."""
    from grid_b200 import synth

    n, b = 200, 4099
    depth = eng.synth_depth(8, n, b)
    depth[11, 100], depth[150, 4098] = 999.99, 660.0
    counts = torch.from_numpy(synth.read_counts(8, n, [42599])).cuda()
    p = eng.Params(num_neighbors=30, n_nbr=20)
    d16 = eng.to_depth16(depth, b)
    a = eng.run_chain(depth, b, counts, p)
    c = eng.run_chain(d16, b, counts, p)
    for x, y, name in zip(a[1:], c[1:], ("idx", "d2", "cn")):
        _same(x, y, name)
    for bits in (13, 16):
        packed = eng.pack_depth(depth, b, chunk_rows=64, bits=bits)
        out = eng.Depth16(torch.full(d16.codes.shape, 7, dtype=torch.int16, device="cuda").view(torch.uint16), b)
        eng.upload_packed16(packed, out, chunk_rows=77)
        torch.cuda.synchronize()
        assert torch.equal(out.codes.view(torch.int16)[:, :b], d16.codes.view(torch.int16)[:, :b])
        assert torch.equal(out.esc_key, d16.esc_key) and torch.equal(out.esc_h, d16.esc_h)
