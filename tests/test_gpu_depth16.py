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


@pytest.mark.parametrize("bits", ["best", 10, 11, 12])
def test_relative_host_format_round_trips_into_both_device_forms(eng, bits):
    """Residuals against the rank-one prediction (sample depth x bin bias) in 9..15 bits: the same fp32 matrix and the same
    2-byte matrix + escape table come back, whatever the width -- a width that is too narrow only lengthens the escape
    list.  Absent cells, cells far from the prediction, depths beyond the 2-byte range, a sample without data and a bin
    without data are all in the cohort."""
    n, b = 150, 4099
    depth = eng.synth_depth(31, n, b)
    depth[3, 40] = 0.0                    # absent
    depth[4, 41] = 0.01                   # far below the prediction
    depth[5, 42] = 399.5                  # far above it (an escape at every width), still a 2-byte value
    depth[6, 4098] = 12345.67             # beyond the 2-byte range too
    depth[9, :] = 0.0                     # a sample without data: row_q = 0
    depth[:, 77] = 0.0                    # a bin without data: col_q = 0
    depth[20, 77] = 31.25                 # ... but for one cell
    want = depth.clone()
    d16 = eng.to_depth16(depth, b)
    if bits == "best":
        packed = eng.pack_depth(depth, b, chunk_rows=64, bits="best")
        assert packed.relative and packed.bits <= 12, "a 30x cohort is smaller as residuals (11-12 bits against 13)"
    else:
        packed = eng.pack_depth(depth, b, chunk_rows=64, bits=bits, tables=eng.rank_one_tables(depth, b))
        assert packed.relative and packed.bits == bits
    assert packed.row_q.dtype == torch.int32 and packed.col_q.numel() == b and int(packed.row_q[9]) == 0
    cells = set(packed.esc_cell.tolist())
    assert {5 * b + 42, 6 * b + 4098} <= cells and (3 * b + 40) not in cells
    assert packed.n_big == 1
    out = torch.full_like(depth, -1.0)
    eng.upload_packed(packed, out, chunk_rows=37)
    torch.cuda.synchronize()
    assert torch.equal(out[:, :b].view(torch.int32), want[:, :b].view(torch.int32))
    o16 = eng.Depth16(torch.full(d16.codes.shape, 7, dtype=torch.int16, device="cuda").view(torch.uint16), b)
    eng.upload_packed16(packed, o16, chunk_rows=77)
    torch.cuda.synchronize()
    assert torch.equal(o16.codes.view(torch.int16)[:, :b], d16.codes.view(torch.int16)[:, :b])
    assert torch.equal(o16.esc_key, d16.esc_key) and torch.equal(o16.esc_h, d16.esc_h)
    assert torch.equal(o16.esc_cols, d16.esc_cols)
    # without a cell beyond the 2-byte range every escape folds back into the matrix and no table is left
    depth[6, 4098] = 30.0
    p2 = eng.pack_depth(depth, b, bits=packed.bits, tables=(packed.row_q, packed.col_q))
    assert p2.n_big == 0 and p2.esc_cell.numel() > 0
    eng.upload_packed16(p2, o16)
    torch.cuda.synchronize()
    assert o16.esc_key is None and torch.equal(o16.codes.view(torch.int16)[:, :b], eng.to_depth16(depth, b).codes.view(torch.int16)[:, :b])


def _decode_row(words, b, bits, row_q, col_q):
    """The bit-packed host format restated in plain Python from include/gridb200.h: groups of 32 cells, `bits` 32-bit
    little-endian words per group, cell c at bit c * bits; 0 = absent, all-ones = escape (returned as -1); relative codes
    add pred = (row_q * col_q[j] + 2048) >> 12 and subtract the bias 2^(bits-1)."""
    mask = (1 << bits) - 1
    out = []
    for g in range((b + 31) // 32):
        stream = 0
        for w in range(bits):
            stream |= (int(words[g * bits + w]) & 0xFFFFFFFF) << (32 * w)
        for c in range(32):
            j = g * 32 + c
            if j >= b:
                break
            v = (stream >> (c * bits)) & mask
            if v == 0 or v == mask:
                out.append(0 if v == 0 else -1)
            elif row_q is None:
                out.append(v)
            else:
                out.append(((row_q * int(col_q[j]) + 2048) >> 12) + v - (1 << (bits - 1)))
    return out


@pytest.mark.parametrize("relative", [False, True])
def test_bit_packed_codes_decode_as_the_header_says(eng, relative):
    """Not a round trip through our own unpacker: the packed words are decoded by the plain-Python restatement of the
    documented layout and must give the cohort's hundredths (escapes aside, which must be exactly the listed cells)."""
    from grid_b200 import synth

    n, b = 6, 1000
    depth = eng.synth_depth(31, n, b)
    depth[2, 7] = 0.0
    depth[3, 999] = 400.0
    h = synth.depth_hundredths(31, n, b)
    h[2, 7], h[3, 999] = 0, 40000
    if relative:
        packed = eng.pack_depth(depth, b, bits=11, tables=eng.rank_one_tables(depth, b))
        rq, cq = packed.row_q.tolist(), packed.col_q.tolist()
    else:
        packed = eng.pack_depth(depth, b, bits=13)
        rq, cq = [None] * n, None
    esc = set(packed.esc_cell.tolist())
    words = packed.codes.numpy()
    for i in range(n):
        got = _decode_row(words[i], b, packed.bits, rq[i], cq)
        for j in range(b):
            if got[j] < 0:
                assert i * b + j in esc, (i, j)
            else:
                assert i * b + j not in esc and got[j] == int(h[i, j]), (i, j, got[j], int(h[i, j]))
    assert 3 * b + 999 in esc and abs(float(packed.esc_val[packed.esc_cell.tolist().index(3 * b + 999)]) - 400.0) < 1e-3
