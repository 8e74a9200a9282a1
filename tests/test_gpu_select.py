"""gridb200_select / gridb200_knn_col_filter through the C ABI against numpy: the radix select's last four passes run on the
keys collected in pass 3 (one block, one launch); when more than 65 536 keys share a target's 24-bit prefix the collection
overflows and the full-vector passes take over.  Both ways, and the allocation-free entry point (which never collects),
must give the order statistics numpy gives (normalize_mosdepth.py:460-468, :479-499; find_neighbors.py:128-175)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _select(r: np.ndarray, top_frac: float):
    from grid_b200 import _lib
    from grid_b200._lib import call, ptr

    b = r.size
    ratio = torch.from_numpy(r).cuda()
    st = torch.cuda.current_stream().cuda_stream
    sel = torch.full((b,), -7, dtype=torch.int32, device="cuda")
    info = torch.full((4,), -1.0, dtype=torch.float64, device="cuda")
    ws = torch.empty(_lib.load().gridb200_select_ws_bytes(b), dtype=torch.uint8, device="cuda")
    call("gridb200_select", ptr(ratio), b, float(top_frac), ptr(sel), ptr(info), ptr(ws), ws.numel(), st)
    torch.cuda.synchronize()
    return ratio, sel, info, ws


def _check_against_numpy(r, top_frac, sel, info):
    fin = np.sort(r[~np.isnan(r)])
    thr = fin[int(top_frac * fin.size)]
    med = float(np.median(fin))
    want = np.nonzero(~np.isnan(r) & (r > thr))[0]
    got = sel.cpu().numpy()
    np.testing.assert_array_equal(np.nonzero(got >= 0)[0], want)
    np.testing.assert_array_equal(got[want], np.arange(want.size))
    h = info.cpu().numpy()
    assert h[1] == thr and int(h[2]) == fin.size and int(h[3]) == want.size
    assert h[0] == (1.0 / np.sqrt(med / 100.0) if med > 0 else 1.0)


CASES = {
    "gamma": lambda rng, b: rng.gamma(2.0, 40.0, size=b),
    # 150 000 copies of the median value and 90 000 of the threshold value: far more than 65 536 keys agree with a target
    # beyond the 24-bit prefix, the collection overflows and all eight passes run over the vector
    "plateaus": lambda rng, b: np.concatenate([rng.gamma(2.0, 40.0, size=b - 240_000), np.full(150_000, 61.25),
                                               np.full(90_000, 173.0625)]),
    # every value in one binade with a common 20-bit mantissa prefix: pass 3 collects (nearly) everything -> overflow
    "narrow": lambda rng, b: 64.0 + rng.random(b) * 2.0 ** -12,
    # values a few ulps apart: only the very last passes tell them apart
    "ulps": lambda rng, b: np.float64(100.0) + np.arange(b) % 977 * np.spacing(np.float64(100.0)),
}


@pytest.mark.parametrize("case", sorted(CASES))
@pytest.mark.parametrize("b,top_frac", [(400_003, 0.1), (70_001, 0.5), (513, 0.25)])
def test_select_matches_numpy_with_and_without_the_collected_tail(case, b, top_frac):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200._lib import call, ptr

    rng = np.random.default_rng(sum(map(ord, case)) * 1000003 + b)
    if case == "plateaus" and b < 300_000:
        pytest.skip("the plateaus need a long vector")
    r = CASES[case](rng, b).astype(np.float64)
    rng.shuffle(r)
    r[rng.random(b) < 0.02] = np.nan
    ratio, sel, info, _ = _select(r, top_frac)
    _check_against_numpy(r, top_frac, sel, info)
    # the allocation-free entry point keeps all eight passes on the whole vector: same answer bit for bit
    sel2 = torch.full((b,), -9, dtype=torch.int32, device="cuda")
    info2 = torch.full((4,), -2.0, dtype=torch.float64, device="cuda")
    call("gridb200_select_static", ptr(ratio), b, float(top_frac), ptr(sel2), ptr(info2), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert torch.equal(sel, sel2) and torch.equal(info.view(torch.int64), info2.view(torch.int64))


@pytest.mark.parametrize("frac_r", [1.0, 0.8, 0.35])
def test_knn_filter_threshold_matches_numpy(frac_r):
    """sigma2_min = sorted(%.3f ratios of the selected bins)[int(R * (1 - frac_r))] (find_neighbors.py:150-160)."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200._lib import call, ptr

    rng = np.random.default_rng(7 + int(frac_r * 100))
    b, top_frac, s2max = 250_007, 0.1, 1000.0
    r = rng.gamma(2.0, 40.0, size=b)
    r[rng.random(b) < 0.02] = np.nan
    ratio, sel, info, ws = _select(r, top_frac)
    keep = torch.zeros(b, dtype=torch.uint8, device="cuda")
    info2 = torch.zeros(2, dtype=torch.float64, device="cuda")
    call("gridb200_knn_col_filter", ptr(ratio), ptr(sel), b, float(frac_r), s2max, ptr(keep), ptr(info), ptr(info2), ptr(ws),
         ws.numel(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    chosen = np.nonzero(sel.cpu().numpy() >= 0)[0]
    rq = np.array([float("%.3f" % v) for v in r[chosen]])
    srt = np.sort(rq)
    idx = min(int(chosen.size * (1.0 - frac_r)), chosen.size - 1)
    smin = srt[idx]
    want = (rq >= smin) & (rq <= s2max)
    got = keep.cpu().numpy()[chosen].astype(bool)
    np.testing.assert_array_equal(got, want)
    h2 = info2.cpu().numpy()
    assert int(h2[0]) == int(want.sum()) and h2[1] == smin
