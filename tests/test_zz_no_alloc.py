"""gridb200_select_static (the allocation-free variant of gridb200_select: scratch in a static __device__ array) through
the C ABI against gridb200_select with a caller-provided workspace: identical outputs (integer positions and the four info
doubles bit for bit -- the tolerance of every selection test here), on two calls in a row (the array keeps the first
call's contents), and the size limit is refused with a clear error.  Collected last on purpose (file name)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _both(ratio, top_frac):
    from grid_b200 import _lib
    from grid_b200._lib import call, ptr

    b = ratio.numel()
    st = torch.cuda.current_stream().cuda_stream
    sel_a = torch.full((b,), -7, dtype=torch.int32, device="cuda")
    sel_b = torch.full((b,), -9, dtype=torch.int32, device="cuda")
    info_a = torch.full((4,), -1.0, dtype=torch.float64, device="cuda")
    info_b = torch.full((4,), -2.0, dtype=torch.float64, device="cuda")
    ws = torch.empty(_lib.load().gridb200_select_ws_bytes(b), dtype=torch.uint8, device="cuda")
    call("gridb200_select", ptr(ratio), b, float(top_frac), ptr(sel_a), ptr(info_a), ptr(ws), ws.numel(), st)
    call("gridb200_select_static", ptr(ratio), b, float(top_frac), ptr(sel_b), ptr(info_b), st)
    torch.cuda.synchronize()
    return sel_a, info_a, sel_b, info_b


def test_static_scratch_select_matches_the_workspace_entry_point_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200 import _lib
    from grid_b200._lib import call, ptr

    rng = np.random.default_rng(20251105)
    # two different problems back to back: the second call finds the first one's state, histograms and tile sums
    for b, top_frac in ((300_001, 0.1), (70_000, 0.5)):
        r = rng.gamma(2.0, 40.0, size=b)
        r[rng.random(b) < 0.03] = np.nan                     # invalid bins
        r[:5] = r[5:10]                                       # exact ties around anywhere
        ratio = torch.from_numpy(r).cuda()
        sel_a, info_a, sel_b, info_b = _both(ratio, top_frac)
        assert torch.equal(sel_a, sel_b)
        assert torch.equal(info_a.view(torch.int64), info_b.view(torch.int64))
        # and against numpy (normalize_mosdepth.py:460-468, :479-499), so that "equal" is not "equally wrong"
        fin = np.sort(r[~np.isnan(r)])
        thr = fin[int(top_frac * fin.size)]
        want = np.nonzero(~np.isnan(r) & (r > thr))[0]
        got = sel_b.cpu().numpy()
        np.testing.assert_array_equal(np.nonzero(got >= 0)[0], want)
        np.testing.assert_array_equal(got[want], np.arange(want.size))
        assert int(info_b[3].item()) == want.size and float(info_b[1].item()) == thr
    # the limit is stated and enforced before anything is launched
    lim = 4194304
    big = torch.zeros(8, dtype=torch.float64, device="cuda")
    with pytest.raises(_lib.GridB200Error, match="GRIDB200_SELECT_STATIC_MAX_BINS"):
        call("gridb200_select_static", ptr(big), lim + 1, 0.1, ptr(big), ptr(big), None)
