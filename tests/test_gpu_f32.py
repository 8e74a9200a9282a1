"""The general fp32 neighbour search (north_star's 3xTF32 path, gridb200_split_tf32 / gridb200_gram_f32 /
gridb200_topk_merge_f32) against an fp64 brute force: distances to fp32 accuracy, ranking identical except inside
groups of candidates whose distances agree to a few 1e-6 relative (north_star's tie window is 1e-6; an fp32 accumulation
over tens of thousands of same-sign terms -- an exact duplicate's dot product -- is itself only good to ~3e-6 of
|x|^2 + |y|^2, whatever the hardware)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

REL = 3e-6


@pytest.mark.parametrize("n,r,k", [(700, 5000, 60), (300, 131, 299), (1030, 40000, 25)])
def test_knn_f32_matches_fp64_brute_force(n, r, k):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200 import engine

    rng = np.random.default_rng(n + r)
    x = (rng.standard_normal((n, r)) * 1.3).astype(np.float32)          # unrounded z-scores
    x[5] = x[17]                                                         # an exact duplicate (distance 0 both ways)
    idx, d2 = engine.knn_f32(torch.from_numpy(x).cuda(), k)
    idx, d2 = idx.cpu().numpy(), d2.cpu().numpy()
    xd = x.astype(np.float64)
    sq = (xd * xd).sum(1)
    want = sq[:, None] + sq[None, :] - 2.0 * (xd @ xd.T)
    np.fill_diagonal(want, np.inf)
    kk = min(k, n - 1)
    scale = sq[:, None] + sq[None, :]                                    # what the error of the expansion is relative to
    for i in range(n):
        assert len(set(idx[i])) == kk and i not in idx[i]
        got = want[i, idx[i]]
        # values: fp32-level agreement with the exact distance
        assert np.all(np.abs(d2[i] - got) <= REL * scale[i, idx[i]] + 1e-12), i
        # ranking: ascending up to the tie window, and nothing closer than the k-th was left out
        assert np.all(np.diff(got) >= -2 * REL * scale[i, idx[i]][1:]), i
        kth = np.sort(want[i])[kk - 1]
        assert got.max() <= kth + 2 * REL * scale[i].max(), i
    assert idx[5, 0] == 17 and idx[17, 0] == 5 and d2[5, 0] <= REL * scale[5, 17]


def test_split_is_exact_and_3xtf32_beats_1xtf32():
    """hi + lo reproduces x bit for bit; the three-product Gram is far closer to fp64 than hi.hi alone."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200._lib import call, ptr

    rng = np.random.default_rng(9)
    n, r = 256, 4096
    x = rng.standard_normal((n, r)).astype(np.float32)
    tx = torch.from_numpy(x).cuda()
    hi = torch.empty((r // 32, n, 32), dtype=torch.float32, device="cuda")
    lo = torch.empty_like(hi)
    norms = torch.empty(n, dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    call("gridb200_split_tf32", ptr(tx), n, r, r, ptr(hi), ptr(lo), n, ptr(norms), st)
    back = (hi + lo).permute(1, 0, 2).reshape(n, r)
    assert torch.equal(back, tx)
    assert int((hi.view(torch.int32) & 0x1FFF).abs().max()) == 0          # hi carries 10 mantissa bits
    np.testing.assert_allclose(norms.cpu().numpy(), (x.astype(np.float64) ** 2).sum(1), rtol=1e-13)
    dot = torch.empty((n, n), dtype=torch.float32, device="cuda")
    call("gridb200_gram_f32", ptr(hi), ptr(lo), n, n, ptr(hi), ptr(lo), n, n, r, ptr(dot), n, st)
    exact = x.astype(np.float64) @ x.astype(np.float64).T
    err3 = np.abs(dot.cpu().numpy() - exact).max()
    zero = torch.zeros_like(lo)
    call("gridb200_gram_f32", ptr(hi), ptr(zero), n, n, ptr(hi), ptr(zero), n, n, r, ptr(dot), n, st)
    err1 = np.abs(dot.cpu().numpy() - exact).max()
    assert err3 < 8e-3 and err1 > 20 * err3, (err1, err3)   # the diagonal (4096 same-sign terms) has ulp 4.9e-4
