"""The fused Gram + k-best kernel (gridb200_gram_topk) against numpy: exact int64 distances formed in the epilogue,
threshold filter against the lists' current k-th entries, candidates for the query rows AND (the half ring's reading of
the transpose) for the reference rows, int32 -> int64 carries inside the pair, overflow and not-ready refusals."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _operand(codes):
    from grid_b200 import engine

    zq = torch.from_numpy(codes.astype(np.int16)).cuda().contiguous()
    return engine.prepare_from_zq(zq, codes.shape[1], None, 200)        # (zh k-blocked fp16, kdim, exact norms)


def _brute(q, refs, q_gid0, ref_gids, k):
    """k best (d2, gid) per row of q among the rows of refs, self excluded by gid."""
    d2 = (q * q).sum(1)[:, None] + (refs * refs).sum(1)[None, :] - 2 * (q @ refs.T)
    out = []
    for i in range(q.shape[0]):
        out.append(sorted((int(d2[i, j]), int(g)) for j, g in enumerate(ref_gids) if g != q_gid0 + i)[:k])
    return out


def _lists(d2_t, idx_t):
    return [[(int(d), int(j)) for d, j in zip(a, b) if j >= 0] for a, b in zip(d2_t.tolist(), idx_t.tolist())]


@pytest.mark.parametrize("kdim,nq,nr", [(1024, 300, 520), (65536, 130, 260)])
def test_fused_block_matches_numpy(kdim, nq, nr):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200 import engine

    rng = np.random.default_rng(kdim)
    k = 40
    A = rng.integers(-200, 201, size=(200, kdim)).astype(np.int64)      # what the query rows have already seen
    C = rng.integers(-200, 201, size=(150, kdim)).astype(np.int64)      # what the reference rows have already seen
    Q = rng.integers(-200, 201, size=(nq, kdim)).astype(np.int64)
    B = rng.integers(-200, 201, size=(nr, kdim)).astype(np.int64)
    # rows whose dot products leave the int32 range at kdim = 65536 (carries inside the pair), exact ties, a duplicate
    Q[0], B[0], B[1] = 200, 200, -200
    B[7] = Q[5]
    B[9] = B[8]
    gq, ga, gb, gc = 1000, 5000, 2000, 7000                              # global ids of Q, A, B, C rows
    zq, kd, nQ = _operand(Q)
    za, _, nA = _operand(A)
    zb, _, nB = _operand(B)
    zc, _, nC = _operand(C)
    bq_d2, bq_idx = engine.knn_init(nq, k, "cuda")
    engine.knn_block(zq, nQ, gq, za, nA, ga, kd, 200, bq_d2, bq_idx)    # bootstrap: lists are full, thresholds finite
    bb_d2, bb_idx = engine.knn_init(nr, k, "cuda")
    engine.knn_block(zb, nB, gb, zc, nC, gc, kd, 200, bb_d2, bb_idx)
    tau_r = bb_d2[:, k - 1].contiguous()
    t_d2, t_idx = engine.knn_init(nr, k, "cuda")
    before = engine.COUNTERS["fused_blocks"]
    ok = engine.knn_block_fused(zq, nQ, gq, zb, nB, gb, kd, 200, bq_d2, bq_idx, tau_r=tau_r, t_lists=(t_d2, t_idx))
    assert ok and engine.COUNTERS["fused_blocks"] == before + 1
    want_q = _brute(Q, np.concatenate([A, B]), gq, list(range(ga, ga + 200)) + list(range(gb, gb + nr)), k)
    assert _lists(bq_d2, bq_idx) == want_q
    engine.merge_lists(t_d2, t_idx, bb_d2, bb_idx)                       # what the shard's owner does with the lists it gets
    want_b = _brute(B, np.concatenate([C, Q]), gb, list(range(gc, gc + 150)) + list(range(gq, gq + nq)), k)
    assert _lists(bb_d2, bb_idx) == want_b


def test_fused_sub_ranges_and_refusals(monkeypatch):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200 import engine

    rng = np.random.default_rng(3)
    k, kdim = 30, 512
    A = rng.integers(-200, 201, size=(300, kdim)).astype(np.int64)
    Q = rng.integers(-200, 201, size=(280, kdim)).astype(np.int64)
    B = rng.integers(-200, 201, size=(270, kdim)).astype(np.int64)
    zq, kd, nQ = _operand(Q)
    za, _, nA = _operand(A)
    zb, _, nB = _operand(B)
    d2, idx = engine.knn_init(280, k, "cuda")
    # lists still empty: no threshold to filter against -> refused, nothing changed
    assert not engine.knn_block_fused(zq, nQ, 0, zb, nB, 1000, kd, 200, d2, idx)
    assert int((idx >= 0).sum()) == 0
    engine.knn_block(zq, nQ, 0, za, nA, 5000, kd, 200, d2, idx)
    snap = (d2.clone(), idx.clone())
    # every row of a block of small-magnitude samples is closer than anything in the lists (d2 ~ |q|^2 against ~ 2|q|^2):
    # 270 survivors per row do not fit 64 slots -> refused, nothing changed
    near = rng.integers(-10, 11, size=(270, kdim)).astype(np.int64)
    zn, _, nN = _operand(near)
    monkeypatch.setattr(engine, "FUSED_CAP", 64)
    over = engine.COUNTERS["fused_overflow"]
    assert not engine.knn_block_fused(zq, nQ, 0, zn, nN, 3000, kd, 200, d2, idx)
    assert engine.COUNTERS["fused_overflow"] == over + 1 and torch.equal(d2, snap[0]) and torch.equal(idx, snap[1])
    monkeypatch.setattr(engine, "FUSED_CAP", 1024)
    # the even-world middle hop: query rows 140.. against reference rows ..135 only
    assert engine.knn_block_fused(zq, nQ, 0, zb, nB, 1000, kd, 200, d2, idx, q_range=(140, 280), r_range=(0, 135))
    got = _lists(d2, idx)
    assert got[:140] == _lists(*snap)[:140]
    want = _brute(Q[140:], np.concatenate([A, B[:135]]), 140, list(range(5000, 5300)) + list(range(1000, 1135)), k)
    assert got[140:] == want
