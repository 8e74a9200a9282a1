"""Row-sharded (ring) chain on GPUs against the oracle: one rank (every kernel and collective call of the path, no
hop) on any box; two and more ranks when the box has the GPUs (NCCL)."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

from grid_b200 import synth  # noqa: E402
from oracle import grid_oracle as O  # noqa: E402

DIPCN_REL = 1e-5


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker_bins(rank, world, port, seed, n, b, params, out):
    """Bin-sharded chain: every rank holds all samples x a bin slice; Gram partials are reduce-scattered."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), NCCL_DEBUG="WARN")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from grid_b200 import engine

        h = synth.depth_hundredths(seed, n, b)
        counts = synth.read_counts(seed, n, [42599, 1000])
        want = O.chain_arrays(h / 100.0, counts, **params)
        per = -(-(-(-b // world)) // 64) * 64
        lo, hi = min(b, rank * per), min(b, (rank + 1) * per)
        depth = engine.synth_depth(seed, n, hi - lo, 0, lo)                      # all samples, this rank's bins
        res, idx, d2, cn = engine.run_chain(depth, hi - lo, torch.from_numpy(counts).cuda(), engine.Params(**params),
                                            group=dist.group.WORLD)
        np.testing.assert_array_equal(idx.cpu().numpy(), want["nbr_idx"])
        np.testing.assert_array_equal(d2.cpu().numpy(), want["nbr_d2"])
        np.testing.assert_allclose(cn.cpu().numpy(), want["dipcn"], rtol=DIPCN_REL)
        out.put((rank, "ok"))
    except Exception as exc:
        out.put((rank, f"{type(exc).__name__}: {str(exc)[:400]}"))
    finally:
        dist.destroy_process_group()


def _worker(rank, world, port, seed, n, b, params, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), NCCL_DEBUG="WARN")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from grid_b200 import engine, ring

        h = synth.depth_hundredths(seed, n, b)
        counts = synth.read_counts(seed, n, [42599, 1000])
        want = O.chain_arrays(h / 100.0, counts, **params)
        cnt, off = ring.shard_rows(n, world)
        lo, hi = off[rank], off[rank] + cnt[rank]
        depth = engine.synth_depth(seed, cnt[rank], b, row0=lo)                  # this rank's samples, all bins
        res, idx, d2, cn = ring.run_chain_rows(depth, b, torch.from_numpy(counts[lo:hi]).cuda(), engine.Params(**params),
                                               dist.group.WORLD)
        np.testing.assert_array_equal(idx.cpu().numpy(), want["nbr_idx"][lo:hi])  # global ids, (d2, index) order
        np.testing.assert_array_equal(d2.cpu().numpy(), want["nbr_d2"][lo:hi])
        np.testing.assert_allclose(cn.cpu().numpy(), want["dipcn"][lo:hi], rtol=DIPCN_REL)
        np.testing.assert_array_equal(res.scale_q.cpu().numpy(), want["scale_q"][lo:hi])
        assert res.rw == len(want["sel"]) and res.r_use == want["r_use"]
        out.put((rank, "ok"))
    except Exception as exc:
        out.put((rank, f"{type(exc).__name__}: {str(exc)[:400]}"))
    finally:
        dist.destroy_process_group()


def _worker_blocked(rank, world, port, seed, n, b, params, out):
    """Out-of-core schedule: no rank ever holds its rows; blocks are regenerated from the synthetic source on demand."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), NCCL_DEBUG="WARN")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from grid_b200 import engine, ring

        params = dict(params)
        block_rows, resident = params.pop("block_rows"), params.pop("resident_blocks", 0)
        h = synth.depth_hundredths(seed, n, b)
        counts = synth.read_counts(seed, n, [42599, 1000])
        want = O.chain_arrays(h / 100.0, counts, **params)
        cnt, off = ring.shard_rows(n, world)
        lo, hi = off[rank], off[rank] + cnt[rank]
        asked = []

        def source(r0, nr):
            asked.append((r0, nr))
            return engine.synth_depth(seed, nr, b, row0=lo + r0)

        res, idx, d2, cn = ring.run_chain_rows_blocked(source, cnt[rank], b, torch.from_numpy(counts[lo:hi]).cuda(),
                                                       engine.Params(**params), dist.group.WORLD, block_rows,
                                                       resident_blocks=resident)
        rounds = res.extra["rounds"]
        assert rounds >= 3 and max(nr for _, nr in asked) <= block_rows
        # two statistics passes + one operand per (query block, round) the half ring runs (with fewer than three ranks
        # the rounds below the diagonal are somebody's mirror image and are not even regenerated), minus the re-uses of
        # the pinned prefix
        pinned, made = set(), 0
        for qi in range(rounds):
            for bi in [qi] + [ri for ri in range(rounds) if ri != qi and not (world < 3 and ri < qi)]:
                if bi not in pinned:
                    made += 1
                    if bi < resident:
                        pinned.add(bi)
        assert len(asked) == 2 * rounds + made
        np.testing.assert_array_equal(idx.cpu().numpy(), want["nbr_idx"][lo:hi])
        np.testing.assert_array_equal(d2.cpu().numpy(), want["nbr_d2"][lo:hi])
        np.testing.assert_allclose(cn.cpu().numpy(), want["dipcn"][lo:hi], rtol=DIPCN_REL)
        np.testing.assert_array_equal(res.scale_q.cpu().numpy(), want["scale_q"][lo:hi])
        assert res.rw == len(want["sel"]) and res.r_use == want["r_use"]
        out.put((rank, "ok"))
    except Exception as exc:
        out.put((rank, f"{type(exc).__name__}: {str(exc)[:400]}"))
    finally:
        dist.destroy_process_group()


def _run(world, seed, n, b, params, worker=None):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=worker or _worker, args=(r, world, port, seed, n, b, params, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(out.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {r: "ok" for r in range(world)}


def test_rows_chain_single_rank():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    _run(1, 31, 130, 3000, dict(num_neighbors=20, n_nbr=8))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_rows_chain_ring(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    # uneven shards (n % world != 0), k larger than a shard, several 256-row tiles per shard at world = 2
    _run(world, 32, 601, 4000, dict(num_neighbors=350, n_nbr=300))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_bins_chain_reduce_scatter(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    # n not divisible by world (padded reduce-scatter rows), k > rows per rank
    _run(world, 33, 601, 4000, dict(num_neighbors=350, n_nbr=300), worker=_worker_bins)


@pytest.mark.parametrize("resident", [0, 2, 99])
def test_blocked_chain_single_rank(resident):
    """Query blocks x rounds on one GPU (every kernel of the out-of-core path; the ring degenerates to hop 0)."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    _run(1, 34, 530, 3000, dict(num_neighbors=300, n_nbr=40, block_rows=150, resident_blocks=resident), worker=_worker_blocked)


@pytest.mark.parametrize("fused", ["1", "0"])
def test_blocked_chain_single_rank_fused_epilogue(fused, monkeypatch):
    """The same schedule with the fused Gram + k-best kernel forced on every eligible block (k = 60 < block rows, so the
    lists fill and later blocks are filtered in the epilogue), and with it switched off: identical results."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    monkeypatch.setenv("GRIDB200_FUSED", fused)          # spawned workers inherit the environment
    _run(1, 36, 530, 3000, dict(num_neighbors=60, n_nbr=40, block_rows=150, resident_blocks=0), worker=_worker_blocked)


@pytest.mark.parametrize("world", [2, 4])
def test_rows_chain_ring_fused_epilogue(world, monkeypatch):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    monkeypatch.setenv("GRIDB200_FUSED", "1")
    _run(world, 37, 901, 4000, dict(num_neighbors=100, n_nbr=80))


@pytest.mark.parametrize("world", [2, 4])
def test_blocked_chain_ring(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    # (at least three rounds per rank: 1001 rows over `world` ranks in blocks of at most block_rows)
    _run(world, 35, 1001, 4000, dict(num_neighbors=350, n_nbr=300, block_rows=260 // world, resident_blocks=1),
         worker=_worker_blocked)


def test_transposed_and_list_merges_match_numpy():
    """gridb200_topk_merge_t reads a Gram block by columns; gridb200_topk_merge_lists merges two k-best lists."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200 import engine
    from grid_b200._lib import call, ptr

    rng = np.random.default_rng(5)
    nq, nr, kdim, k = 300, 170, 128, 40
    q = rng.integers(-200, 201, size=(nq, kdim)).astype(np.int64)
    r = rng.integers(-200, 201, size=(nr, kdim)).astype(np.int64)
    r[7] = q[11]                                                                   # exact tie partners
    dot = torch.from_numpy(q @ r.T).cuda().contiguous()
    nqv, nrv = torch.from_numpy((q * q).sum(1)).cuda(), torch.from_numpy((r * r).sum(1)).cuda()
    d2 = (q * q).sum(1)[:, None] + (r * r).sum(1)[None, :] - 2 * (q @ r.T)
    st = torch.cuda.current_stream().cuda_stream
    # lists of the reference rows (global ids 1000..), candidates = query rows (global ids 50..), in two candidate chunks
    t_d2, t_idx = engine.knn_init(nr, k, "cuda")
    for c0 in (0, 128):
        n_c = min(128, nq - c0) if c0 == 0 else nq - c0
        call("gridb200_topk_merge_t", dot.data_ptr() + c0 * nr * 8, nr, n_c, nr, nqv.data_ptr() + c0 * 8, ptr(nrv), 50 + c0,
             1000, k, ptr(t_d2), ptr(t_idx), st)
    for j in range(nr):
        want = sorted((int(d2[i, j]), 50 + i) for i in range(nq))[:k]
        assert [(int(a), int(b)) for a, b in zip(t_d2[j].tolist(), t_idx[j].tolist())] == want
    # merging two disjoint halves of the candidates through lists gives the same lists
    a_d2, a_idx = engine.knn_init(nr, k, "cuda")
    b_d2, b_idx = engine.knn_init(nr, k, "cuda")
    call("gridb200_topk_merge_t", ptr(dot), nr, 150, nr, ptr(nqv), ptr(nrv), 50, 1000, k, ptr(a_d2), ptr(a_idx), st)
    call("gridb200_topk_merge_t", dot.data_ptr() + 150 * nr * 8, nr, nq - 150, nr, nqv.data_ptr() + 150 * 8, ptr(nrv), 200,
         1000, k, ptr(b_d2), ptr(b_idx), st)
    engine.merge_lists(b_d2, b_idx, a_d2, a_idx)
    assert torch.equal(a_d2, t_d2) and torch.equal(a_idx, t_idx)
    with pytest.raises(ValueError, match="num_neighbors"):
        engine.check_k(5000)


def _worker_bins_ties(rank, world, port, seed, n, b, params, out):
    """Bin-sharded step 4 on rows built to sit EXACTLY on %.2f ties: the scales must be numpy's (pairwise sum over the
    whole row), as on one GPU -- no rank holds a whole row."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), NCCL_DEBUG="WARN")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from grid_b200 import engine

        rng = np.random.default_rng(seed)
        h = rng.integers(2500, 5000, size=(n, b)).astype(np.int64)
        h[rng.random((n, b)) < 0.02] = 0
        h[1, 3], h[4, b - 2] = 70000, 123456                                   # cells beyond 16 bits (escape table) in tie rows
        ties = 0
        for i in range(n):
            cols = np.nonzero(h[i] > 0)[0]
            cnt = len(cols)
            if cnt % 2 or i % 3 == 2:
                continue
            h[i, cols[-1]] += (cnt // 2 - int(h[i, cols].sum())) % cnt      # the nudged cell lives on the LAST rank
            ties += 1
        assert ties > n // 4
        x = np.where(h > 0, h / 100.0, np.nan)
        want_mean = np.nanmean(np.ascontiguousarray(x), axis=1)
        want_q = np.array([float("%.2f" % v) for v in want_mean])
        per = -(-(-(-b // world)) // 64) * 64
        lo, hi = min(b, rank * per), min(b, (rank + 1) * per)
        ld = (hi - lo + 3) // 4 * 4
        depth = torch.zeros((n, ld), dtype=torch.float32, device="cuda")
        depth[:, :hi - lo] = torch.from_numpy((h[:, lo:hi] / 100.0).astype(np.float32)).cuda()
        params = dict(params)
        if params.pop("u16", False):     # the 2-byte device matrix: the tie rows are rebuilt from codes + escape table
            depth = engine.to_depth16(depth, hi - lo)
        res = engine.normalize(depth, hi - lo, engine.Params(**params), want_zq=False, want_zh=False, group=dist.group.WORLD)
        assert int(res.valid.sum()) == hi - lo                                # every bin passes the depth filter
        np.testing.assert_array_equal(res.scale_q.cpu().numpy(), want_q)
        tie_rows = np.array([(2 * int(s)) % (2 * int(c)) == int(c) for s, c in zip(h.sum(1), (h > 0).sum(1))])
        assert tie_rows.sum() == ties
        np.testing.assert_array_equal(res.raw_mean.cpu().numpy()[tie_rows], want_mean[tie_rows])
        out.put((rank, "ok"))
    except Exception as exc:
        out.put((rank, f"{type(exc).__name__}: {str(exc)[:400]}"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,u16", [(1, False), (2, False), (4, False), (1, True), (2, True)])
def test_bin_sharded_scales_on_exact_ties(world, u16):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    _run(world, 38, 48, 1000, dict(min_depth=20.0, max_depth=100.0, u16=u16), worker=_worker_bins_ties)
