"""The row-sharded ring schedule on CPU: world_size 2 and 3 over gloo, with a numpy block kernel standing in for the
CUDA Gram/top-k kernels (which need a GPU).  What is under test is the host logic of grid_b200/ring.py: shard
layout, which shard is resident at which hop, global row offsets, double-buffered send/recv, uneven shards."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

from grid_b200 import ring  # noqa: E402


def test_shard_rows_and_ring_source():
    assert ring.shard_rows(10, 3) == ([4, 3, 3], [0, 4, 7])
    assert ring.shard_rows(3202, 8)[0] == [401, 401] + [400] * 6 and sum(ring.shard_rows(3202, 8)[0]) == 3202
    assert ring.shard_rows(2, 4) == ([1, 1, 0, 0], [0, 1, 2, 2])
    for world in (1, 2, 5, 8):
        for rank in range(world):
            seen = [ring.ring_source(rank, t, world) for t in range(world)]
            assert seen[0] == rank and sorted(seen) == list(range(world))       # every shard visits every rank once
            # what rank receives at hop t+1 is what rank-1 held at hop t
            assert all(ring.ring_source(rank, t + 1, world) == ring.ring_source((rank - 1) % world, t, world)
                       for t in range(world - 1))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_total, kdim, k, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(7)
        full = rng.integers(-200, 201, size=(n_total, kdim)).astype(np.int64)     # every rank can rebuild the cohort
        counts, offsets = ring.shard_rows(n_total, world)
        mine = full[offsets[rank]:offsets[rank] + counts[rank]]
        shard = torch.from_numpy(mine.astype(np.float32).copy())
        best = [[] for _ in range(counts[rank])]
        hops = []

        def block(step, src, flat, n_ref, r_gid0):
            hops.append((step, src, n_ref, r_gid0))
            ref = flat.view(n_ref, kdim).numpy().astype(np.int64)
            np.testing.assert_array_equal(ref, full[r_gid0:r_gid0 + n_ref])       # the right shard arrived intact
            d2 = ((mine[:, None, :] - ref[None, :, :]) ** 2).sum(axis=2)
            for i in range(counts[rank]):
                cand = best[i] + [(int(d2[i, j]), r_gid0 + j) for j in range(n_ref) if r_gid0 + j != offsets[rank] + i]
                best[i] = sorted(cand)[:k]

        ring.ring_knn(shard, max(counts), counts, offsets, dist.group.WORLD, block, flat_per_row=kdim)
        assert [h[0] for h in hops] == list(range(world))
        assert sorted(h[1] for h in hops) == list(range(world))
        # the union over hops equals the global brute-force answer for this rank's rows
        for i in range(counts[rank]):
            g = offsets[rank] + i
            dd = ((full[g][None, :] - full) ** 2).sum(axis=1)
            want = sorted((int(dd[j]), j) for j in range(n_total) if j != g)[:k]
            assert best[i] == want
        # unequal row blocks gather in rank order
        t = torch.arange(counts[rank], dtype=torch.int64).reshape(-1, 1) + 100 * rank
        got = ring._all_gather_rows(t, counts, dist.group.WORLD)
        want = torch.cat([torch.arange(c, dtype=torch.int64).reshape(-1, 1) + 100 * r for r, c in enumerate(counts)])
        assert torch.equal(got, want)
        out.put((rank, "ok"))
    except Exception as exc:  # surface the failure in the parent
        out.put((rank, f"{type(exc).__name__}: {exc}"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_total", [(2, 9), (3, 10)])
def test_ring_over_gloo_matches_global_knn(world, n_total):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, 48, 4, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(out.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {r: "ok" for r in range(world)}


# ---- query blocks x rounds x ring hops (cohorts larger than aggregate HBM) -------------------------------------------
def test_block_layout():
    rounds, blocks = ring.block_layout([10, 9, 9], 4)
    assert rounds == 3 and blocks[0] == ([4, 3, 3], [0, 4, 7]) and blocks[1] == ([3, 3, 3], [0, 3, 6])
    assert ring.block_layout([5, 5], 100)[0] == 1
    with pytest.raises(ValueError):
        ring.block_layout([40, 2], 10)          # a rank with fewer rows than rounds cannot take part in every round


def _worker_blocked(rank, world, port, n_total, kdim, k, block_rows, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(11)
        full = rng.integers(-200, 201, size=(n_total, kdim)).astype(np.int64)
        counts, offsets = ring.shard_rows(n_total, world)
        rounds, blocks = ring.block_layout(counts, block_rows)
        my_cnt, my_off = blocks[rank]
        best = [[] for _ in range(counts[rank])]
        state, calls, loads = {}, [], []

        def rows_of(r, bi):
            lo = offsets[r] + blocks[r][1][bi]
            return lo, lo + blocks[r][0][bi]

        def load_query(qi):
            state["qi"] = qi

        def load_ref(qi, ri):
            loads.append((qi, ri))
            lo, hi = rows_of(rank, ri)
            return torch.from_numpy(full[lo:hi].astype(np.float32).copy())      # "regenerated" on demand

        def block(qi, ri, step, src, flat, n_ref, r_gid0):
            assert qi == state["qi"]
            calls.append((qi, ri, step, src))
            lo, hi = rows_of(src, ri)
            assert (r_gid0, n_ref) == (lo, hi - lo)
            ref = flat.view(n_ref, kdim).numpy().astype(np.int64)
            np.testing.assert_array_equal(ref, full[lo:hi])                      # the right block arrived intact
            qlo, qhi = rows_of(rank, qi)
            d2 = ((full[qlo:qhi, None, :] - ref[None, :, :]) ** 2).sum(axis=2)
            for i in range(qhi - qlo):
                li = qlo - offsets[rank] + i
                cand = best[li] + [(int(d2[i, j]), r_gid0 + j) for j in range(n_ref) if r_gid0 + j != qlo + i]
                best[li] = sorted(cand)[:k]

        ring.ring_knn_blocked(rounds, blocks, offsets, dist.group.WORLD, load_query, load_ref, block, flat_per_row=kdim)
        assert loads == [(qi, ri) for qi in range(rounds) for ri in range(rounds)]        # one regeneration per round
        assert len(calls) == rounds * rounds * world and len(set(calls)) == len(calls)     # W Gram blocks per round
        for i in range(counts[rank]):
            g = offsets[rank] + i
            dd = ((full[g][None, :] - full) ** 2).sum(axis=1)
            assert best[i] == sorted((int(dd[j]), j) for j in range(n_total) if j != g)[:k]
        out.put((rank, "ok"))
    except Exception as exc:
        out.put((rank, f"{type(exc).__name__}: {exc}"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_total,block_rows", [(1, 11, 4), (2, 23, 5), (3, 31, 4)])
def test_blocked_ring_over_gloo_matches_global_knn(world, n_total, block_rows):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_blocked, args=(r, world, port, n_total, 32, 5, block_rows, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(out.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {r: "ok" for r in range(world)}


# ---- the HALF ring: every pair of samples computed once, lists for the peer's rows read off the block's transpose ------
INF = np.iinfo(np.int64).max


class _NumpyLists:
    """Stand-in for the CUDA k-best kernels: sorted (d2, idx) lists as torch CPU tensors (they travel through gloo)."""

    def __init__(self, k):
        self.k = k

    def new(self, m):
        return torch.full((m, self.k), INF, dtype=torch.int64), torch.full((m, self.k), -1, dtype=torch.int32)

    def merge_rows(self, d2_t, idx_t, rows, cand):
        """cand[i] = list of (d2, gid) for list row rows[i]."""
        for i, row in enumerate(rows):
            have = [(int(d), int(j)) for d, j in zip(d2_t[row].tolist(), idx_t[row].tolist()) if j >= 0]
            assert not {j for _, j in have} & {j for _, j in cand[i]}, "a pair of samples met twice"
            best = sorted(have + cand[i])[:self.k]
            d2_t[row] = torch.tensor([d for d, _ in best] + [INF] * (self.k - len(best)), dtype=torch.int64)
            idx_t[row] = torch.tensor([j for _, j in best] + [-1] * (self.k - len(best)), dtype=torch.int32)

    def merge_lists(self, dst_d2, dst_idx, lo, src_d2, src_idx):
        cand = [[(int(d), int(j)) for d, j in zip(src_d2[i].tolist(), src_idx[i].tolist()) if j >= 0]
                for i in range(src_d2.shape[0])]
        self.merge_rows(dst_d2, dst_idx, range(lo, lo + src_d2.shape[0]), cand)


def _block_numpy(L, full, q_gid0, nq_all, r_gid0, n_ref, q_range, r_range, best, t_lists, symmetric):
    qa, qz = q_range or (0, nq_all)
    ra, rz = r_range or (0, n_ref)
    q, r = full[q_gid0 + qa:q_gid0 + qz], full[r_gid0 + ra:r_gid0 + rz]
    d2 = ((q[:, None, :] - r[None, :, :]) ** 2).sum(axis=2)
    cand = [[(int(d2[i, j]), r_gid0 + ra + j) for j in range(rz - ra) if r_gid0 + ra + j != q_gid0 + qa + i]
            for i in range(qz - qa)]
    L.merge_rows(best[0], best[1], range(qa, qz), cand)
    if t_lists is not None:
        assert not symmetric
        cand_t = [[(int(d2[i, j]), q_gid0 + qa + i) for i in range(qz - qa)] for j in range(rz - ra)]
        L.merge_rows(t_lists[0], t_lists[1], range(ra, rz), cand_t)


def _worker_half(rank, world, port, n_total, kdim, k, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(17)
        full = rng.integers(-200, 201, size=(n_total, kdim)).astype(np.int64)
        full[3] = full[n_total - 2]                                                # an exact duplicate across ranks (ties)
        counts, offsets = ring.shard_rows(n_total, world)
        L = _NumpyLists(k)
        best = L.new(counts[rank])
        shard = torch.from_numpy(full[offsets[rank]:offsets[rank] + counts[rank]].astype(np.float32).copy())
        hops = []

        def block(step, src, flat, n_ref, r_gid0, q_range, r_range, t_lists):
            hops.append((step, src, q_range, r_range))
            np.testing.assert_array_equal(flat.view(n_ref, kdim).numpy().astype(np.int64), full[r_gid0:r_gid0 + n_ref])
            _block_numpy(L, full, offsets[rank], counts[rank], r_gid0, n_ref, q_range, r_range, best, t_lists, step == 0)

        ops = ring.HalfRingOps(L.new, lambda lo, d2, idx: L.merge_lists(best[0], best[1], lo, d2, idx))
        ring.ring_knn(shard, max(counts), counts, offsets, dist.group.WORLD, block, flat_per_row=kdim, half=ops)
        assert [h[0] for h in hops] == list(range(world // 2 + 1))                 # half the hops
        for i in range(counts[rank]):
            g = offsets[rank] + i
            dd = ((full[g][None, :] - full) ** 2).sum(axis=1)
            want = sorted((int(dd[j]), j) for j in range(n_total) if j != g)[:k]
            got = [(int(d), int(j)) for d, j in zip(best[0][i].tolist(), best[1][i].tolist()) if j >= 0]
            assert got == want, (rank, i)
        out.put((rank, "ok"))
    except Exception as exc:
        import traceback

        out.put((rank, f"{type(exc).__name__}: {exc} {traceback.format_exc()[-800:]}"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_total", [(1, 7), (2, 9), (3, 10), (4, 14), (5, 11)])
def test_half_ring_over_gloo_matches_global_knn(world, n_total):
    """Odd and even worlds (the even ones split the opposite rank's block), uneven shards, a shard of one row."""
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_half, args=(r, world, port, n_total, 24, 4, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(out.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {r: "ok" for r in range(world)}


def _worker_half_blocked(rank, world, port, n_total, kdim, k, block_rows, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(19)
        full = rng.integers(-200, 201, size=(n_total, kdim)).astype(np.int64)
        counts, offsets = ring.shard_rows(n_total, world)
        rounds, blocks = ring.block_layout(counts, block_rows)
        my_cnt, my_off = blocks[rank]
        L = _NumpyLists(k)
        best = L.new(counts[rank])                                                 # all my rows' lists stay resident
        state, calls, loads = {}, [], []

        def rows_of(r, bi):
            lo = offsets[r] + blocks[r][1][bi]
            return lo, lo + blocks[r][0][bi]

        def load_query(qi):
            state["qi"] = qi

        def load_ref(qi, ri):
            loads.append((qi, ri))
            lo, hi = rows_of(rank, ri)
            return torch.from_numpy(full[lo:hi].astype(np.float32).copy())

        def block(qi, ri, step, src, flat, n_ref, r_gid0, q_range, r_range, t_lists):
            assert qi == state["qi"]
            calls.append((qi, ri, step, src))
            lo, hi = rows_of(src, ri)
            assert (r_gid0, n_ref) == (lo, hi - lo)
            qlo, qhi = rows_of(rank, qi)
            view = (best[0][my_off[qi]:my_off[qi] + my_cnt[qi]], best[1][my_off[qi]:my_off[qi] + my_cnt[qi]])
            _block_numpy(L, full, qlo, qhi - qlo, r_gid0, n_ref, q_range, r_range, view, t_lists, step == 0 and qi == ri)

        ops = ring.HalfRingOps(L.new, lambda ri, lo, d2, idx: L.merge_lists(best[0], best[1], my_off[ri] + lo, d2, idx))
        ring.ring_knn_blocked(rounds, blocks, offsets, dist.group.WORLD, load_query, load_ref, block, flat_per_row=kdim,
                              half=ops)
        full_ring_blocks = rounds * rounds * world
        assert len(set(calls)) == len(calls) and len(calls) <= full_ring_blocks // 2 + rounds * rounds
        if world < 3:                                                              # lower-triangle rounds are not even regenerated
            assert loads == [(qi, ri) for qi in range(rounds) for ri in range(qi, rounds)]
        for i in range(counts[rank]):
            g = offsets[rank] + i
            dd = ((full[g][None, :] - full) ** 2).sum(axis=1)
            want = sorted((int(dd[j]), j) for j in range(n_total) if j != g)[:k]
            got = [(int(d), int(j)) for d, j in zip(best[0][i].tolist(), best[1][i].tolist()) if j >= 0]
            assert got == want, (rank, i)
        out.put((rank, "ok"))
    except Exception as exc:
        import traceback

        out.put((rank, f"{type(exc).__name__}: {exc} {traceback.format_exc()[-800:]}"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_total,block_rows", [(1, 11, 4), (2, 23, 5), (3, 31, 4), (4, 41, 4)])
def test_half_blocked_ring_over_gloo_matches_global_knn(world, n_total, block_rows):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_half_blocked, args=(r, world, port, n_total, 16, 5, block_rows, out))
             for r in range(world)]
    for p in procs:
        p.start()
    res = dict(out.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {r: "ok" for r in range(world)}
