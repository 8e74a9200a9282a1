"""oracle/cohort_check.py (the chunked fp64 checker bench.py uses at full size) against oracle.chain_arrays on
cohorts small enough for both: one process, and world_size 2 over gloo with both partitions."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

from grid_b200 import synth  # noqa: E402
from oracle import cohort_check as C  # noqa: E402
from oracle import grid_oracle as O  # noqa: E402


def _source(seed):
    return lambda r0, nr, c0, nc: torch.from_numpy(synth.depth_matrix(seed, nr, nc, r0, c0))


def _compare(out, want, qrows, counts, params, lo_hi=None):
    np.testing.assert_array_equal(out["sel"], want["valid"][want["sel"]])
    assert out["r_use"] == want["r_use"]
    np.testing.assert_array_equal(out["keep"], want["keep"])
    ties = set(out["tie_rows"].tolist())
    for i in range(len(want["scale_q"])):
        if i not in ties:
            assert out["scale_q"][i] == want["scale_q"][i]
    np.testing.assert_array_equal(out["idx"], want["nbr_idx"][qrows])
    np.testing.assert_array_equal(out["d2"], want["nbr_d2"][qrows])
    a, z = out["sel_local"]
    zc_full = np.zeros((len(qrows), len(want["sel"])), dtype=np.int64)
    zc_full[:, want["keep"]] = want["zc"][qrows]
    np.testing.assert_array_equal(out["qcodes"].numpy().astype(np.int64), zc_full[:, a:z])
    cn = C.dipcn_rows(counts, out["scale_q"], qrows, out["idx"], params["n_nbr"])
    np.testing.assert_allclose(cn, want["dipcn"][qrows], rtol=1e-12)


@pytest.mark.parametrize("seed,n,b,extra", [(41, 90, 1800, {}), (42, 61, 700, dict(top_frac=0.5, zmax=1.0)),
                                            (43, 40, 600, dict(min_depth=30.0, max_depth=38.0, sigma2_max=3.0))])
def test_single_process_matches_chain_arrays(seed, n, b, extra):
    params = dict(num_neighbors=25, n_nbr=10, **extra)
    counts = synth.read_counts(seed, n, [42599, 1000])
    want = O.chain_arrays(synth.depth_hundredths(seed, n, b) / 100.0, counts, **params)
    qrows = C.spread_rows(n, 3, 4)
    chk = C.CohortCheck(_source(seed), n, b, params, torch.device("cpu"), chunk_cells=n * 512)
    out = chk.run(qrows)
    assert out["stats"]["chunks"] >= 2
    _compare(out, want, qrows, counts, params)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, sharding, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        seed, n, b = 44, 75, 1500
        params = dict(num_neighbors=30, n_nbr=12)
        counts = synth.read_counts(seed, n, [42599])
        want = O.chain_arrays(synth.depth_hundredths(seed, n, b) / 100.0, counts, **params)
        qrows = C.spread_rows(n, world, 3)
        if sharding == "bins":
            per = -(-b // world)
            kw = dict(bins=(rank * per, min(b, (rank + 1) * per)))
        else:
            base = n // world
            kw = dict(rows=(rank * base, n if rank == world - 1 else (rank + 1) * base))
        chk = C.CohortCheck(_source(seed), n, b, params, torch.device("cpu"), group=dist.group.WORLD, sharding=sharding,
                            chunk_cells=n * 256, **kw)
        res = chk.run(qrows)
        _compare(res, want, qrows, counts, params)
        out.put((rank, "ok"))
    except Exception as exc:
        import traceback

        out.put((rank, f"{type(exc).__name__}: {exc}\n{traceback.format_exc()[-1500:]}"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("sharding", ["bins", "rows"])
def test_two_ranks_over_gloo(sharding):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, sharding, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(out.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {0: "ok", 1: "ok"}
