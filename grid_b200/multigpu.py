"""Several GPUs behind the drop-in step functions: the additive config key

    gpu:
      devices: [0, 1, 2, 3]        # or "all"; absent / one entry = the single current device, exactly as before
      sharding: auto | bins | rows # step 5's partition: bins = every rank all samples x a bin slice (N << bins),
                                   # rows = a sample block per rank + the half ring (large N); auto picks by N

`grid wgs` runs the three steps inside ONE Python process (pipeline.py:66-93).  A step that is asked for several
devices spawns one worker process per device (torch.multiprocessing, spawn), joins them in an NCCL group over
127.0.0.1, hands them its host-resident inputs through shared memory, and gets rank 0's (or every rank's) pieces back
through a queue; the files are written by the calling process, so the on-disk contract is untouched.  The workers run the
very schedules bench.py measures: engine.normalize / engine.knn with `group` (bin-sharded) and ring.knn_rows (row-sharded).
"""
from __future__ import annotations

import os
import socket
import traceback

import numpy as np
import torch


def configured_devices(config) -> list[int] | None:
    """Device indices named by config['gpu']['devices'], or None when the step should run on the current device."""
    gpu = (config or {}).get("gpu") or {}
    devs = gpu.get("devices")
    if devs is None:
        return None
    if isinstance(devs, str):
        if devs != "all":
            raise ValueError(f"gpu.devices must be a list of device indices or 'all', not {devs!r}")
        devs = list(range(torch.cuda.device_count()))
    devs = [int(d) for d in devs]
    if len(set(devs)) != len(devs) or any(d < 0 or d >= torch.cuda.device_count() for d in devs):
        raise ValueError(f"gpu.devices = {devs}: indices must be distinct and below the {torch.cuda.device_count()} visible devices")
    return devs if len(devs) > 1 else None


def knn_sharding(config, n: int) -> str:
    mode = ((config or {}).get("gpu") or {}).get("sharding", "auto")
    if mode not in ("auto", "bins", "rows"):
        raise ValueError(f"gpu.sharding must be auto, bins or rows, not {mode!r}")
    if mode == "auto":  # the bin-sharded exchange is an N x N int64 block per rank: fine while that is a few hundred MB
        return "bins" if n <= 12000 else "rows"
    return mode


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, devices, port, worker_name, payload, out):
    import torch.distributed as dist

    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        torch.cuda.set_device(devices[rank])
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", devices[rank]))
        try:
            res = globals()[worker_name](rank, world, dist.group.WORLD, payload)
            torch.cuda.synchronize()
            out.put((rank, "ok", res))
        finally:
            dist.destroy_process_group()
    except Exception as exc:  # surfaced in the parent with the rank and the traceback
        out.put((rank, "error", f"{type(exc).__name__}: {exc}\n{traceback.format_exc()[-2000:]}"))


def run_workers(worker_name: str, devices: list[int], payload: dict, timeout: float = 24 * 3600.0) -> list:
    """Run globals()[worker_name](rank, world, group, payload) on every device; returns the per-rank results in rank
    order.  CPU tensors inside `payload` are moved to shared memory first (the workers see them without a copy)."""
    import torch.multiprocessing as mp

    for v in payload.values():
        if isinstance(v, torch.Tensor) and not v.is_cuda:
            v.share_memory_()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    world = len(devices)
    procs = [ctx.Process(target=_entry, args=(r, world, devices, port, worker_name, payload, out)) for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    try:
        for _ in procs:
            rank, status, res = out.get(timeout=timeout)
            if status != "ok":
                raise RuntimeError(f"multi-GPU worker {rank} (cuda:{devices[rank]}) failed: {res}")
            got[rank] = res
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()
    return [got[r] for r in range(world)]


def bin_slices(b: int, world: int, quantum: int = 64) -> list[tuple[int, int]]:
    per = -(-(-(-b // world)) // quantum) * quantum
    return [(min(b, r * per), min(b, (r + 1) * per)) for r in range(world)]


# ---- step 4 ----------------------------------------------------------------------------------------------------------
def _w_normalize(rank, world, group, payload):
    """Bin-sharded normalize_matrix + selection + codes: this rank's bin slice of the packed fp32 matrix."""
    from . import engine

    depth_host, b, p = payload["depth"], payload["b"], payload["params"]
    lo, hi = bin_slices(b, world, 4)[rank]
    nb = hi - lo
    ld = (max(nb, 1) + 3) // 4 * 4
    dev = torch.device("cuda", torch.cuda.current_device())
    depth = torch.zeros((depth_host.shape[0], ld), dtype=torch.float32, device=dev)
    if nb:
        depth[:, :nb] = depth_host[:, lo:hi].to(dev)
    sizes = [z - a for a, z in bin_slices(b, world, 4)]
    res = engine.normalize(depth, max(nb, 1), p, want_zq=True, want_zh=False, group=group, bin_sizes=sizes) if nb else None
    if res is None:
        raise ValueError(f"gpu.devices: {world} devices for {b} bins leaves a rank without bins; use fewer devices")
    sel = torch.nonzero(res.sel_pos >= 0).flatten()
    piece = {"lo": lo, "rw": res.rw, "rw_local": res.rw_local, "sel": (sel + lo).cpu().numpy(),
             "mu": res.mu[sel].cpu().numpy(), "ratio": res.ratio[sel].cpu().numpy(),
             "zq": res.zq[:, :res.rw_local].cpu() if res.zq is not None else torch.zeros((res.n, 0), dtype=torch.int16)}
    if rank == 0:
        piece.update(rowcnt=res.extra["rowcnt"].cpu().numpy(), raw_mean=res.raw_mean.cpu().numpy())
    return piece


def normalize(depth_host: torch.Tensor, b: int, params, devices: list[int]):
    """Step 4 over several GPUs.  depth_host: fp32 [N, >= b] on the host.  Returns (zq int16 [N, Rw] host tensor,
    rw, mu_sel, ratio_sel, raw_mean, rowcnt) -- what the single-GPU path hands the writer."""
    pieces = run_workers("_w_normalize", devices, {"depth": depth_host, "b": int(b), "params": params})
    rw = pieces[0]["rw"]
    if rw < 0:
        raise IndexError("list index out of range")
    assert sum(p["rw_local"] for p in pieces) == rw
    zq = torch.cat([p["zq"] for p in pieces], dim=1) if rw else torch.zeros((depth_host.shape[0], 0), dtype=torch.int16)
    return (zq, rw, np.concatenate([p["mu"] for p in pieces]), np.concatenate([p["ratio"] for p in pieces]),
            pieces[0]["raw_mean"], pieces[0]["rowcnt"])


# ---- step 5 ----------------------------------------------------------------------------------------------------------
def _w_knn(rank, world, group, payload):
    import torch.distributed as dist

    from . import engine, ring

    zq_host, keep_host, k, zclip, mode = payload["zq"], payload["keep"], payload["k"], payload["zclip"], payload["sharding"]
    n, rw = zq_host.shape
    dev = torch.device("cuda", torch.cuda.current_device())
    if mode == "bins":
        lo, hi = bin_slices(rw, world)[rank]
        if hi > lo:
            zq = zq_host[:, lo:hi].contiguous().to(dev)
            zh, kdim, norms = engine.prepare_from_zq(zq, hi - lo, keep_host[lo:hi].to(dev), zclip)
        else:  # more devices than 64-bin blocks: a zero partial
            zh, kdim = torch.zeros((1, n, 64), dtype=torch.float16, device=dev), 64
            norms = torch.zeros(n, dtype=torch.int64, device=dev)
        dist.all_reduce(norms, group=group)
        idx, d2 = engine.knn(zh, kdim, norms, k, zclip, group=group)
        return (idx.cpu(), d2.cpu()) if rank == 0 else None
    cnt, off = ring.shard_rows(n, world)
    if min(cnt) == 0:
        raise ValueError(f"gpu.devices: {world} devices for {n} samples leaves a rank without rows; use fewer devices")
    zq = zq_host[off[rank]:off[rank] + cnt[rank]].contiguous().to(dev)
    zh, kdim, norms = engine.prepare_from_zq(zq, rw, keep_host.to(dev), zclip)
    idx, d2 = ring.knn_rows(zh, kdim, norms, cnt, off, k, zclip, group)
    return idx.cpu(), d2.cpu()


def knn(zq_host: torch.Tensor, keep_host: torch.Tensor, k: int, zclip: int, devices: list[int], sharding: str):
    """Step 5's search over several GPUs from the hundredths matrix of a step-4 file (host int16 [N, Rw], keep uint8
    [Rw]).  Returns (idx int32 [N, k'], d2 int64 [N, k']) on the host, global sample ids, (d2, index) order."""
    res = run_workers("_w_knn", devices, {"zq": zq_host, "keep": keep_host, "k": int(k), "zclip": int(zclip),
                                          "sharding": sharding})
    if sharding == "bins":
        return res[0]
    return torch.cat([r[0] for r in res]), torch.cat([r[1] for r in res])
