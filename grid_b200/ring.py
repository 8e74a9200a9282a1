"""Row-sharded multi-GPU schedule: every rank owns a contiguous block of SAMPLES (all bins) and the reference
shards travel around a ring (rank r -> r+1) while the Gram/top-k kernels work on the shard that has arrived.

This is the schedule for cohorts whose N x N Gram partials no longer fit (BASELINE configs 3/4); for N << bins
(configs 1/2) the bin-sharded path in engine.py moves far fewer bytes (DESIGN.md "Multi-GPU").

What crosses the ring / the collectives (torch.distributed; NCCL over NVLink on GPUs, gloo in the CPU tests):
  * per-bin raw sums and counts, per-bin normalised sums: all-reduce (B x 8..12 bytes each)
  * the clipped operand shards zh[kdim/64, n_r, 64] (fp16): W-1 ring hops, double-buffered so that hop t+1 is in
    flight while hop t is being multiplied; every rank keeps its own running k-best lists -> no merge collective
  * row norms, %.2f scales, read counts: all-gather (N x a few bytes)

The ring driver (`ring_knn`) is device-agnostic: the per-hop work is a callable, so the CPU tests drive the very
same schedule over gloo with a numpy block kernel.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_rows(n_total: int, world: int) -> tuple[list[int], list[int]]:
    """Balanced contiguous row blocks: (counts, offsets) per rank; the first n_total % world ranks get one more."""
    base, extra = divmod(n_total, world)
    counts = [base + (1 if r < extra else 0) for r in range(world)]
    offsets = [sum(counts[:r]) for r in range(world)]
    return counts, offsets


def ring_source(rank: int, step: int, world: int) -> int:
    """Rank whose shard is resident on `rank` at hop `step` (hop 0 = its own)."""
    return (rank - step) % world


class HalfRingOps:
    """What the half ring needs from the device besides the block kernel (CUDA: engine.knn_init / engine.merge_lists;
    the CPU tests pass numpy stand-ins).

    new_lists(nrows) -> (d2 int64 [nrows, k], idx int32 [nrows, k]), empty (d2 = +inf, idx = -1)
    merge(row_lo, d2, idx)    merge candidate lists for the rows row_lo.. of the CURRENT REFERENCE BLOCK of this rank
                              (= all its rows in the in-core ring) into its running lists
    before_hop(step)          optional: called at the start of every hop BEFORE the next shard's transfer is issued --
                              the place for any small collective the hop's block needs (a collective issued later
                              queues behind the multi-GB transfer on the communicator's stream and the block waits for
                              the whole shard: 23 ms per hop at 25 000 x 1e6 on 4 GPUs, 50 ms at config 3)
    """

    def __init__(self, new_lists, merge, before_hop=None):
        self.new_lists, self.merge, self.before_hop = new_lists, merge, before_hop


def half_ring_hops(world: int) -> int:
    """Last hop of the half ring: every unordered pair of ranks meets once within hops 0..world // 2."""
    return world // 2


def _tick(label: str) -> None:
    from . import engine as E

    E._tick(label)


def ring_knn(shard: torch.Tensor, shard_rows_max: int, counts: list[int], offsets: list[int], group, block_fn,
             flat_per_row: int, bufs=None, half: HalfRingOps | None = None, rel: int = 0) -> None:
    """Drive the ring.

    shard          this rank's reference shard, a contiguous tensor whose storage is `counts[rank] * flat_per_row`
                   elements (any layout the block kernel understands -- it is only ever moved, never indexed here)
    shard_rows_max max(counts): the two ring buffers are sized for the largest shard
    bufs           optional pair of flat receive buffers of at least shard_rows_max * flat_per_row elements

    half = None (the full ring): block_fn(step, src_rank, ref_shard_view, n_ref, r_gid0) is called for all `world` hops;
    every rank multiplies its rows with every shard, so each pair of samples is computed twice across the ranks.

    half given (the HALF ring, what the CUDA path runs): the squared distance is symmetric, so the block rank a computes
    against the shard of rank b also holds, read by columns, everything b needs to know about a's rows.  Hops run only
    to world // 2; on hop t >= 1 the block kernel is asked to ALSO fill k-best lists for the shard's rows
    (block_fn(step, src, view, n_ref, r_gid0, q_range, r_range, t_lists)), those lists (n_ref x k x 12 bytes instead of
    a 33 GB shard) go straight to the shard's owner, and the lists this rank's rows earned on rank + t arrive and are
    merged (`half.merge`).  With an even world the two ranks of a pair meet on the same (last) hop: the lower rank
    takes all its rows x the first half of the peer's, the higher one its second half x all of the peer's rows.
    rel: sign(reference block index - query block index) in the blocked schedule (0 in core): blocks of the same rank
    (hop 0) and of a rank's opposite (last hop of an even world) are computed only for rel >= 0 -- the mirror image is
    the peer's (or this rank's own) rel > 0 case -- and read both ways.  block_fn must not keep a reference past the
    following hop.
    """
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    cur = shard.reshape(-1)
    if half is None:
        if world == 1:
            block_fn(0, rank, cur, counts[rank], offsets[rank])
            return
        hops = world - 1
    else:
        hops = half_ring_hops(world)
        even = world % 2 == 0
        if rel < 0 and even:
            hops -= 1  # the opposite rank's block is its rel > 0 case
    if world > 1 and hops > 0 and bufs is None:  # (a caller that runs many rings passes its two buffers in)
        bufs = [torch.empty(shard_rows_max * flat_per_row, dtype=shard.dtype, device=shard.device) for _ in range(2)]
    nxt_rank = dist.get_global_rank(group, (rank + 1) % world) if world > 1 else rank
    prv_rank = dist.get_global_rank(group, (rank - 1) % world) if world > 1 else rank
    pending = []  # list exchanges in flight: (works, row_lo, recv_d2, recv_idx, keep-alive)
    for step in range(hops + 1):
        src = ring_source(rank, step, world)
        works = []
        if half is not None and half.before_hop is not None:
            half.before_hop(step)
        if step < hops:
            # pass the resident shard on and receive the next one into the other buffer WHILE this hop computes;
            # the collective library orders the transfer after everything already queued on the current stream,
            # i.e. after the previous hop's kernels have finished with the buffer being overwritten
            incoming = ring_source(rank, step + 1, world)
            recv_view = bufs[(step + 1) % 2][: counts[incoming] * flat_per_row]
            send_view = cur[: counts[src] * flat_per_row]
            ops = [dist.P2POp(dist.isend, send_view, nxt_rank, group), dist.P2POp(dist.irecv, recv_view, prv_rank, group)]
            works = dist.batch_isend_irecv(ops)
            _tick(f"hop {step}: shard transfer issued")
        view = cur[: counts[src] * flat_per_row]
        if half is None:
            block_fn(step, src, view, counts[src], offsets[src])
        elif step == 0:
            if rel == 0:
                block_fn(0, src, view, counts[src], offsets[src], None, None, None)       # own block: symmetric kernel
            elif rel > 0:  # another block of my own rows: both readings stay here
                t = half.new_lists(counts[src])
                block_fn(0, src, view, counts[src], offsets[src], None, None, t)
                half.merge(0, t[0], t[1])
        else:
            dst = (rank + step) % world                          # the rank that meets MY shard on this hop
            middle = even and step == world // 2
            q_range = r_range = None
            send_rows, recv_rows = (0, counts[src]), (0, counts[rank])
            if middle and rel == 0:
                if rank < src:   # lower rank of the pair: all my rows x the first half of the peer's
                    r_range = send_rows = (0, counts[src] // 2)
                else:            # higher rank: my second half x all of the peer's rows
                    q_range = (counts[rank] // 2, counts[rank])
                    recv_rows = (0, counts[rank] // 2)
            t = half.new_lists(counts[src])
            block_fn(step, src, view, counts[src], offsets[src], q_range, r_range, t)
            out = [x[send_rows[0]:send_rows[1]] for x in t]
            got = [x[recv_rows[0]:recv_rows[1]] for x in half.new_lists(counts[rank])]
            ops = []
            if send_rows[1] > send_rows[0]:
                ops += [dist.P2POp(dist.isend, x, dist.get_global_rank(group, src), group) for x in out]
            if recv_rows[1] > recv_rows[0]:
                ops += [dist.P2POp(dist.irecv, x, dist.get_global_rank(group, dst), group) for x in got]
            if ops:
                pending.append((dist.batch_isend_irecv(ops), recv_rows, got, out))
        _tick(f"hop {step}: block enqueued")
        for w in works:
            w.wait()
        _tick(f"hop {step}: next shard here")
        # lists that arrived for the PREVIOUS hop are merged now, behind this hop's kernels: nobody waits for a peer
        while len(pending) > (1 if step < hops else 0):
            lw, rows, got, _keep = pending.pop(0)
            for w in lw:
                w.wait()
            if rows[1] > rows[0]:
                half.merge(rows[0], got[0], got[1])
        if step < hops:
            cur = bufs[(step + 1) % 2]


def _all_reduce(t, group):
    dist.all_reduce(t, group=group)


def _all_gather_rows(t: torch.Tensor, counts: list[int], group) -> torch.Tensor:
    """Concatenate per-rank row blocks of unequal length (dim 0) on every rank."""
    world = dist.get_world_size(group)
    mx = max(counts)
    pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    pad[: t.shape[0]] = t
    out = torch.empty((world * mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    return torch.cat([out[r * mx: r * mx + counts[r]] for r in range(world)])


def knn_rows(zh: torch.Tensor, kdim: int, norms: torch.Tensor, counts: list[int], offsets: list[int], num_neighbors: int,
             vmax: int, group, impl: int = 0, extra: dict | None = None):
    """Exact k nearest neighbours of this rank's rows among ALL ranks' rows (self excluded) by the half ring.
    zh [kdim/64, n_local, 64]: this rank's k-blocked operand, norms [n_local] its exact row norms; counts/offsets: the
    row blocks of all ranks.  Returns (idx [n_local, k] int32 GLOBAL ids, d2 [n_local, k] int64), (d2, index) order."""
    from . import engine as E
    from ._lib import call

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n, dev = zh.shape[1], zh.device
    n_total = sum(counts)
    extra = extra if extra is not None else {}
    norms_all = _all_gather_rows(norms, counts, group)
    k = max(1, min(num_neighbors, n_total - 1)) if n_total > 1 else 1
    best_d2, best_idx = E.knn_init(n, k, dev)

    def block(step, src, flat, n_ref, r_gid0, q_range=None, r_range=None, t_lists=None):
        ref = zh if step == 0 else flat.view(kdim // 64, n_ref, 64)  # hop 0 is the symmetric (same-operand) case
        tau_r = None
        if t_lists is not None and tau_all[0] is not None:
            # the shard owner's current k-th best distances: what the fused epilogue filters the peer's candidates against
            tau_r = tau_all[0][r_gid0:r_gid0 + n_ref]
        E.knn_block_auto(zh, norms, offsets[rank], ref, norms_all[r_gid0:r_gid0 + n_ref], r_gid0, kdim, vmax, best_d2,
                         best_idx, impl=impl, q_range=q_range, r_range=r_range, t_lists=t_lists, tau_r=tau_r,
                         symmetric=step == 0)

    def merge(lo, d2_in, idx_in):
        with E._mark("merge_lists"):
            E.merge_lists(d2_in, idx_in, best_d2[lo:lo + d2_in.shape[0]], best_idx[lo:lo + d2_in.shape[0]])

    tau_all = [None]

    def before_hop(step):
        # every rank's current k-th best distances, gathered BEFORE this hop's shard transfer is queued (every rank is at
        # the same hop; a value that is a hop stale would only let a few more candidates through)
        tau_all[0] = None
        if step >= 1 and E.fused_mode() != "0":
            tau_all[0] = _all_gather_rows(best_d2[:, k - 1].contiguous(), counts, group)

    # The Gram kernel shares the GPU with NCCL's send/recv kernel while a transfer is in flight.  Its default
    # schedule (cluster launch control) adapts to the SMs it gets.  GRIDB200_RING_COMM_SMS=<n> instead selects the
    # static schedule on all but n SMs (A/B aid).  GRIDB200_RING=full runs all W hops (every pair twice: A/B aid).
    import os

    comm_sms = int(os.environ.get("GRIDB200_RING_COMM_SMS", "0")) if world > 1 else 0
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    full_ring = os.environ.get("GRIDB200_RING", "half") == "full"
    half = None if full_ring else HalfRingOps(lambda m: E.knn_init(m, k, dev), merge, before_hop)
    extra.update(ring_schedule="full" if full_ring else "half", ring_hops=(world - 1) if full_ring else half_ring_hops(world))
    call("gridb200_gram_set_max_pairs", max(1, (sms - comm_sms) // 2) if comm_sms else 0)
    try:
        ring_knn(zh, max(counts), counts, offsets, group, block, flat_per_row=kdim, half=half)
    finally:
        call("gridb200_gram_set_max_pairs", 0)
    return best_idx, best_d2


def run_chain_rows(depth: torch.Tensor, b: int, counts_local: torch.Tensor, p, group, impl: int = 0,
                   release_depth: bool = False):
    """normalize -> kNN -> dipCN for the row block this rank holds (depth [n_local, ld], counts_local [n_local, G]).

    release_depth: free the depth matrix's device memory as soon as step 4 has consumed it (the caller's tensor is left
    with empty storage).  The fp32 matrix is twice the size of the fp16 operand, and the ring needs room for two
    incoming shards: at BASELINE config 3 (6 250 samples x 2.9e6 bins per rank) that is 72 GB handed back in time.

    Returns (NormResult for the local rows, idx [n_local, k] GLOBAL sample ids, d2 [n_local, k], dipcn [n_local, G]).
    Same arithmetic and the same kernels as engine.run_chain; only the exchange steps differ.
    """
    from . import engine as E
    from ._lib import call, load, ptr

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n, ld = E.check_depth(depth, b)
    dev, st = depth.device, E._stream()
    i64, f64 = dict(dtype=torch.int64, device=dev), dict(dtype=torch.float64, device=dev)
    nloc = torch.zeros(world, **i64)
    nloc[rank] = n
    _all_reduce(nloc, group)
    counts = [int(x) for x in nloc.tolist()]
    offsets = [sum(counts[:r]) for r in range(world)]
    n_total = sum(counts)

    # ---- step 4: per-bin statistics are summed over ranks, everything per-sample is local ------------------
    colsum, colcnt = torch.empty(b, **i64), torch.empty(b, dtype=torch.int32, device=dev)
    valid = torch.empty(b, dtype=torch.uint8, device=dev)
    rowsum_all, rowcnt_all = torch.empty(n, **i64), torch.empty(n, dtype=torch.int32, device=dev)
    with E._mark("col_raw"):
        call("gridb200_col_row_raw_stats", ptr(depth), n, b, ld, ptr(colsum), ptr(colcnt), ptr(rowsum_all),
             ptr(rowcnt_all), st)
    _all_reduce(colsum, group)
    _all_reduce(colcnt, group)
    call("gridb200_depth_filter", ptr(colsum), ptr(colcnt), b, float(p.min_depth), float(p.max_depth), ptr(valid), st)
    rowsum, rowcnt = torch.empty(n, **i64), torch.empty(n, dtype=torch.int32, device=dev)
    ws0 = E._ws(load().gridb200_row_stats_fused_ws_bytes(b), dev)
    with E._mark("row_stats"):
        call("gridb200_row_stats_fused", ptr(depth), n, b, ld, ptr(valid), ptr(colcnt), ptr(rowsum_all), ptr(rowcnt_all),
             ptr(rowsum), ptr(rowcnt), ptr(ws0), ws0.numel(), st)
    raw_mean, scale_q, row_factor = torch.empty(n, **f64), torch.empty(n, **f64), torch.empty(n, **f64)
    call("gridb200_row_scales", ptr(rowsum), ptr(rowcnt), n, ptr(raw_mean), ptr(scale_q), ptr(row_factor), st)
    call("gridb200_row_means_exact", ptr(depth), n, b, ld, ptr(valid), ptr(rowsum), ptr(rowcnt), ptr(raw_mean),
         ptr(scale_q), st)  # rows are whole on their rank here
    sd, sd2 = torch.empty(b, **f64), torch.empty(b, **f64)
    ws = E._ws(load().gridb200_col_norm_stats_ws_bytes(n, b), dev)
    with E._mark("col_norm"):
        call("gridb200_col_norm_stats", ptr(depth), n, b, ld, ptr(valid), ptr(row_factor), ptr(sd), ptr(sd2), ptr(ws),
             ws.numel(), st)
    _all_reduce(sd, group)
    _all_reduce(sd2, group)
    n_eff_t = (rowcnt > 0).sum().to(torch.int64).reshape(1)
    _all_reduce(n_eff_t, group)
    n_eff = int(n_eff_t.item())
    mu, s2, ratio = torch.empty(b, **f64), torch.empty(b, **f64), torch.empty(b, **f64)
    call("gridb200_col_finalize", ptr(sd), ptr(sd2), ptr(colcnt), ptr(valid), b, n_eff, ptr(mu), ptr(s2), ptr(ratio), st)
    sel_pos = torch.empty(b, dtype=torch.int32, device=dev)
    keep = torch.empty(b, dtype=torch.uint8, device=dev)
    info, info2 = torch.empty(4, **f64), torch.empty(2, **f64)
    ws2 = E._ws(load().gridb200_select_ws_bytes(b), dev)
    with E._mark("select"):
        call("gridb200_select", ptr(ratio), b, float(p.top_frac), ptr(sel_pos), ptr(info), ptr(ws2), ws2.numel(), st)
        call("gridb200_knn_col_filter", ptr(ratio), ptr(sel_pos), b, float(p.frac_r), float(p.sigma2_max), ptr(keep),
             ptr(info), ptr(info2), ptr(ws2), ws2.numel(), st)
    host = torch.cat([info, info2]).cpu()
    zscale, rw, r_use = float(host[0]), int(host[3]), int(host[4])
    if rw < 0:
        raise IndexError("list index out of range")
    if rw == 0:
        raise ValueError("no bins selected")
    res = E.NormResult(n=n, b=b, colsum=colsum, colcnt=colcnt, valid=valid, raw_mean=raw_mean, scale_q=scale_q,
                       row_factor=row_factor, mu=mu, s2=s2, ratio=ratio, sel_pos=sel_pos, keep=keep, zscale=zscale,
                       rw=rw, rw_local=rw, r_use=r_use)
    res.extra.update(n_eff=n_eff, rowcnt=rowcnt, row_counts=counts, row_offsets=offsets)
    kdim = (rw + 63) // 64 * 64
    zh = torch.empty((kdim // 64, n, 64), dtype=torch.float16, device=dev)
    if kdim > rw:
        zh[-1, :, rw % 64:] = 0
    norms = torch.zeros(n, **i64)
    sat = torch.zeros(1, **i64)
    with E._mark("transform"):
        call("gridb200_transform", ptr(depth), n, b, ld, ptr(row_factor), ptr(mu), ptr(sel_pos), ptr(keep), ptr(info),
             p.zclip, None, 0, ptr(zh), n, ptr(norms), ptr(sat), st)
    res.zh, res.kdim, res.norms = zh, kdim, norms
    if release_depth:
        torch.cuda.current_stream(dev).synchronize()
        depth.untyped_storage().resize_(0)
        del depth

    # ---- step 5: ring of reference shards ------------------------------------------------------------------
    best_idx, best_d2 = knn_rows(zh, kdim, norms, counts, offsets, p.num_neighbors, p.zclip, group, impl=impl, extra=res.extra)

    # ---- step 6: scaled counts are tiny -> all-gather, then a local gather-reduce ----------------------------
    counts_all = _all_gather_rows(counts_local.contiguous(), counts, group)
    scale_all = _all_gather_rows(scale_q, counts, group)
    has = torch.ones(n_total, dtype=torch.uint8, device=dev)
    cn = E.dipcn(counts_all, has, scale_all, best_idx, p.n_nbr, q_gid0=offsets[rank])
    return res, best_idx, best_d2, cn


# ---- cohorts larger than aggregate HBM (BASELINE config 4): query blocks x ring rounds -------------------------------
#
# With N = 5e5 samples x 2.6e6 selected bins the fp16 operand is 2.6 TB against 8 x 180 GB, so no rank can even hold
# the operand of its own rows.  Every rank therefore cuts its rows into `rounds` blocks that are REGENERATED from the
# source (a synthetic generator, or a loader of mosdepth files) whenever they are needed:
#
#   for each query block qi of mine (operand resident):              rounds
#       for each round ri:                                           rounds
#           transform my block ri (source -> depth -> fp16 operand), send it round the ring;     1 regeneration
#           on every hop multiply the resident query block with the block that has arrived       W Gram blocks
#
# Regeneration costs one source + transform pass per W Gram blocks, the ring moves each block W - 1 times over NVLink
# while the Gram kernel (10x longer than the transfer at these shapes) runs, and the per-bin statistics of step 4 are
# accumulated over the blocks in two streaming passes.  Nothing is ever resident but one query block, one reference
# block, the two ring buffers and one fp32 source block.
def block_layout(counts: list[int], block_rows: int):
    """(rounds, per-rank (block counts, block offsets)) -- every rank cuts its rows into the same number of blocks."""
    rounds = max(1, -(-max(counts) // max(1, int(block_rows))))
    if min(counts) < rounds:
        raise ValueError(f"a rank holds fewer rows ({min(counts)}) than there are rounds ({rounds})")
    return rounds, [shard_rows(c, rounds) for c in counts]


def ring_knn_blocked(rounds: int, blocks, rank_offsets: list[int], group, load_query, load_ref, block_fn,
                     flat_per_row: int, dtype=None, device=None, half: HalfRingOps | None = None) -> None:
    """The schedule above, device-agnostic (the CPU tests drive it over gloo with a numpy block kernel).

    blocks[r] = (counts, offsets) of rank r's blocks within its own rows; rank_offsets[r] = its first global row.
    load_query(qi)                     make query block qi resident
    load_ref(qi, ri) -> flat tensor    this rank's block ri in ring layout (counts[ri] * flat_per_row elements)
    block_fn(qi, ri, step, src, flat, n_ref, r_gid0[, q_range, r_range, t_lists])   as ring_knn's, told which query
                                       block and round it serves
    half: run every (query block, round) as a HALF ring (see ring_knn): block pairs are computed once -- (qi, ri) and
          (ri, qi) meet the same pairs of samples from both sides, so blocks of the same rank and of opposite ranks are
          taken for ri >= qi only -- and read both ways.  `half.merge(ri, row_lo, d2, idx)` gets the round as well.
    """
    world = dist.get_world_size(group)
    biggest = max(max(b[0]) for b in blocks)
    bufs = None
    for qi in range(rounds):
        load_query(qi)
        for ri in range(rounds):
            rel = (ri > qi) - (ri < qi)
            if half is not None and rel < 0 and world < 3:
                continue  # hop 0 and the opposite rank's hop are somebody's rel > 0 case; no other hop exists
            shard = load_ref(qi, ri)
            if bufs is None and world > 1:
                bufs = [torch.empty(biggest * flat_per_row, dtype=shard.dtype, device=shard.device) for _ in range(2)]
            cnts = [blocks[r][0][ri] for r in range(world)]
            offs = [rank_offsets[r] + blocks[r][1][ri] for r in range(world)]
            ops = None if half is None else HalfRingOps(half.new_lists, lambda lo, d2, idx, ri=ri: half.merge(ri, lo, d2, idx))
            ring_knn(shard, max(cnts), cnts, offs, group,
                     lambda step, src, flat, n_ref, r_gid0, *more, qi=qi, ri=ri: block_fn(qi, ri, step, src, flat, n_ref,
                                                                                        r_gid0, *more),
                     flat_per_row, bufs=bufs, half=ops, rel=rel)


def block_operand(res, depth: torch.Tensor, row0: int, p, out: torch.Tensor | None = None):
    """(zh [kdim/64, nr, 64] fp16, norms [nr] int64) of the local rows row0.. whose depth block is `depth`, with the
    cohort statistics held in `res` (a NormResult of run_chain_rows_blocked).  out: optional flat fp16 buffer of at
    least nr * kdim elements to build the operand in."""
    from . import engine as E
    from ._lib import call, ptr

    nr, ld = E.check_depth(depth, res.b)
    dev = depth.device
    kdim = (res.rw + 63) // 64 * 64
    if out is None:
        zh = torch.empty((kdim // 64, nr, 64), dtype=torch.float16, device=dev)
    else:
        assert out.dtype == torch.float16 and out.is_contiguous() and out.numel() >= nr * kdim
        zh = out.view(-1)[: nr * kdim].view(kdim // 64, nr, 64)
    if kdim > res.rw:
        zh[-1, :, res.rw % 64:] = 0
    norms = torch.zeros(nr, dtype=torch.int64, device=dev)
    sat = torch.zeros(1, dtype=torch.int64, device=dev)
    with E._mark("transform"):
        call("gridb200_transform", ptr(depth), nr, res.b, ld, ptr(res.row_factor[row0:row0 + nr]), ptr(res.mu),
             ptr(res.sel_pos), ptr(res.keep), ptr(res.extra["info"]), p.zclip, None, 0, ptr(zh), nr, ptr(norms), ptr(sat),
             E._stream())
    return zh, norms


def run_chain_rows_blocked(source, n_local: int, b: int, counts_local: torch.Tensor, p, group, block_rows: int,
                           impl: int = 0, resident_blocks: int = 0):
    """normalize -> kNN -> dipCN for the n_local rows this rank owns WITHOUT holding them: `source(row0, nrows)` returns
    the fp32 depth block [nrows, ld] of local rows row0.. (on the device) each time it is asked.

    Same kernels and arithmetic as run_chain_rows; per-bin sums are accumulated block by block (fp64 partials are added
    in block order, then all-reduced).  Returns (NormResult without operand, idx [n_local, k] global ids, d2, dipcn).

    resident_blocks: keep the operands of my first m blocks in HBM once they have been made (they are asked for once per
    query block; a cyclic access pattern, so pinning a prefix beats any replacement policy).  m = rounds is the in-core
    schedule of run_chain_rows; 0 regenerates everything.
    """
    from . import engine as E
    from ._lib import call, load, ptr

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev, st = counts_local.device, E._stream()
    n = int(n_local)
    i64, f64 = dict(dtype=torch.int64, device=dev), dict(dtype=torch.float64, device=dev)
    i32 = dict(dtype=torch.int32, device=dev)
    nloc = torch.zeros(world, **i64)
    nloc[rank] = n
    _all_reduce(nloc, group)
    counts = [int(x) for x in nloc.tolist()]
    offsets = [sum(counts[:r]) for r in range(world)]
    n_total = sum(counts)
    rounds, blocks = block_layout(counts, block_rows)
    my_cnt, my_off = blocks[rank]

    def fetch(bi):
        with E._mark("source"):
            depth = source(my_off[bi], my_cnt[bi])
        nr, ld = E.check_depth(depth, b)
        if nr != my_cnt[bi]:
            raise ValueError(f"source returned {nr} rows for a block of {my_cnt[bi]}")
        return depth, nr, ld

    # ---- step 4, pass 1: per-bin raw sums (-> depth filter) and per-sample totals --------------------------------
    colsum, colcnt = torch.zeros(b, **i64), torch.zeros(b, **i32)
    t_sum, t_cnt = torch.empty(b, **i64), torch.empty(b, **i32)
    rowsum_all, rowcnt_all = torch.empty(n, **i64), torch.empty(n, **i32)
    for bi in range(rounds):
        depth, nr, ld = fetch(bi)
        r0 = my_off[bi]
        with E._mark("col_raw"):
            call("gridb200_col_row_raw_stats", ptr(depth), nr, b, ld, ptr(t_sum), ptr(t_cnt), ptr(rowsum_all[r0:r0 + nr]),
                 ptr(rowcnt_all[r0:r0 + nr]), st)
        colsum += t_sum
        colcnt += t_cnt
        del depth
    _all_reduce(colsum, group)
    _all_reduce(colcnt, group)
    valid = torch.empty(b, dtype=torch.uint8, device=dev)
    call("gridb200_depth_filter", ptr(colsum), ptr(colcnt), b, float(p.min_depth), float(p.max_depth), ptr(valid), st)

    # ---- pass 2: per-sample means over the valid bins, per-bin normalised sums ----------------------------------
    rowsum, rowcnt = torch.empty(n, **i64), torch.empty(n, **i32)
    raw_mean, scale_q, row_factor = torch.empty(n, **f64), torch.empty(n, **f64), torch.empty(n, **f64)
    sd, sd2 = torch.zeros(b, **f64), torch.zeros(b, **f64)
    t_sd, t_sd2 = torch.empty(b, **f64), torch.empty(b, **f64)
    ws0 = E._ws(load().gridb200_row_stats_fused_ws_bytes(b), dev)
    for bi in range(rounds):
        depth, nr, ld = fetch(bi)
        sl = slice(my_off[bi], my_off[bi] + nr)
        with E._mark("row_stats"):
            call("gridb200_row_stats_fused", ptr(depth), nr, b, ld, ptr(valid), ptr(colcnt), ptr(rowsum_all[sl]),
                 ptr(rowcnt_all[sl]), ptr(rowsum[sl]), ptr(rowcnt[sl]), ptr(ws0), ws0.numel(), st)
        call("gridb200_row_scales", ptr(rowsum[sl]), ptr(rowcnt[sl]), nr, ptr(raw_mean[sl]), ptr(scale_q[sl]),
             ptr(row_factor[sl]), st)
        call("gridb200_row_means_exact", ptr(depth), nr, b, ld, ptr(valid), ptr(rowsum[sl]), ptr(rowcnt[sl]),
             ptr(raw_mean[sl]), ptr(scale_q[sl]), st)
        ws = E._ws(load().gridb200_col_norm_stats_ws_bytes(nr, b), dev)
        with E._mark("col_norm"):
            call("gridb200_col_norm_stats", ptr(depth), nr, b, ld, ptr(valid), ptr(row_factor[sl]), ptr(t_sd), ptr(t_sd2),
                 ptr(ws), ws.numel(), st)
        sd += torch.nan_to_num(t_sd, nan=0.0)      # (bins outside `valid` carry no sum)
        sd2 += torch.nan_to_num(t_sd2, nan=0.0)
        del depth
    _all_reduce(sd, group)
    _all_reduce(sd2, group)
    n_eff_t = (rowcnt > 0).sum().to(torch.int64).reshape(1)
    _all_reduce(n_eff_t, group)
    n_eff = int(n_eff_t.item())
    mu, s2, ratio = torch.empty(b, **f64), torch.empty(b, **f64), torch.empty(b, **f64)
    call("gridb200_col_finalize", ptr(sd), ptr(sd2), ptr(colcnt), ptr(valid), b, n_eff, ptr(mu), ptr(s2), ptr(ratio), st)
    sel_pos = torch.empty(b, **i32)
    keep = torch.empty(b, dtype=torch.uint8, device=dev)
    info, info2 = torch.empty(4, **f64), torch.empty(2, **f64)
    ws2 = E._ws(load().gridb200_select_ws_bytes(b), dev)
    with E._mark("select"):
        call("gridb200_select", ptr(ratio), b, float(p.top_frac), ptr(sel_pos), ptr(info), ptr(ws2), ws2.numel(), st)
        call("gridb200_knn_col_filter", ptr(ratio), ptr(sel_pos), b, float(p.frac_r), float(p.sigma2_max), ptr(keep),
             ptr(info), ptr(info2), ptr(ws2), ws2.numel(), st)
    host = torch.cat([info, info2]).cpu()
    zscale, rw, r_use = float(host[0]), int(host[3]), int(host[4])
    if rw < 0:
        raise IndexError("list index out of range")
    if rw == 0:
        raise ValueError("no bins selected")
    res = E.NormResult(n=n, b=b, colsum=colsum, colcnt=colcnt, valid=valid, raw_mean=raw_mean, scale_q=scale_q,
                       row_factor=row_factor, mu=mu, s2=s2, ratio=ratio, sel_pos=sel_pos, keep=keep, zscale=zscale,
                       rw=rw, rw_local=rw, r_use=r_use)
    res.extra.update(n_eff=n_eff, rowcnt=rowcnt, row_counts=counts, row_offsets=offsets, rounds=rounds, blocks=blocks)
    res.extra["info"] = info
    kdim = (rw + 63) // 64 * 64

    pinned = {}
    # two reusable operand buffers (query block, reference block) sized for my largest block: blocks differ by a row,
    # and tens of GB allocated afresh at a new size every round is how a caching allocator runs out of room
    slots = {}

    def operand(bi, slot):
        """fp16 k-blocked operand and exact row norms of my block bi, from a fresh copy of its source rows."""
        if bi in pinned:
            return pinned[bi]
        depth, nr, ld = fetch(bi)
        if bi < resident_blocks:
            pinned[bi] = block_operand(res, depth, my_off[bi], p)
            return pinned[bi]
        if slot not in slots:
            slots[slot] = torch.empty(max(my_cnt) * kdim, dtype=torch.float16, device=dev)
        return block_operand(res, depth, my_off[bi], p, out=slots[slot])

    # ---- step 5: query blocks x rounds x ring hops -----------------------------------------------------------------
    # The running lists of ALL my rows stay resident (n x k x 12 bytes): the half ring merges candidates into the rows of
    # the reference block as well as into those of the query block.
    import os

    k = max(1, min(p.num_neighbors, n_total - 1)) if n_total > 1 else 1
    d2_all, idx_all = E.knn_init(n, k, dev)
    cur = {}
    full_ring = os.environ.get("GRIDB200_RING", "half") == "full"

    def rows_of(bi):
        return slice(my_off[bi], my_off[bi] + my_cnt[bi])

    def load_query(qi):
        cur.clear()
        zq, nq = operand(qi, "query")
        cur.update(qi=qi, zq=zq, nq=nq, best_d2=d2_all[rows_of(qi)], best_idx=idx_all[rows_of(qi)])

    def load_ref(qi, ri):
        cur["zr"], nr_ = (cur["zq"], cur["nq"]) if ri == qi else operand(ri, "ref")   # own block: the symmetric Gram case
        cnts = [blocks[r][0][ri] for r in range(world)]
        cur["norms_round"] = _all_gather_rows(nr_, cnts, group) if world > 1 else nr_
        cur["round_off"] = [sum(cnts[:r]) for r in range(world)]
        # current k-th best distances of every rank's block ri: the fused epilogue's filter for the reference rows
        tau = d2_all[rows_of(ri), k - 1].contiguous()
        cur["tau_round"] = (_all_gather_rows(tau, cnts, group) if world > 1 else tau) if E.fused_mode() != "0" else None
        return cur["zr"]

    def block_fn(qi, ri, step, src, flat, n_ref, r_gid0, q_range=None, r_range=None, t_lists=None):
        ref = cur["zr"] if step == 0 else flat.view(kdim // 64, n_ref, 64)
        nr_ = cur["norms_round"][cur["round_off"][src]:cur["round_off"][src] + n_ref]
        tau_r = None
        if cur["tau_round"] is not None:
            tau_r = cur["tau_round"][cur["round_off"][src]:cur["round_off"][src] + n_ref]
        E.knn_block_auto(cur["zq"], cur["nq"], offsets[rank] + my_off[qi], ref, nr_, r_gid0, kdim, p.zclip, cur["best_d2"],
                         cur["best_idx"], impl=impl, q_range=q_range, r_range=r_range, t_lists=t_lists, tau_r=tau_r,
                         symmetric=step == 0 and qi == ri)

    def merge(ri, lo, d2_in, idx_in):
        a0 = my_off[ri] + lo
        with E._mark("merge_lists"):
            E.merge_lists(d2_in, idx_in, d2_all[a0:a0 + d2_in.shape[0]], idx_all[a0:a0 + d2_in.shape[0]])

    half = None if full_ring else HalfRingOps(lambda m: E.knn_init(m, k, dev), merge)
    res.extra.update(ring_schedule="full" if full_ring else "half")
    ring_knn_blocked(rounds, blocks, offsets, group, load_query, load_ref, block_fn, flat_per_row=kdim, half=half)
    cur.clear()

    # ---- step 6 --------------------------------------------------------------------------------------------------------
    counts_all = _all_gather_rows(counts_local.contiguous(), counts, group)
    scale_all = _all_gather_rows(scale_q, counts, group)
    has = torch.ones(n_total, dtype=torch.uint8, device=dev)
    cn = E.dipcn(counts_all, has, scale_all, idx_all, p.n_nbr, q_gid0=offsets[rank])
    return res, idx_all, d2_all, cn
