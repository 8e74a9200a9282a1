"""Device pipeline: steps 4-6 on torch CUDA tensors through the C ABI (include/gridb200.h).

torch is plumbing here (device memory, streams, torch.distributed); every arithmetic step is a libgridb200
kernel.  Multi-GPU (``group`` given): the cohort is sharded by BINS -- every rank holds all N samples and a
contiguous slice of the bins -- because for N << bins the only cross-rank data are N-vectors (row sums, row
norms), the ratio vector, and one N x N partial Gram matrix; see DESIGN.md "Multi-GPU".
"""

from __future__ import annotations

from dataclasses import dataclass, field

import torch

from . import _lib
from ._lib import call, ptr


@dataclass
class Params:
    """Config knobs of the three steps with the reference's in-function defaults
    (normalize_mosdepth.py:49-51, find_neighbors.py:31-34, compute_dipcn.py:31)."""

    min_depth: float = 20.0
    max_depth: float = 100.0
    top_frac: float = 0.1
    zmax: float = 2.0
    sigma2_max: float = 1000.0
    frac_r: float = 1.0
    num_neighbors: int = 500
    n_nbr: int = 300

    @property
    def zclip(self) -> int:
        return int(round(self.zmax * 100))


@dataclass
class NormResult:
    n: int
    b: int
    colsum: torch.Tensor
    colcnt: torch.Tensor
    valid: torch.Tensor
    raw_mean: torch.Tensor
    scale_q: torch.Tensor
    row_factor: torch.Tensor
    mu: torch.Tensor
    s2: torch.Tensor
    ratio: torch.Tensor
    sel_pos: torch.Tensor       # local compact position or -1, per local bin
    keep: torch.Tensor
    zscale: float
    rw: int                     # bins selected in the whole cohort
    rw_local: int               # ... of which on this rank
    r_use: int
    zq: torch.Tensor | None = None
    zh: torch.Tensor | None = None
    kdim: int = 0               # padded local operand width (multiple of 64)
    norms: torch.Tensor | None = None
    saturated: int = 0
    extra: dict = field(default_factory=dict)


PROFILE: dict | None = None  # bench.py sets this to {} to collect per-stage CUDA-event timings
GRAM_ELEMS: int = 0          # output elements of the Gram tiles launched so far (flops = 2 * K * elements); bench.py resets it


def gram_tile_elems(nq: int, nr: int, symmetric: bool) -> int:
    """Elements of the 256 x 256 output tiles gridb200_gram computes for an nq x nr block, clipped to the block:
    all of them, or -- same operand on both sides -- only the tiles on or above the diagonal."""
    if not symmetric:
        return nq * nr
    sizes = [min(256, nq - e) for e in range(0, nq, 256)]
    return sum(sizes[i] * sizes[j] for i in range(len(sizes)) for j in range(i, len(sizes)))


class _mark:
    """Record CUDA events around a stage on the current stream when PROFILE is enabled (zero cost otherwise)."""

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        _tick(self.name + " >")
        if PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()

    def __exit__(self, *exc):
        _tick(self.name + " <")
        if PROFILE is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            PROFILE.setdefault(self.name, []).append((self.e0, e1))


TIMELINE: list | None = None  # bench.py --timeline: [(label, CUDA event, host perf_counter)] in program order


def _tick(label: str) -> None:
    """A point of the step's timeline (device event + host clock) when TIMELINE is enabled; free otherwise."""
    if TIMELINE is not None:
        import time

        e = torch.cuda.Event(enable_timing=True)
        e.record()
        TIMELINE.append((label, e, time.perf_counter()))


def timeline_summary() -> list:
    """[(label, device ms since the first tick, host ms since the first tick)] (call after a synchronize)."""
    if not TIMELINE:
        return []
    _, e0, h0 = TIMELINE[0]
    return [(lab, round(e0.elapsed_time(e), 3), round((h - h0) * 1e3, 3)) for lab, e, h in TIMELINE]


def profile_summary(steps: int = 0) -> dict:
    """{stage: {"ms": mean per step, "calls": n}} from the recorded events (call after a synchronize)."""
    out = {}
    for name, evs in (PROFILE or {}).items():
        tot = sum(a.elapsed_time(b) for a, b in evs)
        out[name] = {"ms": tot / (steps or len(evs)), "calls": len(evs)}
    return out


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ws(nbytes: int, dev) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)


def _all_reduce(t: torch.Tensor, group) -> None:
    if group is not None:
        import torch.distributed as dist

        dist.all_reduce(t, group=group)


@dataclass
class Depth16:
    """The depth matrix in 2 bytes per cell on the device: codes [n, ld] uint16 hundredths (0 = absent, 0xFFFF = escape),
    esc_key / esc_h = sorted cell keys (row0 + row) * esc_b + bin and true hundredths of the escape cells (depth >=
    655.35).  Every step-4 entry point has a _u16 twin that reads this and returns bit-identical results with half the
    HBM traffic."""

    codes: torch.Tensor
    b: int
    esc_key: torch.Tensor | None = None
    esc_h: torch.Tensor | None = None
    row0: int = 0
    esc_b: int = 0
    esc_cols: torch.Tensor | None = None   # distinct bins (int32, local) that hold an escape cell

    @property
    def device(self):
        return self.codes.device

    @property
    def shape(self):
        return self.codes.shape

    def stride(self, d):
        return self.codes.stride(d)

    @property
    def n_esc(self) -> int:
        return 0 if self.esc_key is None else int(self.esc_key.numel())

    def set_escapes(self, cell: torch.Tensor, val: torch.Tensor) -> None:
        """Escape list as PackedDepth carries it (cell = global_row * b + bin, fp32 depth) -> the kernels' sorted table."""
        dev = self.codes.device
        if cell.numel() == 0:
            self.esc_key = self.esc_h = self.esc_cols = None
            return
        # (on the device: a relative host format can carry millions of escapes; val * 100 is exact in fp64)
        h = torch.round(val.to(dev, non_blocking=True).double() * 100.0).clamp(max=(1 << 22) - 1).to(torch.int32)
        self.set_escape_table(cell.to(dev, non_blocking=True).to(torch.int64), h)

    def set_escape_table(self, cell: torch.Tensor, h: torch.Tensor) -> None:
        """The same from device tensors: cell keys int64 (unsorted), true hundredths int32."""
        if cell.numel() == 0:
            self.esc_key = self.esc_h = self.esc_cols = None
            return
        key, order = torch.sort(cell)
        self.esc_key, self.esc_h = key, h[order]
        self.esc_cols = torch.unique(key % (self.esc_b or self.b)).to(torch.int32)


def check_depth(depth, b: int) -> tuple[int, int]:
    if isinstance(depth, Depth16):
        c = depth.codes
        if not (c.is_cuda and c.dtype == torch.uint16 and c.dim() == 2 and c.stride(1) == 1 and c.stride(0) % 4 == 0):
            raise ValueError("Depth16.codes must be a 2-D uint16 CUDA tensor with unit column stride and ld % 4 == 0")
        return c.shape[0], c.stride(0)
    if not (depth.is_cuda and depth.dtype == torch.float32 and depth.dim() == 2 and depth.stride(1) == 1):
        raise ValueError("depth must be a 2-D fp32 CUDA tensor with unit column stride")
    return depth.shape[0], depth.stride(0)


def _dcall(name: str, depth, n: int, b: int, ld: int, *rest) -> None:
    """call(name, depth, n, b, ld, ...) for either storage type of the depth matrix."""
    if isinstance(depth, Depth16):
        more = ()
        if name == "gridb200_col_norm_stats":  # its patch kernel re-sums the bins that hold an escape
            more = (ptr(depth.esc_cols), 0 if depth.esc_cols is None else int(depth.esc_cols.numel()))
        call(name + "_u16", ptr(depth.codes), n, b, ld, ptr(depth.esc_key), ptr(depth.esc_h), depth.n_esc, depth.row0,
             depth.esc_b or depth.b, *more, *rest)
    else:
        call(name, ptr(depth), n, b, ld, *rest)


def to_depth16(depth: torch.Tensor, b: int, row0: int = 0, esc_b: int | None = None) -> Depth16:
    """fp32 device matrix -> Depth16 (device to device; escapes collected by the pack kernel)."""
    n, ld = check_depth(depth, b)
    dev = depth.device
    ldo = (b + 7) // 8 * 8
    codes = torch.zeros((n, ldo), dtype=torch.int16, device=dev).view(torch.uint16)
    cap = 1 << 16
    while True:
        esc_cell = torch.empty(cap, dtype=torch.int64, device=dev)
        esc_val = torch.empty(cap, dtype=torch.float32, device=dev)
        cnt = torch.zeros(1, dtype=torch.int64, device=dev)
        for r0 in range(0, n, 65535):
            c = min(65535, n - r0)
            call("gridb200_depth_pack_u16", depth.data_ptr() + r0 * ld * 4, c, b, ld, codes.data_ptr() + r0 * ldo * 2, ldo,
                 row0 + r0, ptr(esc_cell), ptr(esc_val), cap, ptr(cnt), _stream())
        m = int(cnt.item())
        if m <= cap:
            break
        cap = m
    d16 = Depth16(codes, b, row0=row0, esc_b=esc_b or b)
    d16.set_escapes(esc_cell[:m], esc_val[:m])
    return d16


_SIDE: dict = {}


def _read_behind(t: torch.Tensor, ready: "torch.cuda.Event") -> list:
    """t.tolist() without draining the main stream: the copy runs on a side stream that waits only for `ready` (recorded
    when t was complete), so kernels already queued behind it keep the GPU busy while the host reads the few values it
    needs to go on (with several GPUs a step is a few milliseconds and every drained queue shows)."""
    dev = t.device
    side = _SIDE.get(dev)
    if side is None:
        side = _SIDE[dev] = torch.cuda.Stream(device=dev)
    side.wait_event(ready)
    with torch.cuda.stream(side):
        host = t.to("cpu", non_blocking=True)
        t.record_stream(side)
    side.synchronize()
    return host.tolist()


def normalize(depth: torch.Tensor, b: int, p: Params, want_zq: bool = False, want_zh: bool = True,
              group=None, bin_sizes: list[int] | None = None, defer_norms: bool = False) -> NormResult:
    """Step 4 on a device-resident depth matrix [n, ld] (first ``b`` columns are bins; 0 = absent).

    With ``group`` the matrix holds this rank's bin slice; ranks are ordered by bin range.  bin_sizes: every rank's
    number of bins when the caller knows them (saves an all-reduce and a host sync per step).
    defer_norms (bin-sharded): the all-reduce of the row norms is started but not waited for -- the Gram kernel does not
    read them -- and its handle is left in ``extra["norms_pending"]`` for the caller to wait on before the k-best merge.
    """
    n, ld = check_depth(depth, b)
    dev, st = depth.device, _stream()
    i64 = dict(dtype=torch.int64, device=dev)
    f64 = dict(dtype=torch.float64, device=dev)
    colsum, colcnt = torch.empty(b, **i64), torch.empty(b, dtype=torch.int32, device=dev)
    valid = torch.empty(b, dtype=torch.uint8, device=dev)
    # one pass yields the per-bin sums AND the per-sample totals over all bins; the per-sample sums over the VALID
    # bins then only need the few bins that failed the depth filter (gridb200_row_stats_fused decides on the device)
    rowsum_all, rowcnt_all = torch.empty(n, **i64), torch.empty(n, dtype=torch.int32, device=dev)
    with _mark("col_raw"):
        _dcall("gridb200_col_row_raw_stats", depth, n, b, ld, ptr(colsum), ptr(colcnt), ptr(rowsum_all),
             ptr(rowcnt_all), st)
    call("gridb200_depth_filter", ptr(colsum), ptr(colcnt), b, float(p.min_depth), float(p.max_depth), ptr(valid), st)

    # one buffer: the int32 counts ride behind the int64 sums, two per word, so a bin-sharded run sums both in ONE
    # all-reduce (the counts are non-negative and their cohort totals fit 31 bits: no carry between the halves)
    rows_buf = torch.zeros(n + (n + 1) // 2, **i64)
    rowsum, rowcnt = rows_buf[:n], rows_buf[n:].view(torch.int32)[:n]
    ws0 = _ws(_lib.load().gridb200_row_stats_fused_ws_bytes(b), dev)
    with _mark("row_stats"):
        _dcall("gridb200_row_stats_fused", depth, n, b, ld, ptr(valid), ptr(colcnt), ptr(rowsum_all), ptr(rowcnt_all),
             ptr(rowsum), ptr(rowcnt), ptr(ws0), ws0.numel(), st)
    _all_reduce(rows_buf, group)
    _tick("rows all-reduced")
    raw_mean, scale_q, row_factor = torch.empty(n, **f64), torch.empty(n, **f64), torch.empty(n, **f64)
    call("gridb200_row_scales", ptr(rowsum), ptr(rowcnt), n, ptr(raw_mean), ptr(scale_q), ptr(row_factor), st)
    # exact %.2f ties take numpy's own mean: a pairwise fp64 sum over the WHOLE row, replayed for the tie rows only
    n_tie_t = None
    if group is None:
        _dcall("gridb200_row_means_exact", depth, n, b, ld, ptr(valid), ptr(rowsum), ptr(rowcnt), ptr(raw_mean),
             ptr(scale_q), st)
    else:  # bins are sharded: rows are not whole on any rank; ties are known exactly from the all-reduced integer sums
        tie = (rowcnt > 0) & ((2 * rowsum) % (2 * rowcnt.to(torch.int64)).clamp(min=1) == rowcnt)
        n_tie_t = tie.sum()

    # what the host needs next (rows with data, tie rows) is complete here: read it BEHIND the statistics pass
    counts_dev = torch.stack([(rowcnt > 0).sum(), n_tie_t if n_tie_t is not None else rowcnt.new_zeros(())])
    counts_ready = torch.cuda.Event()
    counts_ready.record()
    sd, sd2 = torch.empty(b, **f64), torch.empty(b, **f64)
    ws = _ws(_lib.load().gridb200_col_norm_stats_ws_bytes(n, b), dev)
    with _mark("col_norm"):
        _dcall("gridb200_col_norm_stats", depth, n, b, ld, ptr(valid), ptr(row_factor), ptr(sd), ptr(sd2), ptr(ws),
             ws.numel(), st)
    mu, s2, ratio = torch.empty(b, **f64), torch.empty(b, **f64), torch.empty(b, **f64)
    # samples without any valid bin are dropped by the reference before normalising (filter_empty_samples,
    # normalize_mosdepth.py:576-600); they contribute nothing here either, only N in the ddof=1 divisor changes
    counts_host = _read_behind(counts_dev, counts_ready)
    _tick("row counts on host")
    n_eff = int(counts_host[0])
    if counts_host[1]:
        _sharded_tie_means(depth, b, valid, rowsum, rowcnt, raw_mean, scale_q, tie, group)
    call("gridb200_col_finalize", ptr(sd), ptr(sd2), ptr(colcnt), ptr(valid), b, n_eff, ptr(mu), ptr(s2), ptr(ratio), st)

    # selection runs on the whole cohort's ratio vector (every rank redundantly)
    if group is not None:
        import torch.distributed as dist

        world, rank = dist.get_world_size(group), dist.get_rank(group)
        if bin_sizes is not None:
            sizes = [int(x) for x in bin_sizes]
            assert len(sizes) == world and sizes[rank] == b, "bin_sizes must list every rank's bins"
        else:
            sizes = torch.zeros(world, **i64)
            sizes[rank] = b
            dist.all_reduce(sizes, group=group)
            sizes = sizes.tolist()
        bmax = max(sizes)
        mine = torch.full((bmax,), float("nan"), **f64)
        mine[:b] = ratio
        gathered = torch.empty(world * bmax, **f64)
        dist.all_gather_into_tensor(gathered, mine, group=group)
        ratio_all = torch.cat([gathered[r * bmax:r * bmax + sizes[r]] for r in range(world)])
        _tick("ratios gathered")
        lo = sum(sizes[:rank])
    else:
        ratio_all, lo = ratio, 0
    ball = ratio_all.numel()
    sel_all = torch.empty(ball, dtype=torch.int32, device=dev)
    keep_all = torch.empty(ball, dtype=torch.uint8, device=dev)
    info, info2 = torch.empty(4, **f64), torch.empty(2, **f64)
    ws2 = _ws(_lib.load().gridb200_select_ws_bytes(ball), dev)
    with _mark("select"):
        call("gridb200_select", ptr(ratio_all), ball, float(p.top_frac), ptr(sel_all), ptr(info), ptr(ws2), ws2.numel(), st)
        call("gridb200_knn_col_filter", ptr(ratio_all), ptr(sel_all), ball, float(p.frac_r), float(p.sigma2_max),
             ptr(keep_all), ptr(info), ptr(info2), ptr(ws2), ws2.numel(), st)
    if group is not None:
        mine_sel = sel_all[lo:lo + b]
        before = (sel_all[:lo] >= 0).sum().to(torch.int32)
        sel_pos = torch.where(mine_sel >= 0, mine_sel - before, mine_sel).contiguous()
        keep = keep_all[lo:lo + b].contiguous()
        rw_local_t = (mine_sel >= 0).sum()
    else:
        sel_pos, keep, rw_local_t = sel_all, keep_all, None
    host = torch.cat([info, info2, (rw_local_t.to(torch.float64).reshape(1) if rw_local_t is not None else info[3:4])]).cpu()
    zscale, rw, r_use, rw_local = float(host[0]), int(host[3]), int(host[4]), int(host[6])
    _tick("selection on host")
    if rw < 0:
        raise IndexError("list index out of range")  # what the reference's sorted_ratios[int(top_frac*len)] raises
    res = NormResult(n=n, b=b, colsum=colsum, colcnt=colcnt, valid=valid, raw_mean=raw_mean, scale_q=scale_q,
                     row_factor=row_factor, mu=mu, s2=s2, ratio=ratio, sel_pos=sel_pos, keep=keep, zscale=zscale,
                     rw=rw, rw_local=rw_local, r_use=r_use)
    res.extra.update(thr=float(host[1]), n_finite=int(host[2]), sigma2_min=float(host[5]), bin_offset=lo,
                     rowcnt=rowcnt, rowsum=rowsum, n_eff=n_eff)
    if rw_local == 0 or not (want_zq or want_zh):
        return res

    sat = torch.zeros(1, dtype=torch.int64, device=dev)
    zq = zh = norms = None
    ldq = ldh = 0
    if want_zq:
        ldq = (rw_local + 7) // 8 * 8
        zq = torch.empty((n, ldq), dtype=torch.int16, device=dev)
        if ldq > rw_local:
            zq[:, rw_local:] = -32768
    if want_zh:
        ldh = (rw_local + 63) // 64 * 64
        zh = torch.empty((ldh // 64, n, 64), dtype=torch.float16, device=dev)  # k-blocked: [k-block, row, bin]
        if ldh > rw_local:
            zh[-1, :, rw_local % 64:] = 0
        norms = torch.zeros(n, **i64)
    with _mark("transform"):
        _dcall("gridb200_transform", depth, n, b, ld, ptr(row_factor), ptr(mu), ptr(sel_pos), ptr(keep), ptr(info),
             p.zclip, ptr(zq), ldq, ptr(zh), n, ptr(norms), ptr(sat), st)
    if want_zh and group is not None and defer_norms:
        import torch.distributed as dist

        res.extra["norms_pending"] = dist.all_reduce(norms, group=group, async_op=True)
    elif want_zh:
        _all_reduce(norms, group)
        _tick("norms all-reduced")
    res.zq, res.zh, res.kdim, res.norms = zq, zh, ldh, norms
    res.extra["sat"] = sat
    return res


def _sharded_tie_means(depth, b, valid, rowsum, rowcnt, raw_mean, scale_q, tie, group) -> None:
    """Bin-sharded cohorts: the rows whose exact mean sits on a %.2f tie (known from the all-reduced integer sums) are
    assembled whole on every rank -- a few rows x all bins -- and numpy's pairwise sum is replayed for them
    (gridb200_row_means_exact), so that single- and multi-GPU runs print the same scales.  Ties are rare (about one
    row in `bins`), so the exchange is a few rows of fp32."""
    import torch.distributed as dist

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = depth.device
    rows = torch.nonzero(tie).flatten()
    nt = rows.numel()
    sizes = torch.zeros(world, dtype=torch.int64, device=dev)
    sizes[rank] = b
    dist.all_reduce(sizes, group=group)
    sizes = sizes.tolist()
    bmax, btot = max(sizes), sum(sizes)
    mine = torch.zeros((nt, bmax), dtype=torch.float32, device=dev)
    if isinstance(depth, Depth16):  # (rare path) codes -> the fp32 values the text would have given: h / 100 correctly rounded
        hh = depth.codes.view(torch.int16)[rows, :b].to(torch.int32) & 0xFFFF
        if depth.n_esc:
            keys = ((depth.row0 + rows)[:, None] * (depth.esc_b or depth.b) + torch.arange(b, device=dev)[None, :])
            at = torch.searchsorted(depth.esc_key, keys.reshape(-1)).clamp(max=depth.n_esc - 1).reshape(keys.shape)
            hit = (hh == 0xFFFF) & (depth.esc_key[at] == keys)
            hh = torch.where(hit, depth.esc_h[at], hh)
        mine[:, :b] = hh.to(torch.float32) / torch.full((1,), 100.0, dtype=torch.float32, device=dev)
    else:
        mine[:, :b] = depth[rows, :b]
    vmine = torch.zeros(bmax, dtype=torch.uint8, device=dev)
    vmine[:b] = valid
    got = torch.empty((world, nt, bmax), dtype=torch.float32, device=dev)
    vgot = torch.empty((world, bmax), dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(got, mine, group=group)
    dist.all_gather_into_tensor(vgot, vmine, group=group)
    ldf = (btot + 3) // 4 * 4
    full = torch.zeros((nt, ldf), dtype=torch.float32, device=dev)
    vfull = torch.zeros(btot, dtype=torch.uint8, device=dev)
    at = 0
    for r in range(world):  # ranks are ordered by bin range
        full[:, at:at + sizes[r]] = got[r, :, :sizes[r]]
        vfull[at:at + sizes[r]] = vgot[r, :sizes[r]]
        at += sizes[r]
    rs, rc = rowsum[rows].contiguous(), rowcnt[rows].contiguous()
    rm, sq = raw_mean[rows].contiguous(), scale_q[rows].contiguous()
    call("gridb200_row_means_exact", ptr(full), nt, btot, ldf, ptr(vfull), ptr(rs), ptr(rc), ptr(rm), ptr(sq), _stream())
    raw_mean[rows] = rm
    scale_q[rows] = sq


def prepare_from_zq(zq: torch.Tensor, rw: int, keep: torch.Tensor | None, zclip: int):
    """Operand + exact norms from a hundredths matrix read back from a step-4 file (find_neighbors.py:57-65)."""
    n, ldq = zq.shape[0], zq.stride(0)
    ldh = max(64, (rw + 63) // 64 * 64)
    zh = torch.zeros((ldh // 64, n, 64), dtype=torch.float16, device=zq.device)
    norms = torch.empty(n, dtype=torch.int64, device=zq.device)
    call("gridb200_knn_prepare", ptr(zq), n, rw, ldq, ptr(keep), zclip, ptr(zh), n, ptr(norms), _stream())
    return zh, ldh, norms


def zh_rowmajor(zh: torch.Tensor) -> torch.Tensor:
    """[n, kdim] view-copy of a k-blocked operand (tests and debugging)."""
    return zh.permute(1, 0, 2).reshape(zh.shape[1], zh.shape[0] * 64)


def knn_init(nq: int, k: int, dev):
    """Empty running k-best lists (ascending by (d2, global index)) for nq query rows."""
    best_d2 = torch.empty((nq, k), dtype=torch.int64, device=dev)
    best_idx = torch.empty((nq, k), dtype=torch.int32, device=dev)
    call("gridb200_topk_init", ptr(best_d2), ptr(best_idx), nq, k, _stream())
    return best_d2, best_idx


TOPK_MAX_KEYS = 8192  # candidates + k per gridb200_topk_merge* call (shared-memory sort)


def check_k(k: int) -> None:
    """The k-best merge sorts k + candidates <= 8192 keys in shared memory and takes candidates in chunks of 8192 - k:
    refuse a k that leaves no room instead of looping forever or writing empty lists."""
    if not 0 < k <= TOPK_MAX_KEYS // 2:
        raise ValueError(f"num_neighbors = {k}: the device k-best merge supports 1 <= k <= {TOPK_MAX_KEYS // 2} "
                         "(the reference accepts any value; cohorts that need more are not supported on the GPU path)")


def merge_lists(src_d2: torch.Tensor, src_idx: torch.Tensor, best_d2: torch.Tensor, best_idx: torch.Tensor) -> None:
    """Merge candidate lists [n, src_k] (idx < 0 = empty) into the running k-best lists [n, k] of the same rows."""
    n, k = best_d2.shape
    assert src_d2.shape[0] == n and src_d2.is_contiguous() and src_idx.is_contiguous() and best_d2.is_contiguous()
    if n:
        call("gridb200_topk_merge_lists", ptr(src_d2), ptr(src_idx), n, src_d2.shape[1], k, ptr(best_d2), ptr(best_idx),
             _stream())


def knn_block(zh_q: torch.Tensor, norms_q: torch.Tensor, q_gid0: int, zh_r: torch.Tensor, norms_r: torch.Tensor,
              r_gid0: int, kdim: int, vmax: int, best_d2: torch.Tensor, best_idx: torch.Tensor, impl: int = 0,
              group=None, max_block_bytes: int = 2 << 30, q_range: tuple[int, int] | None = None,
              r_range: tuple[int, int] | None = None, t_lists: tuple[torch.Tensor, torch.Tensor] | None = None,
              padded: tuple[torch.Tensor, torch.Tensor] | None = None, pending=None) -> None:
    """Merge the reference rows of ``zh_r`` (global ids r_gid0..) into the running lists of the query rows of
    ``zh_q`` (global ids q_gid0..).  Operands are k-blocked [kdim/64, rows, 64]; the N x N distances only ever
    exist as query-block x reference-shard Gram tiles (<= max_block_bytes).

    q_range / r_range: restrict the block to query rows [a, b) / reference rows [c, d) of the operands (norms, global
    ids and best_* keep indexing the WHOLE operand's rows).
    t_lists = (t_d2, t_idx) [rows of zh_r, k]: also merge, through the block's transpose, the query rows into the
    running lists of the REFERENCE rows -- the half ring's way of computing every pair of samples once.
    padded = (d2, idx) [ceil(n / world) * world, k] of which best_d2 / best_idx are the first n rows (bin-sharded runs):
    the lists are then all-gathered in place, without staging copies.
    pending: handle of an asynchronous collective the norms still depend on; waited for after the Gram launch."""
    for t in (zh_q, zh_r):
        assert t.dim() == 3 and t.shape[2] == 64 and t.is_contiguous() and t.shape[0] * 64 == kdim, "zh is [kdim/64, n, 64]"
    nq_all, nr_all = zh_q.shape[1], zh_r.shape[1]
    qa, qz = q_range or (0, nq_all)
    ra, rz = r_range or (0, nr_all)
    nr = rz - ra
    if qz <= qa or nr <= 0:
        return
    k = best_d2.shape[1]
    check_k(k)
    dev, st = zh_q.device, _stream()
    if impl == 0 and vmax > 511:
        impl = 1  # |v| > 511: a 64-wide fp32 chunk is no longer exact; the integer SIMT kernel takes over (still GPU)
    ldd = (nr + 1) // 2 * 2
    qb = max(1, min(qz - qa, max_block_bytes // (8 * ldd)))
    rchunk = TOPK_MAX_KEYS - k
    world, rank = 1, 0
    if group is not None:
        import torch.distributed as dist

        world, rank = dist.get_world_size(group), dist.get_rank(group)
        assert q_range is None and r_range is None and t_lists is None, "sub-blocks are a row-sharded (ring) feature"
    zr_ptr = zh_r.data_ptr() + ra * 64 * 2
    global GRAM_ELEMS
    for q0 in range(qa, qz, qb):
        nq = min(qb, qz - q0)
        m = -(-nq // world)  # query rows per rank after the reduce-scatter
        dot = torch.empty((m * world, ldd), dtype=torch.int64, device=dev)
        if m * world > nq:
            dot[nq:].zero_()
        with _mark("gram"):
            call("gridb200_gram", zh_q.data_ptr() + q0 * 64 * 2, nq, nq_all, zr_ptr, nr, nr_all, kdim, vmax, ptr(dot), ldd,
                 0, impl, st)
        GRAM_ELEMS += gram_tile_elems(nq, nr, impl == 0 and q0 == 0 and ra == 0 and nq == nr == nq_all == nr_all
                                      and zh_q.data_ptr() == zh_r.data_ptr())
        if pending is not None:  # (the row norms' all-reduce ran beside the Gram kernel)
            pending.wait()
            pending = None
            _tick("norms all-reduced")
        if world > 1:
            # bin-sharded: every rank holds a partial of the whole block.  Reduce-scatter it by query rows, merge the
            # k-best lists of MY rows only, then all-gather the (small) lists -- instead of all-reducing the block and
            # repeating the whole merge on every rank.
            mine = torch.empty((m, ldd), dtype=torch.int64, device=dev)
            dist.reduce_scatter_tensor(mine, dot, group=group)
            _tick("dots reduce-scattered")
            r_lo = min(nq, rank * m)
            n_mine = min(nq, r_lo + m) - r_lo
        else:
            mine, r_lo, n_mine = dot, 0, nq
        with _mark("topk"):
            for r0 in range(0, nr, rchunk):
                if n_mine <= 0:
                    break
                n_r = min(rchunk, nr - r0)
                call("gridb200_topk_merge", mine.data_ptr() + r0 * 8, ldd, n_mine, n_r,
                     norms_q.data_ptr() + (q0 + r_lo) * 8, norms_r.data_ptr() + (ra + r0) * 8, q_gid0 + q0 + r_lo,
                     r_gid0 + ra + r0, k, best_d2.data_ptr() + (q0 + r_lo) * k * 8, best_idx.data_ptr() + (q0 + r_lo) * k * 4, st)
            if t_lists is not None:  # the same block read by columns: lists of the reference rows, query rows as candidates
                t_d2, t_idx = t_lists
                assert t_d2.shape == (nr_all, k) and t_d2.is_contiguous() and t_idx.is_contiguous()
                for c0 in range(0, nq, rchunk):
                    n_c = min(rchunk, nq - c0)
                    call("gridb200_topk_merge_t", dot.data_ptr() + c0 * ldd * 8, ldd, n_c, nr,
                         norms_q.data_ptr() + (q0 + c0) * 8, norms_r.data_ptr() + ra * 8, q_gid0 + q0 + c0, r_gid0 + ra, k,
                         t_d2.data_ptr() + ra * k * 8, t_idx.data_ptr() + ra * k * 4, st)
        if world > 1 and padded is not None and q0 == 0 and nq == nq_all and padded[0].shape[0] == m * world:
            dist.all_gather_into_tensor(padded[0], padded[0][rank * m:(rank + 1) * m], group=group)
            dist.all_gather_into_tensor(padded[1], padded[1][rank * m:(rank + 1) * m], group=group)
            _tick("lists gathered")
        elif world > 1:
            part_d2 = torch.zeros((m, k), dtype=torch.int64, device=dev)
            part_idx = torch.zeros((m, k), dtype=torch.int32, device=dev)
            if n_mine > 0:
                part_d2[:n_mine] = best_d2[q0 + r_lo:q0 + r_lo + n_mine]
                part_idx[:n_mine] = best_idx[q0 + r_lo:q0 + r_lo + n_mine]
            all_d2 = torch.empty((m * world, k), dtype=torch.int64, device=dev)
            all_idx = torch.empty((m * world, k), dtype=torch.int32, device=dev)
            dist.all_gather_into_tensor(all_d2, part_d2, group=group)
            dist.all_gather_into_tensor(all_idx, part_idx, group=group)
            best_d2[q0:q0 + nq] = all_d2[:nq]
            best_idx[q0:q0 + nq] = all_idx[:nq]
            _tick("lists gathered")


FUSED_CAP = 1024          # candidate slots per row of the fused Gram + k-best kernel
FUSED_MIN_ELEMS = 1 << 22  # "auto": blocks smaller than this keep the dot-matrix path (its split-K fills the SMs better)
COUNTERS = {"fused_blocks": 0, "fused_overflow": 0, "unfused_blocks": 0}
_FUSED_WS: dict = {}


def fused_mode() -> str:
    """GRIDB200_FUSED = auto (default) | 1 (every eligible block, whatever its size: tests) | 0 (never)."""
    import os

    return os.environ.get("GRIDB200_FUSED", "auto")


def knn_block_fused(zh_q: torch.Tensor, norms_q: torch.Tensor, q_gid0: int, zh_r: torch.Tensor, norms_r: torch.Tensor,
                    r_gid0: int, kdim: int, vmax: int, best_d2: torch.Tensor, best_idx: torch.Tensor,
                    q_range=None, r_range=None, tau_r: torch.Tensor | None = None, t_lists=None) -> bool:
    """knn_block through gridb200_gram_topk: the block's distances are filtered against the lists' current k-th entries
    in the Gram epilogue and never reach memory.  tau_r [rows of zh_r] = current k-th best d2 of the REFERENCE rows
    (their owner's lists) -- required with t_lists, which then receive only what can still enter those lists.
    Returns False, having changed nothing, when the block is not eligible (a list without k entries yet: nothing to
    filter against) or a candidate buffer overflowed; the caller then takes knn_block."""
    nq_all, nr_all = zh_q.shape[1], zh_r.shape[1]
    qa, qz = q_range or (0, nq_all)
    ra, rz = r_range or (0, nr_all)
    nq, nr = qz - qa, rz - ra
    k = best_d2.shape[1]
    cap = min(FUSED_CAP, TOPK_MAX_KEYS - k)
    if nq <= 0 or nr <= 0:
        return True
    if vmax > 511 or cap < 64 or (t_lists is not None and tau_r is None):
        return False
    dev, st = zh_q.device, _stream()
    inf = torch.iinfo(torch.int64).max
    tau_q = best_d2[qa:qz, k - 1]
    ready = (tau_q != inf).all()
    if t_lists is not None:
        ready = ready & (tau_r[ra:rz] != inf).all()
    if not bool(ready):  # (one host sync: a list that is not full yet has no threshold -- everything would survive)
        return False
    ws = _FUSED_WS.get(dev)
    if ws is None:
        ws = _FUSED_WS[dev] = _ws(_lib.load().gridb200_gram_topk_ws_bytes(), dev)
    i32 = dict(dtype=torch.int32, device=dev)
    cq_d2 = torch.empty((nq, cap), dtype=torch.int64, device=dev)
    cq_idx, cq_cnt = torch.empty((nq, cap), **i32), torch.zeros(nq, **i32)
    cr_d2 = cr_idx = cr_cnt = None
    if t_lists is not None:
        cr_d2 = torch.empty((nr, cap), dtype=torch.int64, device=dev)
        cr_idx, cr_cnt = torch.empty((nr, cap), **i32), torch.zeros(nr, **i32)
    global GRAM_ELEMS
    with _mark("gram"):
        call("gridb200_gram_topk", zh_q.data_ptr() + qa * 64 * 2, nq, nq_all, zh_r.data_ptr() + ra * 64 * 2, nr, nr_all,
             kdim, vmax, norms_q.data_ptr() + qa * 8, norms_r.data_ptr() + ra * 8, q_gid0 + qa, r_gid0 + ra,
             best_d2.data_ptr() + (qa * k + k - 1) * 8, k, (tau_r.data_ptr() + ra * 8) if t_lists is not None else None, 1,
             cap, ptr(cq_d2), ptr(cq_idx), ptr(cq_cnt), ptr(cr_d2), ptr(cr_idx), ptr(cr_cnt), ptr(ws), ws.numel(), st)
    GRAM_ELEMS += nq * nr
    over = (cq_cnt > cap).any()
    if t_lists is not None:
        over = over | (cr_cnt > cap).any()
    if bool(over):  # a row met more than `cap` samples closer than its k-th best: redo the block with the dot matrix
        COUNTERS["fused_overflow"] += 1
        return False
    with _mark("topk"):
        call("gridb200_topk_merge_cand", ptr(cq_d2), ptr(cq_idx), ptr(cq_cnt), nq, cap, k, best_d2.data_ptr() + qa * k * 8,
             best_idx.data_ptr() + qa * k * 4, st)
        if t_lists is not None:
            t_d2, t_idx = t_lists
            call("gridb200_topk_merge_cand", ptr(cr_d2), ptr(cr_idx), ptr(cr_cnt), nr, cap, k, t_d2.data_ptr() + ra * k * 8,
                 t_idx.data_ptr() + ra * k * 4, st)
    COUNTERS["fused_blocks"] += 1
    return True


def knn_block_auto(zh_q, norms_q, q_gid0, zh_r, norms_r, r_gid0, kdim, vmax, best_d2, best_idx, impl=0, q_range=None,
                   r_range=None, t_lists=None, tau_r=None, symmetric=False) -> None:
    """One block of the ring schedules: the fused Gram + k-best kernel where it applies (a rectangular block between
    different sample sets whose lists already hold k entries), the dot-matrix path otherwise or on overflow."""
    mode = fused_mode()
    nq = (q_range[1] - q_range[0]) if q_range else zh_q.shape[1]
    nr = (r_range[1] - r_range[0]) if r_range else zh_r.shape[1]
    want = impl == 0 and not symmetric and mode != "0" and (mode == "1" or nq * nr >= FUSED_MIN_ELEMS)
    if want and knn_block_fused(zh_q, norms_q, q_gid0, zh_r, norms_r, r_gid0, kdim, vmax, best_d2, best_idx, q_range=q_range,
                                r_range=r_range, tau_r=tau_r, t_lists=t_lists):
        return
    COUNTERS["unfused_blocks"] += 1
    knn_block(zh_q, norms_q, q_gid0, zh_r, norms_r, r_gid0, kdim, vmax, best_d2, best_idx, impl=impl, q_range=q_range,
              r_range=r_range, t_lists=t_lists)


def knn(zh: torch.Tensor, kdim: int, norms: torch.Tensor, k: int, vmax: int, impl: int = 0, group=None,
        max_block_bytes: int = 2 << 30, pending=None):
    """Exact k nearest neighbours of every row among all rows (self excluded): (idx[n,k] int32, d2[n,k] int64).

    zh is the k-blocked operand [kdim/64, n, 64] written by gridb200_transform / gridb200_knn_prepare.
    d2 is in units of 1e-4 (hundredths squared).  Rows with fewer than k candidates are padded with -1.
    Bin-sharded multi-GPU (``group``): each rank contributes the Gram partial of its bins; partials are all-reduced.
    """
    assert zh.dim() == 3 and zh.shape[2] == 64 and zh.is_contiguous() and zh.shape[0] * 64 == kdim, "zh is [kdim/64, n, 64]"
    n = zh.shape[1]
    k = max(1, min(k, n - 1)) if n > 1 else 1
    padded = None
    if group is not None:
        import torch.distributed as dist

        world = dist.get_world_size(group)
        padded = knn_init(-(-n // world) * world, k, zh.device)  # every rank's share of the rows the same size
        best_d2, best_idx = padded[0][:n], padded[1][:n]
    else:
        best_d2, best_idx = knn_init(n, k, zh.device)
    knn_block(zh, norms, 0, zh, norms, 0, kdim, vmax, best_d2, best_idx, impl=impl, group=group,
              max_block_bytes=max_block_bytes, padded=padded, pending=pending)
    return best_idx, best_d2


def knn_f32(x: torch.Tensor, k: int, max_block_bytes: int = 2 << 30):
    """Exact-to-fp32 k nearest neighbours (self excluded) of the rows of an UNQUANTISED fp32 matrix x [n, r] on the tensor
    cores with 3xTF32 split precision (north_star step (c) as stated; gram_f32.cu).  The pipeline's own matrices are
    quantised by the %.2f hand-off and go through knn(); this entry serves callers that skip the text round trip.
    Returns (idx [n, k] int32, d2 [n, k] float64 squared Euclidean distances), ascending by (d2, index)."""
    if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1):
        raise ValueError("x must be a 2-D fp32 CUDA tensor with unit column stride")
    n, r = x.shape
    k = max(1, min(k, n - 1)) if n > 1 else 1
    check_k(k)
    dev, st = x.device, _stream()
    kdim = (r + 31) // 32 * 32
    hi = torch.empty((kdim // 32, n, 32), dtype=torch.float32, device=dev)
    lo = torch.empty_like(hi)
    norms = torch.empty(n, dtype=torch.float64, device=dev)
    call("gridb200_split_tf32", ptr(x), n, r, x.stride(0), ptr(hi), ptr(lo), n, ptr(norms), st)
    best_d2, best_idx = knn_init(n, k, dev)
    ldd = (n + 3) // 4 * 4
    qb = max(1, min(n, max_block_bytes // (4 * ldd)))
    rchunk = TOPK_MAX_KEYS - k
    for q0 in range(0, n, qb):
        nq = min(qb, n - q0)
        dot = torch.empty((nq, ldd), dtype=torch.float32, device=dev)
        with _mark("gram_f32"):
            call("gridb200_gram_f32", hi.data_ptr() + q0 * 32 * 4, lo.data_ptr() + q0 * 32 * 4, nq, n, ptr(hi), ptr(lo), n, n,
                 kdim, ptr(dot), ldd, st)
        with _mark("topk"):
            for r0 in range(0, n, rchunk):
                n_r = min(rchunk, n - r0)
                call("gridb200_topk_merge_f32", dot.data_ptr() + r0 * 4, ldd, nq, n_r, norms.data_ptr() + q0 * 8,
                     norms.data_ptr() + r0 * 8, q0, r0, k, best_d2.data_ptr() + q0 * k * 8, best_idx.data_ptr() + q0 * k * 4, st)
    return best_idx, best_d2.view(torch.float64)


def dipcn(counts: torch.Tensor, has_count: torch.Tensor, scale_q: torch.Tensor, nbr_idx: torch.Tensor, n_nbr: int,
          q_gid0: int = 0, rule: str = "pipeline"):
    """Step 6 on device arrays: counts[n_all,g] int32, has_count[n_all] uint8, scale_q[n_all] f64 for the WHOLE cohort,
    nbr_idx[nq,k] int32 (global ids) for query rows q_gid0.. -> out[nq,g] f64.
    rule: "pipeline" = compute_dipcn.py:62-88; "exon" = compute_dipcn_dir/compute_diploid_cn.py:8-63."""
    rule_id = {"pipeline": 0, "exon": 1}[rule]
    n, g = counts.shape
    dev = counts.device
    nq = nbr_idx.shape[0]
    out = torch.empty((nq, g), dtype=torch.float64, device=dev)
    t_ws = torch.empty(n * g, dtype=torch.float64, device=dev)
    with _mark("dipcn"):
        call("gridb200_dipcn", ptr(counts), n, g, counts.stride(0), ptr(has_count), ptr(scale_q), ptr(nbr_idx),
             nq, nbr_idx.shape[1], int(q_gid0), int(n_nbr), rule_id, ptr(out), g, ptr(t_ws), _stream())
    return out


def synth_depth(seed: int, n: int, b: int, row0: int = 0, col0: int = 0, device="cuda", ld: int | None = None,
                out: torch.Tensor | None = None):
    """Device-generated synthetic depth matrix, bit-identical to grid_b200.synth.depth_matrix.
    out: optional flat fp32 buffer of at least n * ld elements to build the matrix in (callers that regenerate blocks
    of slightly different heights over and over keep one buffer instead of tens of GB of allocator churn)."""
    from . import synth

    ld = ld or (b + 3) // 4 * 4
    c, rc = synth.row_params(seed, n, row0)
    bc, rb = synth.col_params(seed, b, col0)
    dev = torch.device(device) if out is None else out.device
    tc, trc, tb, trb = (torch.from_numpy(a).to(dev) for a in (c, rc, bc, rb))
    if out is None:
        depth = torch.empty((n, ld), dtype=torch.float32, device=dev)
    else:
        assert out.dtype == torch.float32 and out.is_contiguous() and out.numel() >= n * ld
        depth = out.view(-1)[: n * ld].view(n, ld)
    if ld > b:
        depth[:, b:] = 0
    call("gridb200_synth_depth", seed, n, b, row0, col0, ptr(tc), ptr(trc), ptr(tb), ptr(trb), ptr(depth), ld, _stream())
    return depth


def run_chain(depth: torch.Tensor, b: int, counts: torch.Tensor, p: Params, impl: int = 0, group=None,
              bin_sizes: list[int] | None = None):
    """normalize -> kNN -> dipCN with everything resident on the device (the benchmarked path)."""
    res = normalize(depth, b, p, want_zq=False, want_zh=True, group=group, bin_sizes=bin_sizes, defer_norms=True)
    pending = res.extra.pop("norms_pending", None)
    if res.rw == 0:
        raise ValueError("no bins selected")
    if res.zh is None:  # this rank's slice holds no selected bin: contribute a zero partial
        res.zh = torch.zeros((1, res.n, 64), dtype=torch.float16, device=depth.device)
        res.kdim, res.norms = 64, torch.zeros(res.n, dtype=torch.int64, device=depth.device)
        _all_reduce(res.norms, group)
    idx, d2 = knn(res.zh, res.kdim, res.norms, p.num_neighbors, p.zclip, impl=impl, group=group, pending=pending)
    has = torch.ones(res.n, dtype=torch.uint8, device=depth.device)
    cn = dipcn(counts, has, res.scale_q, idx, p.n_nbr)
    return res, idx, d2, cn


# ---- packed host format (hundredths in 16 or fewer bits + escape list): 2 bytes or less per cell over PCIe ---------------
@dataclass
class PackedDepth:
    """Host-resident depth matrix in pinned memory.  bits == 16: codes[n, ld] uint16 hundredths (0 = absent, 0xFFFF = see
    escapes).  bits in 9..15: codes[n, ldw] int32 words, groups of 32 cells bit-packed (gridb200_depth_pack_bits; the
    all-ones code = see escapes).  escapes = (cell index = row * b + bin, fp32 value) for depths that do not fit."""

    codes: torch.Tensor
    n: int
    b: int
    esc_cell: torch.Tensor
    esc_val: torch.Tensor
    bits: int = 16
    row_q: torch.Tensor | None = None   # relative codes (bits < 16 only): int32 [n] and [b] on the host, see
    col_q: torch.Tensor | None = None   # gridb200_depth_pack_bits -- a cell travels as its residual against row_q x col_q

    @property
    def relative(self) -> bool:
        return self.row_q is not None

    @property
    def n_big(self) -> int:
        """Escape cells that do not fit the 2-byte device matrix either (depth >= 655.35); counted once, on the host."""
        if getattr(self, "_n_big", None) is None:
            self._n_big = int((self.esc_val >= 655.345).sum()) if self.esc_val.numel() else 0
        return self._n_big

    @property
    def nbytes(self) -> int:
        tables = (self.row_q.numel() + self.col_q.numel()) * 4 if self.relative else 0
        return self.codes.numel() * self.codes.element_size() + self.esc_cell.numel() * 12 + tables


def choose_bits(depth: torch.Tensor, b: int, max_escape_frac: float = 2e-4) -> int:
    """Smallest code width (13..16 bits) that leaves at most max_escape_frac of the cells to the escape list, judged on
    a sample of rows (a 30x cohort fits 13 bits: depths below 81.91)."""
    n = depth.shape[0]
    rows = torch.linspace(0, n - 1, min(n, 64)).long().to(depth.device)
    h = torch.round(depth[rows, :b].double() * 100.0)
    for bits in (13, 14, 15):
        if float((h >= (1 << bits) - 1).double().mean()) <= max_escape_frac:
            return bits
    return 16


def rank_one_tables(depth: torch.Tensor, b: int):
    """(row_q int32 [n], col_q int32 [b]) of the relative host format from the device fp32 matrix: col_q = the bin's mean
    hundredths over its present cells, row_q = 4096 x (the sample's mean over the cohort mean) -- from the exact integer
    sums of the raw statistics pass, so pred(i, j) = (row_q[i] * col_q[j] + 2048) >> 12 tracks depth_i x bias_j."""
    n, ld = check_depth(depth, b)
    dev = depth.device
    colsum, colcnt = torch.empty(b, dtype=torch.int64, device=dev), torch.empty(b, dtype=torch.int32, device=dev)
    rowsum, rowcnt = torch.empty(n, dtype=torch.int64, device=dev), torch.empty(n, dtype=torch.int32, device=dev)
    _dcall("gridb200_col_row_raw_stats", depth, n, b, ld, ptr(colsum), ptr(colcnt), ptr(rowsum), ptr(rowcnt), _stream())
    cmean = colsum.double() / colcnt.clamp(min=1).double()
    rmean = rowsum.double() / rowcnt.clamp(min=1).double()
    gmean = (rowsum.sum().double() / rowcnt.sum().clamp(min=1).double()).clamp(min=1.0)
    col_q = torch.round(cmean).clamp(0, 65535).to(torch.int32)
    row_q = torch.round(rmean / gmean * 4096.0).clamp(0, 32767).to(torch.int32)
    return row_q, col_q


def choose_format(depth: torch.Tensor, b: int, max_escape_frac: float = 2e-3):
    """(bits, row_q, col_q): the host format with the fewest bytes per cell, judged on a sample of rows -- absolute codes
    (row_q = col_q = None) or residuals against the rank-one prediction; an escape costs 12 bytes, and at most
    max_escape_frac of the cells may be escapes (each one is a scattered write on the device)."""
    n = depth.shape[0]
    rows = torch.linspace(0, n - 1, min(n, 64)).long().to(depth.device)
    h = torch.round(depth[rows, :b].double() * 100.0).to(torch.int64)
    present = h > 0
    best = (16 / 8.0, 16, None, None)
    for bits in range(9, 16):
        frac = float((h >= (1 << bits) - 1).double().mean())
        if frac <= max_escape_frac and bits / 8.0 + 12.0 * frac < best[0]:
            best = (bits / 8.0 + 12.0 * frac, bits, None, None)
    row_q, col_q = rank_one_tables(depth, b)
    pred = (row_q[rows].to(torch.int64)[:, None] * col_q.to(torch.int64)[None, :] + 2048) >> 12
    for bits in range(9, 16):
        r = h - pred + (1 << (bits - 1))
        bad = present & ((r < 1) | (r >= (1 << bits) - 1) | (h >= 0xFFFF))
        frac = float(bad.double().mean())
        if frac <= max_escape_frac and bits / 8.0 + 12.0 * frac < best[0]:
            best = (bits / 8.0 + 12.0 * frac, bits, row_q, col_q)
    return best[1], best[2], best[3]


def pack_depth(depth: torch.Tensor, b: int, chunk_rows: int = 256, pin: bool = True, bits: int | str = 16,
               tables: tuple[torch.Tensor, torch.Tensor] | None = None) -> PackedDepth:
    """Device fp32 depth[n, ld] -> PackedDepth on the host (export / side-car / bench set-up).  bits: 16, 9..15, "auto"
    (choose_bits: absolute codes) or "best" (choose_format: absolute or relative, whichever is smaller); tables =
    (row_q, col_q) on the device forces relative codes of the given width."""
    n, ld = depth.shape[0], depth.stride(0)
    dev = depth.device
    row_q = col_q = None
    if bits == "auto":
        bits = choose_bits(depth, b)
    elif bits == "best":
        bits, row_q, col_q = choose_format(depth, b)
    if tables is not None:
        row_q, col_q = (t.to(dev, torch.int32).contiguous() for t in tables)
        assert row_q.numel() == n and col_q.numel() == b
    bits = int(bits)
    if bits == 16:
        assert row_q is None, "relative codes are a bit-packed format (9..15 bits)"
        ldo, dt = (b + 7) // 8 * 8, torch.uint16
    else:
        ldo, dt = int(_lib.load().gridb200_depth_packed_row_words(b, bits)), torch.int32
        if ldo < 0:
            raise ValueError(f"bits = {bits}: must be 16 or 9..15")
    codes = torch.empty((n, ldo), dtype=dt, pin_memory=pin)
    cap = 1 << 16
    while True:
        esc_cell = torch.empty(cap, dtype=torch.int64, device=dev)
        esc_val = torch.empty(cap, dtype=torch.float32, device=dev)
        cnt = torch.zeros(1, dtype=torch.int64, device=dev)
        stage = torch.empty((min(chunk_rows, n), ldo), dtype=dt, device=dev)
        for r0 in range(0, n, chunk_rows):
            c = min(chunk_rows, n - r0)
            if bits == 16:
                call("gridb200_depth_pack_u16", depth.data_ptr() + r0 * ld * 4, c, b, ld, ptr(stage), ldo, r0, ptr(esc_cell),
                     ptr(esc_val), cap, ptr(cnt), _stream())
            else:
                call("gridb200_depth_pack_bits", depth.data_ptr() + r0 * ld * 4, c, b, ld, bits,
                     (row_q.data_ptr() + r0 * 4) if row_q is not None else None, ptr(col_q), ptr(stage), ldo, r0,
                     ptr(esc_cell), ptr(esc_val), cap, ptr(cnt), _stream())
            codes[r0:r0 + c].copy_(stage[:c])
        m = int(cnt.item())
        if m <= cap:
            def host(t):  # (the escape list and the tables travel with every upload: pinned like the codes)
                out = torch.empty(t.shape, dtype=t.dtype, pin_memory=pin)
                out.copy_(t)
                return out

            return PackedDepth(codes, n, b, host(esc_cell[:m]), host(esc_val[:m]), bits,
                               host(row_q) if row_q is not None else None, host(col_q) if col_q is not None else None)
        cap = m


def upload_packed(packed: PackedDepth, depth: torch.Tensor, chunk_rows: int = 128) -> torch.Tensor:
    """PackedDepth (host) -> device fp32 depth[n, ld] (preallocated), bit-identical to ingesting the text.
    Code chunks are copied on a side stream into two staging buffers while the previous chunk is unpacked."""
    n, b, ld = packed.n, packed.b, depth.stride(0)
    dev = depth.device
    lds = packed.codes.stride(0)
    chunk_rows = max(1, min(chunk_rows, n, 65535))
    key = (dev, chunk_rows, lds, packed.codes.dtype)
    st = _UPLOAD_STATE.get(key)
    if st is None:
        st = {"stage": [torch.empty((chunk_rows, lds), dtype=packed.codes.dtype, device=dev) for _ in range(2)],
              "copy": torch.cuda.Stream(device=dev), "done": [torch.cuda.Event() for _ in range(2)],
              "free": [torch.cuda.Event() for _ in range(2)]}
        _UPLOAD_STATE.clear()
        _UPLOAD_STATE[key] = st
    main = torch.cuda.current_stream(dev)
    n_esc = packed.esc_cell.numel()
    esc_cell = packed.esc_cell.to(dev, non_blocking=True) if n_esc else None
    esc_val = packed.esc_val.to(dev, non_blocking=True) if n_esc else None
    row_q = packed.row_q.to(dev, non_blocking=True) if packed.relative else None
    col_q = packed.col_q.to(dev, non_blocking=True) if packed.relative else None
    for s in range(2):
        st["free"][s].record(main)
    for t, r0 in enumerate(range(0, n, chunk_rows)):
        s, c = t & 1, min(chunk_rows, n - r0)
        with torch.cuda.stream(st["copy"]):
            st["copy"].wait_event(st["free"][s])  # the unpack that last read this staging buffer has finished
            st["stage"][s][:c].copy_(packed.codes[r0:r0 + c], non_blocking=True)
            st["done"][s].record(st["copy"])
        main.wait_event(st["done"][s])
        if packed.bits == 16:
            call("gridb200_depth_unpack_u16", ptr(st["stage"][s]), c, b, lds, depth.data_ptr() + r0 * ld * 4, ld,
                 ptr(esc_cell), ptr(esc_val), n_esc, r0, main.cuda_stream)
        else:
            call("gridb200_depth_unpack_bits", ptr(st["stage"][s]), c, b, lds, packed.bits,
                 (row_q.data_ptr() + r0 * 4) if row_q is not None else None, ptr(col_q), depth.data_ptr() + r0 * ld * 4, ld,
                 ptr(esc_cell), ptr(esc_val), n_esc, r0, main.cuda_stream)
        st["free"][s].record(main)
    return depth


def upload_packed16(packed: PackedDepth, out: Depth16, chunk_rows: int = 128) -> Depth16:
    """PackedDepth (host) -> the 2-byte device matrix `out` (preallocated codes [n, ld], ld = roundup8(b)).  16-bit host
    codes ARE the device format (plain copies); narrower widths are widened by gridb200_depth_unpack_bits_u16 chunk by
    chunk while the next chunk is in flight on a copy stream."""
    n, b = packed.n, packed.b
    codes = out.codes
    dev, ldo = codes.device, codes.stride(0)
    assert codes.shape[0] == n and out.b == b
    main = torch.cuda.current_stream(dev)
    out.row0, out.esc_b = 0, b
    if packed.bits == 16:
        out.set_escapes(packed.esc_cell, packed.esc_val)
        assert packed.codes.stride(0) == ldo, "16-bit host codes are copied as they are: same leading dimension"
        codes.copy_(packed.codes, non_blocking=True)
        return out
    lds = packed.codes.stride(0)
    chunk_rows = max(1, min(chunk_rows, n, 65535))
    key = (dev, chunk_rows, lds, packed.codes.dtype)
    st = _UPLOAD_STATE.get(key)
    if st is None:
        st = {"stage": [torch.empty((chunk_rows, lds), dtype=packed.codes.dtype, device=dev) for _ in range(2)],
              "copy": torch.cuda.Stream(device=dev), "done": [torch.cuda.Event() for _ in range(2)],
              "free": [torch.cuda.Event() for _ in range(2)]}
        _UPLOAD_STATE.clear()
        _UPLOAD_STATE[key] = st
    row_q = packed.row_q.to(dev, non_blocking=True) if packed.relative else None
    col_q = packed.col_q.to(dev, non_blocking=True) if packed.relative else None
    for s_ in range(2):
        st["free"][s_].record(main)
    for t, r0 in enumerate(range(0, n, chunk_rows)):
        s_, c = t & 1, min(chunk_rows, n - r0)
        with torch.cuda.stream(st["copy"]):
            st["copy"].wait_event(st["free"][s_])
            st["stage"][s_][:c].copy_(packed.codes[r0:r0 + c], non_blocking=True)
            st["done"][s_].record(st["copy"])
        main.wait_event(st["done"][s_])
        call("gridb200_depth_unpack_bits_u16", ptr(st["stage"][s_]), c, b, lds, packed.bits,
             (row_q.data_ptr() + r0 * 4) if row_q is not None else None, ptr(col_q), codes.data_ptr() + r0 * ldo * 2, ldo,
             main.cuda_stream)
        st["free"][s_].record(main)
    # escapes of the narrow host width that DO fit 16 bits go back into the matrix (a scatter); the rest -- depths from
    # 655.35 up, counted once on the host -- form the device's escape table
    out.esc_key = out.esc_h = out.esc_cols = None
    if packed.esc_cell.numel():
        cell = packed.esc_cell.to(dev, non_blocking=True)
        h = torch.round(packed.esc_val.to(dev, non_blocking=True).double() * 100.0).clamp(max=(1 << 22) - 1).to(torch.int32)
        flat = codes.view(torch.int16).view(-1)
        if packed.n_big == 0:
            flat[(cell // b) * ldo + (cell % b)] = h.to(torch.int16)   # (wraps: bit pattern)
        else:
            fits = h < 0xFFFF
            k = cell[fits]
            flat[(k // b) * ldo + (k % b)] = h[fits].to(torch.int16)
            out.set_escape_table(cell[~fits], h[~fits])
    return out


_UPLOAD_STATE: dict = {}
