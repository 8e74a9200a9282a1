"""Step 4, drop-in for grid/utils/normalize_mosdepth.py: same entry point, config keys, defaults, output file.

Host work: find and inflate the per-sample mosdepth files, apply the line filters, pack the fp32 depth matrix.
Device work (libgridb200): population-mean depth filter, per-sample means, per-bin statistics, median rescale,
variance selection, transform and printf-exact quantisation.
"""
from __future__ import annotations

import gzip
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

from .. import io as gio
from ..io import (find_bed_gz_for_individual, find_bed_gz_for_individuals, load_repeat_mask,  # noqa: F401  (reference names)
                  map_mosdepth_files_to_samples, norm_chrom)
from .utils import get_samples, log, progress_bar, setup_output_file


def pack_depth_matrix(individuals: dict, mosdepth_dir, chrom, start, end, excluded, threads: int = 1, console=None):
    """Parse every sample once and pack depth[N, B] (fp32, 0 = absent) over the union of bins seen.

    Rows follow sorted sample ids, columns sorted (start, end) -- the order build_matrix_from_regions uses
    (normalize_mosdepth.py:379-416).  Unreadable files contribute nothing (:286-287, :353-355).
    """
    ids = sorted(individuals)
    path_of = dict(zip(ids, find_bed_gz_for_individuals(ids, mosdepth_dir)))

    def one(sid):
        path = path_of[sid]
        if not path.exists():
            return None
        try:
            with gzip.open(path, "rb") as fh:
                return gio.parse_regions_text(fh.read(), chrom, start, end, excluded)
        except Exception:
            return None

    with progress_bar(console, total=len(ids), description="Extracting per-sample regions...") as (progress, task):
        with ThreadPoolExecutor(max_workers=max(1, threads)) as ex:
            parsed = []
            for res in ex.map(one, ids):
                parsed.append(res)
                progress.update(task, advance=1)
    keys = [np.stack([p[0], p[1]], axis=1) for p in parsed if p is not None and len(p[0])]
    if not keys:
        return ids, np.zeros((0, 2), dtype=np.int64), np.zeros((len(ids), 0), dtype=np.float32)
    bins = np.unique(np.concatenate(keys), axis=0)  # sorted by (start, end)
    code = {(int(s), int(e)): j for j, (s, e) in enumerate(bins)}
    depth = np.zeros((len(ids), (len(bins) + 3) // 4 * 4), dtype=np.float32)
    for i, p in enumerate(parsed):
        if p is None:
            continue
        for s, e, h in zip(p[0].tolist(), p[1].tolist(), p[2].tolist()):
            depth[i, code[(s, e)]] = h / 100.0  # later duplicates overwrite: last wins
    return ids, bins, depth


def _inflate(path) -> bytes | None:
    try:
        with gzip.open(path, "rb") as fh:
            return fh.read()
    except Exception:
        return None


def _last_bin_start(text: bytes) -> int:
    """Start coordinate of the last parseable line (to size the bin grid when no window is configured)."""
    for line in reversed(text[-4096:].split(b"\n")):
        f = line.split(b"\t")
        if len(f) >= 4 and f[1].isdigit():
            return int(f[1])
    return -1


def pack_depth_matrix_device(individuals: dict, mosdepth_dir, chrom, start, end, excluded, threads: int = 1,
                             console=None):
    """Device ingest (gridb200_ingest_text): the host only inflates; parsing, filtering and packing run on the GPU.

    Applies when the bins form the regular 1000-bp grid `mosdepth --by 1000` writes (mosdepth.py:193-225).  Returns
    None when they do not (or a file needs Python's int()/float() semantics everywhere) -- the caller then packs
    on the host with pack_depth_matrix; individual irregular files are re-parsed on the host into their row.
    Result: (ids, bins[(start, end)], depth tensor [N, ld] on the GPU).
    """
    import torch

    from .. import _lib
    from .._lib import call, ptr

    ids = sorted(individuals)
    want = norm_chrom(chrom).encode() if chrom else b""
    if len(want) > 32:
        return None
    paths = find_bed_gz_for_individuals(ids, mosdepth_dir)
    cached: dict[int, bytes | None] = {}  # texts inflated ahead of the stream (to size the grid): used once, then dropped

    def inflate(i):
        if i in cached:
            return cached.pop(i)
        return _inflate(paths[i]) if paths[i].exists() else None

    def stream():
        """(i, text) in sample order, at most 2 * threads files inflated ahead: host memory stays bounded however
        large the cohort is (config 2 is 3 202 files of ~90 MB of text each)."""
        ahead = max(2, 2 * int(threads))
        with ThreadPoolExecutor(max_workers=max(1, int(threads))) as ex:
            futs, nxt = [], 0
            for i in range(len(paths)):
                while nxt < len(paths) and len(futs) < ahead:
                    futs.append(ex.submit(inflate, nxt))
                    nxt += 1
                yield i, futs.pop(0).result()

    if start is not None and end is not None:
        bin0 = max(0, -(-(start - 1000) // 1000)) * 1000       # first 1000-aligned bin whose end reaches `start`
        last = (end // 1000) * 1000
    else:
        # no window: the grid ends where the files end.  Every sample of a cohort is binned over the same reference,
        # so the first few files tell; a later file that reaches further sends the cohort to the host packer (below).
        bin0, last = 0, -1
        for i in range(len(paths)):
            if len(cached) >= 4:
                break
            if paths[i].exists():
                cached[i] = _inflate(paths[i])
                if cached[i]:
                    last = max(last, _last_bin_start(cached[i]))
    if last < bin0:
        return None
    b = (last - bin0) // 1000 + 1
    ld = (b + 3) // 4 * 4
    dev = torch.device("cuda")
    depth = torch.zeros((len(ids), ld), dtype=torch.float32, device=dev)
    stats = torch.zeros((len(ids), 8), dtype=torch.int64, device=dev)
    ws = torch.empty(_lib.load().gridb200_ingest_ws_bytes(b), dtype=torch.uint8, device=dev)
    # the mask is looked up by the LINE's chromosome (:281-283); one bitset serves when every kept line carries
    # exactly `chrom` (checked below through stats[4]) -- genome-wide masks go through the host packer
    mask_bits, mask_kb = None, 0
    if excluded:
        if not chrom:
            return None
        kbs = excluded.get(norm_chrom(chrom), set())
        if kbs:
            mask_kb = max(kbs) + 1
            bits = np.zeros((mask_kb + 31) // 32, dtype=np.uint32)
            for kb in kbs:
                bits[kb >> 5] |= np.uint32(1 << (kb & 31))
            mask_bits = torch.from_numpy(bits).to(dev)
    st = torch.cuda.current_stream().cuda_stream
    keep_alive = []
    with progress_bar(console, total=len(ids), description="Extracting per-sample regions...") as (progress, task):
        for i, text in stream():
            if text:
                if _last_bin_start(text) > last and not (start is not None and end is not None):
                    return None  # this file reaches beyond the grid sized from the first ones
                t = torch.frombuffer(bytearray(text), dtype=torch.uint8).to(dev, non_blocking=False)
                keep_alive.append(t)
                call("gridb200_ingest_text", ptr(t), len(text), want, len(want),
                     start if start is not None and end is not None else -1, end if end is not None else -1,
                     ptr(mask_bits), mask_kb, bin0, b, depth.data_ptr() + i * ld * 4, stats.data_ptr() + i * 64,
                     ptr(ws), ws.numel(), st)
                if len(keep_alive) > 4:
                    torch.cuda.current_stream().synchronize()
                    keep_alive.clear()
            progress.update(task, advance=1)
    hs = stats.cpu().numpy()
    redo = np.nonzero((hs[:, 1] > 0) | ((hs[:, 4] > 0) & bool(excluded)))[0]
    for i in redo.tolist():  # exotic file: the reference's own parsing rules, on the host, into the same row
        try:
            s_, e_, h_ = gio.parse_regions_text(_inflate(paths[i]), chrom, start, end, excluded)
        except Exception:
            s_ = e_ = h_ = np.zeros(0, dtype=np.int64)  # the reference drops a file it cannot parse (:286-287)
        rel = s_ - bin0
        if len(s_) and (np.any(rel % 1000) or np.any(rel < 0) or np.any(rel // 1000 >= b) or np.any(e_ - s_ > 1000)
                        or np.any(e_ <= s_)):
            return None  # bins off the regular grid: only the host packer can represent them
        row = np.zeros(ld, dtype=np.float32)
        row[rel // 1000] = h_ / 100.0  # later duplicates overwrite: last wins
        depth[i] = torch.from_numpy(row).to(dev)
    bins = np.stack([bin0 + 1000 * np.arange(b, dtype=np.int64), bin0 + 1000 * np.arange(1, b + 1, dtype=np.int64)], axis=1)
    return ids, bins, depth


def normalize_mosdepth(config, console):
    """Same contract as the reference's normalize_mosdepth(config, console) (normalize_mosdepth.py:23-145)."""
    try:
        samples = get_samples(config["samples_file"])
        chrom = config.get("chrom", None)
        start = config.get("start_bp", None)
        end = config.get("end_bp", None)
        threads = config.get("threads", 1)
        ncfg = config["mosdepth"]["normalize"]
        prefix = config.get("mosdepth", {}).get("normalize", {}).get("output_file_prefix", None)
        ftype = config.get("output_file_type", "tsv")
        output_file = Path(f"{config.get('output_dir', '.')}/{prefix}.{ftype}.gz")
        remove_intermediate = config.get("mosdepth", {}).get("remove_intermediate", False)
        mosdepth_dir = config.get("mosdepth", {}).get("work_dir", None)
        min_depth = ncfg.get("min_depth", 20)
        max_depth = ncfg.get("max_depth", 100)
        top_frac = ncfg.get("top_frac", 0.1)
        repeat_mask = ncfg.get("repeat_mask_file", None)
    except Exception as e:
        log(console, f"[red]Config error: {e}[/red]")
        return

    output_path = setup_output_file(output_file, chrom, start, end)
    individuals = map_mosdepth_files_to_samples(mosdepth_dir, samples)
    if not individuals:
        log(console, f"✗ No mosdepth files found in {mosdepth_dir}", style="danger")
        sys.exit(1)
    excluded = load_repeat_mask(repeat_mask)

    import torch

    from .. import engine

    packed = None
    if config.get("gpu", {}).get("device_ingest", True):  # additive key; default = parse on the GPU
        packed = pack_depth_matrix_device(individuals, mosdepth_dir, chrom, start, end, excluded, threads, console)
    if packed is None:
        ids, bins, host_depth = pack_depth_matrix(individuals, mosdepth_dir, chrom, start, end, excluded, threads, console)
        depth = torch.from_numpy(host_depth).cuda()
    else:
        ids, bins, depth = packed
    b = len(bins)
    if b == 0:
        log(console, "No valid samples with regions found.", style="danger")
        sys.exit(1)
    p = engine.Params(min_depth=min_depth, max_depth=max_depth, top_frac=top_frac)
    from .. import multigpu

    devices = multigpu.configured_devices(config)        # additive key gpu.devices: bin-sharded over several GPUs
    if devices and b >= 4 * len(devices):
        from types import SimpleNamespace

        log(console, f"Normalising on {len(devices)} GPUs (bin-sharded)", style="info")
        zq_h, rw_, mu_h, ratio_h, raw_h, rowcnt_h = multigpu.normalize(depth.cpu(), b, p, devices)
        sel_pos = torch.zeros(max(rw_, 1), dtype=torch.int32)  # only its length of non-negative entries is used below
        zq_d = None
        if rw_:  # same layout as engine.normalize's: rows padded to a multiple of 8 codes with NA
            zq_d = torch.full((zq_h.shape[0], (rw_ + 7) // 8 * 8), gio.ZQ_NA, dtype=torch.int16, device=depth.device)
            zq_d[:, :rw_] = zq_h.to(depth.device)
        res = SimpleNamespace(extra={"rowcnt": torch.from_numpy(rowcnt_h)}, sel_pos=sel_pos[:rw_], rw=rw_,
                              mu=torch.from_numpy(mu_h), ratio=torch.from_numpy(ratio_h), raw_mean=torch.from_numpy(raw_h),
                              zq=zq_d)
    else:
        res = engine.normalize(depth, b, p, want_zq=True, want_zh=False)
    rows = np.nonzero(res.extra["rowcnt"].cpu().numpy() > 0)[0]  # filter_empty_samples (:576-600)
    if len(rows) < len(ids):
        log(console, f"Removed {len(ids) - len(rows)} samples with 0 regions", style="warning")
    if len(rows) == 0:
        log(console, "No valid samples with regions found.", style="danger")
        sys.exit(1)
    sel = np.nonzero(res.sel_pos.cpu().numpy() >= 0)[0]
    mu_sel, ratio_sel = res.mu.cpu().numpy()[sel], res.ratio.cpu().numpy()[sel]
    raw_means = res.raw_mean.cpu().numpy()[rows]
    kept_ids = [ids[i] for i in rows]
    if config.get("gpu", {}).get("device_text", True) and res.zq is not None:  # additive key; default = format on the GPU
        from .. import textio

        zq_dev = res.zq if len(rows) == len(ids) else res.zq[torch.from_numpy(rows).to(res.zq.device)]
        textio.write_normalized_output(output_path, kept_ids, zq_dev, res.rw, mu_sel, ratio_sel, raw_means, threads=threads,
                                       level=config.get("gpu", {}).get("gzip_level"))
    else:
        zq = res.zq.cpu().numpy()[rows][:, : res.rw] if res.zq is not None else np.zeros((len(rows), 0), dtype=np.int16)
        gio.write_normalized_output(output_path, kept_ids, zq, mu_sel, ratio_sel, raw_means)
    from .. import handoff

    if handoff.enabled(config) and res.zq is not None and len(rows) == len(ids):
        # what read_normalized_data would hand step 5: ids, the %.3f ratios, the codes ("-0.00" reads as 0), %.2f scales
        res.zq.masked_fill_(res.zq == gio.ZQ_NEG_ZERO, 0)
        handoff.put(output_path, ids=kept_ids, zq=res.zq, rw=res.rw,
                    ratios=np.array([float("nan") if v != v else float("%.3f" % v) for v in ratio_sel.tolist()]),
                    scales=np.array([float("%.2f" % v) for v in raw_means.tolist()]))
    log(console, f"Mosdepth normalization complete. Results written to {output_path}", style="success")

    if remove_intermediate:  # mosdepth.remove_intermediate_files(dir, console, include_region_bed_gz=True), mosdepth.py:300-326
        suffixes = ("mosdepth.global.dist.txt", "mosdepth.region.dist.txt", "regions.bed.gz.csi", "regions.bed.gz")
        for f in Path(mosdepth_dir).glob("*"):
            if f.name.endswith(suffixes):
                try:
                    f.unlink()
                except Exception as e:
                    log(console, f"Failed to remove intermediate file {f}: {e}", style="warning")
