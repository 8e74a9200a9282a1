"""Region coverage from mosdepth output (the arithmetic of step 3), drop-in for grid/utils/mosdepth.py:264-297.

The reference runs the `mosdepth` binary per CRAM and then reduces the sample's `regions.bed.gz` to one number,
`int(round(100 * overlap-weighted mean depth))`, with a Python loop over every line of the file.
Running the binary is out of scope; the reduction is done here for a whole batch of samples per launch:

    host   : inflate (thread pool), concatenate, upload
    device : gridb200_region_coverage -> per file the EXACT integers  H = sum hundredths * overlap,  covered bp,
             number of overlapping lines, number of lines outside the strict grammar
    host   : the answer is round(H / covered).  The reference forms it through a sequential fp64 sum, whose
             rounding noise matters only when H / covered is within (lines + 8) * 2^-52 relative of a .5 tie;
             exactly those files (and any file with a line outside the grammar, where the reference would raise
             or apply Python's int()/float() rules) are replayed with the reference's own loop.
"""
from __future__ import annotations

import gzip
import io
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

from .normalize_mosdepth import _inflate
from .utils import get_samples, log, progress_bar, setup_output_file


def _replay(text: bytes, chrom: str, start: int, end: int) -> int:
    """The reference's loop (mosdepth.py:279-297) on one inflated file, for the files the exact-integer decision
    cannot settle; raises exactly where the reference raises."""
    region_cov, covered_bp = 0.0, 0
    for line in io.TextIOWrapper(io.BytesIO(text)):
        r_chr, r_start, r_end, mean_cov = line.strip().split("\t")[:4]
        r_start, r_end, mean_cov = int(r_start), int(r_end), float(mean_cov)
        if r_chr != chrom:
            continue
        overlap = min(end, r_end) - max(start, r_start)
        if overlap > 0:
            region_cov += mean_cov * overlap
            covered_bp += overlap
    return int(round(100 * (region_cov / covered_bp))) if covered_bp > 0 else 0


def decide_coverage(weighted: int, covered: int, lines: int) -> int | None:
    """round(weighted / covered) when the reference's fp64 evaluation provably gives the same integer, else None.

    The reference computes fl(100 * fl(S / covered)) with S the sequential fp64 sum of fl(float(depth) * overlap):
    relative error below (lines + 3) * 2^-53.  The result can differ from the exact rounding only if the exact
    quotient q + rem / covered lies within that error of q + 1/2, i.e. unless
    |2 rem - covered| * 2^51 > (lines + 8) * weighted   (integers, evaluated exactly).
    """
    if covered <= 0:
        return 0
    if covered >= 1 << 40:
        return None
    q, rem = divmod(weighted, covered)
    if abs(2 * rem - covered) << 51 > (lines + 8) * weighted:
        return q + (1 if 2 * rem > covered else 0)
    return None


def region_coverage_batch(paths, chrom: str, start: int, end: int, threads: int = 1, batch_bytes: int = 1 << 28) -> list:
    """Coverage of [start, end) on `chrom` for every file of `paths`; an entry is the int the reference returns, or
    the Exception it raises (unreadable file, malformed line)."""
    import torch

    from .._lib import call, load, ptr

    load()  # no CPU fallback: fail before any work if the CUDA library is missing
    paths = [Path(p) for p in paths]
    want = chrom.encode()
    results: list = [None] * len(paths)
    if len(want) > 64:  # the reference compares whole strings; such a name never matches a real file's column
        want = None
    dev = torch.device("cuda")
    st = torch.cuda.current_stream().cuda_stream

    def flush(batch):
        if not batch:
            return
        sizes = [len(t) for _, t in batch]
        off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        host = torch.empty(max(int(off[-1]), 1), dtype=torch.uint8, pin_memory=True)
        view = host.numpy()
        for (_, t), a in zip(batch, off[:-1]):
            view[a:a + len(t)] = np.frombuffer(t, dtype=np.uint8)
        text = host.to(dev, non_blocking=True)
        file_off = torch.from_numpy(off).to(dev)
        out = torch.empty((len(batch), 4), dtype=torch.int64, device=dev)
        call("gridb200_region_coverage", ptr(text), int(off[-1]), ptr(file_off), len(batch), want, len(want), int(start),
             int(end), ptr(out), st)
        for (i, t), (weighted, covered, lines, irregular) in zip(batch, out.cpu().tolist()):
            res = None if irregular else decide_coverage(weighted, covered, lines)
            if res is None:
                try:
                    res = _replay(t, chrom, start, end)
                except Exception as e:  # what the reference raises on this file
                    res = e
            results[i] = res

    def inflate(i):
        t = _inflate(paths[i])
        return t if t is None or t.endswith(b"\n") or not t else t + b"\n"

    ahead = max(2, 2 * int(threads))
    batch, size = [], 0
    with ThreadPoolExecutor(max_workers=max(1, int(threads))) as ex:
        futs, nxt = [], 0
        for i in range(len(paths)):
            while nxt < len(paths) and len(futs) < ahead:
                futs.append(ex.submit(inflate, nxt))
                nxt += 1
            t = futs.pop(0).result()
            if t is None or want is None or b"\r" in t:  # (carriage returns end lines in text mode: rare enough to replay)
                try:  # let the reference's own failure mode surface (missing file, bad gzip, ...)
                    with gzip.open(paths[i], "rb") as fh:
                        results[i] = _replay(fh.read(), chrom, start, end)
                except Exception as e:
                    results[i] = e
                continue
            batch.append((i, t))
            size += len(t)
            if size >= batch_bytes:
                flush(batch)
                batch, size = [], 0
        flush(batch)
    return results


def compute_region_coverage(regions_file, chrom: str, start: int, end: int) -> int:
    """Same name, arguments, result and exceptions as mosdepth.py:264-297."""
    res = region_coverage_batch([regions_file], chrom, start, end)[0]
    if isinstance(res, Exception):
        raise res
    return res


def collect_region_coverage(config, console=None):
    """The part of compute_mosdepth (mosdepth.py:16-108) that follows the mosdepth runs: one `sample\\tcoverage` line per
    sample whose `{sample}*regions.bed.gz` exists in mosdepth.work_dir, under the reference's header, in sample
    order (the reference writes in thread-completion order).  Samples whose file is missing or malformed are
    reported and left out, as the reference does for a failed sample (:93-99)."""
    try:
        samples = get_samples(config["samples_file"])
        chrom, start, end = config.get("chrom", None), config.get("start_bp", None), config.get("end_bp", None)
        md = config.get("mosdepth", {})
        output_file = Path(f"{config.get('output_dir', '.')}/{md.get('output_file_prefix', None)}."
                           f"{config.get('output_file_type', 'tsv')}")
        threads = config.get("threads", 1)
        work_dir = Path(md.get("work_dir", None)).expanduser()
    except Exception as e:
        log(console, f"Config error: {e}", style="danger")
        return
    output_file = setup_output_file(output_file, chrom, start, end)
    found = {}
    for s in samples:
        matches = list(work_dir.glob(f"{s}*regions.bed.gz"))  # wait_for_mosdepth_output's pattern (:244)
        if matches:
            found[s] = matches[0]
        else:
            log(console, f"✗ {s} failed", style="danger")
    ids = list(found)
    with progress_bar(console, total=len(ids), description="Computing region coverage"):
        res = region_coverage_batch([found[s] for s in ids], chrom, start, end, threads=threads)
    with open(output_file, "a", newline="") as fh:
        for s, r in zip(ids, res):
            if isinstance(r, Exception):
                log(console, f"✗ {s} failed", style="danger")
            else:
                fh.write(f"{s}\t{r}\n")
    log(console, f"Mosdepth coverage results written to {output_file}", style="success")
