"""Read counting for target regions (step 2's arithmetic), after grid/utils/count_reads.py:82-108.

The reference walks `bam.fetch(chrom, start, end)` read by read in Python, once per region; batch dipCN over the 734
coding VNTR regions (BASELINE config 5) needs the N x 734 count matrix, i.e. 734 such passes per sample.  Decoding
CRAM/BAM stays on the host (htslib); what is data-parallel -- the flag / MAPQ / mate-reference / position filter and the
binning into regions -- runs on the GPU over columnar records, all regions in ONE pass:

    flag u16 | mapq u8 | tid i32 | next_tid i32 | pos i32      15 bytes per read

`gridb200_count_reads` histograms the passing reads over the elementary intervals between the sorted region
boundaries and turns prefix sums into per-region counts, so overlapping, nested and repeated regions are all exact.
"""
from __future__ import annotations

import glob
import os
from pathlib import Path

import numpy as np

from .utils import get_samples, log, progress_bar, setup_output_file

POS_LIMIT = (1 << 31) - 1


def flag_bitmap(proper_flags) -> np.ndarray:
    """65536-bit table of `read.flag in proper_flags` (count_reads.py:96): an exact-value test, not a mask."""
    bits = np.zeros(2048, dtype=np.uint32)
    for f in proper_flags:
        if isinstance(f, (int, np.integer)) and not isinstance(f, bool) and 0 <= int(f) < 65536:
            bits[int(f) >> 5] |= np.uint32(1 << (int(f) & 31))
    return bits


def region_table(regions):
    """Sorted distinct boundary keys (tid << 32 | position) and, per region, the indices of its start and end key.
    Region = (tid, start, end), counted positions start <= pos < end (count_reads.py:101-105)."""
    keys_lo, keys_hi = [], []
    for tid, start, end in regions:
        start, end = max(0, int(start)), min(int(end), POS_LIMIT)
        if int(tid) < 0 or end <= start:
            start = end = 0
            tid = 0
        keys_lo.append((int(tid) << 32) | start)
        keys_hi.append((int(tid) << 32) | end)
    bounds = np.unique(np.asarray(keys_lo + keys_hi, dtype=np.int64))
    lo = np.searchsorted(bounds, np.asarray(keys_lo, dtype=np.int64)).astype(np.int32)
    hi = np.searchsorted(bounds, np.asarray(keys_hi, dtype=np.int64)).astype(np.int32)
    return bounds, lo, hi


MAX_BUCKETS = 1 << 17  # directory entries (512 KB): small enough to live in L1/L2 next to the read stream


def bucket_directory(bounds: np.ndarray):
    """(first int32[n_tid * per_tid + 1], shift, per_tid, n_tid) over the sorted boundary keys: first[b] = number of keys
    below the first key of bucket b = tid * per_tid + (pos >> shift).  None when there is no boundary at all."""
    if bounds.size == 0:
        return None
    n_tid = int(bounds[-1] >> 32) + 1
    max_pos = int((bounds & 0xFFFFFFFF).max())
    shift = 10
    while n_tid * ((max_pos >> shift) + 2) > MAX_BUCKETS and shift < 30:
        shift += 1
    per_tid = (max_pos >> shift) + 2          # one bucket past the last boundary: every region end lies inside
    if n_tid * per_tid > (1 << 24):
        return None                            # absurd reference ids: plain binary search
    b = np.arange(n_tid * per_tid + 1, dtype=np.int64)
    starts = ((b // per_tid) << 32) | ((b % per_tid) << shift)
    return np.searchsorted(bounds, starts, side="left").astype(np.int32), shift, per_tid, n_tid


def count_reads_columns(cols, sample_off, regions, proper_flags, min_mapq: int = 1, use_directory: bool = True):
    """counts[ns, G] (int64 tensor on the GPU) for ns samples' reads given as five columns (numpy arrays or CUDA
    tensors: flag u16, mapq u8, tid i32, next_tid i32, pos i32), sample s owning reads sample_off[s]:sample_off[s+1]."""
    import torch

    from .. import _lib
    from .._lib import call, ptr

    lib = _lib.load()
    dev = torch.device("cuda")
    want = (torch.uint16, torch.uint8, torch.int32, torch.int32, torch.int32)

    def dev_col(c, dt):
        if not isinstance(c, torch.Tensor):
            c = torch.from_numpy(np.ascontiguousarray(c))
        if c.dtype != dt:
            raise TypeError(f"read column has dtype {c.dtype}, expected {dt}")
        c = c.to(dev).contiguous()
        return c.clone() if c.data_ptr() % 16 else c  # a slice of a larger column: the kernel loads 16 bytes at a time

    flag, mapq, tid, ntid, pos = (dev_col(c, dt) for c, dt in zip(cols, want))
    n = flag.numel()
    if not (mapq.numel() == tid.numel() == ntid.numel() == pos.numel() == n):
        raise ValueError("read columns differ in length")
    off = np.asarray(sample_off, dtype=np.int64)
    ns = off.size - 1
    if ns < 1 or off[0] != 0 or off[-1] != n or np.any(np.diff(off) < 0):
        raise ValueError("sample_off must rise from 0 to the number of reads")
    g = len(regions)
    counts = torch.zeros((ns, g), dtype=torch.int64, device=dev)
    if g == 0:
        return counts
    bounds, lo, hi = region_table(regions)
    mm = int(np.clip(int(min_mapq), -(1 << 31), (1 << 31) - 1))
    d = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    off_d, bits_d, bounds_d, lo_d, hi_d = d(off), d(flag_bitmap(proper_flags)), d(bounds), d(lo), d(hi)
    ws = torch.empty(lib.gridb200_count_reads_ws_bytes(ns, bounds.size), dtype=torch.uint8, device=dev)
    directory = bucket_directory(bounds) if use_directory else None
    first_d, shift, per_tid, n_tid = (d(directory[0]), *directory[1:]) if directory else (None, 0, 0, 0)
    call("gridb200_count_reads", ptr(flag), ptr(mapq), ptr(tid), ptr(ntid), ptr(pos), n, ptr(off_d), ns, ptr(bits_d), mm,
         ptr(bounds_d), int(bounds.size), ptr(lo_d), ptr(hi_d), g, ptr(first_d), shift, per_tid, n_tid, ptr(counts),
         ptr(ws), ws.numel(), torch.cuda.current_stream().cuda_stream)
    return counts


def read_columns_pysam(cram_file, ref_fasta, chrom, start, end):
    """Decode the reads `bam.fetch(chrom, start, end)` yields (count_reads.py:93-94) into the five columns.
    Needs pysam (htslib); there is no substitute decoder in this package."""
    import pysam  # noqa: WPS433 -- optional dependency of the reference as well (pyproject.toml)

    flag, mapq, tid, ntid, pos = [], [], [], [], []
    with pysam.AlignmentFile(cram_file, "rc", reference_filename=ref_fasta) as bam:
        region_tid = bam.get_tid(chrom)
        for read in bam.fetch(chrom, start, end):
            flag.append(read.flag), mapq.append(read.mapq), tid.append(read.reference_id)
            ntid.append(read.next_reference_id), pos.append(read.reference_start)
    return (np.asarray(flag, dtype=np.uint16), np.asarray(mapq, dtype=np.uint8), np.asarray(tid, dtype=np.int32),
            np.asarray(ntid, dtype=np.int32), np.asarray(pos, dtype=np.int32)), region_tid


def count_reads_in_region(cram_file, ref_fasta, chrom, start, end, proper_flags, min_mapq: int = 1, reader=None) -> int:
    """Same name, arguments and result as count_reads.py:82-108; `reader(cram_file, ref_fasta, chrom, start, end)`
    -> (columns, tid of chrom) replaces the pysam decode (tests, other decoders)."""
    cols, region_tid = (reader or read_columns_pysam)(cram_file, ref_fasta, chrom, start, end)
    n = len(cols[0])
    return int(count_reads_columns(cols, [0, n], [(region_tid, start, end)], proper_flags, min_mapq)[0, 0])


def count_reads(config, console=None, reader=None):
    """Drop-in for count_reads(config, console) (count_reads.py:14-79): same config keys, header and `sample\\tcount`
    rows (in sample order; the reference writes in thread-completion order); a sample whose file cannot be decoded
    gets `Error`, as process_single_cram returns (:140-142).  All samples' reads are counted in one launch."""
    try:
        directory_loc = config["directory_loc"]
        samples = get_samples(config["samples_file"])
        chrom, start, end = config.get("chrom", None), config.get("start_bp", None), config.get("end_bp", None)
        flags = config.get("count_reads", {}).get("flags", [])
        min_mapq = config.get("min_mapq", 1)
        output_file = Path(f"{config.get('output_dir', '.')}/{config.get('count_reads', {}).get('output_file_prefix', None)}."
                           f"{config.get('output_file_type', 'tsv')}")
        ref = config.get("reference_genome")
        file_type = config.get("file_type")
    except Exception as e:
        log(console, f"Config error: {e}", style="danger")
        return
    output_path = setup_output_file(output_file, chrom, start, end)
    files = {}
    for s in samples:  # find_file (utils.py:46-53): first match of *{sample}*.{type}; samples without a file are skipped
        matches = glob.glob(os.path.join(directory_loc, f"*{s}*.{file_type}")) if file_type else []
        if matches:
            files[s] = matches[0]
    decoded, failed, region_tid = [], set(), None
    with progress_bar(console, total=len(files), description="Counting reads"):
        for s, path in files.items():
            try:
                cols, t = (reader or read_columns_pysam)(path, ref, chrom, start, end)
                region_tid = t if region_tid is None else region_tid
                decoded.append((s, cols, t))
            except Exception as e:
                log(console, f"Failed to count reads for {Path(path).name}: {e}", style="danger")
                failed.add(s)
    counts = {}
    if decoded:
        off = np.concatenate([[0], np.cumsum([len(c[0]) for _, c, _ in decoded])]).astype(np.int64)
        cat = [np.concatenate([np.asarray(c[k]) for _, c, _ in decoded]) for k in range(5)]
        regions = sorted({t for _, _, t in decoded})
        res = count_reads_columns(cat, off, [(t, start, end) for t in regions], flags, min_mapq).cpu().numpy()
        for row, (s, _, t) in zip(res, decoded):
            counts[s] = int(row[regions.index(t)])
    with open(output_path, "a") as fh:
        for s in files:
            fh.write(f"{s}\t{'Error' if s in failed else counts[s]}\n")
    log(console, f"Read counting completed. Results written to {output_path}", style="success")
