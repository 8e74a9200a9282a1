"""CPU restatement of the two data-parallel producers of the path's inputs.

TEST INFRASTRUCTURE ONLY: imported by tests/, never by grid_b200/.

  compute_region_coverage   mosdepth.py:264-297   overlap-weighted mean depth of a region, int(round(100 * mean))
  count_reads_columns       count_reads.py:82-108 the per-read filter of count_reads_in_region on columnar records

Parity PINNED: tests/golden/step3.json holds what the unmodified reference returned for the seeded inputs of
grid_b200.synth.region_coverage_cases / synth.read_columns (oracle/gen_golden.py step3; pysam's AlignmentFile is
replaced by an in-memory reader that honours fetch()'s overlap contract), and tests/test_step3_oracle.py also
replays the reference's own known answers (test/test_mosdepth.py:23-56).
"""

from __future__ import annotations

import gzip
import io

import numpy as np


def compute_region_coverage_text(text: str, chrom: str, start: int, end: int) -> int:
    """mosdepth.py:279-297 on already-inflated text.  Raises what the reference raises on malformed lines."""
    region_cov = 0.0
    covered_bp = 0
    for line in io.StringIO(text, newline=None):  # text-mode iteration, universal newlines: what gzip.open(.., "rt") gives
        r_chr, r_start, r_end, mean_cov = line.strip().split("\t")[:4]          # :283
        r_start, r_end = int(r_start), int(r_end)                               # :284
        mean_cov = float(mean_cov)                                              # :285
        if r_chr != chrom:                                                      # :287
            continue
        overlap = min(end, r_end) - max(start, r_start)                         # :290-292
        if overlap > 0:
            region_cov += mean_cov * overlap                                    # :295
            covered_bp += overlap
    return int(round(100 * (region_cov / covered_bp))) if covered_bp > 0 else 0  # :297


def compute_region_coverage(regions_file, chrom: str, start: int, end: int) -> int:
    """mosdepth.py:264-297 (file form)."""
    region_cov = 0.0
    covered_bp = 0
    with gzip.open(regions_file, "rt") as f:
        for line in f:
            r_chr, r_start, r_end, mean_cov = line.strip().split("\t")[:4]
            r_start, r_end = int(r_start), int(r_end)
            mean_cov = float(mean_cov)
            if r_chr != chrom:
                continue
            overlap = min(end, r_end) - max(start, r_start)
            if overlap > 0:
                region_cov += mean_cov * overlap
                covered_bp += overlap
    return int(round(100 * (region_cov / covered_bp))) if covered_bp > 0 else 0


def count_reads_columns(flag, mapq, tid, next_tid, pos, region_tid: int, start: int, end: int, proper_flags,
                        min_mapq: int = 1) -> int:
    """count_reads.py:82-108 for the reads of ONE sample given as columns.

    `bam.fetch(chrom, start, end)` yields the reads on `chrom` whose alignment overlaps [start, end); every read
    that passes `start <= reference_start < end` (:103-105) is among them, so the count is the predicate below
    over the reads with tid == region_tid.
    """
    flag = np.asarray(flag).astype(np.int64)
    mapq = np.asarray(mapq).astype(np.int64)
    tid = np.asarray(tid).astype(np.int64)
    next_tid = np.asarray(next_tid).astype(np.int64)
    pos = np.asarray(pos).astype(np.int64)
    ok = np.isin(flag, np.asarray(sorted(set(int(f) for f in proper_flags if f is not None)), dtype=np.int64))  # :96
    ok &= mapq >= min_mapq                                                                                # :97
    ok &= tid == next_tid                                                                                 # :98
    ok &= (flag & 0x400) == 0                                                                             # :99
    ok &= (flag & 0x100) == 0                                                                             # :100
    ok &= (tid == region_tid) & (pos >= start) & (pos < end)                                              # :101-105
    return int(ok.sum())


def count_reads_regions(cols, sample_off, regions, proper_flags, min_mapq: int = 1) -> np.ndarray:
    """counts[ns, G]: count_reads_columns for every sample slice and region (the 734 calls of config 5)."""
    ns = len(sample_off) - 1
    out = np.zeros((ns, len(regions)), dtype=np.int64)
    for s in range(ns):
        a, b = int(sample_off[s]), int(sample_off[s + 1])
        sl = [np.asarray(c)[a:b] for c in cols]
        for g, (rt, lo, hi) in enumerate(regions):
            out[s, g] = count_reads_columns(*sl, rt, lo, hi, proper_flags, min_mapq)
    return out
