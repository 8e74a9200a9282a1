"""CPU restatement of the reference's step 7, haplotype-informed inference (grid/utils/hi_inference.py) -- TEST
INFRASTRUCTURE, like grid_oracle.py: only tests/, smoke() and bench legs may import it.

Pinned against the unmodified reference: tests/golden/hi/ (written by oracle/gen_golden.py hi).
The algorithm is an in-place Gauss-Seidel iteration over samples in file order, so every function here is a plain
Python loop in the reference's operation order; results are compared as the %.2f text the reference writes and as
exact doubles.
"""

from __future__ import annotations

import gzip
import math
from collections import defaultdict
from pathlib import Path


def _open(path, mode="rt"):
    return gzip.open(path, mode) if str(path).endswith(".gz") else open(path, mode)


def read_dip_cn(path):
    """ids, values, id -> row; lines whose second field is not a number (the header) are skipped
    (hi_inference.py:10-31).  A repeated id gets a new row and the map points at the last one."""
    ids, vals, index = [], [], {}
    with _open(path) as fh:
        for raw in fh:
            f = raw.split()
            if len(f) < 2:
                continue
            try:
                v = float(f[1])
            except ValueError:
                continue
            index[f[0]] = len(vals)
            ids.append(f[0])
            vals.append(v)
    return ids, vals, index


def load_ibs(path, index, max_nbr):
    """computeIBSpbwt rows `ID hap nbrInd cMlen cMedge IDnbr hapNbr` after one header line; haplotypes are 1/2;
    the first max_nbr accepted rows per haplotype are kept, weight 1 (hi_inference.py:34-75)."""
    n = len(index)
    out = [[] for _ in range(2 * n)]
    with _open(path) as fh:
        next(fh)
        for raw in fh:
            f = raw.split()
            if len(f) < 7:
                continue
            try:
                h, hn = int(f[1]), int(f[6])
            except ValueError:
                continue
            if h not in (1, 2) or hn not in (1, 2):
                continue
            i, j = index.get(f[0]), index.get(f[5])
            if i is None or j is None:
                continue
            slot = out[2 * i + h - 1]
            if len(slot) < max_nbr:
                slot.append((2 * j + hn - 1, 1.0))
    return out


def load_ibd(path, index, max_nbr, region_start, region_end, min_length=0.5, min_match=0.70, weighted=False,
             weight_scale=1_000_000):
    """iLASH rows `FID1 HAP1 FID2 HAP2 CHR BP1 BP2 SNP1 SNP2 LENGTH MATCH` (tab- or blank-separated, no header);
    HAP = `<fid>_<0|1>`.  Segments shorter than min_length cM or below min_match are dropped; each segment is an
    edge in BOTH directions; per haplotype the max_nbr longest (stable sort) are kept; weight 1 or
    scale/(distance to the region + scale) * match (hi_inference.py:78-176)."""
    n = len(index)
    raw = defaultdict(list)
    with _open(path) as fh:
        for line in fh:
            line = line.strip()
            if not line:
                continue
            f = line.split("\t")
            if len(f) < 11:
                f = line.split()
            if len(f) < 11:
                continue
            try:
                bp1, bp2, length, match = int(f[5]), int(f[6]), float(f[9]), float(f[10])
            except (ValueError, IndexError):
                continue
            if length < min_length or match < min_match:
                continue
            try:
                h1, h2 = int(f[1].rsplit("_", 1)[-1]), int(f[3].rsplit("_", 1)[-1])
            except ValueError:
                continue
            if h1 not in (0, 1) or h2 not in (0, 1):
                continue
            i, j = index.get(f[0]), index.get(f[2])
            if i is None or j is None:
                continue
            if weighted:
                if bp2 < region_start:
                    dist = float(region_start - bp2)
                elif bp1 > region_end:
                    dist = float(bp1 - region_end)
                else:
                    dist = 0.0
                w = (weight_scale / (dist + weight_scale)) * match
            else:
                w = 1.0
            a, b = 2 * i + h1, 2 * j + h2
            raw[a].append((b, w, length))
            raw[b].append((a, w, length))
    out = [[] for _ in range(2 * n)]
    for h, segs in raw.items():
        segs.sort(key=lambda s: -s[2])
        out[h] = [(nb, w) for nb, w, _ in segs[:max_nbr]]
    return out


def _neighbour_means(i, hap, nbrs):
    sums = []
    for h in range(2):
        wsum, wval = 1e-9, 0.0
        for nb, w in nbrs[2 * i + h]:
            v = hap[nb]
            if not math.isnan(v):
                wsum += w
                wval += w * v
        sums.append((wval, wsum))
    return sums


def run_phasing(irr, nbrs, min_nbr, n_iters):
    """(hap values [2N], mean IRR of the phased samples).  A sample takes part when both haplotypes have at least
    min_nbr neighbours; it starts at IRR/2 per haplotype; every sweep visits the samples in order and replaces the
    pair IN PLACE by IRR * m_h / (m_0 + m_1), m_h = weighted mean of the neighbours' current values
    (hi_inference.py:179-233)."""
    n = len(irr)
    hap = [float("nan")] * (2 * n)
    phased, mean = 0, 0.0
    for i in range(n):
        if len(nbrs[2 * i]) >= min_nbr and len(nbrs[2 * i + 1]) >= min_nbr:
            hap[2 * i] = hap[2 * i + 1] = irr[i] / 2
            phased += 1
            mean += irr[i]
    if phased:
        mean /= phased
    for _ in range(n_iters):
        for i in range(n):
            if math.isnan(hap[2 * i]):
                continue
            (v0, s0), (v1, s1) = _neighbour_means(i, hap, nbrs)
            m0, m1 = v0 / s0, v1 / s1
            d = m0 + m1
            if d > 0:
                hap[2 * i] = irr[i] * m0 / d
                hap[2 * i + 1] = irr[i] * m1 / d
    return hap, mean


def impute(i, hap, nbrs, mean):
    """Neighbour means after the last sweep; mean/2 when a haplotype has no phased neighbour (hi_inference.py:236-256)."""
    (v0, s0), (v1, s1) = _neighbour_means(i, hap, nbrs)
    a, b = v0 / s0, v1 / s1
    if s0 <= 1e-9:
        a = mean / 2
    if s1 <= 1e-9:
        b = mean / 2
    return a, b


def hi_inference(config: dict) -> Path | None:
    """File-level restatement of hi_inference(config, console) (hi_inference.py:259-339)."""
    hc = config.get("compute_haploid_genotypes", {})
    ftype = config.get("output_file_type", "tsv")
    od = config.get("output_dir", ".")
    out = Path(f"{od}/{hc.get('output_file_prefix', 'haploid_genotypes')}.{ftype}")
    dip = Path(f"{od}/{config['compute_diploid_genotypes'].get('output_file_prefix')}.{ftype}")
    method = hc.get("method", "ibs").lower()
    min_nbr, max_nbr, n_iters = hc.get("min_neighbors", 1), hc.get("max_neighbors", 10), hc.get("n_iters", 100)
    ids, irr, index = read_dip_cn(dip)
    if method == "ibs":
        if not hc.get("ibs_output"):
            return None
        nbrs = load_ibs(hc["ibs_output"], index, max_nbr)
    elif method == "ibd":
        if not hc.get("ibd_output"):
            return None
        nbrs = load_ibd(hc["ibd_output"], index, max_nbr, config.get("start_bp"), config.get("end_bp"),
                        min_length=hc.get("min_length", 0.5), min_match=hc.get("min_match", 0.70),
                        weighted=hc.get("weighted", False), weight_scale=hc.get("weight_scale", 1_000_000))
    else:
        return None
    hap, mean = run_phasing(irr, nbrs, min_nbr, n_iters)
    with _open(out, "wt") as fh:
        fh.write("ID\tIRRs\thap1phased\thap2phased\thap1imp\thap2imp\n")
        for i in range(len(irr)):
            a, b = impute(i, hap, nbrs, mean)
            fh.write(f"{ids[i]}\t{irr[i]:.2f}\t{hap[2 * i]:.2f}\t{hap[2 * i + 1]:.2f}\t{a:.2f}\t{b:.2f}\n")
    return out
