"""CPU oracle for the GRiD depth-inference hot path (TEST INFRASTRUCTURE ONLY).

This module is a numpy restatement of the reference's steps 4-6
(normalize_mosdepth -> find_neighbors -> compute_dipcn).  It is the checker the
parity tests, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs use.  Nothing under ``grid_b200/`` may import it: the
product path is the CUDA library and fails loudly when that is missing.

Parity pin: ``oracle/gen_golden.py`` imports the real reference from
``/root/reference`` (pysam stubbed the way the reference's own test/conftest.py:9-11
does), runs both on the same seeded inputs and commits the reference's outputs as
fixtures under ``tests/golden/``; ``tests/test_oracle_golden.py`` replays them
against this file on every run (no reference needed at test time).

Third-party arithmetic: the reference's kNN is scikit-learn's
``NearestNeighbors(algorithm="auto", metric="euclidean")`` (reference
grid/utils/find_neighbors.py:207-213; dependency ``scikit-learn>=1.0``,
pyproject.toml:32, un-vendored; 1.9.0 installed in this image).  ``knn_sklearn``
calls that same dependency; ``knn_exact`` restates its published brute-force
algorithm (||x||^2 - 2 x.y + ||y||^2, k smallest, lower index wins ties) in exact
integer arithmetic on the hundredths grid so it is usable where sklearn's fp64
round-off would make tie order arbitrary.

All ``file:line`` citations are into /root/reference/grid/utils/.
"""

from __future__ import annotations

import glob
import gzip
import math
import os
from pathlib import Path

import numpy as np

RATIO_MULT = 100.0


# --------------------------------------------------------------------------------------
# Step 4: normalize_mosdepth
# --------------------------------------------------------------------------------------
def norm_chrom(chrom: str) -> str:
    """'6' -> 'chr6'; names already starting with 'chr' pass through (normalize_mosdepth.py:210-215)."""
    return chrom if chrom.startswith("chr") else "chr" + chrom


def load_repeat_mask(repeat_bed) -> dict:
    """BED -> {chrN: set(kb)} with kb in [start//1000, end//1000] INCLUSIVE (normalize_mosdepth.py:177-207).

    Comment/blank lines, short lines and non-integer coordinates are skipped.  The
    reference opens the path unconditionally (a None path raises TypeError there);
    callers in this repo pass None only when no mask is configured and get {}.
    """
    mask: dict = {}
    if repeat_bed is None:
        return mask
    with open(repeat_bed) as fh:
        for raw in fh:
            if raw.startswith("#") or not raw.strip():
                continue
            tok = raw.strip().split()
            if len(tok) < 3:
                continue
            try:
                lo, hi = int(tok[1]), int(tok[2])
            except ValueError:
                continue
            mask.setdefault(norm_chrom(tok[0]), set()).update(range(lo // 1000, hi // 1000 + 1))
    return mask


def map_files_to_samples(mosdepth_dir, samples) -> dict:
    """Longest '_'-joined prefix of the file stem that is a sample id (normalize_mosdepth.py:148-174)."""
    wanted = set(samples)
    found = {}
    for path in Path(mosdepth_dir).glob("*.regions.bed.gz"):
        pieces = path.name.split(".")[0].split("_")
        for n in range(len(pieces), 0, -1):
            cand = "_".join(pieces[:n])
            if cand in wanted:
                found[cand] = path
                break
    return found


def find_bed_gz(sample_id, mosdepth_dir) -> Path:
    """Substring glob, first match (normalize_mosdepth.py:557-573)."""
    hits = list(Path(mosdepth_dir).glob(f"*{sample_id}*regions.bed.gz"))
    return hits[0] if hits else Path(mosdepth_dir) / f"{sample_id}.regions.bed.gz"


def scan_regions(bed_gz, chrom, start, end, excluded, valid_regions=None):
    """Line filter shared by both passes (normalize_mosdepth.py:262-285 and :321-352).

    Keeps (start, end, depth) when: the line starts with the normalised chromosome
    string (a PREFIX test), has >= 4 tab fields, depth > 0, the bin touches the
    closed window [start, end] (only when both bounds are given), the bin is in
    ``valid_regions`` when that set is given, and none of the kb bins
    [s//1000, e//1000] is masked for the line's chromosome.  Any exception while
    reading discards the whole file (returns None), as the reference does.
    """
    want = norm_chrom(chrom) if chrom else None
    kept = []
    try:
        with gzip.open(bed_gz, "rt") as fh:
            for line in fh:
                if want and not line.startswith(want):
                    continue
                f = line.strip().split("\t")
                if len(f) < 4:
                    continue
                c = norm_chrom(f[0])
                s, e, d = int(f[1]), int(f[2]), float(f[3])
                if not d > 0:
                    continue
                if start is not None and end is not None and not (e >= start and s <= end):
                    continue
                if valid_regions is not None and (s, e) not in valid_regions:
                    continue
                bad = excluded.get(c)
                if bad and any(kb in bad for kb in range(s // 1000, e // 1000 + 1)):
                    continue
                kept.append((s, e, d))
    except Exception:
        return None
    return kept


def population_mean_depths(sample_ids, mosdepth_dir, chrom, start, end, excluded) -> dict:
    """Pass 1: per-bin mean over the samples that HAVE the bin (normalize_mosdepth.py:218-301).

    Within one file a repeated (start,end) keeps its last depth (:285).  Sums are taken
    in ``sample_ids`` order (the reference's order is thread-completion order).
    """
    tot: dict = {}
    cnt: dict = {}
    for sid in sample_ids:
        path = find_bed_gz(sid, mosdepth_dir)
        if not path.exists():
            continue
        rows = scan_regions(path, chrom, start, end, excluded)
        if rows is None:
            continue
        last = {(s, e): d for s, e, d in rows}
        for key, d in last.items():
            tot[key] = tot.get(key, 0.0) + d
            cnt[key] = cnt.get(key, 0) + 1
    return {k: tot[k] / cnt[k] for k in tot if cnt[k] > 0}


def build_matrix(per_sample: dict):
    """rows = sorted ids, cols = sorted (start,end); NaN where a sample lacks a bin (normalize_mosdepth.py:379-416)."""
    ids = sorted(per_sample)
    bins = sorted({(s, e) for sid in ids for s, e, _ in per_sample[sid]})
    col = {b: j for j, b in enumerate(bins)}
    mat = np.full((len(ids), len(bins)), np.nan)
    for i, sid in enumerate(ids):
        for s, e, d in per_sample[sid]:
            mat[i, col[(s, e)]] = d
    return ids, bins, mat


def normalize_matrix(mat: np.ndarray):
    """fp64 within-row then across-column normalisation (normalize_mosdepth.py:419-476).

    Returns (z, ratio, mu, s2):  rows / nanmean(row) (0 -> NaN); mu = nanmean over rows;
    s2 = nansum((x-mu)^2)/(N-1) with N = ALL rows (:446); ratio = 100*s2/mu where mu>0 else
    NaN; z = (x-mu)/sqrt(mu) on mu>0 columns; finally z *= 1/sqrt(median(finite ratio)/100)
    when that median is > 0 (:460-470).  ``ratio`` is the dense array; the reference's dict
    is {j: ratio[j] for finite j}.
    """
    with np.errstate(all="ignore"):
        x = np.array(mat, dtype=np.float64)
        rmean = np.nanmean(x, axis=1)
        rmean = np.where(rmean == 0, np.nan, rmean)
        x = x / rmean[:, None]
        n = x.shape[0]
        mu = np.nanmean(x, axis=0)
        s2 = np.nansum(np.square(x - mu), axis=0) / (n - 1)
        pos = mu > 0
        ratio = np.where(pos, RATIO_MULT * s2 / mu, np.nan)
        x[:, pos] = (x[:, pos] - mu[pos]) / np.sqrt(mu[pos])
        finite = ratio[~np.isnan(ratio)]
        scale = 1.0
        if finite.size:
            med = float(np.median(finite))
            if med > 0:
                scale = 1.0 / np.sqrt(med / RATIO_MULT)
        x *= scale
    return x, ratio, mu, s2


def select_high_variance(ratio: np.ndarray, top_frac: float) -> np.ndarray:
    """Indices (ascending) with ratio STRICTLY above sorted(finite)[int(top_frac*len)] (normalize_mosdepth.py:479-499).

    top_frac=0.1 therefore DROPS the lowest ~10 % and keeps ~90 %.  Raises IndexError when
    int(top_frac*len) == len, like the reference's list indexing would.
    """
    ok = ~np.isnan(ratio)
    vals = np.sort(ratio[ok])
    if vals.size == 0:
        return np.zeros(0, dtype=np.int64)
    thr = vals[int(top_frac * vals.size)]
    return np.nonzero(ok & (ratio > thr))[0]


def fmt(v: float, nd: int) -> str:
    """'NA' for NaN else fixed-point with nd decimals (normalize_mosdepth.py:544-553)."""
    return "NA" if np.isnan(v) else f"{v:.{nd}f}"


def write_normalized(path, ids, z, sel, mu, s2, raw_means) -> None:
    """The .tsv.gz hand-off file (normalize_mosdepth.py:502-554): %.3f headers, %.2f scale and z."""
    sel = list(sel)
    with np.errstate(all="ignore"):
        m = mu[sel]
        r = np.where(m > 0, RATIO_MULT * s2[sel] / m, np.nan)
    with gzip.open(path, "wt") as out:
        out.write(f"{len(ids)}\t{len(sel)}\t" + "\t".join(fmt(v, 3) for v in m) + "\n")
        out.write(f"{len(ids)}\t{len(sel)}\t" + "\t".join(fmt(v, 3) for v in r) + "\n")
        for i, sid in enumerate(ids):
            out.write(f"{sid}\t{raw_means[i]:.2f}\t" + "\t".join(fmt(z[i, j], 2) for j in sel) + "\n")


def normalize_mosdepth(config: dict) -> Path | None:
    """File-level step 4 (normalize_mosdepth.py:23-145).  Returns the output path, or None when
    the reference would sys.exit(1) (no files / no samples with bins)."""
    samples = [ln.strip() for ln in open(config["samples_file"]) if ln.strip()]
    chrom, start, end = config.get("chrom"), config.get("start_bp"), config.get("end_bp")
    ncfg = config["mosdepth"]["normalize"]
    out = Path(f"{config.get('output_dir', '.')}/{ncfg.get('output_file_prefix')}.{config.get('output_file_type', 'tsv')}.gz")
    out.parent.mkdir(parents=True, exist_ok=True)
    work = config["mosdepth"].get("work_dir")
    files = map_files_to_samples(work, samples)
    if not files:
        return None
    excluded = load_repeat_mask(ncfg.get("repeat_mask_file"))
    pop = population_mean_depths(list(files), work, chrom, start, end, excluded)
    lo, hi = ncfg.get("min_depth", 20), ncfg.get("max_depth", 100)
    valid = {k for k, v in pop.items() if lo <= v <= hi}
    per_sample = {}
    for sid in files:
        path = find_bed_gz(sid, work)
        rows = scan_regions(path, chrom, start, end, excluded, valid) if path.exists() else []
        if rows:
            per_sample[sid] = rows
    if not per_sample:
        return None
    ids, _, mat = build_matrix(per_sample)
    raw_means = np.nanmean(mat, axis=1)
    z, ratio, mu, s2 = normalize_matrix(mat)
    sel = select_high_variance(ratio, ncfg.get("top_frac", 0.1))
    write_normalized(out, ids, z, sel, mu, s2, raw_means)
    return out


# --------------------------------------------------------------------------------------
# Step 5: find_neighbors
# --------------------------------------------------------------------------------------
def _num(tok: str) -> float:
    return np.nan if tok in ("NA", "nan") else float(tok)


def read_normalized(path):
    """Parse the step-4 file back (find_neighbors.py:81-124): ids, ratios[Rw], z[N,Rw], {id: scale}."""
    ids, scales, rows = [], {}, []
    with gzip.open(path, "rt") as fh:
        fh.readline()
        ratios = np.array([_num(t) for t in fh.readline().strip().split("\t")[2:]])
        for line in fh:
            p = line.strip().split("\t")
            ids.append(p[0])
            scales[p[0]] = float(p[1])
            rows.append([_num(t) for t in p[2:]])
    return ids, ratios, np.array(rows, dtype=float), scales


def filter_regions_by_variance(ratios: np.ndarray, frac_r: float = 1.0, sigma2_max: float = 1000.0):
    """finite & >= sorted(finite)[min(int(R*(1-frac_r)), n-1)] & <= sigma2_max; all-NaN keeps all (find_neighbors.py:128-175)."""
    r = len(ratios)
    fin = np.isfinite(ratios)
    vals = np.sort(ratios[fin])
    if vals.size == 0:
        return np.arange(r), r
    lo = float(vals[min(int(r * (1.0 - frac_r)), vals.size - 1)])
    with np.errstate(invalid="ignore"):
        keep = np.nonzero(fin & (ratios >= lo) & (ratios <= sigma2_max))[0]
    return keep, len(keep)


def knn_sklearn(x: np.ndarray, n_neighbors: int):
    """The reference's call into scikit-learn (find_neighbors.py:205-225), on index level.

    k' = min(k+1, N); kneighbors; drop j == i wherever it appears; keep the first k; the
    stored value is distance**2.  Returns (idx[N, m], d2[N, m]) with m = min(k, N-1) --
    except that when the row itself is not among its own k' nearest (exact duplicates),
    the reference keeps k' > m entries for that row; that ragged case is returned as lists.
    """
    from sklearn.neighbors import NearestNeighbors

    n = x.shape[0]
    kk = min(n_neighbors + 1, n)
    dist, ind = NearestNeighbors(n_neighbors=kk, algorithm="auto", metric="euclidean").fit(x).kneighbors(x)
    out_i, out_d = [], []
    for i in range(n):
        keep = [(int(j), float(d) ** 2) for j, d in zip(ind[i], dist[i]) if j != i][:n_neighbors]
        out_i.append([j for j, _ in keep])
        out_d.append([d for _, d in keep])
    return out_i, out_d


def knn_exact(zq: np.ndarray, n_neighbors: int, block: int = 256):
    """Exact-integer restatement of sklearn's brute-force ArgKmin for data on the 0.01 grid.

    ``zq`` holds the clipped z-scores in hundredths (integers, |v| <= 32767).  Squared
    distances are exact int64 (units of 1e-4); the k smallest per row are returned sorted by
    (d2, index) -- the order sklearn's max-heap produces when no two candidates tie (lower
    index wins at the k-th boundary, sklearn/utils/_heap.pyx).  Self is excluded by index, as
    find_neighbors.py:218-219 does.  Returns (idx[N,m] int32, d2[N,m] int64), m = min(k, N-1).
    """
    v = np.asarray(zq, dtype=np.int64)
    n = v.shape[0]
    m = min(n_neighbors, n - 1)
    sq = np.einsum("ij,ij->i", v, v)
    idx = np.empty((n, m), dtype=np.int32)
    d2o = np.empty((n, m), dtype=np.int64)
    vf = v.astype(np.float64)
    exact_in_f64 = v.shape[1] * float(np.abs(v).max(initial=0)) ** 2 < 2.0**53
    for a in range(0, n, block):
        b = min(n, a + block)
        g = (vf[a:b] @ vf.T).astype(np.int64) if exact_in_f64 else v[a:b] @ v.T
        d2 = sq[a:b, None] + sq[None, :] - 2 * g
        d2[np.arange(b - a), np.arange(a, b)] = np.iinfo(np.int64).max
        order = np.lexsort((np.broadcast_to(np.arange(n), d2.shape), d2), axis=1)[:, :m]
        idx[a:b] = order
        d2o[a:b] = np.take_along_axis(d2, order, axis=1)
    return idx, d2o


def save_neighbors(path, ids, scales: dict, nbr_idx, nbr_d2, r_use: int) -> None:
    """id, scale(%.2f), then per neighbour id, scale(%.2f), d2/(2*R_use) (%.2f) (find_neighbors.py:231-267)."""
    r_use = r_use or 1
    with gzip.open(path, "wt") as out:
        for i, sid in enumerate(ids):
            line = f"{sid}\t{scales.get(sid, 1.0):.2f}"
            for j, d2 in zip(nbr_idx[i], nbr_d2[i]):
                nid = ids[j]
                line += f"\t{nid}\t{scales.get(nid, 1.0):.2f}\t{d2 / (2 * r_use):.2f}"
            out.write(line + "\n")


def find_neighbors(config: dict) -> Path:
    """File-level step 5 (find_neighbors.py:11-77)."""
    ncfg = config["mosdepth"]["neighbors"]
    zmax = ncfg.get("zmax", 2.0)
    ftype = config.get("output_file_type", "tsv")
    odir = config.get("output_dir", ".")
    src = Path(f"{odir}/{config['mosdepth']['normalize'].get('output_file_prefix')}.{ftype}.gz")
    dst = Path(odir) / f"{ncfg.get('output_file_prefix', 'neighbor_coverage')}.zMax{zmax:.1f}.{ftype}.gz"
    ids, ratios, z, scales = read_normalized(src)
    z = np.nan_to_num(np.clip(z, -zmax, zmax), nan=0.0)
    keep, r_use = filter_regions_by_variance(ratios, ncfg.get("frac_r", 1.0), ncfg.get("sigma2_max", 1000.0))
    idx, d2 = knn_sklearn(z[:, keep], ncfg.get("num_neighbors", 500))
    save_neighbors(dst, ids, scales, idx, d2, r_use)
    return dst


# --------------------------------------------------------------------------------------
# Step 6: compute_dipcn
# --------------------------------------------------------------------------------------
def load_neighbors(path):
    """{id: [(nbr_id, nbr_scale)]}, {id: scale}; the distance column is ignored (compute_dipcn.py:105-152)."""
    nbrs, scales = {}, {}
    with gzip.open(path, "rt") as fh:
        for line in fh:
            p = line.strip().split("\t")
            if len(p) < 2:
                continue
            try:
                scales[p[0]] = float(p[1])
            except ValueError:
                continue
            lst = []
            for t in range(2, len(p) - 1, 3):
                try:
                    lst.append((p[t], float(p[t + 1])))
                except ValueError:
                    pass
            nbrs[p[0]] = lst
    return nbrs, scales


def load_read_counts(path) -> dict:
    """First line is a header; second column coerced to a number, non-numeric rows dropped (compute_dipcn.py:46-49)."""
    import pandas as pd

    df = pd.read_csv(path, sep="\t", header=0, names=["Sample", "Reads"])
    df["Reads"] = pd.to_numeric(df["Reads"], errors="coerce")
    df = df.dropna(subset=["Reads"])
    return df.set_index("Sample")["Reads"].to_dict()


def dipcn(reads: dict, nbrs: dict, scales: dict, n_nbr: int):
    """dipCN_i = (r_i/s_i) / mean over the first n_nbr neighbours THAT HAVE a count of r_j/s_j (compute_dipcn.py:62-88).

    Neighbours without a count are skipped and do not use up the n_nbr budget.  Samples without
    a count, or with no contributing neighbour, are omitted.  Order = neighbour-file order.
    """
    out = []
    for sid, lst in nbrs.items():
        if scales.get(sid) is None or sid not in reads:
            continue
        tot, cnt = 0.0, 0
        for nid, ns in lst:
            if cnt >= n_nbr:
                break
            if nid not in reads:
                continue
            tot += reads[nid] / ns
            cnt += 1
        if cnt:
            out.append((sid, (reads[sid] / scales[sid]) / (tot / cnt)))
    return out


def dipcn_arrays(counts, has_count, scales, nbr_idx, n_nbr: int) -> np.ndarray:
    """Array form of :func:`dipcn` for G regions at once: counts[N,G], has_count[N], scales[N], nbr_idx[N,k] (-1 pads).

    Returns out[N,G] with NaN where the reference would omit the sample.
    """
    counts = np.asarray(counts, dtype=np.float64)
    n, g = counts.shape
    with np.errstate(all="ignore"):
        t = counts / np.asarray(scales, dtype=np.float64)[:, None]
    out = np.full((n, g), np.nan)
    for i in range(n):
        if not has_count[i]:
            continue
        js = [j for j in nbr_idx[i] if j >= 0 and has_count[j]][:n_nbr]
        if js:
            tot = np.zeros(g)
            for j in js:
                tot += t[j]
            with np.errstate(all="ignore"):
                out[i] = t[i] / (tot / len(js))
    return out


def exon_count(counts_dict: dict, exon_type: str) -> int:
    """Count of one exon type from the four stored keys (compute_dipcn_dir/get_exon_count.py:7-44)."""
    g = counts_dict.get
    if exon_type == "1B_KIV3":
        return g("1B_KIV3", 0)
    if exon_type == "1B_notKIV3":
        return g("1B_KIV2", 0) + g("1B_tied", 0)
    if exon_type == "1B":
        return g("1B_KIV3", 0) + g("1B_KIV2", 0) + g("1B_tied", 0)
    if exon_type == "1A":
        return g("1A", 0)
    raise ValueError(f"Unknown exon type: {exon_type}")


EXON_TYPES = ("1B_KIV3", "1B_notKIV3", "1B", "1A")


def dipcn_exon(counts: dict, neighbors: dict, exon_type: str, n_neighbors: int = 200) -> dict:
    """The exon variant of dipCN (compute_dipcn_dir/compute_diploid_cn.py:8-63).  Differs from :func:`dipcn`:
    the neighbour list is cut to its first n entries BEFORE anything is skipped (a neighbour without counts uses up
    a slot); zero counts and non-positive scales do not contribute; a sample with a zero count, a non-positive
    scale or no contributor is omitted.  neighbors[id] = (scale, [(nbr_id, nbr_scale, distance), ...])."""
    out = {}
    for sid, (s_scale, lst) in neighbors.items():
        if sid not in counts:
            continue
        c = exon_count(counts[sid], exon_type)
        if c == 0:
            continue
        tot, num = 0.0, 0
        for nid, ns, _d in lst[:n_neighbors]:
            if nid not in counts:
                continue
            nc = exon_count(counts[nid], exon_type)
            if nc > 0 and ns > 0:
                tot += nc / ns
                num += 1
        if num > 0 and s_scale > 0:
            mean = tot / num
            if mean > 0:
                out[sid] = (c / s_scale) / mean
    return out


def dipcn_arrays_exon(counts, has_count, scales, nbr_idx, n_nbr: int) -> np.ndarray:
    """Array form of :func:`dipcn_exon` for G count columns at once (same argument layout as dipcn_arrays)."""
    counts = np.asarray(counts)
    scales = np.asarray(scales, dtype=np.float64)
    n, g = counts.shape
    out = np.full((n, g), np.nan)
    for i in range(n):
        if not has_count[i]:
            continue
        js = [j for j in nbr_idx[i][:n_nbr] if j >= 0 and has_count[j]]
        for c in range(g):
            if counts[i, c] == 0:
                continue
            tot, num = 0.0, 0
            for j in js:
                if counts[j, c] > 0 and scales[j] > 0:
                    tot += float(counts[j, c]) / float(scales[j])
                    num += 1
            if num > 0 and scales[i] > 0 and tot / num > 0:
                out[i, c] = (float(counts[i, c]) / float(scales[i])) / (tot / num)
    return out


def compute_diploid_genotypes(config: dict) -> Path:
    """File-level step 6 (compute_dipcn.py:10-101); output written the way DataFrame.to_csv does (repr floats)."""
    import pandas as pd

    ftype = config.get("output_file_type", "tsv")
    odir = config.get("output_dir", ".")
    dcfg = config.get("compute_diploid_genotypes", {})
    out = Path(f"{odir}/{dcfg.get('output_file_prefix')}.{ftype}")
    zmax = config["mosdepth"]["neighbors"].get("zmax", 2.0)
    reads = load_read_counts(Path(f"{odir}/{config['count_reads'].get('output_file_prefix')}.{ftype}"))
    nbrs, scales = load_neighbors(Path(f"{odir}/{config['mosdepth']['neighbors'].get('output_file_prefix')}.zMax{zmax:.1f}.{ftype}.gz"))
    rows = dipcn(reads, nbrs, scales, dcfg.get("n_nbr", 300))
    pd.DataFrame(rows, columns=["Sample", "Norm_Reads"]).to_csv(out, sep="\t", index=False)
    return out


# --------------------------------------------------------------------------------------
# Array-level chain used at sizes where text files are infeasible (SURVEY.md 8c "L-array")
# --------------------------------------------------------------------------------------
def q2(x: np.ndarray) -> np.ndarray:
    """float('%.2f' % v) element-wise: what a value becomes after the text hand-off (NaN stays NaN)."""
    flat = np.asarray(x, dtype=np.float64).ravel()
    out = np.array([np.nan if np.isnan(v) else float(f"{v:.2f}") for v in flat])
    return out.reshape(np.shape(x))


def q3(x: np.ndarray) -> np.ndarray:
    flat = np.asarray(x, dtype=np.float64).ravel()
    out = np.array([np.nan if np.isnan(v) else float(f"{v:.3f}") for v in flat])
    return out.reshape(np.shape(x))


def hundredths(x: np.ndarray) -> np.ndarray:
    """Integer n such that '%.2f' % v == repr(n/100) (vectorised; exact for |v| < 2**40).

    printf rounds the EXACT binary value half-to-even; t = v*100 can land on k+0.5 after
    rounding, so the exact residual decides those cases.
    """
    v = np.asarray(x, dtype=np.float64)
    t = v * 100.0
    n = np.rint(t)
    tie = np.abs(t - n) == 0.5
    if tie.any():
        from fractions import Fraction

        for pos in np.argwhere(tie):
            val = Fraction(float(v[tuple(pos)])) * 100
            fl = math.floor(val)
            fr = val - fl
            if fr > Fraction(1, 2) or (fr == Fraction(1, 2) and fl % 2 == 1):
                fl += 1
            n[tuple(pos)] = fl
    return n.astype(np.int64)


def chain_arrays(depth: np.ndarray, counts: np.ndarray, min_depth=20, max_depth=100, top_frac=0.1,
                 zmax=2.0, sigma2_max=1000.0, frac_r=1.0, num_neighbors=500, n_nbr=300, exact_knn=True):
    """Steps 4-6 on in-memory arrays with the text hand-off quantisation applied.

    depth[N,B]: raw depths, 0 or NaN = bin absent for that sample.  counts[N,G]: read counts.
    Returns a dict with every intermediate the CUDA path exposes.
    """
    d = np.array(depth, dtype=np.float64)
    d[~(d > 0)] = np.nan
    with np.errstate(all="ignore"):
        pop = np.nanmean(d, axis=0)
    valid = np.nonzero((pop >= min_depth) & (pop <= max_depth))[0]
    mat = d[:, valid]
    have = ~np.all(np.isnan(mat), axis=1)
    rows = np.nonzero(have)[0]
    mat = np.ascontiguousarray(mat[rows])  # the reference fills a fresh C-order matrix (:404): nanmean over a row is pairwise
    raw_means = np.nanmean(mat, axis=1)
    z, ratio, mu, s2 = normalize_matrix(mat)
    sel = select_high_variance(ratio, top_frac)
    zq = np.full((mat.shape[0], sel.size), -32768, dtype=np.int64)
    zs = z[:, sel]
    ok = ~np.isnan(zs)
    zq[ok] = hundredths(zs[ok])
    ratio_q = q3(ratio[sel])
    keep, r_use = filter_regions_by_variance(ratio_q, frac_r, sigma2_max)
    lim = int(round(zmax * 100))
    zc = np.where(zq == -32768, 0, np.clip(zq, -lim, lim))[:, keep]
    if exact_knn:
        idx, d2 = knn_exact(zc, num_neighbors)
    else:
        li, ld = knn_sklearn(zc / 100.0, num_neighbors)
        idx, d2 = np.array(li, dtype=np.int32), np.rint(np.array(ld) * 1e4).astype(np.int64)
    scale_q = hundredths(raw_means) / 100.0
    cn = dipcn_arrays(counts[rows], np.ones(len(rows), bool), scale_q, idx, n_nbr)
    return dict(valid=valid, rows=rows, raw_means=raw_means, mu=mu, s2=s2, ratio=ratio, sel=sel, zq=zq,
                keep=keep, r_use=r_use, zc=zc, nbr_idx=idx, nbr_d2=d2, scale_q=scale_q, dipcn=cn)
