"""On-disk formats of steps 4-6, byte-compatible with the reference so either implementation can run any step.

Formats (reference file:line, paths under /root/reference/grid/utils/):
  mosdepth input   ``chrom\\tstart\\tend\\tdepth``                                  normalize_mosdepth.py:262-272
  step-4 output    two ``N\\tRw\\t...`` header lines (%.3f) + ``id\\tscale(%.2f)\\tz(%.2f|NA)...`` rows, gzip   :502-554
  step-5 output    ``id\\tscale(%.2f)`` + ``\\tid\\tscale(%.2f)\\td2/(2*Ruse) (%.2f)`` per neighbour, gzip  find_neighbors.py:231-267
  read counts      header line + ``id\\tcount``                                    compute_dipcn.py:46-49
  step-6 output    ``Sample\\tNorm_Reads`` + ``id\\trepr(float)``                  compute_dipcn.py:99-100
"""

from __future__ import annotations

import gzip
import math
import io as _io
from pathlib import Path

import numpy as np

ZQ_NA = -32768        # "NA"
ZQ_NEG_ZERO = -32767  # a negative value that rounds to "-0.00" (numerically 0)
ZQ_LIMIT = 32766


def norm_chrom(chrom: str) -> str:
    """Chromosome names are compared in 'chrN' form (normalize_mosdepth.py:210-215)."""
    return chrom if chrom.startswith("chr") else f"chr{chrom}"


def load_repeat_mask(repeat_bed) -> dict[str, set[int]]:
    """{chrN: masked kb bins}; a BED interval masks every kb from start//1000 through end//1000 inclusive
    (normalize_mosdepth.py:177-207).  Like the reference, the path is opened unconditionally."""
    masked: dict[str, set[int]] = {}
    with open(repeat_bed) as fh:
        for raw in fh:
            if raw.startswith("#") or not raw.strip():
                continue
            cols = raw.strip().split()
            if len(cols) < 3:
                continue
            try:
                first, last = int(cols[1]) // 1000, int(cols[2]) // 1000
            except ValueError:
                continue
            masked.setdefault(norm_chrom(cols[0]), set()).update(range(first, last + 1))
    return masked


def map_mosdepth_files_to_samples(mosdepth_dir, samples) -> dict[str, Path]:
    """sample id -> regions.bed.gz, by the longest '_'-joined stem prefix that is a known id (normalize_mosdepth.py:148-174)."""
    known = set(samples)
    out: dict[str, Path] = {}
    for path in Path(mosdepth_dir).glob("*.regions.bed.gz"):
        parts = path.name.split(".")[0].split("_")
        for cut in range(len(parts), 0, -1):
            sid = "_".join(parts[:cut])
            if sid in known:
                out[sid] = path
                break
    return out


def find_bed_gz_for_individual(individual_id: str, mosdepth_dir) -> Path:
    """First ``*{id}*regions.bed.gz`` match (normalize_mosdepth.py:557-573)."""
    hits = list(Path(mosdepth_dir).glob(f"*{individual_id}*regions.bed.gz"))
    return hits[0] if hits else Path(mosdepth_dir) / f"{individual_id}.regions.bed.gz"


def find_bed_gz_for_individuals(individual_ids, mosdepth_dir) -> list[Path]:
    """find_bed_gz_for_individual for many ids from ONE directory listing: the reference globs the directory once per
    sample (N^2 name matches -- 1.3 s of a 1 000-sample step); the first match in directory order is the same."""
    root = Path(mosdepth_dir)
    listing = list(root.glob("*regions.bed.gz"))              # pathlib's own order: what every per-id glob walks through
    stems = [p.name[: -len("regions.bed.gz")] for p in listing]
    out = []
    for sid in individual_ids:
        if any(c in sid for c in "*?[") or not sid:           # glob metacharacters in an id: let pathlib interpret them
            out.append(find_bed_gz_for_individual(sid, root))
            continue
        hit = next((p for p, stem in zip(listing, stems) if sid in stem), None)
        out.append(hit if hit is not None else root / f"{sid}.regions.bed.gz")
    return out


def parse_regions_text(text: bytes, chrom, start, end, excluded) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Host parser of one sample's decompressed mosdepth text -> (starts, ends, hundredths) after the line filters
    of normalize_mosdepth.py:262-285: chromosome prefix, >= 4 fields, depth > 0, window overlap (closed), kb mask.
    Within a file a repeated (start, end) keeps its LAST depth (:285 / matrix fill order :409-414)."""
    want = norm_chrom(chrom).encode() if chrom else None
    ss, ee, hh = [], [], []
    for line in text.split(b"\n"):
        if want and not line.startswith(want):
            continue
        f = line.strip().split(b"\t")
        if len(f) < 4:
            continue
        s, e, d = int(f[1]), int(f[2]), float(f[3])
        if not d > 0:
            continue
        if start is not None and end is not None and not (e >= start and s <= end):
            continue
        bad = excluded.get(norm_chrom(f[0].decode()))
        if bad and not bad.isdisjoint(range(s // 1000, e // 1000 + 1)):
            continue
        ss.append(s)
        ee.append(e)
        hh.append(round(d * 100))
    return np.array(ss, dtype=np.int64), np.array(ee, dtype=np.int64), np.array(hh, dtype=np.int64)


# ---- step-4 file -------------------------------------------------------------------------------------------
_Z_TABLE = None


def _z_table() -> list[str]:
    """Text of every int16 hundredths code, indexed by code + 32768."""
    global _Z_TABLE
    if _Z_TABLE is None:
        tab = [""] * 65536
        for code in range(-ZQ_LIMIT, ZQ_LIMIT + 1):
            a = abs(code)
            tab[code + 32768] = f"{'-' if code < 0 else ''}{a // 100}.{a % 100:02d}"
        tab[ZQ_NA + 32768] = "NA"
        tab[ZQ_NEG_ZERO + 32768] = "-0.00"
        _Z_TABLE = tab
    return _Z_TABLE


def fixed(v: float, nd: int) -> str:
    return "NA" if v != v else f"{v:.{nd}f}"


def write_normalized_output(path, ids, zq: np.ndarray, mu_sel: np.ndarray, ratio_sel: np.ndarray,
                            raw_means: np.ndarray) -> None:
    """Step-4 hand-off file from the quantised matrix zq[N, Rw] (int16 codes)."""
    n, rw = len(ids), zq.shape[1]
    tab = _z_table()
    with gzip.open(path, "wt", compresslevel=6) as out:
        out.write(f"{n}\t{rw}\t" + "\t".join(fixed(v, 3) for v in mu_sel) + "\n")
        out.write(f"{n}\t{rw}\t" + "\t".join(fixed(v, 3) for v in ratio_sel) + "\n")
        codes = zq.astype(np.int32) + 32768
        for i, sid in enumerate(ids):
            out.write(f"{sid}\t{raw_means[i]:.2f}\t" + "\t".join([tab[c] for c in codes[i].tolist()]) + "\n")


def read_normalized_data(path):
    """Step-4 file -> (ids, ratios[Rw] float64, zq[N,Rw] int16 codes, scales[N] float64) (find_neighbors.py:81-124).

    Fields are two-decimal text, so round(float*100) recovers the integer exactly.  Values beyond the int16 code
    range raise (the device path stores hundredths in 16 bits).
    """
    ids, scales, rows = [], [], []
    with gzip.open(path, "rt") as fh:
        fh.readline()
        ratios = np.array([float("nan") if t in ("NA", "nan") else float(t) for t in fh.readline().strip().split("\t")[2:]])
        for line in fh:
            p = line.strip().split("\t")
            ids.append(p[0])
            scales.append(float(p[1]))
            rows.append(p[2:])
    rw = len(ratios)
    zq = np.full((len(ids), rw), ZQ_NA, dtype=np.int16)
    for i, toks in enumerate(rows):
        arr = np.array([np.nan if t in ("NA", "nan") else float(t) for t in toks])
        ok = ~np.isnan(arr)
        q = np.rint(arr[ok] * 100.0)
        if q.size and np.abs(q).max() > ZQ_LIMIT:
            raise ValueError(f"z-score beyond +-{ZQ_LIMIT / 100} in {path}")
        zq[i, : len(toks)][ok] = q.astype(np.int16)
    return ids, ratios, zq, np.array(scales)


# ---- step-5 file -------------------------------------------------------------------------------------------
def save_neighbors(path, ids, scales: np.ndarray, nbr_idx: np.ndarray, nbr_d2: np.ndarray, r_use: int) -> None:
    """nbr_d2 is exact squared distance in units of 1e-4; the file carries d2/(2*R_use) (find_neighbors.py:231-267)."""
    r_use = r_use or 1
    sc = [f"{v:.2f}" for v in scales]
    with gzip.open(path, "wt", compresslevel=6) as out:
        for i, sid in enumerate(ids):
            parts = [sid, sc[i]]
            for j, d2 in zip(nbr_idx[i].tolist(), nbr_d2[i].tolist()):
                if j < 0:
                    break
                # the reference squares sklearn's sqrt again (find_neighbors.py:222) before normalising; sqrt and the
                # product are the correctly rounded IEEE operations (C sqrt, numpy's square), as on the device
                s = math.sqrt(d2 / 1e4)
                sq = s * s
                parts += [ids[j], sc[j], f"{sq / (2 * r_use):.2f}"]
            out.write("\t".join(parts) + "\n")


def load_neighbors(path):
    """Step-5 file -> (ids, scales[N], lists of (neighbour id, neighbour scale)) in file order (compute_dipcn.py:105-152)."""
    ids, scales, lists = [], [], []
    with gzip.open(path, "rt") as fh:
        for line in fh:
            p = line.strip().split("\t")
            if len(p) < 2:
                continue
            try:
                s = float(p[1])
            except ValueError:
                continue
            cur = []
            for t in range(2, len(p) - 1, 3):
                try:
                    cur.append((p[t], float(p[t + 1])))
                except ValueError:
                    pass
            ids.append(p[0])
            scales.append(s)
            lists.append(cur)
    return ids, np.array(scales), lists


def load_read_counts(path) -> dict[str, float]:
    """pandas semantics of compute_dipcn.py:46-49: first line is a header, non-numeric counts are dropped."""
    import pandas as pd

    df = pd.read_csv(path, sep="\t", header=0, names=["Sample", "Reads"])
    df["Reads"] = pd.to_numeric(df["Reads"], errors="coerce")
    df = df.dropna(subset=["Reads"])
    return df.set_index("Sample")["Reads"].to_dict()


def write_dipcn(path, rows) -> None:
    import pandas as pd

    pd.DataFrame(rows, columns=["Sample", "Norm_Reads"]).to_csv(path, sep="\t", index=False)
