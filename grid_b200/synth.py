"""Deterministic synthetic cohorts, reproducible bit-for-bit in numpy and CUDA.

Everything is a counter-based hash of (seed, stream, row, col) followed by IEEE +,*,/ and sqrt
in fp64 only -- no transcendental functions, no fused multiply-add -- so the device generator in
``csrc/synth.cu`` and :func:`depth_hundredths` below produce identical integers.

Model (per sample i, bin j, latent batch g = i mod 16):

    c_i   = 25 + 20 u(1,i)                       sample coverage
    b_j   = 0.7 + 0.6 u(2,j)                     bin factor
    e_gj  = 1 + 0.05 z(3,g,j)                    batch x bin effect
    lam   = (c_i * b_j) * e_gj
    d_ij  = lam * (1 + z(4,i,j) * (rc_i * rb_j)) rc_i = 1/sqrt(6.7 c_i), rb_j = 1/sqrt(b_j)
    h_ij  = floor(d_ij*100 + 0.5) if d_ij > 0 else 0        depth in hundredths (mosdepth prints %.2f)
    h_ij  = 0 when mix(hash(4,i,j)) % 1000 == 0             0.1 % absent bins

u() is a 53-bit uniform, z() an Irwin-Hall(4) approximate normal built from the four 16-bit
fields of one hash.  Read counts: r_ig = floor(c_i * (L_g/150) * max(0, 1 + 0.25 z(6,i,g))).
"""

from __future__ import annotations

import gzip
from pathlib import Path

import numpy as np

DEFAULT_SEED = 20251104
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
Z_SCALE = 1.0 / 37837.22  # ~ 1/sd of the sum of four 16-bit uniforms
LPA = ("chr6", 160605062, 160647661)


def splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = (np.asarray(x, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15)) & _M64
        x = ((x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        x = ((x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return x ^ (x >> np.uint64(31))


def stream_key(seed: int, stream: int) -> np.uint64:
    return splitmix64(np.uint64((seed * 0x100 + stream) & 0xFFFFFFFFFFFFFFFF))


def hash2(key: np.uint64, a, b) -> np.ndarray:
    a = np.asarray(a, dtype=np.uint64)
    b = np.asarray(b, dtype=np.uint64)
    return splitmix64(key ^ ((a << np.uint64(32)) | b))


def uniform(h: np.ndarray) -> np.ndarray:
    return (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def znorm(h: np.ndarray) -> np.ndarray:
    m = np.uint64(0xFFFF)
    s = (h & m) + ((h >> np.uint64(16)) & m) + ((h >> np.uint64(32)) & m) + (h >> np.uint64(48))
    return (s.astype(np.int64) - 131070).astype(np.float64) * Z_SCALE


def row_params(seed: int, n: int, row0: int = 0):
    """(c_i, rc_i) for rows row0..row0+n."""
    i = np.arange(row0, row0 + n, dtype=np.uint64)
    c = 25.0 + 20.0 * uniform(hash2(stream_key(seed, 1), i, 0))
    return c, 1.0 / np.sqrt(6.7 * c)


def col_params(seed: int, r: int, col0: int = 0):
    """(b_j, rb_j) for columns col0..col0+r."""
    j = np.arange(col0, col0 + r, dtype=np.uint64)
    b = 0.7 + 0.6 * uniform(hash2(stream_key(seed, 2), 0, j))
    return b, 1.0 / np.sqrt(b)


def depth_hundredths(seed: int, n: int, r: int, row0: int = 0, col0: int = 0) -> np.ndarray:
    """h[n, r] int64: depth * 100 for rows row0.., columns col0.. (0 = bin absent)."""
    c, rc = row_params(seed, n, row0)
    b, rb = col_params(seed, r, col0)
    i = np.arange(row0, row0 + n, dtype=np.uint64)[:, None]
    j = np.arange(col0, col0 + r, dtype=np.uint64)[None, :]
    e = 1.0 + 0.05 * znorm(hash2(stream_key(seed, 3), i % np.uint64(16), j))
    lam = (c[:, None] * b[None, :]) * e
    h4 = hash2(stream_key(seed, 4), i, j)
    d = lam * (1.0 + znorm(h4) * (rc[:, None] * rb[None, :]))
    h = np.where(d > 0, np.floor(d * 100.0 + 0.5), 0.0).astype(np.int64)
    h[splitmix64(h4) % np.uint64(1000) == 0] = 0
    return h


def depth_matrix(seed: int, n: int, r: int, row0: int = 0, col0: int = 0) -> np.ndarray:
    """fp32 depth matrix as ingest would pack it (0.0 = absent)."""
    return (depth_hundredths(seed, n, r, row0, col0) / 100.0).astype(np.float32)


def read_counts(seed: int, n: int, region_len, row0: int = 0) -> np.ndarray:
    """counts[n, G] int32 for G target regions of the given lengths (bp)."""
    region_len = np.atleast_1d(np.asarray(region_len, dtype=np.float64))
    c, _ = row_params(seed, n, row0)
    i = np.arange(row0, row0 + n, dtype=np.uint64)[:, None]
    g = np.arange(region_len.size, dtype=np.uint64)[None, :]
    cn = np.maximum(0.0, 1.0 + 0.25 * znorm(hash2(stream_key(seed, 6), i, g)))
    return np.floor((c[:, None] * (region_len[None, :] / 150.0)) * cn).astype(np.int32)


def sample_ids(n: int, row0: int = 0) -> list[str]:
    """Fixed-width ids so the reference's substring glob cannot alias S1/S10."""
    return [f"S{i:07d}" for i in range(row0, row0 + n)]


def bin_starts(r: int, window=None) -> np.ndarray:
    """1 kb bin starts.  window=(chrom,start,end): the bins mosdepth --by 1000 emits that touch it."""
    if window is None:
        return np.arange(r, dtype=np.int64) * 1000
    first = (window[1] // 1000) * 1000
    return first + np.arange(r, dtype=np.int64) * 1000


def write_cohort_files(root, seed: int, n: int, r: int, window=None, region_name="LPA", decoys=True) -> dict:
    """Emit the on-disk inputs of steps 4-6 for a small cohort and return a ready config dict.

    Layout: {root}/work/{id}_{region}.regions.bed.gz, {root}/samples.txt, {root}/mask.bed (empty),
    {root}/out/read_counts.tsv.  window=None -> pseudo-chromosome chr1 with globally unique bins;
    window=LPA -> chr6 bins touching chr6:160605062-160647661 (r is then ignored: 43 bins) plus,
    when ``decoys``, out-of-window and other-chromosome lines the line filter must reject.
    """
    root = Path(root)
    (root / "work").mkdir(parents=True, exist_ok=True)
    (root / "out").mkdir(parents=True, exist_ok=True)
    if window is not None:
        chrom, lo, hi = window
        r = hi // 1000 - lo // 1000 + 1
    else:
        chrom, lo, hi = "chr1", 0, r * 1000
    starts = bin_starts(r, window)
    h = depth_hundredths(seed, n, r)
    ids = sample_ids(n)
    for i, sid in enumerate(ids):
        with gzip.open(root / "work" / f"{sid}_{region_name}.regions.bed.gz", "wt", compresslevel=1) as fh:
            if decoys and window is not None:
                fh.write(f"chr5\t{starts[0]}\t{starts[0] + 1000}\t31.00\n")
                fh.write(f"{chrom}\t{starts[0] - 5000}\t{starts[0] - 4000}\t29.50\n")
            for j in range(r):
                fh.write(f"{chrom}\t{starts[j]}\t{starts[j] + 1000}\t{h[i, j] // 100}.{h[i, j] % 100:02d}\n")
            if decoys and window is not None:
                fh.write(f"{chrom}\t{starts[-1] + 3000}\t{starts[-1] + 4000}\t33.25\n")
    (root / "samples.txt").write_text("\n".join(ids) + "\n")
    (root / "mask.bed").write_text("")
    counts = read_counts(seed, n, hi - lo)
    with open(root / "out" / "read_counts.tsv", "w") as fh:
        fh.write(f"Sample\t{chrom}:{lo}-{hi}\n")
        for sid, c in zip(ids, counts[:, 0]):
            fh.write(f"{sid}\t{c}\n")
    return {
        "samples_file": str(root / "samples.txt"),
        "output_dir": str(root / "out"),
        "threads": 4,
        "chrom": chrom,
        "start_bp": lo,
        "end_bp": hi,
        "output_file_type": "tsv",
        "count_reads": {"run": False, "output_file_prefix": "read_counts"},
        "mosdepth": {
            "run": False,
            "work_dir": str(root / "work"),
            "region_name": region_name,
            "normalize": {"run": True, "min_depth": 20, "max_depth": 100, "top_frac": 0.1,
                          "output_file_prefix": "norm", "repeat_mask_file": str(root / "mask.bed")},
            "neighbors": {"run": True, "output_file_prefix": "nbrs", "num_neighbors": 50, "zmax": 2.0,
                          "sigma2_max": 1000},
        },
        "compute_diploid_genotypes": {"run": True, "output_file_prefix": "dipcn", "n_nbr": 30},
        "compute_haploid_genotypes": {"run": False},
        "index": {"run": False},
    }


# ---- step 7 inputs (hi_inference): a dipCN table plus IBS (computeIBSpbwt) or IBD (iLASH) haplotype neighbours ----
def write_hi_inputs(root, seed: int, n: int, method: str = "ibs", weighted: bool = False, n_iters: int = 100,
                    max_neighbors: int = 10, min_neighbors: int = 1) -> dict:
    """Synthetic step-7 cohort under `root`; returns the config.  Deliberately untidy: haplotypes with no neighbours or
    more than max_neighbors, ids missing from the dipCN table, malformed rows, self-neighbours, duplicate edges."""
    import gzip
    from pathlib import Path

    root = Path(root)
    out = root / "out"
    out.mkdir(parents=True, exist_ok=True)
    rng = np.random.default_rng(seed)
    ids = sample_ids(n)
    irr = np.round(rng.normal(20.0, 4.0, n).clip(2.0, None), 6)
    with open(out / "dipcn.tsv", "w") as fh:
        fh.write("Sample\tNorm_Reads\n")
        for sid, v in zip(ids, irr):
            fh.write(f"{sid}\t{float(v)!r}\n")
    cfg = {"output_dir": str(out), "output_file_type": "tsv", "start_bp": 160_605_062, "end_bp": 160_647_661,
           "compute_diploid_genotypes": {"output_file_prefix": "dipcn"},
           "compute_haploid_genotypes": {"output_file_prefix": "hap", "method": method, "n_iters": n_iters,
                                         "max_neighbors": max_neighbors, "min_neighbors": min_neighbors}}
    hc = cfg["compute_haploid_genotypes"]
    if method == "ibs":
        path = root / "ibs.txt.gz"
        with gzip.open(path, "wt") as fh:
            fh.write("ID\thap\tnbrInd\tcMlen\tcMedge\tIDnbr\thapNbr\n")
            for i in range(n):
                for h in (1, 2):
                    cnt = int(rng.integers(0, max_neighbors + 5)) if rng.random() > 0.06 else 0
                    for t in range(cnt):
                        j = int(rng.integers(0, n))
                        nid = ids[j] if rng.random() > 0.03 else f"X{j:07d}"
                        hn = int(rng.integers(1, 3)) if rng.random() > 0.02 else 3
                        fh.write(f"{ids[i]}\t{h}\t{t}\t{rng.uniform(0.5, 9):.3f}\t{rng.uniform(0, 1):.3f}\t{nid}\t{hn}\n")
                if rng.random() < 0.02:
                    fh.write(f"{ids[i]}\t1\tbroken\n\n")
        hc["ibs_output"] = str(path)
    else:
        path = root / "ilash.match.gz"
        lo, hi = cfg["start_bp"], cfg["end_bp"]
        with gzip.open(path, "wt") as fh:
            for _ in range(n * max_neighbors):
                i, j = int(rng.integers(0, n)), int(rng.integers(0, n))
                f1 = ids[i] if rng.random() > 0.02 else f"X{i:07d}"
                h1, h2 = int(rng.integers(0, 2)), int(rng.integers(0, 2))
                centre = int(rng.integers(lo - 6_000_000, hi + 6_000_000))
                half = int(rng.integers(50_000, 3_000_000))
                length = float(np.round(rng.uniform(0.2, 6.0), 2 if rng.random() > 0.3 else 1))  # ties in length too
                match = float(np.round(rng.uniform(0.6, 1.0), 3))
                sep = "\t" if rng.random() > 0.1 else " "
                row = [f1, f"{f1}_{h1}", ids[j], f"{ids[j]}_{h2}", "6", centre - half, centre + half, centre - half + 7,
                       centre + half - 7, length, match]
                fh.write(sep.join(str(x) for x in row) + "\n")
                if rng.random() < 0.01:
                    fh.write("short\trow\n")
        hc["ibd_output"] = str(path)
        hc["weighted"] = weighted
    return cfg


# ---- step-3 inputs: genome-style regions files for compute_region_coverage, read records for count_reads ----
REGION_QUERIES = [
    ("lpa", "chr6", LPA[1], LPA[2]),                       # both ends cut a bin
    ("inside_one_bin", "chr6", 160_610_100, 160_610_900),
    ("bin_aligned", "chr6", 160_606_000, 160_640_000),
    ("wide", "chr6", 160_500_000, 160_800_000),            # wider than the file's chr6 lines
    ("beyond", "chr6", 170_000_000, 170_100_000),          # no overlap -> 0
    ("other_chrom", "chr60", 160_600_000, 160_650_000),    # exact name test: chr6 lines must not count
    ("absent_chrom", "chrX", 0, 1_000_000),
]


def write_region_files(root, seed: int, n: int, flank_kb: int = 20, by: int = 1000) -> list:
    """n genome-style `{id}_LPA.regions.bed.gz` files under root: chr5 lines, chr6 bins of `by` bp from flank_kb before
    the LPA window to flank_kb after it (0.1 % of them 0.00, a duplicated bin, one line with extra columns), chr60
    lines.  Returns the paths in sample order."""
    root = Path(root)
    root.mkdir(parents=True, exist_ok=True)
    lo = (LPA[1] // 1000 - flank_kb) * 1000
    hi = (LPA[2] // 1000 + 1 + flank_kb) * 1000
    starts = np.arange(lo, hi, by, dtype=np.int64)
    h = depth_hundredths(seed, n, starts.size)
    paths = []
    for i, sid in enumerate(sample_ids(n)):
        p = root / f"{sid}_LPA.regions.bed.gz"
        with gzip.open(p, "wt", compresslevel=1) as fh:
            for j in range(3):
                fh.write(f"chr5\t{lo + j * by}\t{lo + (j + 1) * by}\t{31 + j}.25\n")
            for j, s in enumerate(starts):
                extra = "\tgene7" if j == 5 else ""
                fh.write(f"chr6\t{s}\t{s + by}\t{h[i, j] // 100}.{h[i, j] % 100:02d}{extra}\n")
            fh.write(f"chr6\t{starts[7]}\t{starts[7] + by}\t{h[i, 7] // 100}.{h[i, 7] % 100:02d}\n")  # repeated bin: counted twice
            for j in range(4):
                fh.write(f"chr60\t{LPA[1] - 500 + j * by}\t{LPA[1] - 500 + (j + 1) * by}\t{12 + j}.50\n")
        paths.append(p)
    return paths


READ_FLAG_PALETTE = np.array([83, 99, 147, 163, 83, 99, 147, 163, 83 | 0x400, 99 | 0x100, 65, 4, 0, 67 | 0x800, 163 | 0x400,
                              97], dtype=np.uint16)
READ_REGIONS = [  # (tid, start, end): overlapping, nested, identical, empty, other reference
    (6, LPA[1], LPA[2]), (6, LPA[1] + 10_000, LPA[1] + 20_000), (6, LPA[1], LPA[2]), (6, LPA[1] - 3_000, LPA[1] + 100),
    (6, LPA[2] - 1, LPA[2] + 2_000), (6, LPA[1] + 500, LPA[1] + 500), (5, LPA[1], LPA[2]), (6, 0, 1_000),
    (7, LPA[1] + 7, LPA[1] + 30_007),
]


def read_columns(seed: int, ns: int, mean_reads: int = 4000):
    """Columnar read records for ns samples: (flag u16, mapq u8, tid i32, next_tid i32, pos i32), sample_off[ns+1].
    Positions fall around the LPA window on references 5..7 (mostly 6) with reads exactly on region boundaries;
    every fourth sample is short and sample 2 has no reads at all."""
    key = stream_key(seed, 9)
    s_idx = np.arange(ns, dtype=np.uint64)
    per = (mean_reads * (0.5 + uniform(hash2(key, s_idx, 0)))).astype(np.int64)
    per[3::4] //= 16
    if ns > 2:
        per[2] = 0
    sample_off = np.concatenate([[0], np.cumsum(per)]).astype(np.int64)
    n = int(sample_off[-1])
    r = np.arange(n, dtype=np.uint64)
    hf, hp = hash2(key, r, 1), hash2(key, r, 2)
    flag = READ_FLAG_PALETTE[(hf & np.uint64(15)).astype(np.int64)]
    mapq = ((hf >> np.uint64(8)) % np.uint64(61)).astype(np.uint8)
    mapq[(hf >> np.uint64(20)) % np.uint64(9) == 0] = 0
    sel = (hf >> np.uint64(32)) % np.uint64(20)
    tid = np.where(sel == 0, 5, np.where(sel == 1, 7, 6)).astype(np.int32)
    next_tid = np.where((hf >> np.uint64(40)) % np.uint64(10) == 0, tid + 1, tid).astype(np.int32)
    span = LPA[2] - LPA[1] + 10_000
    pos = (LPA[1] - 5_000 + (hp % np.uint64(span)).astype(np.int64)).astype(np.int64)
    edge = (hp >> np.uint64(40)) % np.uint64(50)
    for code, val in ((0, LPA[1]), (1, LPA[1] - 1), (2, LPA[2] - 1), (3, LPA[2]), (4, LPA[1] + 10_000), (5, LPA[1] + 19_999)):
        pos[edge == code] = val
    unplaced = (hp >> np.uint64(50)) % np.uint64(97) == 0
    tid[unplaced], next_tid[unplaced], pos[unplaced] = -1, -1, -1
    return (flag, mapq, tid, next_tid, pos.astype(np.int32)), sample_off
