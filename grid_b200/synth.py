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
