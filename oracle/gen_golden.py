#!/usr/bin/env python
"""Generate tests/golden/ by running the UNMODIFIED reference (needs /root/reference; dev container only).

    python oracle/gen_golden.py

The reference is imported from /root/reference with pysam stubbed exactly as its own
test/conftest.py:9-11 does.  Inputs come from grid_b200.synth (seeded, regenerated at test time);
what is committed is what the REFERENCE produced:

  c1_lpa/         steps 4-6 through files, N=100, LPA window (BASELINE config 1 shape)
  masked/         steps 4-6 through files, N=24, 80 bins, non-empty repeat mask, tight depth filter,
                  num_neighbors > N-1 (truncation), one sample without a read count
  arrays.npz      normalize_matrix / select_high_variance_regions / filter_regions_by_variance /
                  find_neighbors_sklearn outputs on in-memory arrays (incl. NaN cells)
  exon.npz        compute_diploid_cn_for_exon on a 40-sample cohort with zero counts, samples without counts and
                  non-positive scales, all four exon types (python oracle/gen_golden.py exon regenerates only this)
  step3.json      compute_region_coverage (mosdepth.py:264) and count_reads_in_region (count_reads.py:82, pysam replaced
                  by an in-memory reader) on the seeded step-3 inputs (gen_golden.py step3)
  hi/<case>.tsv   step 7 (hi_inference) outputs for IBS, IBD and weighted IBD neighbour files (gen_golden.py hi)
  manifest.json   shapes, seeds, sha256 of every generated input file
"""

import gzip
import hashlib
import json
import shutil
import sys
import tempfile
from contextlib import contextmanager
from pathlib import Path
from unittest.mock import MagicMock, patch

import numpy as np

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))
sys.path.insert(0, "/root/reference")
sys.modules.setdefault("pysam", MagicMock())

from grid.utils import compute_dipcn as ref_dip  # noqa: E402
from grid.utils import find_neighbors as ref_nbr  # noqa: E402
from grid.utils import normalize_mosdepth as ref_norm  # noqa: E402

from grid_b200 import synth  # noqa: E402

GOLD = REPO / "tests" / "golden"


@contextmanager
def _quiet_progress(*a, **k):
    yield MagicMock(), MagicMock()


def run_reference_steps(cfg):
    with patch.object(ref_norm, "progress_bar", _quiet_progress), patch.object(ref_nbr, "progress_bar", _quiet_progress), \
            patch.object(ref_dip, "progress_bar", _quiet_progress):
        ref_norm.normalize_mosdepth(cfg, None)
        ref_nbr.find_neighbors(cfg, None)
        ref_dip.compute_diploid_genotypes(cfg, None)


def sha(path):
    """sha256 of the (decompressed) content -- gzip headers carry mtimes."""
    path = Path(path)
    data = gzip.open(path, "rb").read() if path.suffix == ".gz" else path.read_bytes()
    return hashlib.sha256(data).hexdigest()


def gunzip_to(src, dst):
    """Store decompressed text (gzip headers carry mtimes; the text is what the format defines)."""
    with gzip.open(src, "rb") as f, open(dst, "wb") as o:
        shutil.copyfileobj(f, o)


def file_case(name, make_cfg, manifest):
    out = GOLD / name
    out.mkdir(parents=True, exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp:
        cfg = make_cfg(Path(tmp))
        run_reference_steps(cfg)
        od = Path(cfg["output_dir"])
        gunzip_to(od / "norm.tsv.gz", out / "norm.tsv")
        gunzip_to(od / "nbrs.zMax2.0.tsv.gz", out / "nbrs.zMax2.0.tsv")
        shutil.copy(od / "dipcn.tsv", out / "dipcn.tsv")
        manifest[name] = {
            "inputs_sha256": {p.name: sha(p) for p in sorted(Path(cfg["mosdepth"]["work_dir"]).glob("*.gz"))[:3]},
            "read_counts_sha256": sha(od / "read_counts.tsv"),
        }


def cfg_c1(tmp):
    return synth.write_cohort_files(tmp, synth.DEFAULT_SEED, 100, 0, window=synth.LPA)


def cfg_masked(tmp):
    cfg = synth.write_cohort_files(tmp, 7, 24, 80)
    # mask two stretches (inclusive kb dilation: 10000-12000 masks kb 10,11,12 -> bins starting 9000..12000)
    Path(cfg["mosdepth"]["normalize"]["repeat_mask_file"]).write_text("# mask\n1\t10000\t12000\nchr1\t40500\t40600\nchrX\t0\t99999999\n")
    cfg["mosdepth"]["normalize"].update(min_depth=28, max_depth=42, top_frac=0.25)
    cfg["mosdepth"]["neighbors"].update(num_neighbors=500, sigma2_max=1000)
    cfg["compute_diploid_genotypes"]["n_nbr"] = 5
    # drop one sample's read count: it must neither be output nor use up a neighbour slot
    rc = Path(cfg["output_dir"]) / "read_counts.tsv"
    lines = rc.read_text().splitlines()
    rc.write_text("\n".join(l for l in lines if not l.startswith("S0000005\t")) + "\n")
    return cfg


def array_case(manifest):
    rng = np.random.default_rng(11)
    depth = synth.depth_hundredths(3, 64, 300).astype(np.float64) / 100.0
    depth[depth == 0] = np.nan
    depth[rng.random(depth.shape) < 0.01] = np.nan
    depth[:, 17] = np.nan            # a bin nobody has would not be a column; keep one sample so it is
    depth[5, 17] = 31.5
    z, vr, mu, s2 = ref_norm.normalize_matrix(depth)
    ratio = np.full(depth.shape[1], np.nan)
    for j, v in vr.items():
        ratio[j] = v
    sel10 = np.array(ref_norm.select_high_variance_regions(vr, 0.1))
    sel50 = np.array(ref_norm.select_high_variance_regions(vr, 0.5))
    # the text hand-off, then find_neighbors' array steps
    with tempfile.TemporaryDirectory() as tmp:
        p = Path(tmp) / "n.tsv.gz"
        ids = synth.sample_ids(64)
        ref_norm.write_normalized_output(z, ids, list(sel10), p, col_means=mu, col_vars=s2, individual_raw_means=np.nanmean(depth, axis=1))
        rid, rratio, rz, rscales = ref_nbr.read_normalized_data(p)
    clipped = np.nan_to_num(np.clip(rz, -2.0, 2.0), nan=0.0)
    keep, r_use = ref_nbr.filter_regions_by_variance(rratio, frac_r=1.0, sigma2_max=1000.0)
    keep_half, _ = ref_nbr.filter_regions_by_variance(rratio, frac_r=0.5, sigma2_max=1000.0)
    res = ref_nbr.find_neighbors_sklearn(clipped[:, keep], rid, n_neighbors=20)
    pos = {s: i for i, s in enumerate(rid)}
    nidx = np.array([[pos[n] for n, _ in res[s]] for s in rid], dtype=np.int32)
    nd2 = np.array([[d for _, d in res[s]] for s in rid])
    np.savez_compressed(
        GOLD / "arrays.npz", depth=depth, z=z, ratio=ratio, mu=mu, s2=s2, sel10=sel10, sel50=sel50,
        read_ratio=rratio, read_z=rz, read_scales=np.array([rscales[s] for s in rid]), keep=keep,
        keep_half=keep_half, r_use=r_use, nbr_idx=nidx, nbr_d2=nd2)
    manifest["arrays"] = {"depth": "synth.depth_hundredths(3,64,300)/100 with NaNs (rng 11)", "k": 20}


def exon_case():
    from grid.utils.compute_dipcn_dir.compute_diploid_cn import compute_diploid_cn_for_exon

    rng = np.random.default_rng(5)
    n, k = 40, 12
    keys = ("1B_KIV3", "1B_KIV2", "1B_tied", "1A")
    raw = rng.integers(0, 400, size=(n, 4))
    raw[rng.random((n, 4)) < 0.15] = 0
    has = rng.random(n) > 0.15
    scales = np.round(rng.uniform(0.6, 1.6, n), 2)
    scales[3], scales[17] = 0.0, -1.0
    idx = np.stack([rng.permutation(np.delete(np.arange(n), i))[:k] for i in range(n)]).astype(np.int32)
    ids = synth.sample_ids(n)
    counts = {ids[i]: dict(zip(keys, (int(v) for v in raw[i]))) for i in range(n) if has[i]}
    nbrs = {ids[i]: (float(scales[i]), [(ids[j], float(scales[j]), 0.5) for j in idx[i]]) for i in range(n)}
    types = ("1B_KIV3", "1B_notKIV3", "1B", "1A")
    res = {}
    for nn in (5, 200):
        out = np.full((n, 4), np.nan)
        for c, t in enumerate(types):
            for sid, v in compute_diploid_cn_for_exon(counts, nbrs, t, n_neighbors=nn).items():
                out[ids.index(sid), c] = v
        res[f"out_n{nn}"] = out
    np.savez_compressed(GOLD / "exon.npz", raw=raw, has=has, scales=scales, idx=idx, **res)


from gen_golden_cases import HI_CASES  # noqa: E402


def hi_case():
    """Step 7 through the reference's hi_inference(config, console): tests/golden/hi/<case>.tsv (inputs are regenerated
    from grid_b200.synth.write_hi_inputs at test time)."""
    from grid.utils import hi_inference as ref_hi

    (GOLD / "hi").mkdir(parents=True, exist_ok=True)
    for name, kw in HI_CASES.items():
        with tempfile.TemporaryDirectory() as tmp:
            cfg = synth.write_hi_inputs(tmp, **kw)
            ref_hi.hi_inference(cfg, None)
            shutil.copy(Path(cfg["output_dir"]) / "hap.tsv", GOLD / "hi" / f"{name}.tsv")


from gen_golden_cases import STEP3_COVERAGE, STEP3_READS  # noqa: E402


class _FakeRead:
    """What count_reads_in_region touches on a pysam.AlignedSegment (count_reads.py:95-105)."""
    __slots__ = ("flag", "mapq", "reference_id", "next_reference_id", "reference_start")

    def __init__(self, flag, mapq, tid, ntid, pos):
        self.flag, self.mapq, self.reference_id, self.next_reference_id, self.reference_start = flag, mapq, tid, ntid, pos

    is_duplicate = property(lambda self: bool(self.flag & 0x400))
    is_secondary = property(lambda self: bool(self.flag & 0x100))


def _fake_alignment_file(samples, read_len=150):
    """Stand-in for pysam.AlignmentFile over in-memory columns; fetch() honours htslib's contract: the reads on that
    reference whose alignment [pos, pos + len) overlaps [start, end)."""
    class Fake:
        def __init__(self, path, mode=None, reference_filename=None):
            self.cols = samples[path]

        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

        def fetch(self, chrom, start, end):
            flag, mapq, tid, ntid, pos = self.cols
            t = int(chrom[3:])
            for i in np.flatnonzero((tid == t) & (pos < end) & (pos + read_len > start)):
                yield _FakeRead(int(flag[i]), int(mapq[i]), int(tid[i]), int(ntid[i]), int(pos[i]))
    return Fake


def step3_case():
    """compute_region_coverage and count_reads_in_region of the unmodified reference on the seeded step-3 inputs:
    tests/golden/step3.json."""
    from grid.utils import count_reads as ref_cr
    from grid.utils import mosdepth as ref_md

    res = {"coverage": {}, "reads": {}}
    for name, kw in STEP3_COVERAGE.items():
        with tempfile.TemporaryDirectory() as tmp:
            paths = synth.write_region_files(tmp, **kw)
            res["coverage"][name] = {q[0]: [ref_md.compute_region_coverage(p, q[1], q[2], q[3]) for p in paths]
                                     for q in synth.REGION_QUERIES}
    for name, kw in STEP3_READS.items():
        cols, off = synth.read_columns(kw["seed"], kw["ns"], kw["mean_reads"])
        samples = {f"s{s}": tuple(c[off[s]:off[s + 1]] for c in cols) for s in range(kw["ns"])}
        with patch.object(ref_cr.pysam, "AlignmentFile", _fake_alignment_file(samples)):
            res["reads"][name] = [[ref_cr.count_reads_in_region(f"s{s}", "ref.fa", f"chr{t}", lo, hi, kw["flags"], kw["min_mapq"])
                                   for (t, lo, hi) in synth.READ_REGIONS] for s in range(kw["ns"])]
    (GOLD / "step3.json").write_text(json.dumps(res, indent=1, sort_keys=True) + "\n")


def main():
    GOLD.mkdir(parents=True, exist_ok=True)
    if sys.argv[1:] == ["step3"]:
        step3_case()
        return
    if sys.argv[1:] == ["exon"]:
        exon_case()
        return
    if sys.argv[1:] == ["hi"]:
        hi_case()
        return
    manifest = {"reference": "caterer-z-t/GRiD @ /root/reference", "sklearn": __import__("sklearn").__version__,
                "numpy": np.__version__}
    file_case("c1_lpa", cfg_c1, manifest)
    file_case("masked", cfg_masked, manifest)
    array_case(manifest)
    exon_case()
    hi_case()
    step3_case()
    (GOLD / "manifest.json").write_text(json.dumps(manifest, indent=1, sort_keys=True) + "\n")
    print("golden written to", GOLD)


if __name__ == "__main__":
    main()
