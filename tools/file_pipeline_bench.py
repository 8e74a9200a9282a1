"""Steps 4-6 through their FILE interfaces (the call `grid wgs` makes): python tools/file_pipeline_bench.py [N] [R] [cpu]
Same shape as BASELINE.md section 2's measurement of the unmodified reference (N=1000, R=1000, k=500, n_nbr=30).
Times our three step functions (second run, after a warm-up on a small cohort), and -- with cpu=1 -- the oracle's
file-level restatement of the reference on the same inputs (one core, like-for-like with the reference's loops), then
checks that the output files agree.  One JSON line."""
import gzip
import json
import sys
import tempfile
import time
from pathlib import Path

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from grid_b200 import synth  # noqa: E402
from grid_b200.utils.compute_dipcn import compute_diploid_genotypes  # noqa: E402
from grid_b200.utils.find_neighbors import find_neighbors  # noqa: E402
from grid_b200.utils.normalize_mosdepth import normalize_mosdepth  # noqa: E402

n, r, cpu = (int(x) for x in (sys.argv[1:4] + ["1000", "1000", "1"][len(sys.argv) - 1:]))


def cfg_for(root, n, r):
    cfg = synth.write_cohort_files(root, synth.DEFAULT_SEED, n, r)
    cfg["threads"] = 8
    cfg["mosdepth"]["neighbors"]["num_neighbors"] = 500
    return cfg


def run(cfg):
    t = {}
    for name, fn in (("normalize_mosdepth", normalize_mosdepth), ("find_neighbors", find_neighbors),
                     ("compute_diploid_genotypes", compute_diploid_genotypes)):
        t0 = time.perf_counter()
        fn(cfg, None)
        t[name] = round(time.perf_counter() - t0, 3)
    return t


with tempfile.TemporaryDirectory() as tmp:
    run(cfg_for(Path(tmp) / "warm", 64, 200))  # CUDA context, library load, allocator
    cfg = cfg_for(Path(tmp) / "c", n, r)
    t = run(cfg)
    total = sum(t.values())
    out = {"workload": f"steps 4-6 through files: N={n}, R={r} one-kb bins, k=500, n_nbr=30, threads=8",
           "seconds": t, "total_s": round(total, 3), "samples_per_s": round(n / total, 1)}
    if cpu:
        from helpers import assert_dipcn_files_equal, assert_neighbor_files_equal, assert_norm_files_equal
        from oracle import grid_oracle as O  # checker / CPU leg only

        od = Path(cfg["output_dir"])
        mine = {}
        for f in ("norm.tsv.gz", "nbrs.zMax2.0.tsv.gz"):
            mine[f] = od / ("mine_" + f[:-3])
            mine[f].write_text(gzip.open(od / f, "rt").read())
        mine["dipcn.tsv"] = od / "mine_dipcn.tsv"
        mine["dipcn.tsv"].write_text((od / "dipcn.tsv").read_text())
        tc = {}
        for name, fn in (("normalize_mosdepth", O.normalize_mosdepth), ("find_neighbors", O.find_neighbors),
                         ("compute_diploid_genotypes", O.compute_diploid_genotypes)):
            t0 = time.perf_counter()
            fn(cfg)
            tc[name] = round(time.perf_counter() - t0, 3)
        assert_norm_files_equal(mine["norm.tsv.gz"], od / "norm.tsv.gz")
        _, ratios, _, _ = O.read_normalized(od / "norm.tsv.gz")
        _, r_use = O.filter_regions_by_variance(ratios, 1.0, cfg["mosdepth"]["neighbors"].get("sigma2_max", 1000.0))
        assert_neighbor_files_equal(mine["nbrs.zMax2.0.tsv.gz"], od / "nbrs.zMax2.0.tsv.gz", r_use=int(r_use))
        assert_dipcn_files_equal(mine["dipcn.tsv"], od / "dipcn.tsv", rel=1e-5)
        out.update({"cpu_port_seconds": tc, "cpu_port_total_s": round(sum(tc.values()), 2),
                    "cpu_port_samples_per_s": round(n / sum(tc.values()), 1), "files_agree": True})
print(json.dumps(out))
