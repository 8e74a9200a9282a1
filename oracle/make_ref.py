#!/usr/bin/env python
"""Recipe: stage the UNMODIFIED reference modules of the hot path under oracle/_ref/ (TEST INFRASTRUCTURE ONLY).

    python oracle/make_ref.py            # needs /root/reference (dev container); idempotent

The reference is pure Python, so "building" it is copying the modules steps 4-6 import --
grid/__init__.py, grid/utils/{__init__,utils,mosdepth,normalize_mosdepth,find_neighbors,compute_dipcn}.py --
and pipeline.py / utils/hi_inference.py (so that a GPU test can drive the reference's real run_wgs_pipeline with
grid_b200.install()) -- byte for byte from where they lie under /root/reference into oracle/_ref/grid/...  oracle/_ref/ is git-ignored
(no reference source enters the history) but not gpurun-ignored, so it travels to the GPU box, where
`bench.py --impl reference` and the `cpu_baseline` leg import the reference's OWN normalize_matrix,
select_high_variance_regions, filter_regions_by_variance, find_neighbors_sklearn, save_neighbors and
compute_diploid_genotypes from it (`kind: "reference"`); without it they fall back to the port (`kind: "port"`).
A manifest with the sha256 of every staged file is written beside them.
"""
import hashlib
import json
import shutil
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = Path("/root/reference/grid")
DST = HERE / "_ref" / "grid"
FILES = ["__init__.py", "pipeline.py", "utils/__init__.py", "utils/utils.py", "utils/mosdepth.py",
         "utils/normalize_mosdepth.py", "utils/find_neighbors.py", "utils/compute_dipcn.py", "utils/hi_inference.py"]


def make(verbose: bool = True) -> bool:
    if not SRC.exists():
        if verbose:
            print(f"{SRC} is absent: oracle/_ref not (re)built", file=sys.stderr)
        return False
    manifest = {}
    for rel in FILES:
        dst = DST / rel
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(SRC / rel, dst)
        manifest[rel] = hashlib.sha256(dst.read_bytes()).hexdigest()
    (DST.parent / "MANIFEST.json").write_text(json.dumps({"source": str(SRC), "sha256": manifest}, indent=1) + "\n")
    if verbose:
        print(f"staged {len(FILES)} reference modules under {DST.parent}")
    return True


def load():
    """(normalize_mosdepth, find_neighbors, compute_dipcn) modules of the staged reference, or None when absent.
    pysam is stubbed exactly as the reference's own test/conftest.py:9-11 does (it is not installed in this image)."""
    if not (DST / "utils" / "normalize_mosdepth.py").exists():
        return None
    from unittest.mock import MagicMock

    sys.modules.setdefault("pysam", MagicMock())
    root = str(DST.parent)
    if root not in sys.path:
        sys.path.insert(0, root)
    from grid.utils import compute_dipcn, find_neighbors, normalize_mosdepth

    return normalize_mosdepth, find_neighbors, compute_dipcn


if __name__ == "__main__":
    sys.exit(0 if make() else 1)
