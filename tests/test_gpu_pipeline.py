"""The reference's REAL run_wgs_pipeline (grid/pipeline.py:9-103, staged unmodified under oracle/_ref by
oracle/make_ref.py; pysam stubbed as its own test/conftest.py:9-11 does) driving OUR step functions after
grid_b200.install(), on the BASELINE config-1 cohort (100 samples, LPA window): the three files must match the goldens the
unmodified reference produced -- on one GPU, and with the additive key gpu.devices naming two."""
import importlib

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from grid_b200 import synth  # noqa: E402
from helpers import assert_dipcn_files_equal, assert_neighbor_files_equal, assert_norm_files_equal  # noqa: E402


@pytest.mark.parametrize("devices,sharding", [(None, "auto"), ([0, 1], "bins"), ([0, 1], "rows")])
def test_reference_pipeline_runs_our_steps(tmp_path, golden, devices, sharding):
    if not torch.cuda.is_available() or (devices and torch.cuda.device_count() < len(devices)):
        pytest.skip("needs the GPUs named in gpu.devices")
    from oracle import make_ref

    if make_ref.load() is None:
        pytest.skip("oracle/_ref is not staged (python oracle/make_ref.py in the dev container)")
    import yaml

    import grid_b200

    pipeline = importlib.import_module("grid.pipeline")
    steps = [importlib.import_module(m) for m in grid_b200.STEP_TARGETS]
    saved = [{k: getattr(m, k) for k in dir(m) if not k.startswith("__")} for m in steps]
    try:
        grid_b200.install()
        cfg = synth.write_cohort_files(tmp_path, synth.DEFAULT_SEED, 100, 0, window=synth.LPA)
        if devices:
            cfg["gpu"] = {"devices": devices, "sharding": sharding}
        path = tmp_path / "config.yaml"
        path.write_text(yaml.safe_dump(cfg))
        pipeline.run_wgs_pipeline(console=None, config=str(path))
    finally:
        for m, attrs in zip(steps, saved):
            for k, v in attrs.items():
                setattr(m, k, v)
    out, gold = tmp_path / "out", golden / "c1_lpa"
    assert_norm_files_equal(out / "norm.tsv.gz", gold / "norm.tsv")
    assert_neighbor_files_equal(out / "nbrs.zMax2.0.tsv.gz", gold / "nbrs.zMax2.0.tsv")
    assert_dipcn_files_equal(out / "dipcn.tsv", gold / "dipcn.tsv", rel=1e-5)
