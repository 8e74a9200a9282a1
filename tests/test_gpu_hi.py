"""Step 7 (hi_inference) on the GPU against the reference's golden outputs and the oracle's exact doubles."""
import math
import sys
from pathlib import Path

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from grid_b200 import synth  # noqa: E402
from oracle import hi_oracle as H  # noqa: E402

sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "oracle"))
from gen_golden_cases import HI_CASES  # noqa: E402


@pytest.fixture(scope="module")
def hi():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200.utils import hi_inference

    return hi_inference


@pytest.mark.parametrize("case", sorted(HI_CASES))
def test_file_output_equals_the_reference(case, tmp_path, golden, hi):
    cfg = synth.write_hi_inputs(tmp_path, **HI_CASES[case])
    hi.hi_inference(cfg, None)
    assert (Path(cfg["output_dir"]) / "hap.tsv").read_text() == (golden / "hi" / f"{case}.tsv").read_text()


@pytest.mark.parametrize("method,weighted,n", [("ibs", False, 4000), ("ibd", True, 3000)])
def test_doubles_equal_the_sequential_loop(method, weighted, n, tmp_path, hi):
    cfg = synth.write_hi_inputs(tmp_path, seed=n, n=n, method=method, weighted=weighted, n_iters=12)
    hc = cfg["compute_haploid_genotypes"]
    ids, irr, index = hi._read_dip_cn_file(Path(cfg["output_dir"]) / "dipcn.tsv")
    if method == "ibs":
        lists = hi._load_ibs_neighbors(hc["ibs_output"], index, 10)
        assert lists == H.load_ibs(hc["ibs_output"], index, 10)
    else:
        lists = hi._load_ibd_neighbors(hc["ibd_output"], index, 10, cfg["start_bp"], cfg["end_bp"], weighted=True)
        assert lists == H.load_ibd(hc["ibd_output"], index, 10, cfg["start_bp"], cfg["end_bp"], weighted=True)
    want, want_mean = H.run_phasing(irr, lists, 1, 12)
    got, mean, imp = hi._run_phasing(irr, lists, 1, 12, None)
    assert mean == want_mean
    want = np.array(want)
    assert np.array_equal(np.isnan(got), np.isnan(want)) and np.array_equal(got[~np.isnan(want)], want[~np.isnan(want)])
    want_imp = np.array([v for i in range(n) for v in H.impute(i, want.tolist(), lists, want_mean)])
    assert np.array_equal(imp, want_imp)


def test_edge_cases(tmp_path, hi):
    # nobody has enough neighbours: everything NaN, imputation falls back to mean/2 = 0
    got, mean, imp = hi._run_phasing([10.0, 12.0], [[], [], [], []], 1, 5, None)
    assert np.isnan(got).all() and mean == 0.0 and (imp == 0.0).all()
    # a haplotype that neighbours its own sample, zero iterations, a NaN IRR
    lists = [[(1, 1.0)], [(0, 2.0)], [(0, 1.0)], [(1, 1.0)], [(2, 1.0)], [(3, 1.0)]]
    for iters in (0, 3):
        want, wm = H.run_phasing([8.0, 6.0, float("nan")], lists, 1, iters)
        got, mean, _ = hi._run_phasing([8.0, 6.0, float("nan")], lists, 1, iters, None)
        assert all((a == b) or (math.isnan(a) and math.isnan(b)) for a, b in zip(got.tolist(), want))
        assert math.isnan(mean) and math.isnan(wm)
