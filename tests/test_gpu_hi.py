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


def test_device_level_schedule_equals_the_host_schedule(tmp_path, hi):
    cfg = synth.write_hi_inputs(tmp_path, seed=77, n=6000, method="ibs", n_iters=3)
    ids, irr, index = hi._read_dip_cn_file(Path(cfg["output_dir"]) / "dipcn.tsv")
    lists = hi._load_ibs_neighbors(cfg["compute_haploid_genotypes"]["ibs_output"], index, 10)
    nbrs = hi.HapNeighbours(lists)
    takes_part = (nbrs.count[0::2] >= 1) & (nbrs.count[1::2] >= 1)
    takes_part[::11] = False
    want_order, want_ptr = hi.level_schedule(nbrs, takes_part)
    order, level_ptr = hi.level_schedule_device(torch.from_numpy(nbrs.ptr).cuda(), torch.from_numpy(nbrs.hap).cuda(), takes_part)
    assert np.array_equal(level_ptr, want_ptr) and np.array_equal(order.cpu().numpy(), want_order)
    assert len(level_ptr) > 5
    # nobody takes part
    order, level_ptr = hi.level_schedule_device(torch.from_numpy(nbrs.ptr).cuda(), torch.from_numpy(nbrs.hap).cuda(),
                                                np.zeros_like(takes_part))
    assert order.numel() == 0 and level_ptr.tolist() == [0]


def test_one_cooperative_launch_equals_one_launch_per_level(tmp_path, hi, monkeypatch):
    cfg = synth.write_hi_inputs(tmp_path, seed=5, n=20000, method="ibs", n_iters=30)
    ids, irr, index = hi._read_dip_cn_file(Path(cfg["output_dir"]) / "dipcn.tsv")
    lists = hi._load_ibs_neighbors(cfg["compute_haploid_genotypes"]["ibs_output"], index, 10)
    from grid_b200 import _lib

    before = _lib.load().gridb200_launch_count()
    coop = hi._run_phasing(irr, lists, 1, 30, None)
    mid = _lib.load().gridb200_launch_count()
    monkeypatch.setenv("GRIDB200_HI_SCHED", "levels")
    per_level = hi._run_phasing(irr, lists, 1, 30, None)
    after = _lib.load().gridb200_launch_count()
    assert np.array_equal(coop[0], per_level[0], equal_nan=True) and np.array_equal(coop[2], per_level[2], equal_nan=True)
    assert (after - mid) - (mid - before) > 30 * 5     # the per-level path really launched once per level and sweep
