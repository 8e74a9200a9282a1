"""Step 7 (hi_inference) oracle against the reference's outputs (tests/golden/hi, written by oracle/gen_golden.py hi)
and against the reference's own reader tests (test/test_hi_inference.py)."""
import sys
from pathlib import Path

import pytest

from grid_b200 import synth
from oracle import hi_oracle as H

sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "oracle"))
from gen_golden_cases import HI_CASES  # noqa: E402


@pytest.mark.parametrize("case", sorted(HI_CASES))
def test_oracle_reproduces_the_reference_output(case, tmp_path, golden):
    cfg = synth.write_hi_inputs(tmp_path, **HI_CASES[case])
    out = H.hi_inference(cfg)
    assert out.read_text() == (golden / "hi" / f"{case}.tsv").read_text()


def test_readers_known_answers(tmp_path):  # reference test/test_hi_inference.py:13-103
    f = tmp_path / "dipcn.tsv"
    f.write_text("Sample Norm\n1001 2.5\n\n1002 3.0\n1003 1.8\n")
    ids, irr, index = H.read_dip_cn(f)
    assert ids == ["1001", "1002", "1003"] and irr == [2.5, 3.0, 1.8] and index["1003"] == 2
    hdr = "ID\thap\tnbr_ind\tcM_len\tcM_edge\tid_nbr\thap_nbr\n"
    g = tmp_path / "n.txt"
    g.write_text(hdr + "1001\t1\t0\t1.5\t0.1\t1002\t2\n")
    assert H.load_ibs(g, {"1001": 0, "1002": 1}, 10)[0] == [(3, 1.0)]
    g.write_text(hdr + "-1\t1\t0\t1.0\t0.0\t1002\t1\n")
    assert all(l == [] for l in H.load_ibs(g, {"1001": 0, "1002": 1}, 10))
    g.write_text(hdr + "1001\t1\t0\t1.0\t0.0\t1002\t1\n1001\t1\t0\t1.0\t0.0\t1003\t1\n")
    assert len(H.load_ibs(g, {"1001": 0, "1002": 1, "1003": 2}, 1)[0]) == 1
    g.write_text(hdr + "1001\t1\t0\t1.0\t0.0\t9999\t1\n")
    assert H.load_ibs(g, {"1001": 0}, 10)[0] == []
    g.write_text(hdr)
    assert H.load_ibs(g, {"1001": 0}, 10) == [[], []]


def test_level_schedule_reproduces_the_in_place_sweep(tmp_path):
    """Host logic of the product (no GPU): executing the samples level by level -- neighbours with a lower sample index
    read from the live array, the others from the snapshot taken at the start of the sweep -- gives exactly the values
    of the reference's sequential in-place loop."""
    import math

    import numpy as np

    from grid_b200.utils import hi_inference as P

    cfg = synth.write_hi_inputs(tmp_path, seed=7, n=300, method="ibd", weighted=True, n_iters=6, min_neighbors=2)
    hc = cfg["compute_haploid_genotypes"]
    ids, irr, index = P._read_dip_cn_file(Path(cfg["output_dir"]) / "dipcn.tsv")
    lists = P._load_ibd_neighbors(hc["ibd_output"], index, hc["max_neighbors"], cfg["start_bp"], cfg["end_bp"], weighted=True)
    assert lists == H.load_ibd(hc["ibd_output"], index, hc["max_neighbors"], cfg["start_bp"], cfg["end_bp"], weighted=True)
    want, _ = H.run_phasing(irr, lists, 2, 6)
    nb = P.HapNeighbours(lists)
    part = (nb.count[0::2] >= 2) & (nb.count[1::2] >= 2)
    order, level_ptr = P.level_schedule(nb, part)
    assert sorted(order.tolist()) == np.nonzero(part)[0].tolist() and len(level_ptr) > 3
    hap = [irr[i >> 1] / 2 if part[i >> 1] else float("nan") for i in range(2 * len(irr))]
    for _ in range(6):
        old = list(hap)
        for lv in range(len(level_ptr) - 1):
            new_vals = {}
            for i in order[level_ptr[lv]:level_ptr[lv + 1]].tolist():  # "in parallel": results applied after the level
                m = []
                for h in range(2):
                    wsum, wval = 1e-9, 0.0
                    for nbh, w in lists[2 * i + h]:
                        v = hap[nbh] if (nbh >> 1) < i else old[nbh]
                        if not math.isnan(v):
                            wsum += w
                            wval += w * v
                    m.append(wval / wsum)
                if m[0] + m[1] > 0:
                    new_vals[i] = (irr[i] * m[0] / (m[0] + m[1]), irr[i] * m[1] / (m[0] + m[1]))
            for i, (a, b) in new_vals.items():
                hap[2 * i], hap[2 * i + 1] = a, b
    assert all((x == y) or (math.isnan(x) and math.isnan(y)) for x, y in zip(hap, want))
