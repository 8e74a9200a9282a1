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
