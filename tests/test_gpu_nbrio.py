"""Step-5 neighbours file written on the device (gridb200_format_nbrs_*) against the host writer, byte for byte."""
import gzip

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from grid_b200 import io as gio  # noqa: E402


@pytest.fixture(scope="module")
def textio():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from grid_b200 import textio

    return textio


def cohort(seed, n, k, r_use, ragged=True):
    rng = np.random.default_rng(seed)
    ids = [f"S{i:07d}" if i % 5 else f"sample-{i}-{'x' * int(rng.integers(0, 30))}" for i in range(n)]
    scales = np.round(rng.uniform(0.5, 180.0, n), 2)
    scales[::7] = np.round(scales[::7])          # "31.00"
    idx = rng.integers(0, n, (n, k)).astype(np.int32)
    d2 = rng.integers(0, 160_000 * max(r_use, 1), (n, k)).astype(np.int64)
    if k:
        d2[:, 0] = (2 * rng.integers(0, 500, n) + 1) * 100 * max(r_use, 1)   # exact %.2f ties
        d2[0, :] = 0
    if ragged and k > 2:
        for i in range(0, n, 3):
            idx[i, int(rng.integers(0, k)):] = -1
        idx[1, :] = -1
    return ids, scales, idx, d2


@pytest.mark.parametrize("n,k,r_use", [(1, 0, 5), (1, 1, 1), (37, 1, 0), (100, 50, 38), (257, 500, 2_609_999), (3000, 131, 1000)])
def test_device_writer_equals_host_writer(n, k, r_use, tmp_path, textio):
    ids, scales, idx, d2 = cohort(n + k, n, k, r_use)
    gio.save_neighbors(tmp_path / "host.gz", ids, scales, idx, d2, r_use)
    textio.save_neighbors(tmp_path / "dev.gz", ids, scales, torch.from_numpy(idx).cuda(), torch.from_numpy(d2).cuda(), r_use, threads=3)
    want = gzip.open(tmp_path / "host.gz", "rb").read()
    got = gzip.open(tmp_path / "dev.gz", "rb").read()
    assert got == want
    a, b = gio.load_neighbors(tmp_path / "dev.gz"), gio.load_neighbors(tmp_path / "host.gz")
    assert a[0] == b[0] and a[2] == b[2] and (a[1] == b[1]).all()


def test_small_batches_and_members(tmp_path, textio, monkeypatch):
    """Rows split over many device batches and gzip members: same bytes, and a strided (padded) list layout."""
    ids, scales, idx, d2 = cohort(9, 900, 60, 77)
    gio.save_neighbors(tmp_path / "host.gz", ids, scales, idx, d2, 77)
    monkeypatch.setattr(textio, "BATCH_BYTES", 200_000)
    monkeypatch.setattr(textio, "MEMBER_BYTES", 30_000)
    pad_i = torch.full((900, 64), -7, dtype=torch.int32, device="cuda")
    pad_d = torch.zeros((900, 64), dtype=torch.int64, device="cuda")
    pad_i[:, :60], pad_d[:, :60] = torch.from_numpy(idx).cuda(), torch.from_numpy(d2).cuda()
    textio.save_neighbors(tmp_path / "dev.gz", ids, scales, pad_i[:, :60], pad_d[:, :60], 77, threads=4)
    assert gzip.open(tmp_path / "dev.gz", "rb").read() == gzip.open(tmp_path / "host.gz", "rb").read()
    raw = (tmp_path / "dev.gz").read_bytes()
    assert len(textio._split_members(raw)) > 10


def _expected_lists(idx):
    out = []
    for row in idx.tolist():
        cut = next((t for t, v in enumerate(row) if v < 0), len(row))
        out.append(row[:cut])
    return out


@pytest.mark.parametrize("n,k,r_use", [(1, 0, 5), (37, 1, 0), (100, 50, 38), (257, 500, 1000), (3000, 131, 1000)])
def test_device_loader_round_trip(n, k, r_use, tmp_path, textio):
    ids, scales, idx, d2 = cohort(n + 3 * k, n, k, r_use)
    textio.save_neighbors(tmp_path / "dev.gz", ids, scales, torch.from_numpy(idx).cuda(), torch.from_numpy(d2).cuda(), r_use, threads=2)
    got_ids, got_scales, got_idx = textio.load_neighbors_device(tmp_path / "dev.gz", threads=2)
    assert got_ids == ids
    assert got_scales.tolist() == [float(f"{v:.2f}") for v in scales.tolist()]
    want = _expected_lists(idx)
    assert _expected_lists(got_idx.cpu().numpy()) == want
    assert got_idx.shape[1] == max(1, max(len(w) for w in want))
    # the host reader (the reference's rules) sees the same lists
    h_ids, h_scales, h_lists = gio.load_neighbors(tmp_path / "dev.gz")
    assert h_ids == ids and [[ids.index(a) for a, _ in lst] for lst in h_lists] == want if n <= 300 else h_ids == ids


def test_device_loader_reads_the_host_writers_file_in_small_batches(tmp_path, textio, monkeypatch):
    ids, scales, idx, d2 = cohort(21, 700, 40, 12)
    gio.save_neighbors(tmp_path / "host.gz", ids, scales, idx, d2, 12)   # single-member gzip, as the reference writes
    monkeypatch.setattr(textio, "BATCH_BYTES", 50_000)
    _, _, got_idx = textio.load_neighbors_device(tmp_path / "host.gz")
    assert _expected_lists(got_idx.cpu().numpy()) == _expected_lists(idx)


@pytest.mark.parametrize("damage", ["foreign_neighbour", "other_scale", "two_fields_short", "duplicate_id", "crlf", "blank_line",
                                    "bad_scale", "leading_blank"])
def test_device_loader_defers_on_foreign_files(damage, tmp_path, textio):
    lines = ["A\t1.00\tB\t2.00\t0.10\tC\t3.00\t0.20", "B\t2.00\tA\t1.00\t0.10", "C\t3.00"]
    if damage == "foreign_neighbour":
        lines[1] = "B\t2.00\tZ\t9.00\t0.10"
    elif damage == "other_scale":
        lines[1] = "B\t2.00\tA\t1.50\t0.10"
    elif damage == "two_fields_short":
        lines[0] = "A\t1.00\tB\t2.00\t0.10\tC\t3.00"
    elif damage == "duplicate_id":
        lines.append("A\t1.00")
    elif damage == "crlf":
        lines = [ln + "\r" for ln in lines]
    elif damage == "blank_line":
        lines.insert(1, "")
    elif damage == "bad_scale":
        lines[2] = "C\tabc"
    elif damage == "leading_blank":
        lines[2] = " C\t3.00"
    with gzip.open(tmp_path / "f.gz", "wt", newline="") as fh:
        fh.write("\n".join(lines) + "\n")
    with pytest.raises(textio._Fallback):
        textio.load_neighbors_device(tmp_path / "f.gz")


def test_device_loader_accepts_a_clean_foreign_file(tmp_path, textio):
    lines = ["A\t1.00\tB\t2\t0.10\tC\t3e0\tnot-a-number", "B\t2\tA\t1.00\t0.10", "C\t3e0"]   # last line without newline
    with gzip.open(tmp_path / "f.gz", "wt", newline="") as fh:
        fh.write("\n".join(lines))
    ids, scales, idx = textio.load_neighbors_device(tmp_path / "f.gz")
    assert ids == ["A", "B", "C"] and scales.tolist() == [1.0, 2.0, 3.0]
    assert idx.cpu().tolist() == [[1, 2], [0, -1], [-1, -1]]
