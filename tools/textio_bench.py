"""Throughput of the step-4 hand-off text path: python tools/textio_bench.py [rows] [rw] [threads]
Device formatter / parser alone (GB/s of text, cells/s), then whole-file write and read through gzip, and the
pure-Python writer/reader of grid_b200.io on a small slice for scale.  One JSON line."""
import gzip
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from grid_b200 import io as gio, textio  # noqa: E402
from grid_b200._lib import call, load, ptr  # noqa: E402

rows, rw, threads = (int(x) for x in (sys.argv[1:4] + ["96", "2609999", str(os.cpu_count())][len(sys.argv) - 1:]))
gen = torch.Generator(device="cuda").manual_seed(1)
ld = (rw + 7) // 8 * 8
zq = (torch.randn((rows, ld), device="cuda", generator=gen) * 100).clamp_(-200, 200).round_().to(torch.int16)
zq[torch.rand((rows, ld), device="cuda", generator=gen) < 0.001] = -32768
st = torch.cuda.current_stream().cuda_stream
per = max(1, min(rows, (256 << 20) // (8 * rw)))
ws = torch.empty(load().gridb200_format_z_ws_bytes(per, rw), dtype=torch.uint8, device="cuda")
row_off = torch.empty(per + 1, dtype=torch.int64, device="cuda")
text = torch.empty(per * rw * 8, dtype=torch.uint8, device="cuda")


def fmt():
    call("gridb200_format_z_measure", ptr(zq), per, rw, ld, ptr(row_off), ptr(ws), ws.numel(), st)
    call("gridb200_format_z_write", ptr(zq), per, rw, ld, ptr(row_off), ptr(ws), ptr(text), text.numel(), st)


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


ms_fmt = timed(fmt)
nbytes = int(row_off[per].item())
# parse the text just produced (no id/scale prefix: skip_fields = 0)
back = torch.empty((per, ld), dtype=torch.int16, device="cuda")
stats = torch.zeros(4, dtype=torch.int64, device="cuda")
max_line = int((row_off[1:per + 1] - row_off[:per]).max().item())
ws2 = torch.empty(load().gridb200_parse_z_ws_bytes(per, max_line), dtype=torch.uint8, device="cuda")
ms_parse = timed(lambda: call("gridb200_parse_z_rows", ptr(text), ptr(row_off), per, max_line, 0, rw, ptr(back), ld,
                              ptr(stats), ptr(ws2), ws2.numel(), st))
expect = zq[:per, :rw].clone()
ok = bool(torch.equal(back[:, :rw], expect)) and stats.cpu().tolist()[:3] == [per * rw, 0, 0]

out = {"workload": f"{rows} rows x {rw} cells, batches of {per} rows", "format_ms": round(ms_fmt, 3),
       "format_text_GBps": round(nbytes / ms_fmt / 1e6, 1), "format_cells_per_s": round(per * rw / ms_fmt * 1e3),
       "parse_ms": round(ms_parse, 3), "parse_text_GBps": round(nbytes / ms_parse / 1e6, 1), "round_trip_exact": ok}

with tempfile.TemporaryDirectory() as tmp:
    ids = [f"S{i:07d}" for i in range(rows)]
    mu, ratio, raw = np.ones(rw), np.full(rw, 100.0), np.full(rows, 30.0)
    t0 = time.perf_counter()
    textio.write_normalized_output(f"{tmp}/a.tsv.gz", ids, zq, rw, mu, ratio, raw, threads=threads)
    t_w = time.perf_counter() - t0
    size = os.path.getsize(f"{tmp}/a.tsv.gz")
    t0 = time.perf_counter()
    ids2, _, zq2, _ = textio.read_normalized_data(f"{tmp}/a.tsv.gz", threads=threads)
    t_r = time.perf_counter() - t0
    same = ids2 == ids and bool(torch.equal(zq2[:, :rw], zq[:, :rw]))
    out.update({"file_write_s": round(t_w, 2), "file_write_cells_per_s": round(rows * rw / t_w), "gzip_threads": threads,
                "file_bytes": size, "file_read_s": round(t_r, 2), "file_read_cells_per_s": round(rows * rw / t_r),
                "file_round_trip_exact": same})
    # the pure-Python path on a slice
    sl = min(rw, 200_000)
    small = zq[:4, :sl].cpu().numpy()
    t0 = time.perf_counter()
    gio.write_normalized_output(f"{tmp}/b.tsv.gz", ids[:4], small, mu[:sl], ratio[:sl], raw[:4])
    t_hw = time.perf_counter() - t0
    t0 = time.perf_counter()
    gio.read_normalized_data(f"{tmp}/b.tsv.gz")
    t_hr = time.perf_counter() - t0
    out.update({"python_write_cells_per_s": round(4 * sl / t_hw), "python_read_cells_per_s": round(4 * sl / t_hr)})
print(json.dumps(out))
