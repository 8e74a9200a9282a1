"""Ingest of mosdepth regions.bed.gz files (reference pass 1 + pass 2 + matrix build, normalize_mosdepth.py:218-416):
    python tools/ingest_bench.py [samples] [bins] [threads]
Device parser alone (text resident), the whole packer (inflate on host threads -> upload -> parse), and the Python
line parser of grid_b200.io (same filters as the reference's loop) on a slice for scale.  One JSON line."""
import gzip
import json
import os
import sys
import tempfile
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, ".")
from grid_b200 import io as gio, synth  # noqa: E402
from grid_b200._lib import call, load, ptr  # noqa: E402
from grid_b200.utils import normalize_mosdepth as NM  # noqa: E402

samples, bins, threads = (int(x) for x in (sys.argv[1:4] + ["48", "2900000", str(os.cpu_count())][len(sys.argv) - 1:]))
h = synth.depth_hundredths(5, 1, bins)[0]
text = "".join([f"chr1\t{1000 * j}\t{1000 * j + 1000}\t{v // 100}.{v % 100:02d}\n" for j, v in enumerate(h.tolist())]).encode()
out = {"workload": f"{samples} samples x {bins} lines ({len(text) / 1e6:.0f} MB of text each)", "threads": threads}

# (a) parser alone
dev = torch.device("cuda")
t_text = torch.frombuffer(bytearray(text), dtype=torch.uint8).to(dev)
ld = (bins + 3) // 4 * 4
row = torch.zeros(ld, dtype=torch.float32, device=dev)
stats = torch.zeros(8, dtype=torch.int64, device=dev)
ws = torch.empty(load().gridb200_ingest_ws_bytes(bins), dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream


def parse():
    call("gridb200_ingest_text", ptr(t_text), len(text), b"chr1", 4, -1, -1, None, 0, 0, bins, ptr(row), ptr(stats), ptr(ws),
         ws.numel(), st)


parse()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    parse()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
ok = bool(torch.equal(row[:bins], torch.from_numpy((h / 100.0).astype(np.float32)).to(dev))) and int(stats[0]) == int((h > 0).sum())
out.update({"parse_ms_per_sample": round(ms, 3), "parse_text_GBps": round(len(text) / ms / 1e6, 1),
            "parse_lines_per_s": round(bins / ms * 1e3), "parse_exact": ok})

# (b) the whole packer
with tempfile.TemporaryDirectory() as tmp:
    work = Path(tmp)
    first = work / "S0000000_X.regions.bed.gz"
    t0 = time.perf_counter()
    with gzip.open(first, "wb", compresslevel=1) as f:
        f.write(text)
    ids = synth.sample_ids(samples)
    for sid in ids[1:]:
        os.link(first, work / f"{sid}_X.regions.bed.gz")
    t0 = time.perf_counter()
    res = NM.pack_depth_matrix_device({sid: None for sid in ids}, work, "chr1", None, None, {}, threads=threads)
    torch.cuda.synchronize()
    t_pack = time.perf_counter() - t0
    same = res is not None and bool(torch.equal(res[2][samples - 1, :bins], row[:bins]))
    out.update({"packer_s": round(t_pack, 2), "packer_cells_per_s": round(samples * bins / t_pack),
                "packer_text_MBps": round(samples * len(text) / t_pack / 1e6), "gz_bytes_per_sample": os.path.getsize(first),
                "packer_exact": same})
    # (c) the Python line parser on a slice
    sl = text[: text.index(b"\n", 6_000_000) + 1]
    t0 = time.perf_counter()
    s_, e_, h_ = gio.parse_regions_text(sl, "chr1", None, None, {})
    t_py = time.perf_counter() - t0
    out.update({"python_lines_per_s": round(len(s_) / t_py)})
print(json.dumps(out))
