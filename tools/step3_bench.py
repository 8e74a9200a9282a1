"""The two step-3 producers on the device, with the reference's Python loops on a slice for scale:
    python tools/step3_bench.py [files] [lines_per_file] [reads] [regions]
(a) gridb200_region_coverage over a resident text batch: GB/s of text against the HBM peak (compulsory traffic = the text);
(b) gridb200_count_reads: reads/s and GB/s at 15 B per read, G regions in one pass.  One JSON line."""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "oracle")
import step3_oracle as so  # noqa: E402  (CPU leg only)

from grid_b200 import synth  # noqa: E402
from grid_b200._lib import call, ptr  # noqa: E402
from grid_b200.utils import count_reads as CR  # noqa: E402

files, lines, reads, g_regions = (int(x) for x in (sys.argv[1:5] + ["24", "2900000", "400000000", "734"][len(sys.argv) - 1:]))
peaks = json.loads(Path("MEASURED_PEAKS.json").read_text()) if Path("MEASURED_PEAKS.json").exists() else {}
hbm = float(peaks.get("hbm_gbs", 6553.3))
dev = torch.device("cuda")
st = torch.cuda.current_stream().cuda_stream


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


# (a) region coverage: one genome-wide file replicated (the kernel does not care), LPA query
h = synth.depth_hundredths(5, 1, lines)[0]
one = "".join([f"chr6\t{1000 * j}\t{1000 * j + 1000}\t{v // 100}.{v % 100:02d}\n" for j, v in enumerate(h.tolist())]).encode()
text = torch.frombuffer(bytearray(one), dtype=torch.uint8).to(dev).repeat(files)
off = torch.arange(files + 1, dtype=torch.int64, device=dev) * len(one)
out_rc = torch.empty((files, 4), dtype=torch.int64, device=dev)
lo, hi = synth.LPA[1], synth.LPA[2]
ms = timed(lambda: call("gridb200_region_coverage", ptr(text), files * len(one), ptr(off), files, b"chr6", 4, lo, hi, ptr(out_rc), st))
j = np.arange(lines)
ov = np.clip(np.minimum(hi, 1000 * (j + 1)) - np.maximum(lo, 1000 * j), 0, None)
exact = bool((out_rc.cpu().numpy() == np.array([int((h * ov).sum()), int(ov.sum()), int((ov > 0).sum()), 0])).all())
sl = one[: one.index(b"\n", 8_000_000) + 1].decode()
t0 = time.perf_counter()
so.compute_region_coverage_text(sl, "chr6", lo, hi)
t_py = time.perf_counter() - t0
res = {"region_coverage": {"workload": f"{files} files x {lines} lines ({files * len(one) / 1e9:.2f} GB of text), region chr6:{lo}-{hi}",
                           "ms": round(ms, 3), "text_GBps": round(files * len(one) / ms / 1e6, 1),
                           "frac_hbm": round(files * len(one) / ms / 1e6 / hbm, 3), "lines_per_s": round(files * lines / ms * 1e3),
                           "exact": exact, "python_loop_lines_per_s": round(sl.count("\n") / t_py)}}
del text

# (b) read counting: `reads` reads in 64 samples, G random overlapping regions of 1-50 kb on 24 references
rng = np.random.default_rng(2)
gen = torch.Generator(device=dev).manual_seed(3)
pal = torch.tensor(synth.READ_FLAG_PALETTE.astype(np.int32), device=dev)
flag = pal[torch.randint(0, 16, (reads,), generator=gen, device=dev)].to(torch.int16).view(torch.uint16)
mapq = torch.randint(0, 61, (reads,), generator=gen, device=dev, dtype=torch.int32).to(torch.uint8)
tid = torch.randint(0, 24, (reads,), generator=gen, device=dev, dtype=torch.int32)
ntid = torch.where(torch.rand(reads, generator=gen, device=dev) < 0.95, tid, tid + 1)
pos = torch.randint(0, 100_000_000, (reads,), generator=gen, device=dev, dtype=torch.int32)
ns = 64
sample_off = np.linspace(0, reads, ns + 1).astype(np.int64)
regions = []
for _ in range(g_regions):
    a = int(rng.integers(0, 99_900_000))
    regions.append((int(rng.integers(0, 24)), a, a + int(rng.integers(1_000, 50_000))))
flags = [83, 99, 147, 163]
cols = (flag, mapq, tid, ntid, pos)
counts = CR.count_reads_columns(cols, sample_off, regions, flags, 1)
ms_call = timed(lambda: CR.count_reads_columns(cols, sample_off, regions, flags, 1), reps=3)  # with the host-side tables
# the two kernels alone, tables resident: with the bucket directory and with the plain binary search
from grid_b200._lib import load  # noqa: E402
DEFAULT_CAP = CR.MAX_BUCKETS
bounds, lo_i, hi_i = CR.region_table(regions)
first, shift, per_tid, n_tid = CR.bucket_directory(bounds)
up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
t_off, t_bits, t_bounds, t_lo, t_hi, t_first = (up(a) for a in (sample_off, CR.flag_bitmap(flags), bounds, lo_i, hi_i, first))
ws = torch.empty(load().gridb200_count_reads_ws_bytes(ns, bounds.size), dtype=torch.uint8, device=dev)
out_k = torch.empty((ns, g_regions), dtype=torch.int64, device=dev)


def kernels(directory):
    call("gridb200_count_reads", ptr(flag), ptr(mapq), ptr(tid), ptr(ntid), ptr(pos), reads, ptr(t_off), ns, ptr(t_bits), 1,
         ptr(t_bounds), int(bounds.size), ptr(t_lo), ptr(t_hi), g_regions, ptr(t_first) if directory else None, shift, per_tid,
         n_tid, ptr(out_k), ptr(ws), ws.numel(), st)


sweep = {}
for cap in (1 << 11, 1 << 13, 1 << 15, 1 << 17, 1 << 19):  # directory size against L1/L2 residency
    CR.MAX_BUCKETS = cap
    first, shift, per_tid, n_tid = CR.bucket_directory(bounds)
    t_first = up(first)
    sweep[str(first.size)] = round(timed(lambda: kernels(True)), 3)
CR.MAX_BUCKETS = DEFAULT_CAP
first, shift, per_tid, n_tid = CR.bucket_directory(bounds)
t_first = up(first)
ms = timed(lambda: kernels(True))
same = bool((out_k == counts).all())
ms_plain = timed(lambda: kernels(False))
same = same and bool((out_k == counts).all())
# check one sample against the oracle, and time the oracle's G passes over it
s = 5
sl_cols = [c[sample_off[s]:sample_off[s + 1]].cpu().numpy() for c in (flag.view(torch.int16), mapq, tid, ntid, pos)]
sl_cols[0] = sl_cols[0].view(np.uint16)
t0 = time.perf_counter()
want = so.count_reads_regions(sl_cols, [0, len(sl_cols[0])], regions[:32], flags, 1)
t_np = time.perf_counter() - t0
res["count_reads"] = {"workload": f"{reads} reads in {ns} samples x {g_regions} regions, one pass", "ms": round(ms, 3),
                      "ms_plain_binary_search": round(ms_plain, 3), "ms_python_call_with_table_build": round(ms_call, 3),
                      "directory": {"shift": shift, "buckets": int(first.size), "ms_by_buckets": sweep}, "kernels_agree": same,
                      "reads_per_s": round(reads / ms * 1e3), "GBps_at_15B_per_read": round(15 * reads / ms / 1e6, 1),
                      "frac_hbm": round(15 * reads / ms / 1e6 / hbm, 3), "total_counted": int(counts.sum()),
                      "exact_on_sample": bool((counts[s, :32].cpu().numpy() == want[0]).all()),
                      "numpy_port_read_region_pairs_per_s": round(len(sl_cols[0]) * 32 / t_np)}
res["hbm_peak_GBps"] = hbm
print(json.dumps(res))
