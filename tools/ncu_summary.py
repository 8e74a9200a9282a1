"""Summarise an .ncu-rep (read here, no GPU needed) into the few numbers DESIGN.md / bench.py quote.
    python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x_summary.json"""
import csv
import io
import json
import re
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
KEEP = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed": "tensor_pipe_active_pct",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed": "fp64_pipe_active_pct",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "xu_pipe_inst_pct",
    "lts__t_sector_hit_rate.pct": "l2_hit_rate_pct",
    "l1tex__m_xbar2l1tex_read_bytes.sum": "l2_to_sm_read",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "launch__registers_per_thread": "registers_per_thread",
    "launch__grid_size": "grid_size",
    "launch__block_size": "block_size",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "sm__cycles_elapsed.avg.per_second": "sm_clock",
    "smsp__inst_executed.sum": "warp_instructions",
}
out = []
for r in rows[2:]:
    d = {"kernel": re.sub(r"\(.*", "", r[hdr.index("Kernel Name")])}
    for h, u, v in zip(hdr, units, r):
        if h in KEEP:
            try:
                d[KEEP[h]] = {"value": float(v.replace(",", "")), "unit": u}
            except ValueError:
                pass
    out.append(d)
print(json.dumps(out, indent=1))
