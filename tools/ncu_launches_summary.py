"""Per-kernel totals and shares of an ncu launch list (--metrics gpu__time_duration.sum --csv):
    python tools/ncu_launches_summary.py gpurun_out/x_launches.csv "<command that was profiled>" > profiles/x_launches_summary.json"""
import csv
import json
import re
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
hdr = rows[0]
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot = OrderedDict()
for r in rows[1:]:
    if len(r) <= iv:
        continue
    name = re.sub(r"\(.*", "", r[ik]).strip()
    v = float(r[iv].replace(",", ""))
    ms = v / 1e6 if r[iu] in ("ns", "nsecond") else v / 1e3 if r[iu] in ("us", "usecond") else v
    t = tot.setdefault(name, [0, 0.0])
    t[0] += 1
    t[1] += ms
total = sum(t[1] for t in tot.values())
out = {"command": sys.argv[2] if len(sys.argv) > 2 else "", "note": "cold-cache, serialised: compare SHARES, not absolutes; "
       "k_synth_depth is input generation (outside bench.py's timed region)", "total_ms": round(total, 3),
       "kernels": [{"kernel": k, "launches": v[0], "total_ms": round(v[1], 3), "share": round(v[1] / total, 4)}
                   for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1])]}
print(json.dumps(out, indent=1))
