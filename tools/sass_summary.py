"""SASS evidence for the tcgen05 / TMA / cluster-launch-control paths: python tools/sass_summary.py [object] [out]

Disassembles grid_b200/csrc/build/gram_tc.cu.o (cuobjdump -sass) and writes, per kernel, an opcode histogram and the lines
of the instructions that prove the Blackwell paths: UTCHMMA (tcgen05.mma), LDTM (tcgen05.ld), UTMALDG (TMA), UTCBAR
(tcgen05.commit), UGETNEXTWORKID (clusterlaunchcontrol.try_cancel)."""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
obj = Path(sys.argv[1]).resolve() if len(sys.argv) > 1 else ROOT / "grid_b200/csrc/build/gram_tc.cu.o"
dst = Path(sys.argv[2]).resolve() if len(sys.argv) > 2 else ROOT / "profiles/r02_sass_gram_tc.txt"
txt = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True, check=True).stdout
KEYS = ("UTCHMMA", "LDTM", "UTMALDG", "UTCBAR", "UGETNEXTWORKID", "UTCATOMSWS", "SYNCS", "ATOMG", "RED", "STL", "LDL", "F2I",
        "ERRBAR", "MEMBAR", "UCGABAR")
out = [f"# SASS summary of {obj.relative_to(ROOT)} (cuobjdump -sass, sm_100a; written by tools/sass_summary.py)",
       "# per kernel: opcode histogram, then the lines of tcgen05.mma (UTCHMMA), tcgen05.ld (LDTM), TMA (UTMALDG),",
       "# tcgen05.commit (UTCBAR) and clusterlaunchcontrol.try_cancel (UGETNEXTWORKID).", ""]
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    ops, lines = collections.Counter(), []
    for ln in f.split("\n"):
        m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if not m:
            continue
        op = m.group(2)
        ops[op.split(".")[0]] += 1
        if op.startswith(("UTCHMMA", "LDTM", "UTMALDG", "UTCBAR", "UGETNEXTWORKID")):
            lines.append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/", "", ln).strip())
    out += [f"## {f.splitlines()[0].strip()}", f"total instructions: {sum(ops.values())}",
            "key opcodes: " + ", ".join(f"{k}={ops[k]}" for k in KEYS if ops.get(k)),
            "top opcodes: " + ", ".join(f"{k}={v}" for k, v in ops.most_common(14)), *lines[:40], ""]
dst.write_text("\n".join(out))
print(f"wrote {dst} ({len(out)} lines)")
