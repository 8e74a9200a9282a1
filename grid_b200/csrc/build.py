#!/usr/bin/env python
"""Build libgridb200.so in-tree with nvcc for sm_100a (no GPU needed: nvcc cross-compiles)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
SOURCES = ["api.cu", "normalize.cu", "select.cu", "knn_simt.cu", "gram_tc.cu", "gram_f32.cu", "topk.cu", "dipcn.cu", "synth.cu", "ingest.cu", "textio.cu", "hi.cu", "region.cu", "nbrio.cu"]
LIB = HERE / "libgridb200.so"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
         "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr"]


def _stale(obj: Path, src: Path) -> bool:
    deps = [src, HERE / "common.cuh", HERE.parents[1] / "include" / "gridb200.h", Path(__file__)]
    deps += list(HERE.glob("*.cuh"))
    return not obj.exists() or any(d.stat().st_mtime > obj.stat().st_mtime for d in deps)


def _compile(src: str, verbose: bool) -> Path:
    obj = HERE / "build" / (src + ".o")
    obj.parent.mkdir(exist_ok=True)
    if _stale(obj, HERE / src):
        cmd = [NVCC, *FLAGS, "-c", str(HERE / src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        res = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or res.returncode:
            sys.stderr.write(res.stdout + res.stderr)
        if res.returncode:
            raise RuntimeError(f"nvcc failed on {src}")
    return obj


def build(verbose: bool = False, force: bool = False) -> Path:
    if force:
        for o in (HERE / "build").glob("*.o"):
            o.unlink()
    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), SOURCES))
    if force or not LIB.exists() or any(o.stat().st_mtime > LIB.stat().st_mtime for o in objs):
        cmd = [NVCC, "-shared", "-cudart", "static", "-o", str(LIB), *map(str, objs)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("link failed")
    return LIB


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
