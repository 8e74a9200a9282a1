"""GPU micro-benchmark of gridb200_gram (debugging aid): python tools/gram_bench.py n kdim vmax [reps]
Prints ms per launch and TFLOP/s of the tiles actually computed (upper triangle, padded to 256)."""
import sys
import torch
sys.path.insert(0, ".")
from grid_b200._lib import call, ptr

n, k, vmax = (int(x) for x in sys.argv[1:4])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
g = torch.Generator(device="cuda").manual_seed(1)
t = (torch.randint(-vmax, vmax + 1, (k // 64, n, 64), device="cuda", generator=g, dtype=torch.int32)).to(torch.float16)
dot = torch.empty((n, n + (n % 2)), dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def run():
    call("gridb200_gram", ptr(t), n, n, ptr(t), n, n, k, vmax, ptr(dot), dot.stride(0), 0, 0, st)
for _ in range(2):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
tiles = (n + 255) // 256
import os
sym = os.environ.get("GRIDB200_GRAM_NOSYM") != "1"
ntiles = tiles * (tiles + 1) // 2 if sym else tiles * tiles
flops = 2.0 * ntiles * 256 * 256 * k
print(f"n={n} k={k} vmax={vmax} sym={sym}: {ms:.3f} ms  {flops / ms / 1e9:.1f} TFLOP/s issued ({ntiles} tiles)  operand {n*k*2/1e6:.0f} MB", flush=True)
