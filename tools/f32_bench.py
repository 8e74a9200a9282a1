"""GPU micro-benchmark (measurement aid) of the 3xTF32 general path: python tools/f32_bench.py n r [reps]
Times gridb200_gram_f32 on an n x n self block over r bins; prints ms, useful TFLOP/s (2 n^2 r) and the tensor-pipe
rate actually issued (3 tf32 products per useful one), plus the max error against an fp64 product on a sample."""
import json
import sys

import torch

sys.path.insert(0, ".")
from grid_b200._lib import call, ptr  # noqa: E402

n, r = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
kdim = (r + 31) // 32 * 32
g = torch.Generator(device="cuda").manual_seed(2)
x = torch.randn((n, r), device="cuda", generator=g, dtype=torch.float32)
hi = torch.empty((kdim // 32, n, 32), dtype=torch.float32, device="cuda")
lo = torch.empty_like(hi)
norms = torch.empty(n, dtype=torch.float64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
call("gridb200_split_tf32", ptr(x), n, r, r, ptr(hi), ptr(lo), n, ptr(norms), st)
dot = torch.empty((n, n), dtype=torch.float32, device="cuda")


def run():
    call("gridb200_gram_f32", ptr(hi), ptr(lo), n, n, ptr(hi), ptr(lo), n, n, kdim, ptr(dot), n, st)


run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
m = min(n, 256)
exact = x[:m].double() @ x[:m].double().T
err = float((dot[:m, :m].double() - exact).abs().max())
tiles = ((n + 255) // 256) ** 2
print(json.dumps({"n": n, "r": r, "ms": round(ms, 3), "useful_tflops": round(2.0 * n * n * r / ms / 1e9, 1),
                  "issued_tf32_tflops": round(3 * 2.0 * tiles * 65536 * kdim / ms / 1e9, 1), "max_abs_err_sample": err,
                  "diag_magnitude": float(exact.diagonal().mean())}), flush=True)
