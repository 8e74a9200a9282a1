"""GPU micro-benchmark (measurement aid): one rectangular ring-hop block, dot-matrix path vs fused Gram + k-best.

    python tools/fused_bench.py nq nr kdim k [reps]

Random integer operands in [-200, 200]; lists are bootstrapped against a first block so that thresholds are finite, then
the same second block is merged through (a) gridb200_gram + gridb200_topk_merge (+ the transposed reading) and
(b) gridb200_gram_topk + gridb200_topk_merge_cand.  Prints ms per block for both, TFLOP/s, and whether the lists agree."""
import json
import sys

import torch

sys.path.insert(0, ".")
from grid_b200 import engine  # noqa: E402

nq, nr, kdim, k = (int(x) for x in sys.argv[1:5])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
g = torch.Generator(device="cuda").manual_seed(1)


def operand(n):
    """k-blocked fp16 operand [kdim/64, n, 64] of random integers and its exact row norms (built slab by slab)."""
    zh = torch.empty((kdim // 64, n, 64), dtype=torch.float16, device="cuda")
    norms = torch.zeros(n, dtype=torch.int64, device="cuda")
    step = max(1, (1 << 26) // (n * 64))
    for b0 in range(0, kdim // 64, step):
        z = torch.randint(-200, 201, (min(step, kdim // 64 - b0), n, 64), device="cuda", generator=g, dtype=torch.int32)
        zh[b0:b0 + z.shape[0]] = z.to(torch.float16)
        norms += (z.to(torch.int64) ** 2).sum(dim=(0, 2))
    return zh, norms


# the lists are bootstrapped against blocks as large as the ones under test: what the own-shard hop of a ring leaves behind
zq, nQ = operand(nq)
zb, nB = operand(nr)
za, nA = operand(nr)
zc, nC = operand(nq)


def boot():
    q = engine.knn_init(nq, k, "cuda")
    engine.knn_block(zq, nQ, 0, za, nA, 10_000_000, kdim, 200, q[0], q[1])
    b = engine.knn_init(nr, k, "cuda")
    engine.knn_block(zb, nB, 1_000_000, zc, nC, 20_000_000, kdim, 200, b[0], b[1])
    return q, b


def timed(fn):
    out = None
    ms = []
    for _ in range(reps + 1):
        q, b = boot()
        t = engine.knn_init(nr, k, "cuda")
        tau = b[0][:, k - 1].contiguous()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(q, t, tau)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
        out = (q, t)
    return min(ms[1:]), out


def unfused(q, t, tau):
    engine.knn_block(zq, nQ, 0, zb, nB, 1_000_000, kdim, 200, q[0], q[1], t_lists=t)


def fused(q, t, tau):
    assert engine.knn_block_fused(zq, nQ, 0, zb, nB, 1_000_000, kdim, 200, q[0], q[1], tau_r=tau, t_lists=t)


engine.PROFILE = {}
ms_u, (qu, tu) = timed(unfused)
st_u = engine.profile_summary(reps + 1)
engine.PROFILE = {}
ms_f, (qf, tf) = timed(fused)
st_f = engine.profile_summary(reps + 1)
engine.PROFILE = None
# the fused transposed lists hold only what beats the owner's threshold: compare after the owner's merge
_, b1 = boot()
_, b2 = boot()
engine.merge_lists(tu[0], tu[1], b1[0], b1[1])
engine.merge_lists(tf[0], tf[1], b2[0], b2[1])
same = bool(torch.equal(qu[0], qf[0]) and torch.equal(qu[1], qf[1]) and torch.equal(b1[0], b2[0]) and torch.equal(b1[1], b2[1]))
flops = 2.0 * nq * nr * kdim
print(json.dumps({"nq": nq, "nr": nr, "kdim": kdim, "k": k, "unfused_ms": round(ms_u, 3), "fused_ms": round(ms_f, 3),
                  "unfused_tflops": round(flops / ms_u / 1e9, 1), "fused_tflops": round(flops / ms_f / 1e9, 1),
                  "unfused_stages_ms": {a: round(v["ms"], 3) for a, v in st_u.items()},
                  "fused_stages_ms": {a: round(v["ms"], 3) for a, v in st_f.items()}, "lists_equal": same}), flush=True)
