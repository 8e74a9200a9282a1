"""Batch dipCN over many regions (BASELINE config 5 shape): python tools/dipcn_batch_bench.py [N] [G] [k] [n_nbr]
Prints ms per call, compulsory-byte bandwidth (N*G*4 + N*8 + N*n*4 + N*G*8) and gather bandwidth
(N*n*G*8 bytes of scaled counts summed).  Neighbour lists are random; a sampled check against the oracle runs first."""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from grid_b200 import engine  # noqa: E402
from oracle import grid_oracle as O  # noqa: E402  (checker only)

n, g, k, nn = (int(x) for x in (sys.argv[1:5] + ["50000", "734", "300", "300"][len(sys.argv) - 1:]))
gen = torch.Generator(device="cuda").manual_seed(3)
counts = torch.randint(1, 4000, (n, g), device="cuda", dtype=torch.int32, generator=gen)
has = (torch.rand(n, device="cuda", generator=gen) > 0.02).to(torch.uint8)
scales = torch.round((0.5 + torch.rand(n, device="cuda", dtype=torch.float64, generator=gen)) * 100) / 100
idx = torch.randint(0, n, (n, k), device="cuda", dtype=torch.int32, generator=gen)
out = {}
for rule in ("pipeline", "exon"):
    for _ in range(2):
        cn = engine.dipcn(counts, has, scales, idx, nn, rule=rule)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        cn = engine.dipcn(counts, has, scales, idx, nn, rule=rule)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    # sampled exactness: 16 query rows against the oracle restricted to those rows
    rows = np.linspace(0, n - 1, 16).astype(int)
    fn = O.dipcn_arrays if rule == "pipeline" else O.dipcn_arrays_exon
    c_h, h_h, s_h, i_h = counts.cpu().numpy(), has.cpu().numpy(), scales.cpu().numpy(), idx.cpu().numpy()
    ok = True
    for r in rows:
        want = np.full(g, np.nan)
        # direct restatement for one row (the array oracle loops over all rows)
        js = [j for j in i_h[r] if h_h[j]][:nn] if rule == "pipeline" else [j for j in i_h[r][:nn] if h_h[j]]
        if h_h[r] and js:
            if rule == "pipeline":
                tot = np.zeros(g)
                for j in js:
                    tot += c_h[j] / s_h[j]
                want = (c_h[r] / s_h[r]) / (tot / len(js))
            else:
                tot = np.zeros(g)
                for j in js:
                    tot += c_h[j] / s_h[j]  # all counts > 0 and scales > 0 here
                want = (c_h[r] / s_h[r]) / (tot / len(js))
        got = cn[r].cpu().numpy()
        ok &= bool(np.array_equal(np.isnan(got), np.isnan(want)) and np.allclose(got[~np.isnan(want)], want[~np.isnan(want)], rtol=1e-13))
    comp = n * g * 4 + n * 8 + n * nn * 4 + n * g * 8
    out[rule] = {"ms": round(ms, 3), "compulsory_GBps": round(comp / ms / 1e6, 1),
                 "gather_GBps": round(n * min(nn, k) * g * 8 / ms / 1e6, 1), "sampled_rows_match": ok,
                 "region_genotypes_per_s": round(n * g / ms * 1e3)}
print(json.dumps({"workload": f"dipCN batch N={n} G={g} k={k} n_nbr={nn}", **out}))
