"""Step 7 (hi_inference) phasing at cohort scale: python tools/hi_bench.py [N] [iters] [cpu_N]
GPU: level-scheduled exact Gauss-Seidel sweeps (gridb200_hi_phase, one cooperative launch; schedule from
gridb200_hi_level_pass); CPU: the oracle's sequential loop on cpu_N samples
(same neighbour statistics), scaled linearly in N for the comparison.  One JSON line; the GPU result of the small
cohort is checked against the loop bit for bit."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from grid_b200.utils import hi_inference as P  # noqa: E402
from oracle import hi_oracle as H  # noqa: E402  (checker / CPU leg only)

n, iters, cpu_n = (int(x) for x in (sys.argv[1:4] + ["50000", "100", "2000"][len(sys.argv) - 1:]))


def cohort(n, seed):
    rng = np.random.default_rng(seed)
    irr = rng.normal(20, 4, n).clip(2).tolist()
    lists = []
    for h in range(2 * n):
        cnt = int(rng.integers(0, 11))
        lists.append([(int(x), 1.0) for x in rng.integers(0, 2 * n, cnt)])
    return irr, lists


irr_s, lists_s = cohort(cpu_n, 1)
t0 = time.perf_counter()
want, wm = H.run_phasing(irr_s, lists_s, 1, iters)
cpu_s = time.perf_counter() - t0
got, mean, _ = P._run_phasing(irr_s, lists_s, 1, iters, None)
w = np.array(want)
exact = bool(np.array_equal(np.isnan(got), np.isnan(w)) and np.array_equal(got[~np.isnan(w)], w[~np.isnan(w)]) and mean == wm)

irr, lists = cohort(n, 2)
nb = P.HapNeighbours(lists)
part = (nb.count[0::2] >= 1) & (nb.count[1::2] >= 1)
t0 = time.perf_counter()
P.level_schedule(nb, part)
host_sched_s = time.perf_counter() - t0
P._run_phasing(irr, nb, 1, 2, None)  # warm-up
t_ptr, t_hap = torch.from_numpy(nb.ptr).cuda(), torch.from_numpy(nb.hap).cuda()
torch.cuda.synchronize()
t0 = time.perf_counter()
order, level_ptr = P.level_schedule_device(t_ptr, t_hap, part)
torch.cuda.synchronize()
sched_s = time.perf_counter() - t0
torch.cuda.synchronize()
t0 = time.perf_counter()
P._run_phasing(irr, nb, 1, iters, None)
torch.cuda.synchronize()
gpu_s = time.perf_counter() - t0
print(json.dumps({"workload": f"hi_inference phasing: N={n} samples, <=10 neighbours per haplotype, {iters} sweeps",
                  "levels": len(level_ptr) - 1, "gpu_total_s": round(gpu_s, 3), "of_which_level_schedule_s": round(sched_s, 3),
                  "host_level_schedule_s": round(host_sched_s, 3),
                  "sweep_launches": 1, "grid_barriers": iters * len(level_ptr), "samples_per_s": round(n / gpu_s),
                  "cpu_loop_s_at_cpu_N": round(cpu_s, 2), "cpu_N": cpu_n,
                  "cpu_loop_s_scaled_to_N": round(cpu_s * n / cpu_n, 1), "small_cohort_bit_exact": exact}))
