"""Out-of-core schedule (query blocks x rounds x ring hops, grid_b200.ring.run_chain_rows_blocked) on a cohort shaped like
BASELINE config 4, scaled to the GPUs at hand: timing, where the time goes, and a sampled exact check.

    torchrun --nproc-per-node W tools/blocked_cohort_check.py --samples 40000 --bins 2900000 --block-rows 5000 --k 200

No rank ever holds its rows: every block is regenerated from the synthetic source when it is needed (two statistics
passes, then once per (query block, round) the half ring runs).  After the timed step `--check` query rows PER RANK
(spread over every rank's ownership range) are verified with oracle/cohort_check.py -- a chunked fp64 restatement of the
oracle that regenerates the cohort from the synthetic source: ids and d2 bit for bit, %.2f scales of every row, step-4
statistics, dipCN within 1e-5.  One JSON line on rank 0; exit status 1 on a mismatch."""
import argparse
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from grid_b200 import engine, ring, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--samples", type=int, default=9606)
ap.add_argument("--bins", type=int, default=2_900_000)
ap.add_argument("--block-rows", type=int, default=3202)
ap.add_argument("--k", type=int, default=200)
ap.add_argument("--n-nbr", type=int, default=200)
ap.add_argument("--steps", type=int, default=1)
ap.add_argument("--warmup", type=int, default=0)
ap.add_argument("--check", type=int, default=8, help="query rows verified per rank")
ap.add_argument("--resident-blocks", type=int, default=0, help="operands of the first m blocks stay in HBM")
a = ap.parse_args()

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29577")
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
group = dist.group.WORLD
cnt, off = ring.shard_rows(a.samples, world)
SEED = 20251104
p = engine.Params(num_neighbors=a.k, n_nbr=a.n_nbr)
counts = torch.from_numpy(synth.read_counts(SEED, a.samples, [42599])[off[rank]:off[rank] + cnt[rank]]).to(dev)


_, _layout = ring.block_layout(cnt, a.block_rows)
_depth_buf = torch.empty(max(_layout[rank][0]) * ((a.bins + 3) // 4 * 4), dtype=torch.float32, device=dev)


def source(r0, nr):
    """One reusable fp32 block (the chain consumes a block before it asks for the next one)."""
    return engine.synth_depth(SEED, nr, a.bins, off[rank] + r0, 0, out=_depth_buf)


def barrier():
    dist.barrier(group=group)
    torch.cuda.synchronize()


def one_step():
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    out = ring.run_chain_rows_blocked(source, cnt[rank], a.bins, counts, p, group, a.block_rows,
                                      resident_blocks=a.resident_blocks)
    ev1.record()
    barrier()
    return out, ev0.elapsed_time(ev1)


for _ in range(a.warmup):
    one_step()
engine.PROFILE = {}
engine.GRAM_ELEMS = 0
total_ms = 0.0
for _ in range(a.steps):
    (res, idx, d2, cn), t_ms = one_step()
    total_ms += t_ms
ms = torch.tensor([total_ms / a.steps], dtype=torch.float64, device=dev)
dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
stages = engine.profile_summary(a.steps)
engine.PROFILE = None
peak_gb = torch.cuda.max_memory_allocated() / 1e9

# ---- sampled check: `--check` query rows of EVERY rank against the fp64 checker (cohort regenerated from the source) ----
import numpy as np  # noqa: E402

from oracle import cohort_check as C  # noqa: E402

rounds, blocks = res.extra["rounds"], res.extra["blocks"]
t0 = time.perf_counter()
counts_all = synth.read_counts(SEED, a.samples, [42599])
qrows = C.spread_rows(a.samples, world, a.check, off, cnt)
chk = C.CohortCheck(lambda r0, nr, c0, nc: engine.synth_depth(SEED, nr, nc, r0, c0, device=dev), a.samples, a.bins,
                    dict(num_neighbors=a.k, n_nbr=a.n_nbr), dev, group=group, sharding="rows",
                    rows=(off[rank], off[rank] + cnt[rank]))
out = chk.run(qrows)
mine = [t for t, q in enumerate(qrows) if off[rank] <= q < off[rank] + cnt[rank]]
loc = torch.tensor([qrows[t] - off[rank] for t in mine], dtype=torch.int64, device=dev)
ids_ok = bool(np.array_equal(idx[loc].cpu().numpy(), out["idx"][mine]))
d2_ok = bool(np.array_equal(d2[loc].cpu().numpy(), out["d2"][mine]))
want_cn = C.dipcn_rows(counts_all, out["scale_q"], [qrows[t] for t in mine], out["idx"][mine], a.n_nbr)
rel = float(np.nanmax(np.abs(cn[loc].cpu().numpy() - want_cn) / np.abs(want_cn))) if mine else 0.0
ties = set(out["tie_rows"].tolist())
sq = res.scale_q.cpu().numpy()
bad = [i for i in np.nonzero(sq != out["scale_q"][off[rank]:off[rank] + cnt[rank]])[0].tolist() if i + off[rank] not in ties]
sel_ok = res.rw == len(out["sel"]) and res.r_use == out["r_use"]
flags = torch.tensor([float(ids_ok), float(d2_ok), float(not bad), float(sel_ok), -rel], dtype=torch.float64, device=dev)
dist.all_reduce(flags, op=dist.ReduceOp.MIN, group=group)
fl = flags.tolist()
ok = all(x == 1.0 for x in fl[:4]) and -fl[4] <= 1e-5
check_s = time.perf_counter() - t0
if rank == 0:
    t_s = float(ms.item()) / 1e3
    gram_ms = stages.get("gram", {}).get("ms", 0.0)
    executed = 2.0 * engine.GRAM_ELEMS / max(1, a.steps) * res.rw              # what the tensor pipe did on this rank
    useful = 2.0 * cnt[0] * a.samples * res.rw                                   # this rank's share of 2 N^2 Ru
    peak = 1378.5
    try:
        peak = json.loads(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                            "MEASURED_PEAKS.json")).read()).get("bf16_tflops_sustained", peak)
    except Exception:
        pass
    line = {"workload": f"blocked out-of-core ring: {a.samples} synthetic samples x {a.bins} bins, k={a.k}, "
                        f"block_rows={a.block_rows} ({rounds} rounds)", "n_gpus": world, "s_per_step": round(t_s, 3),
            "samples_per_s": round(a.samples / t_s, 1), "selected_bins": res.rw, "schedule": res.extra.get("ring_schedule"),
            "stages_rank0_ms": {k2: round(v["ms"], 1) for k2, v in stages.items()},
            "gram_tflops_rank0": round(executed / (gram_ms * 1e-3) / 1e12, 1) if gram_ms else None,
            "gram_frac_of_tensor_peak": round(executed / (gram_ms * 1e-3) / 1e12 / peak, 3) if gram_ms else None,
            "gram_tflops_rank0_full_equiv": round(useful / (gram_ms * 1e-3) / 1e12, 1) if gram_ms else None,
            "regeneration_share": round((stages.get("source", {}).get("ms", 0.0) + stages.get("transform", {}).get("ms", 0.0))
                                        / (t_s * 1e3), 3),
            "resident_blocks": a.resident_blocks, "hbm_gb_peak_rank0": round(peak_gb, 1),
            "checked_rows": len(qrows), "checked_rows_per_rank": a.check, "ids_exact": fl[0] == 1.0, "d2_exact": fl[1] == 1.0,
            "scale_q_exact": fl[2] == 1.0, "selected_bins_equal": fl[3] == 1.0, "dipcn_max_rel": -fl[4],
            "exact_match": ok, "check_seconds": round(check_s, 1)}
    print(json.dumps(line), flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
