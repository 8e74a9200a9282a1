"""Out-of-core schedule (query blocks x rounds x ring hops, grid_b200.ring.run_chain_rows_blocked) on a cohort shaped like
BASELINE config 4, scaled to the GPUs at hand: timing, where the time goes, and a sampled exact check.

    torchrun --nproc-per-node W tools/blocked_cohort_check.py --samples 40000 --bins 2900000 --block-rows 5000 --k 200

No rank ever holds its rows: every block is regenerated from the synthetic source when it is needed (two statistics
passes, then once per (query block, round)).  After the timed step `--check` query rows of rank 0 are verified against
exact brute-force distances to ALL samples (every rank regenerates its blocks' operands once more and measures them with
fp64 matmuls of small integers): ids and d2 must match bit for bit.  One JSON line on rank 0."""
import argparse
import json
import os
import sys

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
ap.add_argument("--check", type=int, default=4)
ap.add_argument("--resident-blocks", type=int, default=0, help="operands of the first m blocks stay in HBM")
a = ap.parse_args()

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29577")
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ["NCCL_DEBUG"] = os.environ.get("GRIDB200_NCCL_DEBUG", "WARN")
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
total_ms = 0.0
for _ in range(a.steps):
    (res, idx, d2, cn), t_ms = one_step()
    total_ms += t_ms
ms = torch.tensor([total_ms / a.steps], dtype=torch.float64, device=dev)
dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
stages = engine.profile_summary(a.steps)
engine.PROFILE = None
peak_gb = torch.cuda.max_memory_allocated() / 1e9

# ---- sampled exact check: rank 0's query rows against every row of the cohort ---------------------------------------
rounds, blocks = res.extra["rounds"], res.extra["blocks"]
my_cnt, my_off = blocks[rank]
nq = min(a.check, my_cnt[0]) if rank == 0 else min(a.check, blocks[0][0][0])
kdim = (res.rw + 63) // 64 * 64
rows = torch.linspace(0, blocks[0][0][0] - 1, nq).long().to(dev)          # inside rank 0's first block
q = torch.empty((kdim // 64, nq, 64), dtype=torch.float16, device=dev)
qn = torch.empty(nq, dtype=torch.int64, device=dev)
if rank == 0:
    zh0, n0 = ring.block_operand(res, source(my_off[0], my_cnt[0]), my_off[0], p)
    q.copy_(zh0[:, rows, :])
    qn.copy_(n0[rows])
    del zh0
dist.broadcast(q, src=0, group=group)
dist.broadcast(qn, src=0, group=group)
qd = q.permute(1, 0, 2).reshape(nq, -1).double()
d2_local = torch.empty((nq, cnt[rank]), dtype=torch.int64, device=dev)
for bi in range(rounds):
    zh, nrm = ring.block_operand(res, source(my_off[bi], my_cnt[bi]), my_off[bi], p)
    for r0 in range(0, my_cnt[bi], 512):                                   # fp64 matmul of small integers: exact
        blk = zh[:, r0:r0 + 512, :].permute(1, 0, 2).reshape(-1, kdim).double()
        dots = qd @ blk.T
        d2_local[:, my_off[bi] + r0:my_off[bi] + r0 + blk.shape[0]] = \
            (qn.double()[:, None] + nrm[r0:r0 + blk.shape[0]].double()[None, :] - 2.0 * dots).round().long()
    del zh
d2_all = ring._all_gather_rows(d2_local.T.contiguous(), cnt, group).T       # [nq, N]
ok, detail = True, ""
if rank == 0:
    big = torch.iinfo(torch.int64).max
    for t in range(nq):
        row = d2_all[t].clone()
        row[off[0] + int(rows[t])] = big                                     # self excluded
        key = row * (a.samples + 1) + torch.arange(a.samples, device=dev)    # (d2, index) order
        want = torch.argsort(key)[: idx.shape[1]]
        if not (torch.equal(want.int(), idx[rows[t]]) and torch.equal(row[want], d2[rows[t]])):
            ok, detail = False, f"query row {int(rows[t])} differs"
            break
    t_s = float(ms.item()) / 1e3
    gram_ms = stages.get("gram", {}).get("ms", 0.0)
    useful = 2.0 * cnt[0] * a.samples * res.rw                               # this rank's share of 2 N^2 Ru
    line = {"workload": f"blocked out-of-core ring: {a.samples} synthetic samples x {a.bins} bins, k={a.k}, "
                        f"block_rows={a.block_rows} ({rounds} rounds)", "n_gpus": world, "s_per_step": round(t_s, 3),
            "samples_per_s": round(a.samples / t_s, 1), "selected_bins": res.rw,
            "stages_rank0_ms": {k2: round(v["ms"], 1) for k2, v in stages.items()},
            "gram_tflops_rank0_full_equiv": round(useful / (gram_ms * 1e-3) / 1e12, 1) if gram_ms else None,
            "regeneration_share": round((stages.get("source", {}).get("ms", 0.0) + stages.get("transform", {}).get("ms", 0.0))
                                        / (t_s * 1e3), 3),
            "resident_blocks": a.resident_blocks, "hbm_gb_peak_rank0": round(peak_gb, 1),
            "checked_rows": nq, "exact_match": ok, "detail": detail}
    print(json.dumps(line), flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
