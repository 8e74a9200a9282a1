"""Row-sharded ring at cohort scale (BASELINE config 3 shape, scaled to fit): timing + sampled exact check.

    torchrun --nproc-per-node W tools/ring_cohort_check.py --samples 50000 --bins 1000000 --k 200

Every rank holds samples/W synthetic samples x all bins.  After the timed steps, `--check` query rows of rank 0 are
verified against exact brute-force distances to ALL samples (each rank measures its own rows with fp64 matmuls --
integers below 2^53 are exact -- and the distances are gathered): neighbour ids and d2 must match bit for bit.
Prints one JSON line on rank 0.  (Measurement/validation aid; the product path is grid_b200.ring.)"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from grid_b200 import engine, ring, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--samples", type=int, default=50000)
ap.add_argument("--bins", type=int, default=1_000_000)
ap.add_argument("--k", type=int, default=200)
ap.add_argument("--n-nbr", type=int, default=200)
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--warmup", type=int, default=1)
ap.add_argument("--check", type=int, default=6)
ap.add_argument("--regen", action="store_true",
                help="regenerate the depth matrix before every step (outside the timed region) and let the chain free it "
                     "after step 4: what a one-shot pipeline run does, and what makes BASELINE config 3's full width fit")
a = ap.parse_args()

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
group = dist.group.WORLD
cnt, off = ring.shard_rows(a.samples, world)
SEED = 20251104
p = engine.Params(num_neighbors=a.k, n_nbr=a.n_nbr)
depth = engine.synth_depth(SEED, cnt[rank], a.bins, off[rank], 0, device=dev)
counts = torch.from_numpy(synth.read_counts(SEED, a.samples, [42599])[off[rank]:off[rank] + cnt[rank]]).to(dev)


def barrier():
    dist.barrier(group=group)
    torch.cuda.synchronize()


def one_step():
    global depth
    if a.regen and depth.untyped_storage().size() == 0:
        depth = engine.synth_depth(SEED, cnt[rank], a.bins, off[rank], 0, device=dev)
        barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    out = ring.run_chain_rows(depth, a.bins, counts, p, group, release_depth=a.regen)
    ev1.record()
    barrier()
    return out, ev0.elapsed_time(ev1)


for _ in range(a.warmup):
    res = idx = d2 = cn = None  # drop the previous step's operand before the next one allocates its own
    one_step()
barrier()
engine.PROFILE = {}
total_ms = 0.0
for _ in range(a.steps):
    res = idx = d2 = cn = None
    (res, idx, d2, cn), t_ms = one_step()
    total_ms += t_ms


ms = torch.tensor([total_ms / a.steps], dtype=torch.float64, device=dev)
dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
stages = engine.profile_summary(a.steps)
engine.PROFILE = None

# ---- sampled exact check ------------------------------------------------------------------------------------
zh = res.zh                                              # [kdim/64, n_local, 64] fp16 integers
nq = min(a.check, cnt[0])
rows = torch.linspace(0, cnt[0] - 1, nq).long().to(dev)  # rank 0's local rows under test
q = torch.empty((zh.shape[0], nq, 64), dtype=torch.float16, device=dev)
if rank == 0:
    q.copy_(zh[:, rows, :])
dist.broadcast(q, src=0, group=group)
qd = q.permute(1, 0, 2).reshape(nq, -1).double()
dots = torch.empty((nq, cnt[rank]), dtype=torch.float64, device=dev)
step_rows = 512
for r0 in range(0, cnt[rank], step_rows):                 # fp64 matmul of small integers: exact
    blk = zh[:, r0:r0 + step_rows, :].permute(1, 0, 2).reshape(-1, zh.shape[0] * 64).double()
    dots[:, r0:r0 + blk.shape[0]] = qd @ blk.T
qn = (qd * qd).sum(dim=1, keepdim=True)
d2_local = (qn + res.norms.double()[None, :] - 2.0 * dots).round().long()
d2_all = ring._all_gather_rows(d2_local.T.contiguous(), cnt, group).T    # [nq, N]
ok, detail = True, ""
if rank == 0:
    big = torch.iinfo(torch.int64).max
    for t in range(nq):
        row = d2_all[t].clone()
        row[off[0] + int(rows[t])] = big                                   # self excluded
        key = row * (a.samples + 1) + torch.arange(a.samples, device=dev)  # (d2, index) order
        want = torch.argsort(key)[: idx.shape[1]]
        if not (torch.equal(want.int(), idx[rows[t]]) and torch.equal(row[want], d2[rows[t]])):
            ok, detail = False, f"query row {int(rows[t])} differs"
            break
    flops = 2.0 * cnt[0] * a.samples * res.rw
    line = {"workload": f"row-sharded ring: {a.samples} synthetic samples x {a.bins} bins, k={a.k}, n_nbr={a.n_nbr}",
            "n_gpus": world, "ms_per_step": float(ms.item()), "samples_per_s": a.samples / (float(ms.item()) / 1e3),
            "selected_bins": res.rw, "stages_rank0_ms": {k2: round(v["ms"], 2) for k2, v in stages.items()},
            "gram_tflops_rank0_full_equiv": flops / (stages["gram"]["ms"] * 1e-3) / 1e12 if "gram" in stages else None,
            "checked_rows": nq, "exact_match": ok, "detail": detail,
            "hbm_gb_rank0": round(torch.cuda.max_memory_allocated() / 1e9, 1)}
    print(json.dumps(line), flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
