#!/usr/bin/env python
"""bench.py -- samples/sec of normalize + kNN + dipCN on a synthetic cohort (BASELINE.json metric).

    python bench.py [--gpus N --steps K --warmup W]            our CUDA path (one process per GPU under torchrun)
    python bench.py --impl reference [...]                     the reference algorithm (oracle port) on host cores

Workload (default = BASELINE config 2): 3,202 synthetic samples x 2,900,000 one-kb bins, k = 500 neighbours,
n_nbr = 300, one target region; a "step" is one pass of steps 4-6 over the whole cohort.
`value`  : samples / second with the depth matrix already resident in HBM.
`e2e`    : same, through the public API with the depth matrix in pinned HOST memory (H2D + results D2H timed).
Multi-GPU: the cohort is sharded by bins (strong scaling: the cohort is fixed as N grows), see DESIGN.md.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "samples/sec normalize+kNN+dipCN"
SEED = 20251104
LPA_LEN = 42599


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--samples", type=int, default=3202)
    ap.add_argument("--bins", type=int, default=2_900_000)
    ap.add_argument("--k", type=int, default=500)
    ap.add_argument("--n-nbr", type=int, default=300)
    ap.add_argument("--cpu-bins", type=int, default=16384, help="bins in the bounded CPU sample")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--host-format", choices=["u16", "f32"], default="u16",
                    help="e2e host buffer: packed uint16 hundredths + escape list (the product's host format) or raw fp32")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--gram-impl", type=int, default=0, help="debug: 1 = SIMT cross-check kernel")
    ap.add_argument("--norm-only", action="store_true", help="debug: time step 4 only (not a valid bench line)")
    ap.add_argument("--sharding", default="bins", choices=["bins", "rows"],
                    help="multi-GPU partition: 'bins' (default; all samples x a bin slice per rank, N x N partials "
                         "all-reduced) or 'rows' (a sample block per rank, reference shards ring-passed; the schedule "
                         "for cohorts whose N x N partials do not fit)")
    return ap.parse_args()


def workload_name(a):
    tag = "config2: " if (a.samples, a.bins) == (3202, 2_900_000) else "custom: "
    return f"{tag}{a.samples} synthetic samples x {a.bins} 1 kb bins, k={a.k}, n_nbr={a.n_nbr}, 1 region"


# ------------------------------------------------------------------------------------------------ CPU arms
def cpu_sample_step(a, bins):
    """One pass of the oracle chain on `bins` bins of the workload's cohort; returns seconds."""
    import numpy as np

    from grid_b200 import synth
    from oracle import grid_oracle as O

    depth = synth.depth_hundredths(SEED, a.samples, bins) / 100.0  # the decimals the reference would parse
    counts = synth.read_counts(SEED, a.samples, [LPA_LEN])
    t0 = time.perf_counter()
    O.chain_arrays(depth, counts, num_neighbors=a.k, n_nbr=a.n_nbr, exact_knn=False)
    return time.perf_counter() - t0


def cpu_baseline(a, steps=1):
    bins = min(a.cpu_bins, a.bins)
    secs = min(cpu_sample_step(a, bins) for _ in range(steps))
    full = secs * (a.bins / bins)  # every stage is linear in the number of bins
    cores = os.cpu_count() or 1
    return {
        "value": a.samples / full, "unit": "samples/s", "cores": cores, "kind": "port",
        "sample": f"oracle port (numpy fp64 + scikit-learn kNN, BLAS/OpenMP threads = all {cores} cores) on all "
                  f"{a.samples} samples x {bins} of {a.bins} bins: {secs:.2f} s, scaled linearly in bins",
        "sample_seconds": secs,
    }


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    for _ in range(a.warmup if a.warmup < 2 else 1):
        cpu_sample_step(a, min(a.cpu_bins, a.bins))
    bins = min(a.cpu_bins, a.bins)
    t = [cpu_sample_step(a, bins) for _ in range(a.steps)]
    per = sum(t) / len(t)
    full = per * (a.bins / bins)
    cores = os.cpu_count() or 1
    val = a.samples / full
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "samples/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": full * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a), "l2": "n/a (host)"},
        "cpu_baseline": {"value": val, "unit": "samples/s", "cores": cores, "kind": "port",
                         "sample": f"oracle port on {a.samples} samples x {bins} of {a.bins} bins per step "
                                   f"({per:.2f} s), scaled linearly in bins"},
        "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (NVML every ~10 ms; nvidia-smi as a fallback)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.index, self.sm, self.max_sm, self.reasons, self._stop = index, [], None, set(), threading.Event()
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.source = "nvml"
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            # NVML enumerates physical devices; honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = int(vis.split(",")[index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv, self.source = None, "nvidia-smi"

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                if nv is not None:
                    self.sm.append(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM))
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                    for nm, bit in (("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown),
                                    ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                                    ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown),
                                    ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap)):
                        if r & bit:
                            self.reasons.add(nm)
                    self._stop.wait(0.01)
                    continue
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                row = [x.strip() for x in out.strip().split(",")]
                if row and row[0].isdigit():
                    self.sm.append(int(row[0]))
                    self.max_sm = int(row[1]) if len(row) > 1 and row[1].isdigit() else self.max_sm
                    for i, nm in enumerate(self.NAMES):
                        if len(row) > 2 + i and row[2 + i].lower().startswith("active"):
                            self.reasons.add(nm)
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        self.thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self.thread.join(timeout=6)

    def summary(self):
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_sm,
                "reasons": [nm for nm in self.NAMES if nm in self.reasons], "samples": len(sm), "source": self.source}


# ------------------------------------------------------------------------------------------------ our arm
def run_ours(a):
    import torch
    import torch.distributed as dist

    from grid_b200 import _lib, engine, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    group = None
    if world > 1:
        os.environ["NCCL_DEBUG"] = os.environ.get("GRIDB200_NCCL_DEBUG", "WARN")  # keep NCCL's banner off stdout
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD
    lib = _lib.load()

    # bin slice of this rank (contiguous, multiples of 1024 except the last)
    per = -(-a.bins // world)
    per = -(-per // 1024) * 1024
    lo, hi = min(a.bins, rank * per), min(a.bins, (rank + 1) * per)
    b_local = hi - lo
    p = engine.Params(num_neighbors=a.k, n_nbr=a.n_nbr)
    rows = a.sharding == "rows" and world > 1
    if rows:
        from grid_b200 import ring

        rcnt, roff = ring.shard_rows(a.samples, world)
        b_local = a.bins
        depth = engine.synth_depth(SEED, rcnt[rank], a.bins, roff[rank], 0, device=dev)
        counts = torch.from_numpy(synth.read_counts(SEED, a.samples, [LPA_LEN])[roff[rank]:roff[rank] + rcnt[rank]]).to(dev)
    else:
        depth = engine.synth_depth(SEED, a.samples, b_local, 0, lo, device=dev)
        counts = torch.from_numpy(synth.read_counts(SEED, a.samples, [LPA_LEN])).to(dev)

    def barrier():
        if group is not None:
            dist.barrier(group=group)
        torch.cuda.synchronize()

    def step():
        if a.norm_only:
            r = engine.normalize(depth, b_local, p, want_zq=False, want_zh=True, group=group)
            z = torch.zeros((a.samples, 1), dtype=torch.int32, device=dev)
            return r, z, z, z.double()
        if rows:
            return ring.run_chain_rows(depth, b_local, counts, p, group, impl=a.gram_impl)
        return engine.run_chain(depth, b_local, counts, p, impl=a.gram_impl, group=group)

    for _ in range(a.warmup):
        res, idx, d2, cn = step()
    barrier()
    engine.PROFILE = {}
    l0 = lib.gridb200_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        ev0.record()
        for _ in range(a.steps):
            res, idx, d2, cn = step()
        ev1.record()
        barrier()
    ms = ev0.elapsed_time(ev1) / a.steps
    launches = (lib.gridb200_launch_count() - l0) // a.steps
    stages = engine.profile_summary(a.steps)
    engine.PROFILE = None
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if group is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    ms = float(t.item())

    # ---- end to end: depth matrix in pinned host memory, results back on the host --------------------------
    e2e = None
    if not a.no_e2e:
        try:
            if a.host_format == "u16":
                host = engine.pack_depth(depth, b_local)
                host_bytes = host.nbytes
            else:
                host = torch.empty(depth.shape, dtype=torch.float32, pin_memory=True)
                host.copy_(depth)
                host_bytes = host.numel() * 4
            outs = [torch.empty((idx.shape[0], idx.shape[1]), dtype=torch.int32, pin_memory=True),
                    torch.empty((cn.shape[0], 1), dtype=torch.float64, pin_memory=True)]

            def e2e_step():
                if a.host_format == "u16":
                    engine.upload_packed(host, depth)
                else:
                    depth.copy_(host, non_blocking=True)
                r, i, d, c = step()
                if rank == 0 or rows:  # bin-sharded ranks all hold the same lists; row-sharded ranks hold their own rows
                    outs[0].copy_(i, non_blocking=True)
                    outs[1].copy_(c, non_blocking=True)

            e2e_step()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            nrep = max(1, min(a.steps, 3))
            for _ in range(nrep):
                e2e_step()
            e1.record()
            barrier()
            te = torch.tensor([e0.elapsed_time(e1) / nrep], dtype=torch.float64, device=dev)
            if group is not None:
                dist.all_reduce(te, op=dist.ReduceOp.MAX, group=group)
            h2d = torch.tensor([host_bytes], dtype=torch.float64, device=dev)
            if group is not None:
                dist.all_reduce(h2d, group=group)
            e2e = {"value": a.samples / (float(te.item()) / 1e3), "unit": "samples/s",
                   "h2d_bytes_per_step": int(h2d.item()), "d2h_bytes_per_step": int(sum(o.numel() * o.element_size() for o in outs)),
                   "ms_per_step": float(te.item()),
                   "host_format": "uint16 hundredths + escape list, unpacked on the device (bit-identical fp32 matrix)"
                   if a.host_format == "u16" else "fp32"}
            del host
        except Exception as exc:  # e.g. not enough pinned host memory on this box
            e2e = {"value": None, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                   "error": f"{type(exc).__name__}: {exc}"[:200]}

    if rank == 0:
        peaks = {}
        pk = ROOT / "MEASURED_PEAKS.json"
        if pk.exists():
            peaks = json.loads(pk.read_text())
        which = "measured (MEASURED_PEAKS.json, sustained)" if peaks else "fallback (B200_PROFILING.md)"
        tensor_peak = peaks.get("bf16_tflops_sustained", 1400.0)
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        kdim_total = res.rw  # bins that enter the distance (useful K)
        gram_ms = stages.get("gram", {}).get("ms")
        # The self-Gram is symmetric: the kernel computes the 256x256 tiles on or above the diagonal only (then
        # mirrors them), so the flops it needs are 2*K*(elements of those tiles), not 2*K*N^2.  `achieved` counts
        # only what the tensor pipe has to do; `full_matrix_equivalent` is the SURVEY figure 2*N*N*K over the same time.
        sym = os.environ.get("GRIDB200_GRAM_NOSYM") != "1" and a.gram_impl == 0
        nq0 = rcnt[0] if rows else a.samples  # rank 0's query rows
        edges = list(range(0, nq0, 256)) + [nq0]
        sizes = [edges[i + 1] - edges[i] for i in range(len(edges) - 1)]
        elems = sum(sizes[i] * sizes[j] for i in range(len(sizes)) for j in range(i if sym else 0, len(sizes)))
        if rows:  # hop 0 is the symmetric own-shard block; the other hops are full rectangular blocks
            elems += nq0 * (a.samples - nq0)
        flops = 2.0 * elems * res.rw_local  # this rank's launch
        full_flops = 2.0 * nq0 * a.samples * res.rw_local
        # dram__bytes_read.sum + dram__bytes_write.sum of the Gram kernel from the committed `ncu --set full` capture of
        # this same workload (profiles/*_ncu_full_summary.json, written by tools/ncu_summary.py); single-GPU launch
        traffic, traffic_note = None, None
        try:
            caps = sorted((ROOT / "profiles").glob("*_ncu_full_summary.json"))
            for k in json.loads(caps[-1].read_text()) if caps and world == 1 else []:
                if "k_gram_tc_pair" in k["kernel"]:
                    traffic = (k["dram_read"]["value"] + k["dram_write"]["value"]) * 1e9
                    traffic_note = (f"bytes per launch from {caps[-1].name}; the operand itself is "
                                    f"{a.samples * res.kdim * 2 / 1e9:.1f} GB (read once per wave of tiles), dot matrix 0.08 GB")
        except Exception:
            traffic = None
        roof = None
        if gram_ms:
            ach = flops / (gram_ms * 1e-3) / 1e12
            roof = {"bound": "tensor", "kernel": "k_gram_tc_pair (fp16 tcgen05 cta_group::2, fp32 TMEM chunks drained to int64)",
                    "achieved": ach, "peak": tensor_peak, "unit": "TFLOP/s", "frac": ach / tensor_peak,
                    "traffic": traffic, "traffic_note": traffic_note, "peak_source": which, "launch_ms": gram_ms,
                    "algorithmic": (f"2*Rw_local*(elements of the {'upper-triangle ' if sym else ''}256x256 tiles, clipped to N) = "
                                    f"2*{res.rw_local}*{elems} flop per launch"),
                    "full_matrix_equivalent_tflops": full_flops / (gram_ms * 1e-3) / 1e12, "symmetric": sym}
        cells = (rcnt[0] if rows else a.samples) * b_local
        norm_ms = sum(stages.get(s, {}).get("ms", 0.0) for s in ("col_raw", "row_stats", "col_norm", "transform"))
        stage_roof = {}
        for s, byts in (("col_raw", 4), ("row_stats", 4), ("col_norm", 4), ("transform", 6)):
            if s in stages:
                g = cells * byts / (stages[s]["ms"] * 1e-3) / 1e9
                stage_roof[s] = {"ms": stages[s]["ms"], "GB/s": g, "frac_hbm": g / hbm_peak, "bytes_per_cell": byts}
        line = {
            "metric": METRIC, "value": a.samples / (ms / 1e3), "unit": "samples/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64 (normalise) + f16 x f16 -> exact int64 (kNN)", "data": "synthetic",
            "config": {"workload": workload_name(a), "sharding": f"{'rows (ring)' if rows else 'bins'} x{world}", "l2": "inputs_exceed_l2 (depth matrix "
                       f"{cells * 4 / 1e9:.1f} GB per GPU)", "selected_bins": res.rw, "gram_impl": a.gram_impl, "norm_only_debug": a.norm_only},
            "clocks": clocks.summary(), "e2e": e2e, "gpu_launches": int(launches), "roofline": roof,
            "stages": stages, "normalize_roofline": {"algorithmic_bytes_per_cell": 16, "ms": norm_ms,
                                                     "GB/s": (16 * cells / (norm_ms * 1e-3) / 1e9) if norm_ms else None,
                                                     "frac": (16 * cells / (norm_ms * 1e-3) / 1e9 / hbm_peak) if norm_ms else None,
                                                     "peak": hbm_peak, "per_kernel": stage_roof},
        }
        if world == 1 and not a.no_cpu:
            line["cpu_baseline"] = cpu_baseline(a)
        emit(line)
    if group is not None:
        dist.destroy_process_group()


_RESULT_FD = 1


def emit(line: dict) -> None:
    """The ONE JSON line, on the process's original stdout."""
    os.write(_RESULT_FD, (json.dumps(line) + "\n").encode())


def main():
    global _RESULT_FD
    a = parse_args()
    # Libraries write banners to file descriptor 1 (NCCL prints its version there at NCCL_DEBUG=WARN): keep the real
    # stdout for the result line only and send everything else that lands on fd 1 to stderr.
    sys.stdout.flush()
    try:
        saved = os.dup(1)
        os.dup2(2, 1)
        _RESULT_FD = saved
    except OSError:  # no usable stderr: leave stdout as it is
        _RESULT_FD = 1
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
