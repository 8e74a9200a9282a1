#!/usr/bin/env python
"""bench.py -- samples/sec of normalize + kNN + dipCN on a synthetic cohort (BASELINE.json metric).

    python bench.py [--gpus N --steps K --warmup W]            our CUDA path (one process per GPU under torchrun)
    python bench.py --impl reference [...]                     the reference's own CPU functions on the host cores

Workload (default = BASELINE config 2): 3,202 synthetic samples x 2,900,000 one-kb bins, k = 500 neighbours,
n_nbr = 300, one target region; a "step" is one pass of steps 4-6 over the whole cohort.
`value`  : samples / second with the depth matrix already resident in HBM.
`e2e`    : same, through the public API with the depth matrix in pinned HOST memory (H2D + results D2H timed).
Multi-GPU: the cohort is sharded by bins (strong scaling: the cohort is fixed as N grows); at N > 1
the line also carries a `ring` sub-record: one step of the ROW-sharded ring schedule (north_star's partition) on a
config-3-shaped cohort that fits the GPUs at hand.

What is timed is also CHECKED (outside the timed region): `parity_check` compares neighbour ids, squared distances,
dipCN, the %.2f scales and the step-4 codes of >= 64 query rows spread over every rank's ownership range with
oracle/cohort_check.py, a chunked fp64 restatement of the oracle that regenerates the cohort from the synthetic
source (nothing is taken from the product's intermediates).  A mismatch makes the process exit non-zero.
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
DIPCN_REL = 1e-5      # north_star: dipCN within 1e-5 relative


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
    ap.add_argument("--cpu-bins", type=int, default=0, help="bins of the smallest CPU sample (0 = sized from --cpu-budget)")
    ap.add_argument("--cpu-budget", type=float, default=0.0,
                    help="seconds of CPU work the CPU arm may use (default: 150 for --impl reference, 25 for the cpu_baseline leg)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--host-format", choices=["packed", "abs", "u16", "f32"], default="packed",
                    help="e2e host buffer: the product's host format -- bit-packed codes in the smallest form that fits the "
                         "cohort: absolute hundredths (13 bits for a 30x cohort) or residuals against sample depth x bin bias "
                         "(11 bits), + escape list --, 'abs' = absolute codes only, its 16-bit variant, or raw fp32")
    ap.add_argument("--device-format", choices=["u16", "f32"], default="u16",
                    help="resident form of the depth matrix: uint16 hundredths + escape table (2 bytes per cell, the default; "
                         "bit-identical results) or the packed fp32 matrix (4 bytes per cell)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--timeline", action="store_true",
                    help="debug: one extra step after the timed region with device/host timestamps of every stage, collective "
                         "and host read (rank 0's view, key 'timeline')")
    ap.add_argument("--no-check", action="store_true", help="debug: skip the parity check (the line says so)")
    ap.add_argument("--check-rows", type=int, default=64, help="query rows verified after the timed region (>= 64 by default)")
    ap.add_argument("--no-ring", action="store_true", help="skip the row-sharded ring sub-record at N > 1")
    ap.add_argument("--ring-samples-per-gpu", type=int, default=6250)
    ap.add_argument("--ring-bins", type=int, default=1_000_000)
    ap.add_argument("--ring-k", type=int, default=200)
    ap.add_argument("--ring-regen", action="store_true",
                    help="ring sub-record: regenerate the depth matrix before every step (outside the timed region) and let "
                         "the chain free it after step 4 -- what makes BASELINE config 3's full width (--ring-bins 2900000) fit")
    ap.add_argument("--gram-impl", type=int, default=0, help="debug: 1 = SIMT cross-check kernel")
    ap.add_argument("--norm-only", action="store_true", help="debug: time step 4 only (not a valid bench line)")
    ap.add_argument("--sharding", default="bins", choices=["bins", "rows"],
                    help="multi-GPU partition of the MAIN measurement: 'bins' (default; all samples x a bin slice per rank, "
                         "N x N partials reduce-scattered) or 'rows' (a sample block per rank, reference shards ring-passed)")
    return ap.parse_args()


def workload_name(a):
    tag = "config2: " if (a.samples, a.bins) == (3202, 2_900_000) else "custom: "
    return f"{tag}{a.samples} synthetic samples x {a.bins} 1 kb bins, k={a.k}, n_nbr={a.n_nbr}, 1 region"


# ------------------------------------------------------------------------------------------------ CPU arms
def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def pin_host_threads() -> int:
    """All host cores for BLAS/OpenMP, whatever the launcher exported (torchrun sets OMP_NUM_THREADS=1).
    Must run before numpy / scikit-learn are imported."""
    n = host_threads()
    for v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS", "NUMEXPR_NUM_THREADS"):
        os.environ[v] = str(n)
    return n


def thread_report():
    try:
        from threadpoolctl import threadpool_info

        return [{"api": i.get("user_api"), "lib": i.get("internal_api"), "threads": i.get("num_threads")} for i in threadpool_info()]
    except Exception as exc:  # pragma: no cover
        return [{"error": str(exc)[:80]}]


class CpuArm:
    """One pass of steps 4-6 on the host for a bin slice of the workload's cohort.

    kind "reference": the reference's OWN functions, staged unmodified under oracle/_ref by oracle/make_ref.py --
    normalize_matrix, select_high_variance_regions (normalize_mosdepth.py:419-499), filter_regions_by_variance,
    find_neighbors_sklearn, save_neighbors (find_neighbors.py:128-267), compute_diploid_genotypes (compute_dipcn.py:10-101,
    through its real files).  Not timed, in the reference's favour: parsing the mosdepth text and writing / re-reading the
    N x Rw normalised text matrix (95 % of its wall time on real files); the %.2f/%.3f quantisation of that hand-off is
    applied in memory.  kind "port": oracle.chain_arrays (when oracle/_ref is absent).
    """

    def __init__(self, a):
        from oracle import make_ref

        self.a = a
        self.mods = make_ref.load()
        self.kind = "reference" if self.mods else "port"

    def step(self, bins: int, tail: bool = False) -> dict:
        """Stage seconds of one pass over `bins` bins.  tail: also run the part whose cost does not depend on the number
        of bins (neighbours file + step 6: N x k entries) -- timed once and added as a constant, so that its run-to-run
        noise (seconds) does not swamp the slope of the bins-dependent stages at the bounded widths."""
        import numpy as np

        from grid_b200 import synth
        from oracle import grid_oracle as O

        a = self.a
        depth = synth.depth_hundredths(SEED, a.samples, bins) / 100.0          # the decimals the reference would parse
        counts = synth.read_counts(SEED, a.samples, [LPA_LEN])
        st = {}
        t0 = time.perf_counter()
        if self.mods is None:
            O.chain_arrays(depth, counts, num_neighbors=a.k, n_nbr=a.n_nbr, exact_knn=False)
            st["chain"] = time.perf_counter() - t0
            return st
        import tempfile
        from contextlib import contextmanager
        from unittest.mock import MagicMock, patch

        ref_norm, ref_nbr, ref_dip = self.mods
        ids = synth.sample_ids(a.samples)
        d = np.array(depth, dtype=np.float64)
        d[~(d > 0)] = np.nan
        with np.errstate(all="ignore"):                                             # depth filter (compute_population_mean_depths)
            pop = np.nanmean(d, axis=0)
        mat = np.ascontiguousarray(d[:, (pop >= 20.0) & (pop <= 100.0)])
        raw_means = np.nanmean(mat, axis=1)
        st["filter"] = time.perf_counter() - t0
        t1 = time.perf_counter()
        with np.errstate(all="ignore"):
            z, vr, mu, s2 = ref_norm.normalize_matrix(mat)
        sel = ref_norm.select_high_variance_regions(vr, 0.1)
        st["normalize"] = time.perf_counter() - t1
        t1 = time.perf_counter()
        zs = z[:, sel]                                                              # the text hand-off: %.2f cells, %.3f ratios
        ok = ~np.isnan(zs)
        zq = np.full(zs.shape, np.nan)
        zq[ok] = O.hundredths(zs[ok]) / 100.0
        ratios = O.q3(np.array([vr[j] for j in sel]))
        st["handoff"] = time.perf_counter() - t1
        t1 = time.perf_counter()
        keep, r_use = ref_nbr.filter_regions_by_variance(ratios, 1.0, 1000.0, None)
        data = np.nan_to_num(np.clip(zq, -2.0, 2.0), nan=0.0)[:, keep]              # find_neighbors.py:57-65
        nbrs = ref_nbr.find_neighbors_sklearn(data, ids, a.k)
        st["knn"] = time.perf_counter() - t1
        st["chain"] = time.perf_counter() - t0
        if not tail:
            return st
        t1 = time.perf_counter()

        @contextmanager
        def quiet(*_a, **_k):
            yield MagicMock(), MagicMock()

        with tempfile.TemporaryDirectory() as tmp:
            scales = {sid: float(m) for sid, m in zip(ids, raw_means)}
            cfg = {"output_dir": tmp, "output_file_type": "tsv", "count_reads": {"output_file_prefix": "reads"},
                   "mosdepth": {"neighbors": {"zmax": 2.0, "output_file_prefix": "nbrs"}},
                   "compute_diploid_genotypes": {"output_file_prefix": "dipcn", "n_nbr": a.n_nbr}}
            ref_nbr.save_neighbors(nbrs, scales, Path(tmp) / "nbrs.zMax2.0.tsv.gz", 2.0, r_use)
            with open(Path(tmp) / "reads.tsv", "w") as fh:
                fh.write("Sample\tReads\n" + "".join(f"{s}\t{int(c)}\n" for s, c in zip(ids, counts[:, 0])))
            with patch.object(ref_dip, "progress_bar", quiet), patch.object(ref_dip, "log", lambda *x, **y: None):
                ref_dip.compute_diploid_genotypes(cfg, None)
        st["save+dipcn"] = time.perf_counter() - t1
        return st

    def measure(self, budget_s: float, steps: int, warmup: int, base_bins: int = 0) -> dict:
        """Time the chain at three slice widths (w, 2w, 4w) plus `steps` timed passes at w, fit seconds = c0 + c1 * bins
        and extrapolate to the workload's width.  w is sized from a calibration pass so the whole thing fits `budget_s`."""
        a = self.a
        cal_bins = min(2048, a.bins)
        cal_st = self.step(cal_bins, tail=True)                                     # also the warm-up (imports, thread pools)
        cal, tail_s = cal_st["chain"], cal_st.get("save+dipcn", 0.0)
        per_bin = cal / cal_bins
        w = base_bins or 1024
        if not base_bins:
            while w * 2 <= 65536 and (steps + max(0, warmup - 1) + 7) * per_bin * (w * 2) <= budget_s:
                w *= 2
        w = min(w, a.bins)
        widths = sorted({min(a.bins, w), min(a.bins, 2 * w), min(a.bins, 4 * w)})
        for _ in range(max(0, warmup - 1)):
            self.step(widths[0])
        pts, stages = [], {}
        for wd in widths:
            s = self.step(wd)
            pts.append((wd, s["chain"]))
            stages[str(wd)] = {k: round(v, 3) for k, v in s.items()}
        timed = [self.step(widths[0])["chain"] for _ in range(steps)]
        pts += [(widths[0], t) for t in timed]
        n = len(pts)
        sx, sy = sum(p[0] for p in pts), sum(p[1] for p in pts)
        sxx, sxy = sum(p[0] * p[0] for p in pts), sum(p[0] * p[1] for p in pts)
        den = n * sxx - sx * sx
        c1 = (n * sxy - sx * sy) / den if den else sy / max(sx, 1)
        c0 = (sy - c1 * sx) / n if den else 0.0
        resid = max(abs(c0 + c1 * x - y) / y for x, y in pts)
        full = c0 + c1 * a.bins + tail_s
        naive = (sum(timed) / len(timed)) * a.bins / widths[0]
        return {"kind": self.kind, "widths": widths, "stage_seconds": stages, "timed_bins": widths[0],
                "timed_seconds": [round(t, 3) for t in timed],
                "fit": {"c0_s": c0, "c1_s_per_bin": c1, "max_rel_residual": resid, "bins_independent_tail_s": tail_s},
                "full_seconds": full, "proportional_full_seconds": naive,
                "cpu_seconds_spent": sum(p[1] for p in pts) + cal + tail_s}


def cpu_line_fields(a, m: dict) -> dict:
    cores = host_threads()
    what = ("the reference's own normalize_matrix / select_high_variance_regions / filter_regions_by_variance / "
            "find_neighbors_sklearn / save_neighbors / compute_diploid_genotypes (oracle/_ref, unmodified)"
            if m["kind"] == "reference" else "oracle port (numpy fp64 + scikit-learn kNN)")
    return {"value": a.samples / m["full_seconds"], "unit": "samples/s", "cores": cores, "kind": m["kind"],
            "threads": thread_report(), "extrapolated": True,
            "sample": f"{what} on all {a.samples} samples x {m['widths']} of {a.bins} bins ({m['cpu_seconds_spent']:.1f} s of CPU "
                      f"work); seconds = c0 + c1*bins fitted (max residual {m['fit']['max_rel_residual']:.1%}) and evaluated at "
                      f"{a.bins} bins, plus the bins-independent neighbours-file + step-6 tail timed once "
                      f"({m['fit']['bins_independent_tail_s']:.1f} s); text parsing / formatting of the hand-off files not included",
            "fit": m["fit"], "stage_seconds": m["stage_seconds"], "timed_seconds": m["timed_seconds"],
            "sample_seconds": sum(m["timed_seconds"]) / max(1, len(m["timed_seconds"]))}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm(a)
    m = arm.measure(a.cpu_budget or 150.0, a.steps, a.warmup, a.cpu_bins)
    cb = cpu_line_fields(a, m)
    val = cb["value"]
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "samples/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": m["full_seconds"] * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "extrapolated": True,
        "sample_ms_per_step": cb["sample_seconds"] * 1e3,
        "config": {"workload": workload_name(a), "l2": "n/a (host)"},
        "cpu_baseline": cb,
        "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------ clocks
def bind_to_gpu_numa_node(index: int):
    """Multi-GPU runs: pin this rank's threads to the CPUs NVML names as local to its GPU, so the pinned host buffers of
    the end-to-end leg are first-touched on the GPU's own NUMA node (eight ranks pulling their inputs across the socket
    link held the 8-GPU e2e at 21 GB/s per GPU where one GPU alone reaches 55).  Returns a description or None."""
    try:
        import pynvml

        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        phys = int(vis.split(",")[index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else index
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + bit for w, word in enumerate(words) for bit in range(64) if (int(word) >> bit) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"cpus": len(cpus), "first": min(cpus), "last": max(cpus)}
    except Exception as exc:  # no NVML / not permitted: run unbound, say so
        return {"error": f"{type(exc).__name__}: {exc}"[:120]}


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


# ------------------------------------------------------------------------------------------------ parity check
def parity_check(n_total, b_total, params, dev, group, sharding, rows, bins, per_rank, res, idx, d2, cn, counts_all, row0=0):
    """Compare the LAST timed step's results with oracle/cohort_check.py on query rows spread over every rank's range.

    res/idx/d2/cn are this rank's product outputs; idx/d2/cn cover global rows row0 .. row0 + idx.shape[0].
    Returns the `parity_check` record (identical on every rank)."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from grid_b200 import engine
    from oracle import cohort_check as C

    t0 = time.perf_counter()
    world = dist.get_world_size(group) if group is not None else 1
    if sharding == "rows":
        counts_r, offs_r = res.extra["row_counts"], res.extra["row_offsets"]
        qrows = C.spread_rows(n_total, world, per_rank, offs_r, counts_r)
    else:
        qrows = C.spread_rows(n_total, world, per_rank)

    def source(r0, nr, c0, nc):
        return engine.synth_depth(SEED, nr, nc, r0, c0, device=dev)

    chk = C.CohortCheck(source, n_total, b_total, params, dev, group=group, sharding=sharding, rows=rows, bins=bins)
    out = chk.run(qrows)
    flags = {"ids_exact": True, "d2_exact": True, "scale_q_exact": True, "step4_codes_exact": True, "norms_exact": True}
    rel = 0.0
    # rows of `qrows` this rank can judge: all of them when every rank holds the full lists (bins), its own otherwise
    mine = [t for t, q in enumerate(qrows) if row0 <= q < row0 + idx.shape[0]]
    loc = torch.tensor([qrows[t] - row0 for t in mine], dtype=torch.int64, device=dev)
    if mine:
        got_idx, got_d2, got_cn = idx[loc].cpu().numpy(), d2[loc].cpu().numpy(), cn[loc].cpu().numpy()
        flags["ids_exact"] = bool(np.array_equal(got_idx, out["idx"][mine]))
        flags["d2_exact"] = bool(np.array_equal(got_d2, out["d2"][mine]))
        want_cn = C.dipcn_rows(counts_all, out["scale_q"], [qrows[t] for t in mine], out["idx"][mine], params["n_nbr"])
        with np.errstate(all="ignore"):
            r = np.abs(got_cn - want_cn) / np.abs(want_cn)
        r[np.isnan(got_cn) & np.isnan(want_cn)] = 0.0
        rel = float(np.nanmax(r)) if r.size else 0.0
        if np.isnan(r).any():
            rel = float("inf")
        # step 4: the product's operand (clipped codes at compact positions) and exact norms for those rows
        a_sel, z_sel = out["sel_local"]
        if res.zh is not None and res.rw_local == z_sel - a_sel:
            op = res.zh[:, loc, :].permute(1, 0, 2).reshape(len(mine), -1)[:, :res.rw_local]      # k-blocked -> row-major
            flags["step4_codes_exact"] = bool(torch.equal(op.to(torch.int16), out["qcodes"][mine]))
            flags["norms_exact"] = bool(np.array_equal(res.norms[loc].cpu().numpy(), out["qnorms"][mine]))
        else:
            flags["step4_codes_exact"] = res.rw_local == z_sel - a_sel
    # every row's %.2f scale (exact k + 0.005 ties are numpy's pairwise sum's to decide: listed, not judged)
    sq = res.scale_q.cpu().numpy()
    want_sq = out["scale_q"][row0:row0 + sq.shape[0]]
    ties = set(out["tie_rows"].tolist())
    bad = [i for i in np.nonzero(sq != want_sq)[0].tolist() if (i + row0) not in ties]
    flags["scale_q_exact"] = not bad
    examples = [{"row": int(i + row0), "got": float(sq[i]), "want": float(want_sq[i])} for i in bad[:4]]
    flags["selected_bins_equal"] = bool(res.rw == len(out["sel"]) and res.r_use == out["r_use"])
    vec = torch.tensor([float(all(flags.values())), -rel] + [float(flags[k]) for k in sorted(flags)], dtype=torch.float64, device=dev)
    if group is not None:
        dist.all_reduce(vec, op=dist.ReduceOp.MIN, group=group)
    v = vec.tolist()
    rec = {"rows": len(qrows), "rows_per_rank": per_rank, "ranks_covered": world}
    rec.update({k: bool(v[2 + i]) for i, k in enumerate(sorted(flags))})
    rec.update({"dipcn_max_rel": -v[1], "dipcn_tol": DIPCN_REL, "scale_rows_checked": n_total, "scale_tie_rows": len(ties),
                "codes_checked_per_rank": len(mine) * (out["sel_local"][1] - out["sel_local"][0]) if sharding != "rows" else None,
                "scale_mismatches_this_rank": len(bad), "scale_mismatch_examples": examples,
                "ok": bool(v[0]) and -v[1] <= DIPCN_REL,
                "checker": "oracle/cohort_check.py: chunked fp64 restatement of oracle.chain_arrays, cohort regenerated from the "
                           "synthetic source, exact int64 distances of the query rows to ALL rows",
                "checker_stats": out["stats"], "seconds": round(time.perf_counter() - t0, 1)})
    return rec


# ------------------------------------------------------------------------------------------------ our arm
def ring_subrecord(a, dev, group, peaks):
    """One step of the row-sharded ring (north_star's partition) on a config-3-shaped cohort that fits, checked."""
    import torch
    import torch.distributed as dist

    from grid_b200 import engine, ring, synth

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n_total, b = a.ring_samples_per_gpu * world, a.ring_bins
    torch.cuda.empty_cache()   # (the main workload's cached blocks: config 3 at full width needs the room)
    p = engine.Params(num_neighbors=a.ring_k, n_nbr=min(a.ring_k, 200))
    cnt, off = ring.shard_rows(n_total, world)
    depth = engine.synth_depth(SEED, cnt[rank], b, off[rank], 0, device=dev)
    counts_all = synth.read_counts(SEED, n_total, [LPA_LEN])
    counts = torch.from_numpy(counts_all[off[rank]:off[rank] + cnt[rank]]).to(dev)

    def barrier():
        dist.barrier(group=group)
        torch.cuda.synchronize()

    def step():
        nonlocal depth
        if a.ring_regen and depth.untyped_storage().size() == 0:   # (released by the previous step)
            depth = engine.synth_depth(SEED, cnt[rank], b, off[rank], 0, device=dev)
        return ring.run_chain_rows(depth, b, counts, p, group, release_depth=a.ring_regen)

    step()                                                                          # warm-up (NCCL channels, allocator)
    barrier()
    if a.ring_regen:
        depth = engine.synth_depth(SEED, cnt[rank], b, off[rank], 0, device=dev)    # outside the timed region
        barrier()
    engine.PROFILE = {}
    engine.GRAM_ELEMS = 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    res, idx, d2, cn = step()
    ev1.record()
    barrier()
    stages = engine.profile_summary(1)
    engine.PROFILE = None
    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    ms = float(t.item())
    gram_ms = stages.get("gram", {}).get("ms", 0.0)
    executed = 2.0 * engine.GRAM_ELEMS * res.rw
    full = 2.0 * cnt[rank] * n_total * res.rw
    rec = {"workload": f"config3-shaped: {n_total} synthetic samples x {b} bins, k={a.ring_k}, row-sharded ring x{world}",
           "ms_per_step": ms, "samples_per_s": n_total / (ms / 1e3), "selected_bins": res.rw, "hops": res.extra.get("ring_hops"),
           "schedule": res.extra.get("ring_schedule"),
           "stages_rank0_ms": {k: round(v["ms"], 2) for k, v in stages.items()},
           "gram_tflops_rank0": executed / (gram_ms * 1e-3) / 1e12 if gram_ms else None,
           "gram_frac_of_tensor_peak": (executed / (gram_ms * 1e-3) / 1e12 / peaks["tensor"]) if gram_ms else None,
           "gram_tflops_rank0_full_equiv": full / (gram_ms * 1e-3) / 1e12 if gram_ms else None,
           "hbm_gb_peak_rank0": round(torch.cuda.max_memory_allocated() / 1e9, 1),
           "depth_regenerated_outside_timed_region": bool(a.ring_regen)}
    if a.timeline:
        res = idx = d2 = cn = None
        if a.ring_regen:
            depth = engine.synth_depth(SEED, cnt[rank], b, off[rank], 0, device=dev)
        barrier()
        engine.TIMELINE = []
        engine._tick("step >")
        res, idx, d2, cn = step()
        engine._tick("step <")
        torch.cuda.synchronize()
        rec["timeline"] = {"columns": ["label", "device ms", "host ms"], "rank": 0, "points": engine.timeline_summary()}
        engine.TIMELINE = None
    if not a.no_check:
        rec["parity_check"] = parity_check(n_total, b, dict(num_neighbors=p.num_neighbors, n_nbr=p.n_nbr), dev, group, "rows",
                                           (off[rank], off[rank] + cnt[rank]), None, max(2, a.check_rows // world), res, idx, d2,
                                           cn, counts_all, row0=off[rank])
        rec["exact"] = rec["parity_check"]["ok"]
    return rec


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
    numa = bind_to_gpu_numa_node(local) if world > 1 else None   # before any pinned allocation (first touch)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)     # NCCL_DEBUG is whatever the launcher set: fd 1 is parked on stderr
        group = dist.group.WORLD
    lib = _lib.load()

    # bin slice of this rank (contiguous, multiples of 1024 except the last)
    per = -(-a.bins // world)
    per = -(-per // 1024) * 1024
    lo, hi = min(a.bins, rank * per), min(a.bins, (rank + 1) * per)
    b_local = hi - lo
    bin_sizes = [min(a.bins, (r + 1) * per) - min(a.bins, r * per) for r in range(world)] if world > 1 else None
    p = engine.Params(num_neighbors=a.k, n_nbr=a.n_nbr)
    rows = a.sharding == "rows" and world > 1
    counts_all = synth.read_counts(SEED, a.samples, [LPA_LEN])
    row0 = 0
    if rows:
        from grid_b200 import ring

        rcnt, roff = ring.shard_rows(a.samples, world)
        b_local, row0 = a.bins, roff[rank]
        depth = engine.synth_depth(SEED, rcnt[rank], a.bins, roff[rank], 0, device=dev)
        counts = torch.from_numpy(counts_all[roff[rank]:roff[rank] + rcnt[rank]]).to(dev)
    else:
        depth = engine.synth_depth(SEED, a.samples, b_local, 0, lo, device=dev)
        counts = torch.from_numpy(counts_all).to(dev)
        if a.device_format == "u16":   # the matrix stays resident in 2 bytes per cell (engine.Depth16)
            d32 = depth
            depth = engine.to_depth16(d32, b_local)
            del d32
            torch.cuda.empty_cache()

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
        return engine.run_chain(depth, b_local, counts, p, impl=a.gram_impl, group=group, bin_sizes=bin_sizes)

    for _ in range(a.warmup):
        res, idx, d2, cn = step()
    barrier()
    engine.PROFILE = {}
    engine.GRAM_ELEMS = 0
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
    elems_per_step = engine.GRAM_ELEMS // max(1, a.steps)
    engine.PROFILE = None
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if group is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    ms = float(t.item())
    timeline = None
    if a.timeline:
        barrier()
        engine.TIMELINE = []
        engine._tick("step >")
        step()
        engine._tick("step <")
        torch.cuda.synchronize()
        timeline = engine.timeline_summary()
        engine.TIMELINE = None

    # ---- what was timed is checked: >= 64 query rows spread over every rank's range, against the fp64 checker --------
    check = None
    if not a.no_check and not a.norm_only:
        per_rank = max(2, -(-a.check_rows // world))
        check = parity_check(a.samples, a.bins, dict(num_neighbors=a.k, n_nbr=a.n_nbr), dev, group, "rows" if rows else "bins",
                             (row0, row0 + idx.shape[0]) if rows else None, (lo, hi), per_rank, res, idx, d2, cn, counts_all,
                             row0=row0)

    # ---- end to end: depth matrix in pinned host memory, results back on the host --------------------------
    e2e = None
    if not a.no_e2e:
        try:
            u16_dev = isinstance(depth, engine.Depth16)
            # (the host buffer is built from the fp32 form of the same synthetic cohort, regenerated for the purpose)
            d32 = engine.synth_depth(SEED, a.samples, b_local, 0, lo, device=dev) if u16_dev else depth
            if a.host_format in ("packed", "abs", "u16"):
                host = engine.pack_depth(d32, b_local, bits={"packed": "best", "abs": "auto", "u16": 16}[a.host_format])
                host_bytes = host.nbytes
            else:
                if u16_dev:
                    raise ValueError("--host-format f32 needs --device-format f32")
                host = torch.empty(depth.shape, dtype=torch.float32, pin_memory=True)
                host.copy_(depth)
                host_bytes = host.numel() * 4
            if u16_dev:
                del d32
                torch.cuda.empty_cache()
            outs = [torch.empty((idx.shape[0], idx.shape[1]), dtype=torch.int32, pin_memory=True),
                    torch.empty((cn.shape[0], 1), dtype=torch.float64, pin_memory=True)]

            def e2e_step():
                if u16_dev:
                    engine.upload_packed16(host, depth)
                elif a.host_format != "f32":
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
                   "host_format": (("residuals against the rank-one prediction (sample depth x bin bias)" if host.relative
                                    else "hundredths") +
                                   f" in {host.bits} bits per cell + escape list ({host.esc_cell.numel()} cells), unpacked "
                                   "on the device into the " + ("2-byte matrix" if u16_dev else "bit-identical fp32 matrix"))
                   if a.host_format != "f32" else "fp32",
                   # the end-to-end steps went through the host format: same cohort, so the same lists as the checked step
                   "results_equal_checked_step": bool(torch.equal(outs[0], idx.cpu()) and torch.equal(outs[1], cn.cpu()))
                   if (rank == 0 or rows) else None}
            if numa is not None:
                e2e["host_threads_bound_to_gpu_numa_node"] = numa
            del host
        except Exception as exc:  # e.g. not enough pinned host memory on this box
            e2e = {"value": None, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                   "error": f"{type(exc).__name__}: {exc}"[:200]}

    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    which = "measured (MEASURED_PEAKS.json, sustained)" if peaks else "fallback (B200_PROFILING.md)"
    tensor_peak = peaks.get("bf16_tflops_sustained", 1400.0)
    hbm_peak = peaks.get("hbm_gbs", 6650.0)

    # ---- main-line quantities that need the main run's tensors; then free them for the ring sub-record -----------------
    rw, rw_local, kdim = res.rw, res.rw_local, res.kdim
    nq0 = idx.shape[0] if rows else a.samples
    cells = nq0 * b_local
    cellb = 2 if isinstance(depth, engine.Depth16) else 4
    ring_rec = None
    if world > 1 and not a.no_ring and not rows and not a.norm_only:
        del depth, res, idx, d2, cn
        torch.cuda.empty_cache()
        try:
            ring_rec = ring_subrecord(a, dev, group, {"tensor": tensor_peak})
        except Exception as exc:
            ring_rec = {"error": f"{type(exc).__name__}: {exc}"[:300]}

    ok = True
    if rank == 0:
        gram_ms = stages.get("gram", {}).get("ms")
        # The self-Gram is symmetric: the kernel computes the 256x256 tiles on or above the diagonal only, so the flops it
        # needs are 2*K*(elements of those tiles), not 2*K*N^2.  `achieved` counts only what the tensor pipe has to do
        # (engine counts the elements of every block it launches); `full_matrix_equivalent` is the SURVEY figure 2*N*N*K.
        sym = os.environ.get("GRIDB200_GRAM_NOSYM") != "1" and a.gram_impl == 0
        elems = elems_per_step
        flops = 2.0 * elems * rw_local  # this rank's launches
        full_flops = 2.0 * nq0 * a.samples * rw_local
        # dram__bytes_read.sum + dram__bytes_write.sum of the Gram kernel from the committed `ncu --set full` capture of
        # this same workload (profiles/*_ncu_full_summary.json, written by tools/ncu_summary.py); single-GPU launch
        traffic, traffic_note = None, None
        try:
            # the newest committed capture of THIS workload's Gram launch (the symmetric, non-fused instantiation)
            for cap in sorted((ROOT / "profiles").glob("*_ncu_full_summary.json"), reverse=True) if world == 1 else []:
                hit = [k for k in json.loads(cap.read_text()) if "k_gram_tc_pair" in k["kernel"] and "1, 1>" not in k["kernel"]]
                if hit:
                    k = hit[0]
                    traffic = (k["dram_read"]["value"] + k["dram_write"]["value"]) * 1e9
                    traffic_note = (f"bytes per launch from {cap.name} (non-fused instantiation: same main loop since); the operand "
                                    f"itself is {a.samples * kdim * 2 / 1e9:.1f} GB (read once per wave of tiles)")
                    break
        except Exception:
            traffic = None
        roof = None
        if gram_ms:
            ach = flops / (gram_ms * 1e-3) / 1e12
            roof = {"bound": "tensor", "kernel": "k_gram_tc_pair (fp16 tcgen05 cta_group::2, fp32 TMEM chunks drained exactly)",
                    "achieved": ach, "peak": tensor_peak, "unit": "TFLOP/s", "frac": ach / tensor_peak,
                    "traffic": traffic, "traffic_note": traffic_note, "peak_source": which, "launch_ms": gram_ms,
                    "algorithmic": (f"2*Rw_local*(elements of the {'upper-triangle ' if sym else ''}256x256 tiles, clipped to N) = "
                                    f"2*{rw_local}*{elems} flop per step"),
                    "full_matrix_equivalent_tflops": full_flops / (gram_ms * 1e-3) / 1e12, "symmetric": sym}
        norm_ms = sum(stages.get(s, {}).get("ms", 0.0) for s in ("col_raw", "row_stats", "col_norm", "transform"))
        stage_roof = {}
        for s, byts in (("col_raw", cellb), ("row_stats", cellb), ("col_norm", cellb), ("transform", cellb + 2)):
            if s in stages:
                g = cells * byts / (stages[s]["ms"] * 1e-3) / 1e9
                stage_roof[s] = {"ms": stages[s]["ms"], "GB/s": g, "frac_hbm": g / hbm_peak, "bytes_per_cell": byts}
        line = {
            "metric": METRIC, "value": a.samples / (ms / 1e3), "unit": "samples/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64 (normalise) + f16 x f16 -> exact int64 (kNN)", "data": "synthetic",
            "config": {"workload": workload_name(a), "sharding": f"{'rows (ring)' if rows else 'bins'} x{world}", "l2": "inputs_exceed_l2 (depth matrix "
                       f"{cells * cellb / 1e9:.1f} GB per GPU)", "device_format": "u16 hundredths + escape table" if cellb == 2 else "fp32",
                       "selected_bins": rw, "gram_impl": a.gram_impl, "norm_only_debug": a.norm_only},
            "clocks": clocks.summary(), "e2e": e2e, "gpu_launches": int(launches), "roofline": roof,
            "parity_check": check if check is not None else {"skipped": True},
            "stages": stages, "normalize_roofline": {
                # three reads of the matrix + one 2-byte write per cell: 16 B with the fp32 matrix (SURVEY 8d's figure),
                # 8 B with the 2-byte matrix -- the fraction is taken on the bytes the chosen storage has to move
                "algorithmic_bytes_per_cell": 3 * cellb + (4 if cellb == 4 else 2), "ms": norm_ms,
                "GB/s": ((3 * cellb + (4 if cellb == 4 else 2)) * cells / (norm_ms * 1e-3) / 1e9) if norm_ms else None,
                "frac": ((3 * cellb + (4 if cellb == 4 else 2)) * cells / (norm_ms * 1e-3) / 1e9 / hbm_peak) if norm_ms else None,
                "fp32_accounting_equivalent_GB/s": (16 * cells / (norm_ms * 1e-3) / 1e9) if norm_ms else None,
                "peak": hbm_peak, "per_kernel": stage_roof},
        }
        if ring_rec is not None:
            line["ring"] = ring_rec
        if timeline is not None:
            line["timeline"] = {"columns": ["label", "device ms", "host ms"], "rank": 0, "points": timeline}
        if world == 1 and not a.no_cpu:
            try:
                m = CpuArm(a).measure(a.cpu_budget or 25.0, 1, 1, a.cpu_bins)
                line["cpu_baseline"] = cpu_line_fields(a, m)
            except Exception as exc:
                line["cpu_baseline"] = {"error": f"{type(exc).__name__}: {exc}"[:200]}
        emit(line)
        ok = (check is None or check["ok"]) and (ring_rec is None or ring_rec.get("exact", True))
    if group is not None:
        dist.destroy_process_group()
    if not ok:
        sys.stderr.write("bench.py: PARITY CHECK FAILED (see parity_check / ring.parity_check in the line)\n")
        sys.exit(1)


_RESULT_FD = 1


def emit(line: dict) -> None:
    """The ONE JSON line, on the process's original stdout."""
    os.write(_RESULT_FD, (json.dumps(line) + "\n").encode())


def main():
    global _RESULT_FD
    a = parse_args()
    # The CPU legs use every host core, whatever the launcher exported (torchrun sets OMP_NUM_THREADS=1 for N > 1):
    # decided here, before numpy / scikit-learn / torch create their thread pools.
    if a.impl == "reference" or int(os.environ.get("WORLD_SIZE", "1")) == 1:
        pin_host_threads()
    # Libraries write banners to file descriptor 1 (NCCL prints its version and, at NCCL_DEBUG=INFO, its topology there):
    # keep the real stdout for the result line only and send everything else that lands on fd 1 to stderr.
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
