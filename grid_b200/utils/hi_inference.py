"""Step 7, drop-in for grid/utils/hi_inference.py: same entry point, config keys, defaults and output file.

Host: the three small text inputs (dipCN table, IBS or IBD haplotype neighbours) -> CSR arrays and the level schedule
that lets the GPU keep the reference's in-place sweep order.  Device: the sweeps and the final imputation
(csrc/hi.cu, gridb200_hi_phase / gridb200_hi_impute), bit-identical doubles.
"""
from __future__ import annotations

import gzip
from pathlib import Path

import numpy as np

from .utils import log


def open_maybe_gz(path, mode="rt"):
    """utils.py:250-253."""
    return gzip.open(path, mode) if str(path).endswith(".gz") else open(path, mode)


def _read_dip_cn_file(dip_cn_file):
    """(IDs, IRRs, IDtoInd); rows whose second field is not a number (the header) are skipped (hi_inference.py:10-31)."""
    ids, vals, index = [], [], {}
    with open_maybe_gz(dip_cn_file) as fh:
        for line in fh:
            parts = line.split()
            if len(parts) < 2:
                continue
            try:
                value = float(parts[1])
            except ValueError:
                continue
            index[parts[0]] = len(vals)
            ids.append(parts[0])
            vals.append(value)
    return ids, vals, index


class HapNeighbours:
    """Per-haplotype neighbour lists in CSR form (list order = the reference's)."""

    def __init__(self, lists):
        self.count = np.fromiter((len(l) for l in lists), dtype=np.int64, count=len(lists))
        self.ptr = np.zeros(len(lists) + 1, dtype=np.int64)
        np.cumsum(self.count, out=self.ptr[1:])
        flat = [e for l in lists for e in l]
        self.hap = np.fromiter((e[0] for e in flat), dtype=np.int32, count=len(flat))
        self.w = np.fromiter((e[1] for e in flat), dtype=np.float64, count=len(flat))


def _load_ibs_neighbors(neighbors_file, IDtoInd, MAX_NBR):
    """computeIBSpbwt output -> lists of (neighbour haplotype, 1.0): first MAX_NBR accepted rows per haplotype,
    haplotypes numbered 1/2 in the file (hi_inference.py:34-75)."""
    lists = [[] for _ in range(2 * len(IDtoInd))]
    with open_maybe_gz(neighbors_file) as fh:
        next(fh)  # header
        for line in fh:
            parts = line.split()
            if len(parts) < 7:
                continue
            try:
                hap, hap_nbr = int(parts[1]), int(parts[6])
            except ValueError:
                continue
            if hap not in (1, 2) or hap_nbr not in (1, 2):
                continue
            i, j = IDtoInd.get(parts[0]), IDtoInd.get(parts[5])
            if i is None or j is None:
                continue
            mine = lists[2 * i + hap - 1]
            if len(mine) < MAX_NBR:
                mine.append((2 * j + hap_nbr - 1, 1.0))
    return lists


def _columns(line: str, need: int):
    """Fields of one iLASH line -- tab-separated, or any whitespace when that yields too few -- or None for a blank or
    short line (what hi_inference.py:117-124 accepts)."""
    line = line.strip()
    for sep in ("\t", None):
        parts = line.split(sep) if line else []
        if len(parts) >= need:
            return parts
    return None


def _load_ibd_neighbors(ilash_file, IDtoInd, MAX_NBR, region_start, region_end, min_length=0.5, min_match=0.70,
                        weighted=False, weight_scale=1_000_000):
    """iLASH output -> lists of (neighbour haplotype, weight): both directions of every kept segment, the MAX_NBR
    longest per haplotype (stable), weight 1 or the Lorentzian distance weight times the match score
    (hi_inference.py:78-176)."""
    cand: dict[int, list] = {}
    with open_maybe_gz(ilash_file) as fh:
        for parts in (_columns(line, 11) for line in fh):
            if parts is None:
                continue
            try:
                bp1, bp2 = int(parts[5]), int(parts[6])
                length, match = float(parts[9]), float(parts[10])
            except (ValueError, IndexError):
                continue
            if length < min_length or match < min_match:
                continue
            try:
                hap1, hap2 = int(parts[1].rsplit("_", 1)[-1]), int(parts[3].rsplit("_", 1)[-1])
            except ValueError:
                continue
            if hap1 not in (0, 1) or hap2 not in (0, 1):
                continue
            i, j = IDtoInd.get(parts[0]), IDtoInd.get(parts[2])
            if i is None or j is None:
                continue
            w = 1.0
            if weighted:
                gap = float(region_start - bp2) if bp2 < region_start else float(bp1 - region_end) if bp1 > region_end else 0.0
                w = (weight_scale / (gap + weight_scale)) * match
            a, b = 2 * i + hap1, 2 * j + hap2
            cand.setdefault(a, []).append((b, w, length))
            cand.setdefault(b, []).append((a, w, length))
    lists = [[] for _ in range(2 * len(IDtoInd))]
    for h, segs in cand.items():
        segs.sort(key=lambda s: -s[2])
        lists[h] = [(nb, w) for nb, w, _ in segs[:MAX_NBR]]
    return lists


def level_schedule(nbrs: HapNeighbours, takes_part: np.ndarray):
    """(order, level_ptr): samples that take part, grouped by level(i) = 1 + max level of the neighbour samples j < i
    that take part -- the dependency depth of the reference's in-place sweep."""
    n = len(takes_part)
    level = np.full(n, -1, dtype=np.int64)
    ptr, hap = nbrs.ptr, nbrs.hap
    for i in np.nonzero(takes_part)[0].tolist():
        js = hap[ptr[2 * i]:ptr[2 * i + 2]] >> 1
        js = js[js < i]
        level[i] = 0 if js.size == 0 else int(level[js].max()) + 1  # level = -1 for samples that never change
    active = np.nonzero(level >= 0)[0]
    order = active[np.argsort(level[active], kind="stable")].astype(np.int32)
    n_levels = int(level.max()) + 1 if active.size else 0
    level_ptr = np.zeros(n_levels + 1, dtype=np.int64)
    if active.size:
        np.cumsum(np.bincount(level[active], minlength=n_levels), out=level_ptr[1:])
    return order, level_ptr


def level_schedule_device(t_ptr, t_hap, takes_part: np.ndarray):
    """level_schedule on the GPU: relaxation passes (gridb200_hi_level_pass) until nothing moves, then a stable sort
    by level.  Returns (order int32 DEVICE tensor, level_ptr int64 numpy) -- the same arrays as level_schedule."""
    import torch

    from .._lib import call, ptr

    dev, n = t_ptr.device, len(takes_part)
    st = torch.cuda.current_stream().cuda_stream
    level = torch.from_numpy(np.where(takes_part, 0, -1).astype(np.int32)).to(dev)
    changed = torch.zeros(1, dtype=torch.int32, device=dev)
    while n:
        changed.zero_()
        for _ in range(8):  # the flag is read once per eight passes; a pass over 5e4 samples takes microseconds
            call("gridb200_hi_level_pass", ptr(t_ptr), ptr(t_hap), n, ptr(level), ptr(changed), st)
        if int(changed.item()) == 0:
            break
    lv = level.to(torch.int64)
    act = lv >= 0
    n_active = int(act.sum().item())
    if n_active == 0:
        return torch.zeros(0, dtype=torch.int32, device=dev), np.zeros(1, dtype=np.int64)
    key = torch.where(act, lv, torch.full_like(lv, 1 << 40))
    order = torch.sort(key, stable=True).indices[:n_active].to(torch.int32).contiguous()
    n_levels = int(lv.max().item()) + 1
    level_ptr = np.zeros(n_levels + 1, dtype=np.int64)
    level_ptr[1:] = torch.cumsum(torch.bincount(lv[act], minlength=n_levels), 0).cpu().numpy()
    return order, level_ptr


def _run_phasing(IRRs, hap_nbrs, MIN_NBR, N_ITERS, console):
    """(hap_IRRs [2N], mean_IRRs, imputed [2N]) on the GPU; same values as the reference's loop (hi_inference.py:179-256)."""
    import torch

    from .._lib import call, ptr

    n = len(IRRs)
    nbrs = hap_nbrs if isinstance(hap_nbrs, HapNeighbours) else HapNeighbours(hap_nbrs)
    irr = np.asarray(IRRs, dtype=np.float64)
    takes_part = (nbrs.count[0::2] >= MIN_NBR) & (nbrs.count[1::2] >= MIN_NBR)
    hap0 = np.full(2 * n, np.nan)
    hap0[0::2] = np.where(takes_part, irr / 2, np.nan)
    hap0[1::2] = hap0[0::2]
    n_phase = int(takes_part.sum())
    mean_irr = 0.0
    for v in irr[takes_part].tolist():  # sequential sum, as the reference accumulates it
        mean_irr += v
    log(console, f"Phasing {n_phase} samples with >={MIN_NBR} neighbors for both haps")
    if n_phase:
        mean_irr /= n_phase
        tot = int(nbrs.count[0::2][takes_part].sum() + nbrs.count[1::2][takes_part].sum())
        log(console, f"Avg neighbors per hap: {tot / (2.0 * n_phase):.2f}")
    if n == 0:
        return hap0, mean_irr, hap0.copy()
    # a NaN IRR never takes part in the sweeps (the reference tests the haplotype value, which is then NaN)
    dev = torch.device("cuda")
    st = torch.cuda.current_stream().cuda_stream
    d = {k: torch.from_numpy(v).to(dev) for k, v in dict(irr=irr, ptr=nbrs.ptr, hap=nbrs.hap, w=nbrs.w, val=hap0).items()}
    d["order"], level_ptr = level_schedule_device(d["ptr"], d["hap"], takes_part & ~np.isnan(irr))
    old = torch.empty(2 * n + len(level_ptr) + 1, dtype=torch.float64, device=dev)  # snapshot + the level table
    imp = torch.empty_like(d["val"])
    call("gridb200_hi_phase", ptr(d["irr"]), n, ptr(d["ptr"]), ptr(d["hap"]), ptr(d["w"]), ptr(d["order"]),
         level_ptr.ctypes.data, len(level_ptr) - 1, int(N_ITERS), ptr(d["val"]), ptr(old), st)
    call("gridb200_hi_impute", ptr(d["ptr"]), ptr(d["hap"]), ptr(d["w"]), ptr(d["val"]), n, float(mean_irr), ptr(imp), st)
    return d["val"].cpu().numpy(), mean_irr, imp.cpu().numpy()


class _Settings:
    """The step's knobs, read with the reference's keys and in-function defaults (hi_inference.py:265-280)."""

    def __init__(self, config):
        sec = config.get("compute_haploid_genotypes", {})
        ext = config.get("output_file_type", "tsv")
        folder = config.get("output_dir", ".")
        self.section = sec
        self.out_path = Path(f"{folder}/{sec.get('output_file_prefix', 'haploid_genotypes')}.{ext}")
        self.dipcn_path = Path(f"{folder}/{config['compute_diploid_genotypes'].get('output_file_prefix')}.{ext}")
        self.method = sec.get("method", "ibs").lower()
        self.min_nbr, self.max_nbr, self.sweeps = sec.get("min_neighbors", 1), sec.get("max_neighbors", 10), sec.get("n_iters", 100)
        self.window = (config.get("start_bp"), config.get("end_bp"))


def _neighbours_for(s: _Settings, index, console):
    """Haplotype neighbour lists by the configured method, or None after logging the reference's config error."""
    source_key = {"ibs": "ibs_output", "ibd": "ibd_output"}.get(s.method)
    if source_key is None:
        log(console, f"Config error: unknown method '{s.method}', must be 'ibs' or 'ibd'", style="danger")
        return None
    source = s.section.get(source_key)
    if not source:
        log(console, f"Config error: {source_key} required for method='{s.method}'", style="danger")
        return None
    if s.method == "ibs":
        log(console, f"Loading IBS neighbors from {source}")
        return _load_ibs_neighbors(source, index, s.max_nbr)
    weighted = s.section.get("weighted", False)
    log(console, f"Loading IBD neighbors from {source} (weighted={weighted})")
    return _load_ibd_neighbors(source, index, s.max_nbr, s.window[0], s.window[1],
                               min_length=s.section.get("min_length", 0.5), min_match=s.section.get("min_match", 0.70),
                               weighted=weighted, weight_scale=s.section.get("weight_scale", 1_000_000))


def hi_inference(config, console):
    """Same contract as the reference's hi_inference(config, console) (hi_inference.py:259-339): same config keys and
    defaults, same log lines, same output table."""
    try:
        s = _Settings(config)
    except Exception as e:
        log(console, f"Config error: {e}", style="danger")
        return

    ids, irr, index = _read_dip_cn_file(s.dipcn_path)
    n = len(irr)
    log(console, f"Read diploid IRR data for {n} samples", style="success")
    if len(index) != n:
        # a repeated id: the reference sizes its neighbour table by distinct ids and then indexes it by row -- it
        # fails with an IndexError; so do we, with a message
        raise IndexError(f"{s.dipcn_path}: {n - len(index)} repeated sample id(s)")

    hap_nbrs = _neighbours_for(s, index, console)
    if hap_nbrs is None:
        return
    phased, _mean, imputed = _run_phasing(irr, hap_nbrs, s.min_nbr, s.sweeps, console)
    h, m = phased.tolist(), imputed.tolist()
    with open_maybe_gz(s.out_path, "wt") as fout:
        fout.write("ID\tIRRs\thap1phased\thap2phased\thap1imp\thap2imp\n")
        fout.writelines(f"{ids[i]}\t{irr[i]:.2f}\t{h[2 * i]:.2f}\t{h[2 * i + 1]:.2f}\t{m[2 * i]:.2f}\t{m[2 * i + 1]:.2f}\n"
                        for i in range(n))
    log(console, f"Haploid genotypes written to {s.out_path}", style="success")
