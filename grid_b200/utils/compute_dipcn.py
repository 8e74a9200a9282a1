"""Step 6, drop-in for grid/utils/compute_dipcn.py: same entry point, config keys, defaults, output file."""
from __future__ import annotations

from pathlib import Path

import numpy as np

from .. import io as gio
from ..io import load_neighbors  # noqa: F401  (reference name)
from .utils import log, progress_bar


def compute_diploid_genotypes(config, console) -> None:
    """Same contract as the reference's compute_diploid_genotypes(config, console) (compute_dipcn.py:10-101)."""
    try:
        dcfg = config.get("compute_diploid_genotypes", {})
        ftype = config.get("output_file_type", "tsv")
        output_dir = config.get("output_dir", ".")
        output_file = Path(f"{output_dir}/{dcfg.get('output_file_prefix', None)}.{ftype}")
        n_nbr = dcfg.get("n_nbr", 300)
        read_counts_file = Path(f"{output_dir}/{config['count_reads'].get('output_file_prefix', None)}.{ftype}")
        zmax = config["mosdepth"]["neighbors"].get("zmax", 2.0)
        nprefix = config["mosdepth"]["neighbors"].get("output_file_prefix", None)
        neighbors_file = Path(f"{output_dir}/{nprefix}.zMax{zmax:.1f}.{ftype}.gz")
    except Exception as e:
        log(console, f"Config error: {e}", style="danger")
        return

    import torch

    from .. import engine

    from .. import handoff

    reads = gio.load_read_counts(read_counts_file)
    cached = handoff.take(neighbors_file) if handoff.enabled(config) else None
    if cached is not None:  # step 5 of this process wrote that very file: ids, %.2f scales and index lists as arrays
        _from_arrays(cached["ids"], cached["scales"], cached["idx"], reads, n_nbr, output_file, console)
        return
    if config.get("gpu", {}).get("device_text", True):  # additive key; default = resolve the lists on the GPU
        from .. import textio

        try:
            d_ids, d_scales, d_idx = textio.load_neighbors_device(neighbors_file, threads=config.get("threads", 1))
        except textio._Fallback:  # not a file step 5 wrote (foreign ids, odd fields): the reference's rules below
            pass
        else:
            _from_arrays(d_ids, d_scales, d_idx, reads, n_nbr, output_file, console)
            return
    ids, scales, lists = load_neighbors(neighbors_file)
    # universe of samples = file rows plus any neighbour id that only appears inside a list
    pos = {}
    for sid in ids:                      # a repeated id keeps its last row, as the reference's dict does
        pos.setdefault(sid, len(pos))
    row_of = {sid: i for i, sid in enumerate(ids)}
    uni = list(pos)
    uscale = {sid: scales[row_of[sid]] for sid in uni}
    for lst in lists:
        for nid, ns in lst:
            if nid not in pos:
                pos[nid] = len(pos)
                uni.append(nid)
    n_all, k = len(uni), max(1, max((len(l) for l in lists), default=1))
    counts = np.zeros((n_all, 1), dtype=np.int32)
    has = np.zeros(n_all, dtype=np.uint8)
    for sid, v in reads.items():
        if sid in pos:
            if v != int(v) or not (0 <= v < 2**31):
                raise ValueError(f"read count of {sid} is not a non-negative 32-bit integer: {v}")
            counts[pos[sid], 0] = int(v)
            has[pos[sid]] = 1
    # a neighbour's scale is whatever its entry in THIS list says (compute_dipcn.py:131-142); the array kernel uses
    # one scale per sample, which is the same thing for files written by step 5 (checked here)
    nscale = np.full(n_all, np.nan)
    for sid in uni:
        if sid in uscale:
            nscale[pos[sid]] = uscale[sid]
    idx = np.full((len(uni), k), -1, dtype=np.int32)
    for sid in uni:
        if sid not in row_of:
            continue
        for t, (nid, ns) in enumerate(lists[row_of[sid]]):
            j = pos[nid]
            if np.isnan(nscale[j]):
                nscale[j] = ns
            elif nscale[j] != ns:
                raise ValueError(f"neighbour {nid} appears with two different scales ({nscale[j]} and {ns})")
            idx[pos[sid], t] = j
    nq = len(pos) if False else sum(1 for sid in uni if sid in row_of)
    missing = {nid for lst in lists for nid, _ in lst if nid not in reads}
    with progress_bar(console, total=len(ids), description="Computing dipCN...") as (progress, task):
        dev = torch.device("cuda")
        out = engine.dipcn(torch.from_numpy(counts).to(dev), torch.from_numpy(has).to(dev),
                           torch.from_numpy(np.where(np.isnan(nscale), 1.0, nscale)).to(dev),
                           torch.from_numpy(idx[:nq]).to(dev), n_nbr).cpu().numpy()[:, 0]
        progress.advance(task, len(ids))
    rows = [(sid, float(out[pos[sid]])) for sid in uni[:nq] if not np.isnan(out[pos[sid]])]
    if missing:
        log(console, f"Warning: {len(missing)} neighbor IDs not found in read counts "
                     f"(showing up to 5: {list(missing)[:5]})", style="warning")
    gio.write_dipcn(output_file, rows)
    log(console, f"Saved {len(rows)} samples → {output_file}", style="success")


def _from_arrays(ids, scales, idx, reads, n_nbr, output_file, console) -> None:
    """Step 6 when the neighbour lists are already index arrays (unique ids, every neighbour is a row of the file)."""
    import torch

    from .. import engine

    n = len(ids)
    counts = np.zeros((n, 1), dtype=np.int32)
    has = np.zeros(n, dtype=np.uint8)
    pos = {sid: i for i, sid in enumerate(ids)}
    for sid, v in reads.items():
        i = pos.get(sid)
        if i is not None:
            if v != int(v) or not (0 <= v < 2**31):
                raise ValueError(f"read count of {sid} is not a non-negative 32-bit integer: {v}")
            counts[i, 0] = int(v)
            has[i] = 1
    with progress_bar(console, total=n, description="Computing dipCN...") as (progress, task):
        dev = torch.device("cuda")
        t_idx = idx if isinstance(idx, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(idx))
        out = engine.dipcn(torch.from_numpy(counts).to(dev), torch.from_numpy(has).to(dev),
                           torch.from_numpy(np.asarray(scales, dtype=np.float64)).to(dev), t_idx.to(dev).contiguous(),
                           n_nbr).cpu().numpy()[:, 0]
        progress.advance(task, n)
    rows = [(sid, float(out[i])) for i, sid in enumerate(ids) if not np.isnan(out[i])]
    missing_idx = np.unique(idx[idx >= 0]) if not isinstance(idx, torch.Tensor) else torch.unique(idx[idx >= 0]).cpu().numpy()
    missing = [ids[j] for j in missing_idx.tolist() if not has[j]]
    if missing:
        log(console, f"Warning: {len(missing)} neighbor IDs not found in read counts "
                     f"(showing up to 5: {missing[:5]})", style="warning")
    gio.write_dipcn(output_file, rows)
    log(console, f"Saved {len(rows)} samples → {output_file}", style="success")
