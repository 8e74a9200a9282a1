"""Step 5, drop-in for grid/utils/find_neighbors.py: same entry point, config keys, defaults, output file.

Host: read the step-4 file, write the neighbours file.  Device: region filter on the %.3f ratios, clip / NA->0,
exact Gram matrix on tcgen05 tensor cores, k-best merge.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np

from .. import io as gio
from ..io import read_normalized_data, save_neighbors  # noqa: F401  (reference names)
from .utils import log, progress_bar


def find_neighbors(config, console):
    """Same contract as the reference's find_neighbors(config, console) (find_neighbors.py:11-77)."""
    try:
        ncfg = config["mosdepth"]["neighbors"]
        zmax = ncfg.get("zmax", 2.0)
        sigma2_max = ncfg.get("sigma2_max", 1000.0)
        n_neighbors = ncfg.get("num_neighbors", 500)
        frac_r = ncfg.get("frac_r", 1.0)
        in_prefix = config["mosdepth"]["normalize"].get("output_file_prefix")
        ftype = config.get("output_file_type", "tsv")
        output_dir = config.get("output_dir", ".")
        input_file = Path(f"{output_dir}/{in_prefix}.{ftype}.gz")
        out_prefix = ncfg.get("output_file_prefix", "neighbor_coverage")
        output_file = Path(output_dir) / f"{out_prefix}.zMax{zmax:.1f}.{ftype}.gz"
    except Exception as e:
        log(console, f"Config error: {e}", style="danger")
        return
    output_file.parent.mkdir(parents=True, exist_ok=True)

    import torch

    from .. import _lib, engine
    from .._lib import call, ptr

    dev = torch.device("cuda")
    from .. import handoff

    cached = handoff.take(input_file) if handoff.enabled(config) else None
    if cached is not None:  # this process wrote that very file a moment ago: the arrays are still on the GPU
        ids, ratios, tzq, scales = cached["ids"], cached["ratios"], cached["zq"], cached["scales"]
    elif config.get("gpu", {}).get("device_text", True):  # additive key; default = parse the cells on the GPU
        from .. import textio

        ids, ratios, tzq, scales = textio.read_normalized_data(input_file, dev, threads=config.get("threads", 1))
    else:
        ids, ratios, zq, scales = read_normalized_data(input_file)
        tzq = torch.from_numpy(np.ascontiguousarray(zq)).to(dev)
    n, rw = len(ids), len(ratios)
    p = engine.Params(zmax=zmax, sigma2_max=sigma2_max, frac_r=frac_r, num_neighbors=n_neighbors)
    tidx = td2 = None
    with progress_bar(console, total=n, description="Finding neighbors...") as (progress, task):
        if rw == 0 or n < 2:
            idx = np.full((n, 0), -1, dtype=np.int32)
            d2 = np.zeros((n, 0), dtype=np.int64)
            r_use = 0
        else:
            # region filter (find_neighbors.py:128-175) on the ratios as the file carries them
            t_ratio = torch.from_numpy(ratios).to(dev)
            sel = torch.arange(rw, dtype=torch.int32, device=dev)
            keep = torch.empty(rw, dtype=torch.uint8, device=dev)
            info = torch.tensor([1.0, 0.0, 0.0, float(rw)], dtype=torch.float64, device=dev)
            info2 = torch.empty(2, dtype=torch.float64, device=dev)
            ws = torch.empty(_lib.load().gridb200_select_ws_bytes(rw), dtype=torch.uint8, device=dev)
            call("gridb200_knn_col_filter", ptr(t_ratio), ptr(sel), rw, float(frac_r), float(sigma2_max), ptr(keep),
                 ptr(info), ptr(info2), ptr(ws), ws.numel(), torch.cuda.current_stream().cuda_stream)
            r_use = int(info2[0].item())
            if r_use == 0:  # the reference hands sklearn an N x 0 matrix here, which raises (find_neighbors.py:207-213)
                raise ValueError(f"Found array with 0 feature(s) (shape=({n}, 0)) while a minimum of 1 is required "
                                 "by NearestNeighbors.")
            if not np.isfinite(ratios).any():
                log(console, "Warning: no finite variance ratios — keeping all regions", style="warning")
            extreme = int(np.sum(ratios > sigma2_max))
            if extreme and console:
                log(console, f"Removed {extreme} / {rw} regions with sigma2ratio > {sigma2_max}", style="warning")
            from .. import multigpu

            devices = multigpu.configured_devices(config)    # additive key gpu.devices: several GPUs for the search
            if devices:
                mode = multigpu.knn_sharding(config, n)
                log(console, f"Neighbour search on {len(devices)} GPUs ({mode}-sharded)", style="info")
                hidx, hd2 = multigpu.knn(tzq[:, :rw].cpu().contiguous(), keep.cpu(), n_neighbors, p.zclip, devices, mode)
                tidx, td2 = hidx.to(dev), hd2.to(dev)
            else:
                engine.check_k(max(1, min(n_neighbors, n - 1)))
                zh, kdim, norms = engine.prepare_from_zq(tzq, rw, keep, p.zclip)
                tidx, td2 = engine.knn(zh, kdim, norms, n_neighbors, p.zclip)
            idx = tidx.cpu().numpy()
        progress.advance(task, n)
    if tidx is not None and config.get("gpu", {}).get("device_text", True):  # entries formatted on the GPU
        from .. import textio

        textio.save_neighbors(output_file, ids, scales, tidx, td2, r_use, threads=config.get("threads", 1),
                              level=config.get("gpu", {}).get("gzip_level"))
    else:
        save_neighbors(output_file, ids, scales, idx, d2 if tidx is None else td2.cpu().numpy(), r_use)
    if handoff.enabled(config) and len(set(ids)) == len(ids) and idx.shape[1] > 0:
        handoff.put(output_file, ids=list(ids), scales=np.array([float("%.2f" % v) for v in np.asarray(scales).tolist()]),
                    idx=idx)
    log(console, f"Saved neighbors to {output_file}", style="success")
