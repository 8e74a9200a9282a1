"""grid_b200 -- B200-native drop-in for GRiD's depth-inference hot path (normalize_mosdepth -> find_neighbors ->
compute_dipcn).  See DESIGN.md / INTEGRATION.md.  The CUDA library is mandatory; there is no CPU fallback."""

__version__ = "0.1.0"

STEP_TARGETS = {
    "grid.utils.normalize_mosdepth": ("normalize_mosdepth", "grid_b200.utils.normalize_mosdepth"),
    "grid.utils.find_neighbors": ("find_neighbors", "grid_b200.utils.find_neighbors"),
    "grid.utils.compute_dipcn": ("compute_diploid_genotypes", "grid_b200.utils.compute_dipcn"),
    "grid.utils.hi_inference": ("hi_inference", "grid_b200.utils.hi_inference"),  # step 7, the path's consumer
}


# Helpers of the steps before the path, rebound at function level: run_mosdepth_single_cram calls
# compute_region_coverage (mosdepth.py:185) and process_single_cram calls count_reads_in_region (count_reads.py:137) through
# their module globals.  Optional: the reference's count_reads module needs pysam to import at all.
HELPER_TARGETS = {
    "grid.utils.mosdepth": ("compute_region_coverage", "grid_b200.utils.mosdepth"),
    "grid.utils.count_reads": ("count_reads_in_region", "grid_b200.utils.count_reads"),
}


def install() -> None:
    """Rebind the step functions on the reference's modules so `grid wgs config.yaml` runs them on the GPU.

    grid/pipeline.py:66-93 imports each step lazily (`from .utils.X import f`) at call time, so replacing the
    attribute on the already-imported module is enough -- the mechanism the reference's own tests use to patch
    the steps (test/test_pipeline.py:67-87).
    """
    import importlib

    from . import _lib

    _lib.load()  # fail now, loudly, if the CUDA library is not built
    for ref_mod, (fn, mine) in STEP_TARGETS.items():
        setattr(importlib.import_module(ref_mod), fn, getattr(importlib.import_module(mine), fn))
    for ref_mod, (fn, mine) in HELPER_TARGETS.items():
        try:
            target = importlib.import_module(ref_mod)
        except ImportError:
            continue
        setattr(target, fn, getattr(importlib.import_module(mine), fn))
