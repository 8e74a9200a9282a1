"""Case tables shared by gen_golden.py (dev container, needs /root/reference) and the tests (anywhere)."""

HI_CASES = {"ibs": dict(seed=41, n=60, method="ibs", n_iters=25),
            "ibd": dict(seed=42, n=50, method="ibd", n_iters=25),
            "ibd_weighted": dict(seed=43, n=50, method="ibd", weighted=True, n_iters=40, min_neighbors=2, max_neighbors=6)}

STEP3_COVERAGE = {"by1000": dict(seed=51, n=6), "by500": dict(seed=53, n=3, flank_kb=8, by=500)}
STEP3_READS = {"mapq1": dict(seed=52, ns=9, mean_reads=3000, flags=[83, 99, 147, 163, None], min_mapq=1),
               "mapq20_two_flags": dict(seed=54, ns=5, mean_reads=1500, flags=[99, 147], min_mapq=20),
               "no_flags": dict(seed=55, ns=3, mean_reads=500, flags=[], min_mapq=0)}
