"""Case tables shared by gen_golden.py (dev container, needs /root/reference) and the tests (anywhere)."""

HI_CASES = {"ibs": dict(seed=41, n=60, method="ibs", n_iters=25),
            "ibd": dict(seed=42, n=50, method="ibd", n_iters=25),
            "ibd_weighted": dict(seed=43, n=50, method="ibd", weighted=True, n_iters=40, min_neighbors=2, max_neighbors=6)}
