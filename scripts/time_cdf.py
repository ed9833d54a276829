"""Timing of the exact CDF kernels for a window that starts at c0 (development aid)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from badger_amcl_b200 import cabi  # noqa: E402

ctx = cabi.Context(0)
n = 1000000
rng = np.random.default_rng(1)
for kind in ("uniform", "random"):
    w = np.full(n, 1.0 / (8 * n)) if kind == "uniform" else rng.random(n) / (4.0 * n)
    aos = np.zeros((n, 4))
    aos[:, 3] = w
    pf = cabi.ParticleFilter(ctx, 1, n)
    pf.upload_samples(aos)
    for c0 in (0.0, 0.126, 0.375, 0.62):
        end = C.c_double()
        ts = []
        for _ in range(6):
            ctx.sync()
            ctx.timer_start()
            cabi._check(pf.L.bamcl_pf_shard_cdf(pf.h, c0, C.byref(end)))
            ts.append(ctx.timer_stop_ms())
        print("%s c0=%.3f: %.1f us (min %.1f)" % (kind, c0, 1e3 * np.median(ts[1:]), 1e3 * min(ts)))
    pf.close()
