"""Timing of bamcl_pf_update_action at the bench size (development aid)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from badger_amcl_b200 import cabi, synth  # noqa: E402

ctx = cabi.Context(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
poses = synth.gaussian_poses((2.0, 1.0, -0.3), n, seed=1)
pf = cabi.ParticleFilter(ctx, n // 10, n)
pf.init_with_poses(poses)
pf.srand48(1)
for mt in range(5):
    ts = []
    for _ in range(6):
        ctx.sync()
        t0 = time.perf_counter()
        pf.update_action(mt, [0.2, 0.3, 0.25, 0.15, 0.1], [1.0, 2.0, 0.7], [0.3, 0.1, 0.2], [0.35, 0.05, 0.25])
        ts.append(time.perf_counter() - t0)
    print("model %d: %.3f ms (min %.3f) for %d samples" % (mt, 1e3 * np.median(ts[1:]), 1e3 * min(ts), n))
pf.close()
