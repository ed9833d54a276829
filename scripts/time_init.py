"""Timing of the free-space list, the uniform initialisation and the pose read-out at the bench size (development aid)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from badger_amcl_b200 import cabi, synth  # noqa: E402

ctx = cabi.Context(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
cells = synth.make_grid(4000, 4000, seed=7)
m = cabi.Map2D(ctx, cells, 0.05)
m.build_distances(2.0)
pf = cabi.ParticleFilter(ctx, n // 10, n)
pf.srand48(1)


def timed(fn, reps=5):
    ts = []
    for _ in range(reps):
        ctx.sync()
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return 1e3 * np.median(ts[1:]), 1e3 * min(ts)


print("free-space list (4000x4000): %.3f ms (min %.3f), %d cells" % (*timed(lambda: m.update_free_space(0.1)), m.free_count))
print("init_uniform %d samples: %.3f ms (min %.3f)" % (n, *timed(lambda: pf.init_uniform(m))))
print("clusters %d; max_weight_pose: %.3f ms (min %.3f)" % (pf.last.cluster_count, *timed(lambda: pf.max_weight_pose())))
pf.close()
m.close()
