"""Quick device timing of the LF update kernel (development aid; bench.py is the contract)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from badger_amcl_b200 import cabi, synth  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1080
W = int(sys.argv[3]) if len(sys.argv) > 3 else 4000
cloud = sys.argv[4] if len(sys.argv) > 4 else "global"
ctx = cabi.Context(0)
cells = synth.make_grid(W, W, seed=7)
m = cabi.Map2D(ctx, cells, 0.05)
t = time.time()
m.build_distances(2.0)
print("edt build %.1f ms (host wall)" % ((time.time() - t) * 1e3))
truth = synth.pick_truth_pose(cells, 0.05, seed=5)
ranges, angles = synth.make_scan(cells, 0.05, truth, n_beams=B)
poses = synth.uniform_free_poses(cells, 0.05, P, seed=1) if cloud == "global" else synth.gaussian_poses(truth, P, seed=2)
pf = cabi.ParticleFilter(ctx, P // 10, P)
aos = np.concatenate([poses, np.full((P, 1), 1.0 / P)], axis=1)
pf.upload_samples(aos)
model = cabi.planar_model(model_type=1, max_beams=B, scanner_pose=(0.1, 0, 0))
pf.stage_scan(m, model, ranges, angles, 30.0)
for _ in range(3):
    pf.update_staged(m, want_total=False)
ctx.sync()
ts = []
for _ in range(10):
    ctx.flush_l2()
    ctx.timer_start()
    pf.update_staged(m, want_total=False)
    ts.append(ctx.timer_stop_ms())
valid = int(((ranges < 30.0)).sum())
ms = float(np.median(ts))
print("layout=%s cloud=%s P=%d B=%d(valid %d) W=%d: update %.3f ms (min %.3f)  %.3e evals/s  algo %.1f GB/s" % (
    os.environ.get("BAMCL_LAYOUT", "0"), cloud, P, B, valid, W, ms, min(ts), P * valid / (ms * 1e-3),
    (P * valid * 4 + P * 40 + B * 16) / (ms * 1e-3) / 1e9))
