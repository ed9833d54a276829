"""Parity figures the contract promises beside the numbers, written to
profiles/r02_parity.json. Runs on the GPU box (needs oracle/_ref for the reference's brushfire; falls back to the port's
restatement of it, pinned bit for bit by tests/test_oracle_port_vs_ref.py, where _ref did not travel).

 (i)  distance caches: share of cells where the device's exact EDT equals the reference's heap-order brushfire, max
      deviation, and what the swap does to the weights (max relative difference of log weights scoring the SAME particles
      against either cache) -- 2000^2 and 4000^2 bench maps; 3-D: a 400 x 400 x 60 volume (the reference's FIFO brushfire
      over the 2001 x 2001 x 61 bench volume does not fit the bench's time budget);
 (ii) end-to-end chain (SURVEY H1b iii): the device resamples from its OWN weights (block-tree total, fixed-point or exact
      terms, rotated end points) -- mismatching resample indices against the reference chain at 5k / 100k / 1M particles
      (60-beam scan of config 1, so the CPU side stays in seconds).
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from badger_amcl_b200 import cabi, synth  # noqa: E402
from oracle import oracle as O  # noqa: E402  (checker)

RES = 0.05
out = {"generated_by": "scripts/parity_report.py", "reference_side": "oracle/_ref (compiled reference)" if O.have_ref()
       else "oracle port (restatement pinned to the compiled reference by tests/)"}
ctx = cabi.Context(0)


def ref_cache(cells):
    if O.have_ref():
        rm = O.RefMap2D(cells, RES)
        rm.update_distances(2.0)
        return rm.get_distances().reshape(cells.shape)
    pm = O.PortMap2D(cells, RES)
    pm.update_distances(2.0)
    return pm.dist.reshape(cells.shape) if hasattr(pm, "dist") else None


# ---- (i) 2-D caches
caches = {}
for W in (2000, 4000):
    cells = synth.make_grid(W, W, seed=7)
    t = time.time()
    ref = ref_cache(cells)
    t_ref = time.time() - t
    m = cabi.Map2D(ctx, cells, RES)
    ctx.timer_start()
    m.build_distances(2.0)
    ms = ctx.timer_stop_ms()
    dev = m.download_distances()
    eq = dev == ref
    row = {"cells": int(W * W), "equal_fraction": float(eq.mean()), "device_le_reference_everywhere": bool((dev <= ref).all()),
           "max_deviation_m": float(np.abs(dev - ref).max()), "device_build_ms": ms, "reference_build_s": t_ref}
    # effect on the weights: same particles, same scan, either cache
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=1080)
    n = 20000
    poses = synth.uniform_free_poses(cells, RES, n, seed=1)
    model = cabi.planar_model(model_type=cabi.MODEL_LF, max_beams=1080, scanner_pose=(0.1, 0.0, 0.0))
    ws = []
    for cache in (dev, ref):
        m.upload_distances(cache, 2.0)
        pf = cabi.ParticleFilter(ctx, n // 10, n)
        pf.init_with_poses(poses)
        pf.set_exact_terms(True)
        tot = pf.score_planar(m, model, ranges, angles, 30.0)
        ws.append(pf.weights() * n)  # un-normalised factor p = 1 + sum pz^3
        pf.close()
    lw = np.abs(np.log(ws[0]) - np.log(ws[1])) / np.maximum(np.abs(np.log(ws[1])), 1e-12)
    row["max_rel_log_weight_diff_exact_vs_reference_cache"] = float(lw.max())
    row["particles_whose_weight_changes"] = float((ws[0] != ws[1]).mean())
    caches["%dx%d" % (W, W)] = row
    print(W, row, flush=True)
    m.close()
out["distance_cache_2d"] = caches

# ---- (i) 3-D cache
keys = synth.make_voxel_world(400, 400, 60, seed=3)
ref3 = (O.RefMap3D if O.have_ref() else O.PortMap3D)(RES, keys, 0.36)
ref3.update_distances()
rd = ref3.dense()
m3 = cabi.Map3D(ctx, RES, keys, 0.36)
m3.build_distances()
dd = m3.download_distances()
out["distance_cache_3d"] = {"voxels": list(m3.shape()), "equal_fraction": float((dd == rd).mean()),
                            "max_deviation_quanta": int(np.abs(dd.astype(int) - rd.astype(int)).max()),
                            "device_le_reference_everywhere": bool((dd <= rd).all())}
print(out["distance_cache_3d"], flush=True)
m3.close()

# ---- (ii) end-to-end chain: the device resamples from its own weights
cells = synth.make_grid(2000, 2000, seed=7)
m = cabi.Map2D(ctx, cells, RES)
ref = ref_cache(cells)
m.upload_distances(ref, 2.0)  # the same cache on both sides: this isolates scoring + total + CDF + select
pm = O.PortMap2D(cells, RES)
pm.set_distances(ref, 2.0)
truth = synth.pick_truth_pose(cells, RES, seed=5)
ranges, angles = synth.make_scan(cells, RES, truth, n_beams=60)
kw = dict(model_type=1, max_beams=60, scanner_pose=(0.1, 0.0, 0.0))
chain = []
for n in (5000, 100000, 1000000):
    poses = synth.uniform_free_poses(cells, RES, n, seed=1)
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    port.srand48(12345)
    port.update_sensor(pm, O.planar_model(**kw), ranges, angles, 30.0)
    want_w = port.samples()[:, 3].copy()
    port.update_resample()
    want_src = port.resample_sources()
    for mode, exact in (("default_fixed_point", False), ("exact_terms", True)):
        pf = cabi.ParticleFilter(ctx, n // 10, n)
        pf.init_with_poses(poses)
        pf.set_exact_terms(exact)
        pf.srand48(12345)
        pf.update_sensor(m, cabi.planar_model(**kw), ranges, angles, 30.0)
        got_w = pf.weights()
        r = pf.update_resample()
        src = pf.resample_sources()
        same_n = r.sample_count == port.sample_count()
        mism = int((src != want_src).sum()) if same_n else None
        row = {"particles": n, "mode": mode, "weights_bit_equal_fraction": float((got_w == want_w).mean()),
               "max_rel_weight_diff": float(np.abs(got_w / want_w - 1.0).max()), "sample_count_equal": bool(same_n),
               "mismatching_resample_indices": mism, "kld_leaf_count_equal": bool(r.leaf_count == port.leaf_count())}
        chain.append(row)
        print(row, flush=True)
        pf.close()
out["end_to_end_chain"] = {"what": "update (device scoring + device total) then resample on the device from its OWN weights, "
                                   "against the reference chain on the same inputs and the same draw; 60-beam scan, 2000^2 map, "
                                   "the same (reference) distance cache on both sides", "rows": chain}
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "r02_parity.json"), "w") as f:
    json.dump(out, f, indent=1)
print(json.dumps(out))
