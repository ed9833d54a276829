"""Device timings of the BASELINE.json configs that are not the bench line (development aid; results quoted in
DESIGN.md): C2 beam model 100k x 1080 on 4000^2, C3 LF + KLD resample sweep 10k..10M, C4 3-D cloud 1M x 4096."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from badger_amcl_b200 import cabi, synth  # noqa: E402

RES = 0.05
which = sys.argv[1:] or ["c2", "c3", "c4"]
ctx = cabi.Context(0)
out = {}


def med(fn, reps=7, warm=2):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        ctx.flush_l2()
        ctx.sync()
        ctx.timer_start()
        fn()
        ts.append(ctx.timer_stop_ms())
    return float(np.median(ts))


if "c2" in which or "c3" in which:
    cells = synth.make_grid(4000, 4000, seed=7)
    m = cabi.Map2D(ctx, cells, RES)
    m.build_distances(2.0)
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=1080)

if "c2" in which:
    P = 100000
    poses = synth.uniform_free_poses(cells, RES, P, seed=1)
    aos = np.concatenate([poses, np.full((P, 1), 1.0 / P)], axis=1)
    pf = cabi.ParticleFilter(ctx, P // 10, P)
    model = cabi.planar_model(model_type=cabi.MODEL_BEAM, max_beams=1080, scanner_pose=(0.1, 0.0, 0.0))
    pf.upload_samples(aos)
    pf.stage_scan(m, model, ranges, angles, 30.0)
    ms = med(lambda: pf.update_staged(m, want_total=False))
    # Bresenham steps of the same rays (the algorithmic unit of the beam model)
    sub = poses[:2000]
    steps = 0
    for b in range(0, 1080, 54):
        _, s = m.calc_range(sub[:, 0], sub[:, 1], sub[:, 2] + angles[b], 30.0, with_steps=True)
        steps += s
    steps_per_ray = steps / (len(sub) * 20)
    out["c2_beam"] = dict(P=P, B=1080, ms=ms, rays_per_s=P * 1080 / (ms * 1e-3), steps_per_ray=steps_per_ray,
                          cell_steps_per_s=P * 1080 * steps_per_ray / (ms * 1e-3))
    print("C2", out["c2_beam"], flush=True)
    pf.close()

if "c3" in which:
    model = cabi.planar_model(model_type=cabi.MODEL_LF, max_beams=1080, scanner_pose=(0.1, 0.0, 0.0))
    rows = []
    for P in (10000, 100000, 1000000, 3000000, 10000000):
        for cloud in ("global", "tracking"):
            poses = (synth.uniform_free_poses(cells, RES, P, seed=1) if cloud == "global"
                     else synth.gaussian_poses(truth, P, seed=2))
            aos = np.concatenate([poses, np.full((P, 1), 1.0 / P)], axis=1)
            pf = cabi.ParticleFilter(ctx, P // 10, P)
            pf.stage_scan(m, model, ranges, angles, 30.0)
            pf.srand48(1)
            up, rs = [], []
            for rep in range(5):
                pf.upload_samples(aos)
                pf.init_with_current()
                pf.w_slow = pf.w_fast = 0.0
                ctx.flush_l2()
                ctx.sync()
                ctx.timer_start()
                pf.update_staged(m, want_total=True)
                up.append(ctx.timer_stop_ms())
                ctx.timer_start()
                r = pf.update_resample()
                rs.append(ctx.timer_stop_ms())
            row = dict(P=P, cloud=cloud, update_ms=float(np.median(up[1:])), resample_ms=float(np.median(rs[1:])),
                       new_count=int(r.sample_count), leaf=int(r.leaf_count), nodes=int(r.node_count),
                       clusters=int(r.cluster_count), fast_tree=int(pf.L.bamcl_version() >= 0))
            rows.append(row)
            print("C3", row, flush=True)
            pf.close()
    out["c3_sweep"] = rows

if "c4" in which:
    keys = synth.make_voxel_world(2000, 2000, 60, seed=3)
    t = time.time()
    m3 = cabi.Map3D(ctx, RES, keys, 0.36)
    ctx.timer_start()
    m3.build_distances()
    edt_ms = ctx.timer_stop_ms()
    truth3 = (1.3, -0.7, 0.4)
    pts = synth.make_cloud(keys, RES, truth3, n_points=4096)
    P = 1000000
    cells2 = synth.make_grid(2000, 2000, seed=7)
    res4 = {}
    for cloud in ("global", "tracking"):
        poses = (synth.uniform_free_poses(cells2, RES, P, seed=1) if cloud == "global"
                 else synth.gaussian_poses(truth3, P, seed=2))
        aos = np.concatenate([poses, np.full((P, 1), 1.0 / P)], axis=1)
        pf = cabi.ParticleFilter(ctx, P // 10, P)
        pf.upload_samples(aos)
        cm = cabi.cloud_model(max_beams=4096, mount_translation=(0.2, 0.0, 1.5), mount_rotation=synth.rpy_quaternion(0, 0.1, 0))
        pf.stage_cloud(m3, cm, pts)
        ms = med(lambda: pf.update_cloud_staged(m3, want_total=False), reps=5, warm=1)
        res4[cloud] = dict(ms=ms, evals_per_s=P * len(pts) / (ms * 1e-3), algo_GBps=(P * len(pts) + P * 40) / (ms * 1e-3) / 1e9)
        pf.close()
    out["c4_cloud"] = dict(P=P, N=int(len(pts)), voxels=list(m3.shape()), edt_build_ms=edt_ms, **res4)
    print("C4", out["c4_cloud"], flush=True)

print(json.dumps(out))
