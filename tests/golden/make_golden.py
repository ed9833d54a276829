"""Generate tests/golden/*.npz from the UNMODIFIED reference compiled into oracle/_ref (run in the container that has
/root/reference; the GPU box does not). Each file holds seeded inputs and the reference's outputs for one slice of
the hot path, small enough to commit. Usage: python tests/golden/make_golden.py [node]  ("node": only the two runs
through the reference's ROS node, node_2d.npz / node_3d.npz)"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from badger_amcl_b200 import synth  # noqa: E402
from oracle import oracle as O  # noqa: E402

RES = 0.05
MODELS = {
    "lf": dict(model_type=1),
    "lf_factors": dict(model_type=1, off_map_factor=0.2, non_free_space_factor=0.3, non_free_space_radius=0.4),
    "gompertz": dict(model_type=3, gompertz=(1.0, 2.0, 6.0, -0.2, 1.3, 0.05)),
    "prob": dict(model_type=2),
    "beam": dict(model_type=0),
}


MOTION_ALPHA = [0.2, 0.3, 0.25, 0.15, 0.1]
MOTION_POSE = [1.0, 2.0, 0.7]
MOTION_DELTAS = [[0.3, 0.1, 0.2], [0.001, 0.0, 1.5], [-0.4, 0.2, -3.0]]
MOTION_ABS = [0.35, 0.05, 0.25]


def world():
    cells = synth.make_grid(240, 200, seed=7, wall_every=60)
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    return cells, truth


def node_runs():
    """one localisation run through the reference's own node (node/node.cpp + node_2d.cpp / node_3d.cpp compiled against
    oracle/shim), recorded step by step: tests/node_chain.py"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import node_chain
    np.savez_compressed(os.path.join(HERE, "node_2d.npz"), **node_chain.record_2d(O))
    np.savez_compressed(os.path.join(HERE, "node_3d.npz"), **node_chain.record_3d(O))
    # the node's default resampler + recovery, a kidnapped robot: random poses injected by the node's own pose function
    g = node_chain.record_2d(O, seed=44, params=node_chain.PARAMS_2D_MN, kidnap_from=3)
    g.pop("brushfire")  # the same map as node_2d.npz
    np.savez_compressed(os.path.join(HERE, "node_2d_mn.npz"), **g)


def main():
    O.build(ref=True)
    assert O.have_ref(), "the compiled reference (oracle/_ref) is required"
    node_runs()
    if sys.argv[1:] == ["node"]:
        return
    cells, truth = world()
    rm = O.RefMap2D(cells, RES)
    brushfire = rm.update_distances(2.0)  # occupancy_map.cpp:138-242, the reference's own cache
    n, n_beams = 1500, 60
    glob = synth.uniform_free_poses(cells, RES, n // 2, seed=1)
    track = synth.gaussian_poses(truth, n - n // 2, seed=2)
    poses = np.concatenate([glob, track])
    poses[0, :2] = (50.0, 3.0)  # off the map
    w0 = np.random.default_rng(1).uniform(0.5, 1.5, n) / n
    samples = np.concatenate([poses, w0[:, None]], axis=1)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=n_beams, dropouts=0.03)
    out = dict(cells=cells, brushfire=brushfire, samples=samples, ranges=ranges, angles=angles, truth=np.array(truth))
    # applyModelToSampleSet for every planar model (planar_scanner.cpp:141-164)
    for name, kw in MODELS.items():
        r = ranges.copy()
        if name != "beam":
            r[3] = float("nan")
        model = O.planar_model(max_beams=n_beams, scanner_pose=(0.1, 0.0, 0.0), **kw)
        total, w = O.RefScanner2D(rm, model).apply(r, angles, 30.0, samples)
        out["total_" + name], out["w_" + name] = total, w
    # calcRange (occupancy_map.cpp:257-364)
    rng = np.random.default_rng(9)
    ox, oy, oa = rng.uniform(-6.5, 6.5, 500), rng.uniform(-5.5, 5.5, 500), rng.uniform(-7, 7, 500)
    out.update(ray_ox=ox, ray_oy=oy, ray_oa=oa, ray_range=rm.calc_ranges(ox, oy, oa, 30.0))
    # (updateSensor, updateSensor, updateResample) x 2, systematic (particle_filter.cpp:223-354, 423-471)
    pf = O.RefPF(150, n)
    pf.init_with_poses(poses)
    pf.srand48(4242)
    model = O.planar_model(model_type=1, max_beams=n_beams, scanner_pose=(0.1, 0.0, 0.0))
    scanner = O.RefScanner2D(rm, model)
    out.update(init_leaf=pf.leaf_count(), init_nodes=pf.node_count(), init_clusters=pf.cluster_count(),
               init_labels=pf.sample_clusters(), init_stats=pf.set_stats())
    for step in range(2):
        for _ in range(2):
            pf.update_sensor(scanner, ranges, angles, 30.0)
        out["seq%d_weights" % step] = pf.samples()[:, 3]
        out["seq%d_averages" % step] = np.array(pf.averages())
        pf.update_resample()
        s = pf.samples()
        out["seq%d_poses" % step] = s[:, :3]
        out["seq%d_count" % step] = pf.sample_count()
        out["seq%d_leaf" % step] = pf.leaf_count()
        out["seq%d_nodes" % step] = pf.node_count()
        out["seq%d_clusters" % step] = pf.cluster_count()
        out["seq%d_labels" % step] = pf.sample_clusters()
        out["seq%d_stats" % step] = pf.set_stats()
        out["seq%d_cluster_stats" % step] = pf.cluster_stats()
        out["seq%d_converged" % step] = int(pf.is_converged())
    np.savez_compressed(os.path.join(HERE, "planar_2d.npz"), **out)

    # pose-bin tree: running leaf count after every insertion + labels (pf_kdtree.cpp:49-194)
    t = O.RefKDTree()
    kposes = np.concatenate([synth.uniform_free_poses(cells, RES, 1200, seed=3), synth.gaussian_poses(truth, 800, seed=4)])
    leaf_run = t.insert_many(kposes, 1.0, leaf_counts=True)
    t.cluster()
    labels = np.array([t.get_cluster(p) for p in kposes], np.int32)
    np.savez_compressed(os.path.join(HERE, "kdtree.npz"), poses=kposes, leaf_running=np.asarray(leaf_run),
                        leaf=t.leaf_count(), nodes=t.node_count(), labels=labels)

    # 3-D cache build (octomap.cpp:174-333) on a small synthetic tree
    rng = np.random.default_rng(3)
    keys = np.unique(np.concatenate([
        np.stack([rng.integers(-30, 30, 400), rng.integers(-25, 25, 400), np.zeros(400, int)], 1),   # floor
        np.stack([np.full(200, 10), rng.integers(-25, 25, 200), rng.integers(0, 12, 200)], 1),       # wall
        np.stack([rng.integers(-30, 30, 60), rng.integers(-25, 25, 60), rng.integers(0, 12, 60)], 1)]), axis=0).astype(np.int32)
    m3 = O.RefMap3D(RES, keys, 0.36)
    m3.update_distances()
    mn, mx = m3.bounds()
    np.savez_compressed(os.path.join(HERE, "octomap_3d.npz"), keys=keys, field=m3.dense(), mn=np.array(mn), mx=np.array(mx))
    # 3-D scorer: the UNMODIFIED sensors/point_cloud_scanner.cpp (applyModelToSampleSet :111-129, calcPointCloudModel /
    # Gompertz :132-203, recalcWeight :205-229) over that cache; its tf2 / pcl calls resolve to oracle/shim/
    cpts = synth.make_cloud(keys, RES, (0.3, -0.2, 0.4), n_points=384)
    rng = np.random.default_rng(5)
    cs = np.zeros((400, 4))
    cs[:, 0], cs[:, 1], cs[:, 2] = rng.uniform(-1.8, 1.8, 400), rng.uniform(-1.5, 1.5, 400), rng.uniform(-3.2, 3.2, 400)
    cs[:150, :3] = np.array((0.3, -0.2, 0.4)) + rng.normal(0.0, 0.1, (150, 3))
    cs[-5:, 0] += 40.0  # off the map: the off-map factor of recalcWeight
    cs[:, 3] = 1.0 / len(cs)
    cout = dict(points=cpts, samples=cs)
    for name, mt in (("plain", 0), ("gompertz", 1)):
        cm = O.cloud_model(model_type=mt, max_beams=384, off_map_factor=0.3, gompertz=(1.0, 2.0, 6.0, -0.2, 1.3, 0.05),
                           mount_translation=(0.2, 0.0, 1.5), mount_rotation=synth.rpy_quaternion(0.0, 0.1, 0.0))
        total, w = m3.cloud_apply(cm, cpts, cs)
        cout["total_" + name], cout["w_" + name] = total, w
    np.savez_compressed(os.path.join(HERE, "cloud_3d.npz"), **cout)
    # motion update (odom.cpp:74-311): five models x three odometry deltas on one small cloud, drand48 seeded per case;
    # `next` = the drand48() value after the update (pins the number of draws consumed)
    mposes = synth.gaussian_poses((1.5, -0.7, 0.4), 400, seed=9)
    out = dict(poses=mposes, alpha=np.array(MOTION_ALPHA), pose=np.array(MOTION_POSE), deltas=np.array(MOTION_DELTAS),
               absolute_motion=np.array(MOTION_ABS))
    for mt in range(5):
        for di, delta in enumerate(MOTION_DELTAS):
            pf = O.RefPF(40, len(mposes))
            pf.init_with_poses(mposes)
            pf.srand48(100 + 10 * mt + di)
            pf.update_action(mt, MOTION_ALPHA, MOTION_POSE, delta, MOTION_ABS)
            out["m%d_d%d" % (mt, di)] = pf.samples()[:, :3]
            out["m%d_d%d_next" % (mt, di)] = np.float64(O.ref_lib().ref_drand48())
    np.savez_compressed(os.path.join(HERE, "motion.npz"), **out)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
