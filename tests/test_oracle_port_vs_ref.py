"""Validate the C restatement (oracle/amcl_oracle.c) against the UNMODIFIED reference compiled into
oracle/_ref (nothing in the reference's tests pins sensor-model, resample or distance-cache
outputs, so the compiled reference itself is the anchor). Bit-exact comparisons throughout. CPU only."""
import ctypes as C
import math

import numpy as np
import pytest

from badger_amcl_b200 import synth

RES = 0.05


@pytest.fixture(scope="module")
def world(oracle_built):
    O = oracle_built
    if not O.have_ref():
        pytest.skip("compiled reference not available")
    cells = synth.make_grid(400, 300, seed=7, wall_every=100)
    pm, rm = O.PortMap2D(cells, RES), O.RefMap2D(cells, RES)
    d_ref = rm.update_distances(2.0)
    d_port = pm.update_distances(2.0)
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    return dict(O=O, cells=cells, pm=pm, rm=rm, d_ref=d_ref, d_port=d_port, truth=truth)


def test_brushfire_bit_exact(world):
    """occupancy_map.cpp:138-242 incl. libstdc++ heap tie order"""
    assert np.array_equal(world["d_ref"], world["d_port"])
    # and on a second, denser random map
    O = world["O"]
    rng = np.random.default_rng(3)
    cells = np.where(rng.random((120, 150)) < 0.03, 1, -1).astype(np.int8)
    for mx in (0.3, 1.0):
        assert np.array_equal(O.RefMap2D(cells, RES).update_distances(mx), O.PortMap2D(cells, RES).update_distances(mx))


def test_exact_edt_is_lower_bound_of_brushfire(world):
    """the brushfire is >= the exact capped EDT, equal in most cells"""
    exact = world["pm"].exact_edt(2.0)
    bf = world["d_ref"]
    assert (exact <= bf).all()
    assert (exact == bf).mean() > 0.9


def test_calc_range_random_rays(world):
    rng = np.random.default_rng(9)
    n = 4000
    ox, oy = rng.uniform(-11, 11, n), rng.uniform(-8.5, 8.5, n)  # some start off the map
    oa = rng.uniform(-7, 7, n)
    for max_range in (0.0, 3.0, 30.0):
        assert np.array_equal(world["rm"].calc_ranges(ox, oy, oa, max_range), world["pm"].calc_ranges(ox, oy, oa, max_range))


def _scan(world, n_beams, dropouts=0.0, nan=False):
    r, a = synth.make_scan(world["cells"], RES, world["truth"], n_beams=n_beams, dropouts=dropouts)
    if nan:
        r[3] = float("nan")
    return r, a


def _cloud(world, n, seed=1):
    O = world["O"]
    glob = synth.uniform_free_poses(world["cells"], RES, n // 2, seed=seed)
    track = synth.gaussian_poses(world["truth"], n - n // 2, seed=seed + 1)
    poses = np.concatenate([glob, track])
    poses[0, :2] = (500.0, 3.0)  # one particle off the map
    w = np.random.default_rng(seed).uniform(0.5, 1.5, n) / n
    return np.concatenate([poses, w[:, None]], axis=1)


MODELS = {
    "lf": dict(model_type=1),
    "lf_factors": dict(model_type=1, off_map_factor=0.2, non_free_space_factor=0.3, non_free_space_radius=0.4),
    "lf_subsample": dict(model_type=1, max_beams=30),
    "gompertz": dict(model_type=3, gompertz=(1.0, 2.0, 6.0, -0.2, 1.3, 0.05)),
    "prob": dict(model_type=2),
    "prob_beamskip": dict(model_type=2, do_beamskip=True),
    "beam": dict(model_type=0),
}


@pytest.mark.parametrize("name", list(MODELS))
def test_planar_models_bit_exact(world, name):
    O = world["O"]
    kw = dict(max_beams=60, scanner_pose=(0.1, 0.02, 0.01))
    kw.update(MODELS[name])
    model = O.planar_model(**kw)
    ranges, angles = _scan(world, 60, dropouts=0.05, nan=(name != "beam"))  # the beam model has no NaN guard
    samples = _cloud(world, 600)
    sc = O.RefScanner2D(world["rm"], model)
    st = O.port_lib().orc_planar_state_create()
    for converged in (0, 1):
        t_ref, w_ref = sc.apply(ranges, angles, 30.0, samples, converged)
        t_port, w_port = world["pm"].apply(model, ranges, angles, 30.0, samples, converged, state=st)
        assert t_ref == t_port
        assert np.array_equal(w_ref, w_port)
        assert t_ref > 0
    O.port_lib().orc_planar_state_destroy(st)


@pytest.mark.parametrize("systematic", [True, False])
@pytest.mark.parametrize("cloud", ["global", "tracking"])
def test_pf_sequence_bit_exact(world, systematic, cloud):
    """updateSensor x2 + updateResample, three rounds (resample_interval=2, node_2d.cpp:69), with a motion
    jitter between rounds. Compares every observable of ParticleFilter with the compiled reference and
    checks the port's recorded source indices by pose equality against the reference's output."""
    O = world["O"]
    n = 3000 if systematic else 800
    model = O.planar_model(model_type=1, max_beams=60, scanner_pose=(0.1, 0.0, 0.0))
    ranges, angles = _scan(world, 60)
    if cloud == "global":
        poses = synth.uniform_free_poses(world["cells"], RES, n, seed=1)
    else:
        poses = synth.gaussian_poses(world["truth"], n, seed=2)
    ref = O.RefPF(n // 10, n, systematic=systematic)
    port = O.PortPF(n // 10, n, systematic=systematic)
    sc = O.RefScanner2D(world["rm"], model)
    ref.init_with_poses(poses)
    port.init_with_poses(poses)
    ref.srand48(42)
    port.srand48(42)
    rnd = synth.uniform_free_poses(world["cells"], RES, n, seed=77)
    ref.set_random_poses(rnd)
    port.set_random_poses(rnd)
    assert ref.leaf_count() == port.leaf_count()
    jitter = np.random.default_rng(4)
    for rnd_i in range(3):
        for _ in range(2):
            assert ref.update_sensor(sc, ranges, angles, 30.0) == port.update_sensor(world["pm"], model, ranges, angles, 30.0)
            assert np.array_equal(ref.samples(), port.samples())
            assert ref.averages() == port.averages()
        if rnd_i == 1:  # force the augmented-MCL random injection path: w_fast < w_slow
            ws, wf = ref.averages()
            ref.set_averages(ws, ws * 0.8)
            port.set_averages(ws, ws * 0.8)
        before = port.samples()
        ref.update_resample()
        port.update_resample()
        a, b = ref.samples(), port.samples()
        assert a.shape == b.shape and np.array_equal(a, b)
        src = port.resample_sources()
        sel = src >= 0
        assert np.array_equal(a[sel, :3], before[src[sel], :3])
        assert (before[src[sel], 3] > 0).all()
        if rnd_i == 1:
            assert (~sel).sum() > 0 and ref.L.ref_pf_random_poses_used(ref.h) == (~sel).sum()
        assert ref.leaf_count() == port.leaf_count() and ref.node_count() == port.node_count()
        assert ref.cluster_count() == port.cluster_count()
        assert np.array_equal(ref.sample_clusters(), port.sample_clusters())
        assert np.array_equal(ref.cluster_stats(), port.cluster_stats(), equal_nan=True)
        assert np.array_equal(ref.set_stats(), port.set_stats(), equal_nan=True)
        assert ref.is_converged() == port.is_converged()
        assert ref.averages() == port.averages()
        # motion jitter so the next resample sees distinct poses (same on both sides)
        s = ref.samples()
        s[:, :3] += jitter.normal(0, [0.05, 0.05, 0.02], (len(s), 3))
        ref.set_samples(s)
        port.set_samples(s)


def test_resample_limit(world):
    O = world["O"]
    ref, port = O.RefPF(100, 5000), O.PortPF(100, 5000)
    for k in list(range(0, 300)) + [1000, 5000, 100000]:
        assert ref.resample_limit(k) == port.resample_limit(k)


def test_kdtree_random_streams(world):
    """leaf count after every insertion, node table and cluster labels (pf_kdtree.cpp)"""
    O = world["O"]
    rng = np.random.default_rng(12)
    for n, spread in ((2000, 3.0), (20000, 30.0), (5000, 0.6)):
        poses = np.stack([rng.normal(0, spread, n), rng.normal(0, spread, n), rng.uniform(-4, 4, n)], axis=1)
        rt, pt = O.RefKDTree(), O.PortKDTree()
        assert np.array_equal(rt.insert_many(poses, 0.1, leaf_counts=True), pt.insert_many(poses, 0.1, leaf_counts=True))
        rt.cluster()
        pt.cluster()
        for a, b in zip(rt.nodes(), pt.nodes()):
            assert np.array_equal(a, b)


def _voxels(seed=3, nx=60, ny=50, nz=12):
    rng = np.random.default_rng(seed)
    keys = []
    for i in range(nx):
        for j in range(ny):
            keys.append((i, j, 0))  # floor
            if i % 20 == 0 or j % 25 == 0:
                for k in range(1, nz):
                    keys.append((i, j, k))
    extra = rng.integers(0, [nx, ny, nz], (150, 3))
    keys = np.unique(np.concatenate([np.array(keys), extra]), axis=0)
    return keys[rng.permutation(len(keys))].astype(np.int32)


def test_octomap_lut_bit_exact(world):
    """octomap.cpp:53-74,174-333 (+ setMapBounds :128-150)"""
    O = world["O"]
    keys = _voxels()
    rm, pm = O.RefMap3D(0.05, keys, 0.36), O.PortMap3D(0.05, keys, 0.36)
    assert np.array_equal(rm.bounds()[0], pm.bounds()[0]) and np.array_equal(rm.bounds()[1], pm.bounds()[1])
    rm.update_distances()
    pm.update_distances()
    a, b = rm.dense(), pm.dense()
    assert a.shape == b.shape and np.array_equal(a, b)
    exact = pm.exact_edt()
    assert (exact <= a).all() and (exact == a).mean() > 0.99
    rng = np.random.default_rng(5)
    mn, mx = pm.bounds()
    for _ in range(200):
        i, j, k = (int(v) for v in rng.integers(mn - 3, mx + 4))
        assert rm.distance(i, j, k) == pm.distance(i, j, k)
    # setMapBounds runs the update itself; the reference drops its octree after one update
    # (octomap.cpp:196), so cropping needs a fresh map
    rm, pm = O.RefMap3D(0.05, keys, 0.36), O.PortMap3D(0.05, keys, 0.36)
    rm.set_bounds((0.5, 0.4), (2.0, 1.9))
    pm.set_bounds((0.5, 0.4), (2.0, 1.9))
    assert np.array_equal(rm.bounds()[0], pm.bounds()[0]) and np.array_equal(rm.bounds()[1], pm.bounds()[1])
    assert np.array_equal(rm.dense(), pm.dense())


@pytest.mark.parametrize("systematic", [False, True])
@pytest.mark.parametrize("thr,mult", [(0.0, 0.0), (3.0, 0.5)])
def test_resample_with_the_nodes_pose_function_bit_exact(world, systematic, thr, mult):
    """resampleMultinomial / resampleSystematic (particle_filter.cpp:269-420, UNMODIFIED in oracle/_ref) with w_diff > 0
    and random_pose_fn_ = Node::uniformPoseGenerator drawing from the same drand48 stream (node.cpp:262-264,
    823-868; restated in the harness and in the port): the interleaving of the resampler's and the generator's draws"""
    O = world["O"]
    cells = synth.make_grid(300, 200, seed=7, wall_every=60)
    pm = O.PortMap2D(cells, 0.05)
    pm.update_distances(2.0)
    rm = O.RefMap2D(cells, 0.05)
    rm.set_distances(pm.dist, 2.0)
    fs = pm.free_space(0.1)
    n = 3000
    poses = synth.gaussian_poses((1.0, 0.5, 0.2), n, seed=1)
    rng = np.random.default_rng(0)
    a, b = O.PortPF(n // 10, n, systematic=systematic), O.RefPF(n // 10, n, systematic=systematic)
    a.init_with_poses(poses)
    b.init_with_poses(poses)
    a.set_uniform_pose_source(pm, fs, thr, mult)
    b.set_uniform_pose_source(rm, fs, thr, mult)
    for step in range(2):
        a.srand48(5 + step)
        b.srand48(5 + step)
        m = a.sample_count()
        w = np.exp(rng.normal(0.0, 1.0, m))
        s = a.samples()
        s[:, 3] = w / w.sum()
        a.set_samples(s)
        b.set_samples(s)
        a.set_averages(1.0, 0.7)
        b.set_averages(1.0, 0.7)  # w_diff = 0.3
        a.update_resample()
        b.update_resample()
        assert a.sample_count() == b.sample_count()
        assert np.array_equal(a.samples(), b.samples())
        assert (a.resample_sources() < 0).sum() > 0
        assert O.port_lib().orc_drand48(C.byref(a.rng)) == O.ref_lib().ref_drand48()


@pytest.mark.parametrize("model_type", [0, 1])
def test_cloud_scorer_bit_exact(world, model_type):
    """sensors/point_cloud_scanner.cpp compiled UNMODIFIED (its tf2 / pcl calls resolve to oracle/shim) == the restatement,
    bit for bit: applyModelToSampleSet :111-129, calcPointCloudModel / Gompertz :132-203, recalcWeight :205-229"""
    O = world["O"]
    from badger_amcl_b200 import synth
    keys = synth.make_voxel_world(240, 200, 40, seed=3, wall_every=60)
    ref = O.RefMap3D(0.05, keys, 0.36)
    ref.update_distances()
    port = O.PortMap3D(0.05, keys, 0.36)
    port.set_dense(ref.dense())
    truth = (0.3, -0.2, 0.4)
    pts = synth.make_cloud(keys, 0.05, truth, n_points=512)
    rng = np.random.default_rng(1)
    n = 300
    s = np.zeros((n, 4))
    s[:, 0], s[:, 1], s[:, 2], s[:, 3] = rng.uniform(-6, 6, n), rng.uniform(-5, 5, n), rng.uniform(-3.2, 3.2, n), 1.0 / n
    s[:100, :3] = np.array(truth) + rng.normal(0, 0.2, (100, 3))
    s[-3:, 0] += 50.0  # off the map
    m = O.cloud_model(model_type=model_type, max_beams=512, mount_translation=(0.2, 0.0, 1.5),
                      mount_rotation=synth.rpy_quaternion(0, 0.1, 0), off_map_factor=0.7, gompertz=(1.0, 2.0, 3.0, -0.5, 1.5, 0.1))
    t_ref, w_ref = ref.cloud_apply(m, pts, s)
    t_port, w_port = port.cloud_apply(m, pts, s)
    assert t_ref == t_port
    assert np.array_equal(w_ref, w_port)
