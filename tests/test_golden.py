"""Golden fixtures: outputs of the UNMODIFIED reference (compiled into oracle/_ref in the build container, see
tests/golden/make_golden.py) committed as small .npz files, because /root/reference does not exist on the GPU box.

CPU tests pin the C restatement (oracle/amcl_oracle.c) to them bit for bit; GPU tests compare the CUDA path through
the C ABI: resample indices / pose bins / labels bit-exact, weights 1e-5 on the log (north_star), statistics 1e-6."""
import os

import numpy as np
import pytest

RES = 0.05
HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MODELS = {
    "lf": dict(model_type=1),
    "lf_factors": dict(model_type=1, off_map_factor=0.2, non_free_space_factor=0.3, non_free_space_radius=0.4),
    "gompertz": dict(model_type=3, gompertz=(1.0, 2.0, 6.0, -0.2, 1.3, 0.05)),
    "prob": dict(model_type=2),
    "beam": dict(model_type=0),
}


@pytest.fixture(scope="module")
def g2d():
    return np.load(os.path.join(HERE, "planar_2d.npz"))


def _ranges(g, name):
    r = g["ranges"].copy()
    if name != "beam":
        r[3] = float("nan")
    return r


# ---------------------------------------------------------------- CPU: restatement == reference
def test_port_brushfire_equals_golden(g2d, oracle_built):
    pm = oracle_built.PortMap2D(g2d["cells"], RES)
    assert np.array_equal(pm.update_distances(2.0), g2d["brushfire"])


@pytest.mark.parametrize("name", list(MODELS))
def test_port_models_equal_golden(g2d, oracle_built, name):
    O = oracle_built
    pm = O.PortMap2D(g2d["cells"], RES)
    pm.set_distances(g2d["brushfire"], 2.0)
    model = O.planar_model(max_beams=60, scanner_pose=(0.1, 0.0, 0.0), **MODELS[name])
    total, w = pm.apply(model, _ranges(g2d, name), g2d["angles"], 30.0, g2d["samples"])
    assert total == float(g2d["total_" + name])
    assert np.array_equal(w, g2d["w_" + name])


def test_port_calc_range_equals_golden(g2d, oracle_built):
    pm = oracle_built.PortMap2D(g2d["cells"], RES)
    assert np.array_equal(pm.calc_ranges(g2d["ray_ox"], g2d["ray_oy"], g2d["ray_oa"], 30.0), g2d["ray_range"])


def test_port_update_resample_sequence_equals_golden(g2d, oracle_built):
    O = oracle_built
    pm = O.PortMap2D(g2d["cells"], RES)
    pm.set_distances(g2d["brushfire"], 2.0)
    n = len(g2d["samples"])
    pf = O.PortPF(150, n)
    pf.init_with_poses(g2d["samples"][:, :3])
    pf.srand48(4242)
    assert (pf.leaf_count(), pf.node_count(), pf.cluster_count()) == (int(g2d["init_leaf"]), int(g2d["init_nodes"]),
                                                                      int(g2d["init_clusters"]))
    assert np.array_equal(pf.sample_clusters(), g2d["init_labels"])
    model = O.planar_model(model_type=1, max_beams=60, scanner_pose=(0.1, 0.0, 0.0))
    for step in range(2):
        for _ in range(2):
            pf.update_sensor(pm, model, g2d["ranges"], g2d["angles"], 30.0)
        assert np.array_equal(pf.samples()[:, 3], g2d["seq%d_weights" % step])
        assert np.array_equal(np.array(pf.averages()), g2d["seq%d_averages" % step])
        pf.update_resample()
        assert pf.sample_count() == int(g2d["seq%d_count" % step])
        assert np.array_equal(pf.samples()[:, :3], g2d["seq%d_poses" % step])
        assert (pf.leaf_count(), pf.node_count(), pf.cluster_count()) == (
            int(g2d["seq%d_leaf" % step]), int(g2d["seq%d_nodes" % step]), int(g2d["seq%d_clusters" % step]))
        assert np.array_equal(pf.sample_clusters(), g2d["seq%d_labels" % step])
        assert np.array_equal(pf.set_stats(), g2d["seq%d_stats" % step], equal_nan=True)
        assert int(pf.is_converged()) == int(g2d["seq%d_converged" % step])


def test_port_kdtree_equals_golden(oracle_built):
    g = np.load(os.path.join(HERE, "kdtree.npz"))
    t = oracle_built.PortKDTree()
    run = t.insert_many(g["poses"], 1.0, leaf_counts=True)
    assert np.array_equal(np.asarray(run), g["leaf_running"])
    assert (t.leaf_count(), t.node_count()) == (int(g["leaf"]), int(g["nodes"]))
    t.cluster()
    assert np.array_equal(np.array([t.get_cluster(p) for p in g["poses"]], np.int32), g["labels"])


def test_port_octomap_cache_equals_golden(oracle_built):
    g = np.load(os.path.join(HERE, "octomap_3d.npz"))
    m = oracle_built.PortMap3D(RES, g["keys"], 0.36)
    m.update_distances()
    mn, mx = m.bounds()
    assert np.array_equal(np.array(mn), g["mn"]) and np.array_equal(np.array(mx), g["mx"])
    assert np.array_equal(m.dense(), g["field"])


# ---------------------------------------------------------------- GPU: CUDA path vs the reference's outputs
@pytest.fixture(scope="module")
def dev(g2d):
    from badger_amcl_b200 import cabi
    ctx = cabi.Context(0)
    m = cabi.Map2D(ctx, g2d["cells"], RES)
    m.upload_distances(g2d["brushfire"], 2.0)  # the reference's own cache (the device builds the exact EDT)
    yield dict(cabi=cabi, ctx=ctx, m=m)
    m.close()
    ctx.close()


def _assert_logw(got, want):
    pos = want > 0
    assert np.array_equal(got > 0, pos)
    lg, lw = np.log(got[pos]), np.log(want[pos])
    assert (np.abs(lg - lw) <= 1e-5 * np.abs(lw) + 1e-300).all()  # north_star: 1e-5 relative on the log-weights
    assert np.allclose(got, want, rtol=2e-5, atol=0.0)


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(MODELS))
def test_gpu_models_vs_golden(g2d, dev, name):
    cabi = dev["cabi"]
    n = len(g2d["samples"])
    pf = cabi.ParticleFilter(dev["ctx"], n // 10, n)
    pf.upload_samples(g2d["samples"])
    model = cabi.planar_model(max_beams=60, scanner_pose=(0.1, 0.0, 0.0), **MODELS[name])
    total = pf.score_planar(dev["m"], model, _ranges(g2d, name), g2d["angles"], 30.0)
    _assert_logw(pf.weights(), g2d["w_" + name])
    assert abs(total - float(g2d["total_" + name])) <= 2e-5 * abs(float(g2d["total_" + name]))
    pf.close()


@pytest.mark.gpu
def test_gpu_calc_range_vs_golden(g2d, dev):
    got = dev["m"].calc_range(g2d["ray_ox"], g2d["ray_oy"], g2d["ray_oa"], 30.0)
    assert np.array_equal(got, g2d["ray_range"])


@pytest.mark.gpu
def test_gpu_exact_edt_bounds_the_reference_brushfire(g2d, dev):
    """the reference's heap-order brushfire is >= the exact capped EDT, equal in most cells"""
    cabi = dev["cabi"]
    m = cabi.Map2D(dev["ctx"], g2d["cells"], RES)
    m.build_distances(2.0)
    exact = m.download_distances()
    assert (exact <= g2d["brushfire"]).all()
    assert (exact == g2d["brushfire"]).mean() > 0.9
    m.close()


@pytest.mark.gpu
def test_gpu_resample_sequence_vs_golden(g2d, dev):
    """staged parity: the reference's normalised weights in, bit-identical resampled set, pose bins and
    labels out; statistics to 1e-6"""
    cabi = dev["cabi"]
    n = len(g2d["samples"])
    pf = cabi.ParticleFilter(dev["ctx"], 150, n)
    r0 = pf.init_with_poses(g2d["samples"][:, :3])
    pf.srand48(4242)
    assert (r0.leaf_count, r0.node_count, r0.cluster_count) == (int(g2d["init_leaf"]), int(g2d["init_nodes"]),
                                                                int(g2d["init_clusters"]))
    assert np.array_equal(pf.sample_clusters(), g2d["init_labels"])
    assert np.allclose(pf.set_stats(), g2d["init_stats"], rtol=1e-6, atol=1e-9)
    model = cabi.planar_model(model_type=1, max_beams=60, scanner_pose=(0.1, 0.0, 0.0))
    for step in range(2):
        for _ in range(2):
            pf.update_sensor(dev["m"], model, g2d["ranges"], g2d["angles"], 30.0)
        _assert_logw(pf.weights(), g2d["seq%d_weights" % step])
        pf.upload_weights(g2d["seq%d_weights" % step])
        pf.w_slow, pf.w_fast = [float(v) for v in g2d["seq%d_averages" % step]]
        r = pf.update_resample()
        assert r.sample_count == int(g2d["seq%d_count" % step])
        assert np.array_equal(pf.samples()[:, :3], g2d["seq%d_poses" % step])
        assert (r.leaf_count, r.node_count, r.cluster_count) == (
            int(g2d["seq%d_leaf" % step]), int(g2d["seq%d_nodes" % step]), int(g2d["seq%d_clusters" % step]))
        assert np.array_equal(pf.sample_clusters(), g2d["seq%d_labels" % step])
        assert np.allclose(pf.set_stats(), g2d["seq%d_stats" % step], rtol=1e-6, atol=1e-9)
        gs, ws = pf.cluster_stats(), g2d["seq%d_cluster_stats" % step]
        assert np.array_equal(gs[:, 0], ws[:, 0]) and np.allclose(gs[:, 1:], ws[:, 1:], rtol=1e-6, atol=1e-9)
        assert int(r.converged) == int(g2d["seq%d_converged" % step])
    pf.close()


@pytest.mark.gpu
def test_gpu_kdtree_vs_golden(dev):
    g = np.load(os.path.join(HERE, "kdtree.npz"))
    cabi = dev["cabi"]
    n = len(g["poses"])
    pf = cabi.ParticleFilter(dev["ctx"], 1, n)
    r = pf.init_with_poses(g["poses"])
    assert (r.leaf_count, r.node_count) == (int(g["leaf"]), int(g["nodes"]))
    assert np.array_equal(pf.sample_clusters(), g["labels"])
    pf.close()


@pytest.mark.gpu
def test_gpu_octomap_cache_vs_golden(dev):
    """octomap.cpp:174-333: the device's exact quantised EDT is <= the reference's FIFO brushfire, equal almost everywhere"""
    g = np.load(os.path.join(HERE, "octomap_3d.npz"))
    cabi = dev["cabi"]
    m = cabi.Map3D(dev["ctx"], RES, g["keys"], 0.36)
    mn, mx = m.bounds()
    assert np.array_equal(np.array(mn), g["mn"]) and np.array_equal(np.array(mx), g["mx"])
    m.build_distances()
    got = m.download_distances()
    assert got.shape == g["field"].shape
    assert (got <= g["field"]).all() and (got == g["field"]).mean() > 0.99
    m.upload_distances(g["field"])
    assert np.array_equal(m.download_distances(), g["field"])
    m.close()


def test_port_cloud_scorer_equals_golden(oracle_built):
    """the restatement against outputs of the UNMODIFIED sensors/point_cloud_scanner.cpp (tests/golden/make_golden.py)"""
    O = oracle_built
    g3, gc = np.load(os.path.join(HERE, "octomap_3d.npz")), np.load(os.path.join(HERE, "cloud_3d.npz"))
    m = O.PortMap3D(RES, g3["keys"], 0.36)
    m.set_dense(g3["field"])
    from badger_amcl_b200 import synth
    for name, mt in (("plain", 0), ("gompertz", 1)):
        cm = O.cloud_model(model_type=mt, max_beams=384, off_map_factor=0.3, gompertz=(1.0, 2.0, 6.0, -0.2, 1.3, 0.05),
                           mount_translation=(0.2, 0.0, 1.5), mount_rotation=synth.rpy_quaternion(0.0, 0.1, 0.0))
        total, w = m.cloud_apply(cm, gc["points"], gc["samples"])
        assert total == float(gc["total_" + name])
        assert np.array_equal(w, gc["w_" + name])


@pytest.mark.gpu
@pytest.mark.parametrize("name,mt", [("plain", 0), ("gompertz", 1)])
def test_gpu_cloud_scorer_vs_golden(dev, name, mt):
    """point_cloud_scanner.cpp:111-229 on the device against the compiled reference's weights (log-weights 1e-5)"""
    from badger_amcl_b200 import synth
    cabi = dev["cabi"]
    g3, gc = np.load(os.path.join(HERE, "octomap_3d.npz")), np.load(os.path.join(HERE, "cloud_3d.npz"))
    m = cabi.Map3D(dev["ctx"], RES, g3["keys"], 0.36)
    m.upload_distances(g3["field"])
    cm = cabi.cloud_model(model_type=mt, max_beams=384, off_map_factor=0.3, gompertz=(1.0, 2.0, 6.0, -0.2, 1.3, 0.05),
                          mount_translation=(0.2, 0.0, 1.5), mount_rotation=synth.rpy_quaternion(0.0, 0.1, 0.0))
    s = gc["samples"]
    pf = cabi.ParticleFilter(dev["ctx"], len(s) // 10, len(s))
    pf.upload_samples(s)
    total = pf.score_cloud(m, cm, gc["points"])
    got, want = pf.weights(), gc["w_" + name]
    lg, lw = np.log(got), np.log(want)
    assert (np.abs(lg - lw) <= 1e-5 * np.abs(lw)).all()
    assert (np.abs(got - want) <= 1e-9 * want).mean() > 0.99
    assert abs(total - float(gc["total_" + name])) <= 1e-6 * abs(float(gc["total_" + name]))
    pf.close()
    m.close()
