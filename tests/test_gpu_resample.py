"""GPU parity: the KLD resampler (K-d1..d4) vs the oracle.

Staged as SURVEY H1b describes: the device resampler is fed the ORACLE's normalised weights, kd-tree state and
drand48 draw and must return bit-identical CDF entries, source indices, sample_count, leaf count, node count and
cluster labels; means / covariances within 1e-6 (BASELINE.json north_star)."""
import numpy as np
import pytest

from badger_amcl_b200 import synth

pytestmark = pytest.mark.gpu
RES = 0.05
STAT_RTOL = 1e-6


@pytest.fixture(scope="module")
def env(oracle_built):
    from badger_amcl_b200 import cabi
    ctx = cabi.Context(0)
    cells = synth.make_grid(2000, 2000, seed=7)
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    yield dict(O=oracle_built, cabi=cabi, ctx=ctx, cells=cells, truth=truth)
    ctx.close()


def _weights(kind, n, rng):
    if kind == "uniform":
        return np.full(n, 1.0 / n)
    if kind == "random":
        w = rng.random(n)
    elif kind == "heavy":
        w = np.exp(rng.normal(0.0, 6.0, n))
    elif kind == "zeros":
        w = rng.random(n)
        w[: n // 3] = 0.0
        w[rng.random(n) < 0.3] = 0.0
        if not w.any():
            w[-1] = 1.0
    elif kind == "dominant":
        w = rng.random(n) * 1e-12
        w[n // 2] = 1.0
    elif kind == "ties":
        # dyadic weights whose low bits sit exactly half way between neighbours of the running sum
        w = (rng.integers(1, 8, n) * 2.0 + 1.0) * 2.0 ** -60 + rng.integers(0, 3, n) * 2.0 ** -20
        return w  # deliberately not normalised: exact dyadic arithmetic
    elif kind == "increasing":
        w = np.exp(np.linspace(-40.0, 0.0, n))
    else:
        raise ValueError(kind)
    return w / w.sum()


@pytest.mark.parametrize("kind", ["uniform", "random", "heavy", "zeros", "dominant", "ties", "increasing"])
@pytest.mark.parametrize("n", [1, 5, 2047, 2048, 2049, 5000, 100000, 1000000])
def test_cdf_is_the_sequential_fp64_prefix_sum(env, kind, n):
    """particle_filter.cpp:280-285: c[i+1] = c[i] + w[i], one rounding at a time"""
    O, cabi = env["O"], env["cabi"]
    rng = np.random.default_rng(n * 7 + len(kind))
    w = _weights(kind, n, rng)
    aos = np.zeros((n, 4))
    aos[:, 3] = w
    port = O.PortPF(1, n)
    port.set_samples(aos)
    want = port.cdf()
    pf = cabi.ParticleFilter(env["ctx"], 1, n)
    pf.upload_samples(aos)
    got = pf.cdf()
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, "first mismatch at %d: %r vs %r (%d total)" % (bad[0], got[bad[0]], want[bad[0]], bad.size)
    pf.close()


def _assert_same_resample(pf, port, r, check_labels=True):
    assert r.sample_count == port.sample_count()
    assert np.array_equal(pf.resample_sources(), port.resample_sources())
    got, want = pf.samples(), port.samples()
    assert np.array_equal(got[:, :3], want[:, :3])
    assert np.array_equal(got[:, 3], want[:, 3])
    assert r.leaf_count == port.leaf_count()
    assert r.node_count == port.node_count()
    assert r.cluster_count == port.cluster_count()
    assert bool(r.converged) == port.is_converged()
    if check_labels:
        assert np.array_equal(pf.sample_clusters(), port.sample_clusters())
    assert np.allclose(pf.set_stats(), port.set_stats(), rtol=STAT_RTOL, atol=1e-9)
    gs, ws = pf.cluster_stats(), port.cluster_stats()
    assert gs.shape == ws.shape
    assert np.array_equal(gs[:, 0], ws[:, 0])
    assert np.allclose(gs[:, 1:], ws[:, 1:], rtol=STAT_RTOL, atol=1e-9)


@pytest.mark.parametrize("cloud", ["global", "tracking"])
def test_update_resample_sequence_matches_oracle(env, cloud):
    """C1: 5k particles x 60 beams, (updateSensor, updateSensor, updateResample) x 3 with the oracle's weights"""
    O, cabi = env["O"], env["cabi"]
    n = 5000
    cells, truth = env["cells"], env["truth"]
    pm = O.PortMap2D(cells, RES)
    pm.set_distances(pm.exact_edt(2.0), 2.0)
    poses = (synth.uniform_free_poses(cells, RES, n, seed=1) if cloud == "global"
             else synth.gaussian_poses(truth, n, seed=2))
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=60)
    model = O.planar_model(model_type=1, max_beams=60, scanner_pose=(0.1, 0.0, 0.0))
    port = O.PortPF(100, n)
    port.init_with_poses(poses)
    port.srand48(4242)
    pf = cabi.ParticleFilter(env["ctx"], 100, n)
    r0 = pf.init_with_poses(poses)
    pf.srand48(4242)
    assert r0.leaf_count == port.leaf_count() and r0.node_count == port.node_count()
    assert r0.cluster_count == port.cluster_count()
    assert np.array_equal(pf.sample_clusters(), port.sample_clusters())
    assert np.allclose(pf.set_stats(), port.set_stats(), rtol=STAT_RTOL, atol=1e-9)
    for _ in range(3):
        for _ in range(2):
            port.update_sensor(pm, model, ranges, angles, 30.0)
        pf.upload_weights(port.samples()[:, 3])
        pf.w_slow, pf.w_fast = port.averages()
        port.update_resample()
        r = pf.update_resample()
        _assert_same_resample(pf, port, r)
    pf.close()


def test_resample_with_random_pose_injection(env):
    """w_diff > 0 (particle_filter.cpp:297-306, 317-327): random poses first, averages reset"""
    O, cabi = env["O"], env["cabi"]
    n = 4000
    cells, truth = env["cells"], env["truth"]
    poses = synth.gaussian_poses(truth, n, seed=3)
    rnd = synth.uniform_free_poses(cells, RES, n, seed=9)
    rng = np.random.default_rng(5)
    w = rng.random(n)
    w /= w.sum()
    port = O.PortPF(500, n)
    port.init_with_poses(poses)
    aos = np.concatenate([poses, w[:, None]], axis=1)
    port.set_samples(aos)
    port.set_averages(0.010, 0.007)
    port.set_random_poses(rnd)
    port.srand48(99)
    pf = cabi.ParticleFilter(env["ctx"], 500, n)
    pf.init_with_poses(poses)
    pf.upload_samples(aos)
    pf.w_slow, pf.w_fast = 0.010, 0.007
    pf.set_random_poses(rnd)
    pf.srand48(99)
    port.update_resample()
    r = pf.update_resample()
    assert r.random_poses_used == port.L.orc_pf_random_poses_used(port.h) > 0
    _assert_same_resample(pf, port, r)
    assert (pf.w_slow, pf.w_fast) == port.averages() == (0.0, 0.0)
    pf.close()


@pytest.mark.parametrize("n,cloud", [(100000, "global"), (100000, "tracking"), (1000000, "global")])
def test_large_resample_matches_oracle(env, n, cloud):
    """C3 sizes: synthetic heavy-tailed weights on global / tracking clouds, two consecutive resamples"""
    O, cabi = env["O"], env["cabi"]
    cells, truth = env["cells"], env["truth"]
    poses = (synth.uniform_free_poses(cells, RES, n, seed=1) if cloud == "global"
             else synth.gaussian_poses(truth, n, seed=2))
    rng = np.random.default_rng(n)
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    port.srand48(7)
    pf = cabi.ParticleFilter(env["ctx"], n // 10, n)
    r0 = pf.init_with_poses(poses)
    pf.srand48(7)
    assert r0.leaf_count == port.leaf_count() and r0.node_count == port.node_count()
    for _ in range(2):
        m = port.sample_count()
        w = np.exp(rng.normal(0.0, 2.0, m))
        w /= w.sum()
        s = port.samples()
        s[:, 3] = w
        port.set_samples(s)
        port.set_averages(1.0, 1.0)
        pf.upload_weights(w)
        pf.w_slow = pf.w_fast = 1.0
        port.update_resample()
        r = pf.update_resample()
        _assert_same_resample(pf, port, r, check_labels=(n <= 100000))
    pf.close()


@pytest.mark.parametrize("order", ["random", "sorted_x"])
@pytest.mark.parametrize("n", [1, 2, 300, 2048, 2049, 5000, 60000])
def test_kdtree_leaf_count_paths(env, n, order):
    """pf_kdtree.cpp:97-150: leaf / node counts and cluster labels of initWithPoseFn's tree. 'sorted_x' makes the
    tree a chain so one bucket of the bucketed build overflows and the general grid kernel takes over."""
    O, cabi = env["O"], env["cabi"]
    cells = env["cells"]
    poses = synth.uniform_free_poses(cells, RES, n, seed=11)
    if order == "sorted_x":
        if n > 20000:
            pytest.skip("a 60k-deep chain is minutes of level steps")
        poses = poses[np.argsort(poses[:, 0], kind="stable")]
    port = O.PortPF(1, n)
    port.init_with_poses(poses)
    pf = cabi.ParticleFilter(env["ctx"], 1, n)
    r = pf.init_with_poses(poses)
    assert (r.leaf_count, r.node_count, r.cluster_count) == (port.leaf_count(), port.node_count(), port.cluster_count())
    assert np.array_equal(pf.sample_clusters(), port.sample_clusters())
    pf.close()


@pytest.mark.parametrize("n", [300, 5000, 150000])
def test_kdtree_hashed_bins_path(env, n, monkeypatch):
    """a world too large for the dense bin grid falls back to hashed bins (BAMCL_TREE_HASH forces that path here)"""
    monkeypatch.setenv("BAMCL_TREE_HASH", "1")
    O, cabi = env["O"], env["cabi"]
    poses = synth.uniform_free_poses(env["cells"], RES, n, seed=13)
    port = O.PortPF(1, n)
    port.init_with_poses(poses)
    pf = cabi.ParticleFilter(env["ctx"], 1, n)
    r = pf.init_with_poses(poses)
    assert (r.leaf_count, r.node_count, r.cluster_count) == (port.leaf_count(), port.node_count(), port.cluster_count())
    assert np.array_equal(pf.sample_clusters(), port.sample_clusters())
    pf.close()


def test_kdtree_leaf_count_large_top_tree(env):
    """sets beyond 1.5M samples (a sharded filter's global set) use the 8192-key top tree of tree_fast.cuh"""
    O, cabi = env["O"], env["cabi"]
    n = 2200000
    poses = synth.uniform_free_poses(env["cells"], RES, n, seed=21)
    port = O.PortPF(1, n)
    port.init_with_poses(poses)
    pf = cabi.ParticleFilter(env["ctx"], 1, n)
    r = pf.init_with_poses(poses)
    assert (r.leaf_count, r.node_count) == (port.leaf_count(), port.node_count())
    assert r.cluster_count == port.cluster_count()
    pf.close()


@pytest.mark.parametrize("n,n_min,cloud", [(3000, 100, "global"), (3000, 100, "tracking"), (20000, 500, "global"),
                                           (20000, 50, "tracking"), (1, 1, "tracking"), (2, 1, "global")])
def test_multinomial_resample_matches_oracle(env, n, n_min, cloud):
    """resampleMultinomial (particle_filter.cpp:356-420) with w_diff = 0: two draws per output, stop as soon as the
    count exceeds resampleLimit(leaf count so far); indices, count, tree and the advanced drand48 state"""
    O, cabi = env["O"], env["cabi"]
    cells, truth = env["cells"], env["truth"]
    poses = (synth.uniform_free_poses(cells, RES, n, seed=1) if cloud == "global"
             else synth.gaussian_poses(truth, n, seed=2, sigma_xy=0.3))
    rng = np.random.default_rng(n + n_min)
    port = O.PortPF(n_min, n, systematic=False)
    port.init_with_poses(poses)
    port.srand48(31)
    pf = cabi.ParticleFilter(env["ctx"], n_min, n, systematic=False)
    pf.init_with_poses(poses)
    pf.srand48(31)
    for _ in range(2):
        m = port.sample_count()
        w = np.exp(rng.normal(0.0, 1.0, m))
        w /= w.sum()
        s = port.samples()
        s[:, 3] = w
        port.set_samples(s)
        port.set_averages(1.0, 1.0)
        pf.upload_weights(w)
        pf.w_slow = pf.w_fast = 1.0
        port.update_resample()
        r = pf.update_resample()
        _assert_same_resample(pf, port, r)
    # both generators consumed the same number of draws: the next systematic start agrees
    port.L.orc_pf_set_resample_model(port.h, 1)
    pf.systematic = True
    s = port.samples()
    pf.w_slow = pf.w_fast = 1.0
    port.set_averages(1.0, 1.0)
    port.update_resample()
    r = pf.update_resample()
    assert np.array_equal(pf.resample_sources(), port.resample_sources())
    pf.close()


@pytest.mark.parametrize("n,n_min,w_fast,thr,mult", [(3000, 100, 0.7, 0.0, 0.0), (20000, 500, 0.95, 3.0, 0.5),
                                                     (20000, 50, 0.3, 0.0, 0.0), (1, 1, 0.5, 0.0, 0.0),
                                                     (200000, 20000, 0.9, 1.5, 0.5), (5000, 100, 0.0, 0.0, 0.0)])
@pytest.mark.parametrize("systematic", [False, True])
def test_resample_with_device_pose_source_matches_oracle(env, n, n_min, w_fast, thr, mult, systematic):
    """resampleMultinomial (and resampleSystematic, :291-324: start draw, then the random poses) with w_diff = 1 - w_fast / w_slow > 0 (particle_filter.cpp:385-388) and the node's pose
    function (uniformPoseGenerator over the free-space list, node.cpp:262-264) drawing from the same drand48 stream:
    sources (-1 = random pose), poses, count, tree, clusters and the stream state, bit for bit"""
    O, cabi = env["O"], env["cabi"]
    cells = synth.make_grid(300, 200, seed=7, wall_every=60)
    pm = O.PortMap2D(cells, RES, (0.5, -0.25))
    pm.update_distances(2.0)
    fs = pm.free_space(0.1)
    m = cabi.Map2D(env["ctx"], cells, RES, origin=(0.5, -0.25))
    m.upload_distances(pm.dist, 2.0)
    assert m.update_free_space(0.1) == len(fs)
    poses = synth.gaussian_poses((0.5, 0.3, 0.2), n, seed=2, sigma_xy=0.3)
    rng = np.random.default_rng(n + n_min)
    port = O.PortPF(n_min, n, systematic=systematic)
    port.init_with_poses(poses)
    port.set_uniform_pose_source(pm, fs, thr, mult)
    port.srand48(31)
    pf = cabi.ParticleFilter(env["ctx"], n_min, n, systematic=systematic)
    pf.init_with_poses(poses)
    pf.set_random_pose_source(m, thr, mult)
    pf.srand48(31)
    for _ in range(2):
        k = port.sample_count()
        w = np.exp(rng.normal(0.0, 1.0, k))
        w /= w.sum()
        s = port.samples()
        s[:, 3] = w
        port.set_samples(s)
        port.set_averages(1.0, w_fast)
        pf.upload_samples(s)
        pf.w_slow, pf.w_fast = 1.0, w_fast
        port.update_resample()
        r = pf.update_resample()
        _assert_same_resample(pf, port, r)
        assert r.random_poses_used == int((port.resample_sources() < 0).sum())
        assert pf.drand48_state == port.rng.x
        if 0.0 < w_fast < 1.0 and n > 100:
            assert r.random_poses_used > 0
    pf.close()
    m.close()


@pytest.mark.parametrize("n,n_min,w_fast,stream_poses", [(3000, 100, 0.7, True), (20000, 500, 0.95, True), (20000, 50, 0.3, False),
                                                         (1, 1, 0.5, True), (5000, 100, 0.0, False)])
def test_multinomial_with_host_drawn_branches_matches_oracle(env, n, n_min, w_fast, stream_poses):
    """bamcl_pf_resample_multinomial_drawn: the caller runs the branch loop of particle_filter.cpp:385-393 with ITS pose
    function - one that draws from the same drand48 stream (the node's generator) or one that does not (a queue) - and hands
    the draws over; sources, poses, count, tree, clusters bit for bit against the sequential loop, and the generator state
    the caller restores (recorded before the first output that was not kept) is where the sequential loop stops"""
    O, cabi = env["O"], env["cabi"]
    C = O.C
    cells = synth.make_grid(300, 200, seed=7, wall_every=60)
    pm = O.PortMap2D(cells, RES, (0.5, -0.25))
    pm.update_distances(2.0)
    fs = pm.free_space(0.1)
    poses = synth.gaussian_poses((0.5, 0.3, 0.2), n, seed=2, sigma_xy=0.3)
    queue = synth.uniform_free_poses(cells, RES, 2 * n + 8, seed=9)
    rng = np.random.default_rng(n + n_min)
    port = O.PortPF(n_min, n, systematic=False)
    port.init_with_poses(poses)
    if stream_poses:
        port.set_uniform_pose_source(pm, fs, 0.0, 0.0)
    port.srand48(31)
    pf = cabi.ParticleFilter(env["ctx"], n_min, n, systematic=False)
    pf.init_with_poses(poses)
    pf.srand48(31)
    L = O.port_lib()
    used = 0
    for _ in range(2):
        k = port.sample_count()
        w = np.exp(rng.normal(0.0, 1.0, k))
        w /= w.sum()
        s = port.samples()
        s[:, 3] = w
        port.set_samples(s)
        port.set_averages(1.0, w_fast)
        pf.upload_samples(s)
        w_diff = max(0.0, 1.0 - w_fast / 1.0)
        if not stream_poses:
            port.set_random_poses(queue[used:])
        # the caller's loop (what the host mirror does), on a copy of the generator
        g = O.Rng(pf.drand48_state)
        draws, rp, states = np.empty(n), [], np.empty(n + 1, np.uint64)
        for i in range(n):
            states[i] = g.x
            if L.orc_drand48(C.byref(g)) < w_diff:
                rp.append(pm.uniform_poses(fs, g, 1)[0] if stream_poses else queue[used + len(rp)])
                draws[i] = -float(len(rp))
            else:
                draws[i] = L.orc_drand48(C.byref(g))
        states[n] = g.x
        port.update_resample()
        r = pf.update_resample_drawn(w_diff, draws, np.array(rp).reshape(-1, 3), states)
        _assert_same_resample(pf, port, r)
        assert r.random_poses_used == int((port.resample_sources() < 0).sum())
        assert pf.drand48_state == port.rng.x
        used += r.random_poses_used
        if 0.0 < w_fast < 1.0 and n > 100:
            assert r.random_poses_used > 0
    pf.close()


def test_multinomial_with_random_injection_is_refused(env):
    cabi = env["cabi"]
    n = 1000
    poses = synth.gaussian_poses(env["truth"], n, seed=2)
    pf = cabi.ParticleFilter(env["ctx"], 100, n, systematic=False)
    pf.init_with_poses(poses)
    pf.w_slow, pf.w_fast = 0.010, 0.007
    pf.set_random_poses(poses)
    with pytest.raises(cabi.BamclError):
        pf.update_resample()
    pf.close()


def test_host_buffer_entry_points_equal_the_resident_ones(env):
    """bamcl_pf_update_planar_host / bamcl_pf_resample_host (pipelined copies) == upload + update + resample + download"""
    cabi = env["cabi"]
    n = 400000  # above the chunking threshold
    cells, truth = env["cells"], env["truth"]
    m = cabi.Map2D(env["ctx"], cells, RES)
    m.build_distances(2.0)
    poses = synth.uniform_free_poses(cells, RES, n, seed=8)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=90)
    model = cabi.planar_model(model_type=cabi.MODEL_LF, max_beams=90, scanner_pose=(0.1, 0.0, 0.0))
    aos = cabi.pinned_array((n, 4))
    aos[:, :3] = poses
    aos[:, 3] = 1.0 / n
    a = cabi.ParticleFilter(env["ctx"], n // 10, n)
    a.init_with_poses(poses)
    a.srand48(5)
    a.upload_samples(aos)
    a.update_sensor(m, model, ranges, angles, 30.0)
    wa = a.weights()
    ra = a.update_resample()
    sa = a.samples()
    b = cabi.ParticleFilter(env["ctx"], n // 10, n)
    b.init_with_poses(poses)
    b.srand48(5)
    b.update_sensor_host(m, model, ranges, angles, 30.0, aos)
    assert np.array_equal(b.weights(), wa)
    out = cabi.pinned_array((n, 4))
    rb = b.update_resample(out=out)
    assert (rb.sample_count, rb.leaf_count, rb.node_count, rb.cluster_count) == (
        ra.sample_count, ra.leaf_count, ra.node_count, ra.cluster_count)
    assert np.array_equal(out[:rb.sample_count], sa)
    assert np.array_equal(b.samples(), sa)
    a.close()
    b.close()
    m.close()
