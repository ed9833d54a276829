"""A whole localisation run as the reference's ROS node sequences it (tests/node_chain.py), recorded from the compiled node
into tests/golden/node_2d.npz / node_3d.npz and replayed here: through the restated chain on the CPU (bit-exact) and through
the CUDA library on the GPU (bit-exact where the arithmetic is integer / copy / identical, 1e-9 where device libm differs
from glibc in the last ulp)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import node_chain as nc  # noqa: E402


def golden(name):
    g = dict(np.load(os.path.join(HERE, "golden", name)))
    if name == "node_2d_mn.npz":   # the same map message as node_2d.npz: its cache is stored once
        g["brushfire"] = np.load(os.path.join(HERE, "golden", "node_2d.npz"))["brushfire"]
    return g


def exact(bad):
    def check(name, got, want):
        if not np.array_equal(np.asarray(got), np.asarray(want)):
            bad.append(name)
    return check


def test_restated_chain_2d_is_the_node_bit_for_bit(oracle_built):
    O, g = oracle_built, golden("node_2d.npz")
    b = nc.PortBackend2D(O, g)
    assert np.array_equal(b.distances(), g["brushfire"])
    assert b.geometry == (float(g["resolution"]), (float(g["origin"][0]), float(g["origin"][1])))
    assert np.array_equal(b.free, g["free"])
    bad = []
    nc.replay_2d(b, g, exact(bad))
    assert bad == []


def test_restated_chain_2d_multinomial_with_recovery_is_the_node_bit_for_bit(oracle_built):
    """the node's default resampler with augmented-MCL recovery and a kidnapped robot: updateResample injects poses from
    the node's own generator, interleaved with the resampler's draws (particle_filter.cpp:385-393)"""
    O, g = oracle_built, golden("node_2d_mn.npz")
    assert (g["s3_averages"] == 0.0).all() and int(g["s3_clusters"]) > 50        # the injection happened
    b = nc.PortBackend2D(O, g, nc.PARAMS_2D_MN)
    bad = []
    nc.replay_2d(b, g, exact(bad), nc.PARAMS_2D_MN)
    assert bad == []


def test_restated_chain_3d_is_the_node_bit_for_bit(oracle_built):
    O, g = oracle_built, golden("node_3d.npz")
    b = nc.PortBackend3D(O, g)
    assert np.array_equal(b.field(), g["field"])
    mn, mx = b.bounds()
    assert np.array_equal(mn, g["mn"]) and np.array_equal(mx, g["mx"])
    bad = []
    nc.replay_3d(b, g, exact(bad))
    assert bad == []


@pytest.mark.parametrize("which", ["2d", "3d", "2d_mn"])
def test_committed_golden_is_what_the_compiled_node_produces(oracle_built, which):
    O = oracle_built
    if not O.have_ref():
        pytest.skip("oracle/_ref not built (needs the reference checkout)")
    g = golden("node_%s.npz" % which)
    fresh = (nc.record_2d(O) if which == "2d" else nc.record_3d(O) if which == "3d" else
             nc.record_2d(O, seed=44, params=nc.PARAMS_2D_MN, kidnap_from=3))
    assert sorted(fresh) == sorted(g)
    for k in g:
        assert np.array_equal(np.asarray(fresh[k]), g[k], equal_nan=np.asarray(g[k]).dtype.kind == "f"), k


def test_the_run_exercises_motion_resampling_and_readout():
    """the fixture is only worth something if the node actually moved, resampled and published"""
    g = golden("node_2d.npz")
    assert int(g["s5_resample_count"]) == 6 and int(g["s5_pose_count"]) == 4
    assert not np.array_equal(g["s1_samples"][:, :3], g["s0_samples"][:, :3])            # updateAction moved the set
    assert abs(g["s5_pose"][0] - nc.truth_pose(5)[0]) < 0.15 and abs(g["s5_pose"][1] - nc.truth_pose(5)[1]) < 0.15
    assert np.unique(g["s5_samples"][:, :3], axis=0).shape[0] < len(g["s5_samples"])  # resampling duplicated particles


# ---------------------------------------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def dev():
    import torch
    assert torch.cuda.is_available()
    from badger_amcl_b200 import build, cabi
    build.build()
    return dict(cabi=cabi, ctx=cabi.Context(0))


def tolerant(bad, weight_rtol, same_rows=1.0):
    """same_rows: share of particles that must be the reference's particle (1.0 = every resampling index identical)"""
    def check(name, got, want):
        got, want = np.asarray(got), np.asarray(want)
        if name.endswith("_samples") and name != "init_samples":
            # poses: Odom::updateAction through device log / sincos (last ulp vs glibc); weights: scorer precision
            ok = got.shape == want.shape
            if ok:
                row = (np.abs(got[:, :3] - want[:, :3]).max(axis=1) <= 1e-9) & np.isclose(got[:, 3], want[:, 3],
                                                                                        rtol=weight_rtol, atol=1e-300)
                ok = row.mean() >= same_rows
        elif name.endswith("_max_weight") or name.endswith("_pose_xy") or name.endswith("_cov"):
            ok = got.shape == want.shape and np.allclose(got, want, rtol=1e-6, atol=1e-9)
        else:   # scan conversion, counts, labels, the generator state: exact
            ok = np.array_equal(got, want)
        if not ok:
            bad.append(name)
    return check


@pytest.mark.gpu
@pytest.mark.parametrize("exact_terms,rtol,same_rows", [(True, 1e-9, 1.0), (False, 2e-5, 0.999)])
def test_gpu_chain_2d_vs_the_node(dev, exact_terms, rtol, same_rows):
    """exact_terms: f64 terms and sums, every resampling index must be the reference's; the default fixed-point table
    (2e-5 on the weights) may move a CDF boundary past a target: at most 1 particle in 1000 may differ"""
    g = golden("node_2d.npz")
    b = nc.CudaBackend2D(dev["cabi"], dev["ctx"], g, exact_terms=exact_terms)
    assert b.map.resolution == float(g["resolution"]) and tuple(np.float64(b.map.origin)) == tuple(g["origin"])
    assert np.array_equal(b.map.free_space(), g["free"])
    bad = []
    nc.replay_2d(b, g, tolerant(bad, rtol, same_rows))
    b.close()
    assert bad == []


@pytest.mark.gpu
def test_gpu_chain_2d_with_the_device_built_cache(dev):
    """the drop-in path: cache from the device's exact EDT instead of the reference's brushfire: same sequence,
    same counts and generator state, the published pose within a few centimetres"""
    g = golden("node_2d.npz")
    b = nc.CudaBackend2D(dev["cabi"], dev["ctx"], g, reference_cache=False)
    seen = {}

    def check(name, got, want):
        seen[name] = (np.asarray(got), np.asarray(want))
    nc.replay_2d(b, g, check)
    b.close()
    for k in range(nc.N_STEPS):
        assert seen["s%d_resample_count" % k][0] == seen["s%d_resample_count" % k][1]
        assert seen["s%d_samples" % k][0].shape == seen["s%d_samples" % k][1].shape
    got, want = seen["s5_pose_xy"]
    assert np.abs(got - want).max() < 0.05


@pytest.mark.gpu
def test_gpu_chain_3d_vs_the_node(dev):
    g = golden("node_3d.npz")
    b = nc.CudaBackend3D(dev["cabi"], dev["ctx"], g)
    mn, mx = b.map.bounds()
    assert np.array_equal(mn, g["mn"]) and np.array_equal(mx, g["mx"])
    bad = []
    nc.replay_3d(b, g, tolerant(bad, 1e-6))
    b.close()
    assert bad == []


@pytest.mark.gpu
@pytest.mark.parametrize("thr,mult", [(0.0, 0.0), (5.0, 0.4)])
def test_gpu_init_uniform_3d_vs_oracle(oracle_built, dev, thr, mult):
    """global localisation with map_type 3 (node_3d.cpp:306-318 + node.cpp:823-883): poses from the OctoMap's cell rectangle"""
    import ctypes as C
    O, cabi = oracle_built, dev["cabi"]
    g = golden("node_3d.npz")
    m = cabi.Map3D(dev["ctx"], 0.1, g["keys"], 0.5)
    m.upload_distances(g["field"])
    n = 5000
    rng = O.Rng(0)
    O.port_lib().orc_srand48(C.byref(rng), 123)
    poses = O.uniform_poses_rect(g["mn"], g["mx"], 0.1, rng, n, thr, mult)
    ref = O.PortPF(100, n)
    ref.init_with_poses(poses)
    pf = cabi.ParticleFilter(dev["ctx"], 100, n)
    pf.srand48(123)
    res = pf.init_uniform(m, thr, mult)
    assert pf.drand48_state == rng.x
    assert np.array_equal(pf.samples(), ref.samples())
    assert (res.leaf_count, res.cluster_count) == (ref.leaf_count(), ref.cluster_count())
    assert np.array_equal(pf.sample_clusters(), ref.sample_clusters())
    pf.close()
    m.close()


@pytest.mark.gpu
@pytest.mark.parametrize("which,seed,n,cd", [("2d", 21, 1500, None), ("3d", 33, 800, None)])
def test_gpu_init_gaussian_reproduces_the_nodes_first_particle_set(dev, which, seed, n, cd):
    """Node::initFromNewMap -> ParticleFilter::initWithGaussian (node.cpp:670-701, particle_filter.cpp:105-133) with the
    node's default initial pose and covariance: the set the compiled node drew from srand48(seed) is in the golden run.
    The node draws it TWICE (initFromNewMap calls initWithGaussian, then setInitialPose -> applyInitialPose calls it again,
    node.cpp:688-708, 980-1002): the golden set is the second one"""
    cabi = dev["cabi"]
    g = golden("node_%s.npz" % which)
    pf = cabi.ParticleFilter(dev["ctx"], 100, n)
    pf.srand48(seed)
    sig = np.sqrt(np.array([0.5 * 0.5, 0.5 * 0.5, (np.pi / 12.0) * (np.pi / 12.0)]))   # pdf_gaussian.cpp:41-43
    pf.init_gaussian(np.zeros(3), np.eye(3), sig)
    res = pf.init_gaussian(np.zeros(3), np.eye(3), sig)
    got, want = pf.samples(), g["init_samples"]
    assert pf.drand48_state == int(g["init_state"])          # the same attempts were accepted
    assert np.array_equal(got[:, 3], want[:, 3])
    assert np.abs(got[:, :3] - want[:, :3]).max() <= 1e-12   # device log / sqrt vs glibc, last ulp
    assert res.sample_count == n and res.converged == 0
    if "init_leaf" in g:
        assert res.leaf_count == int(g["init_leaf"])
        assert np.array_equal(pf.sample_clusters(), g["init_labels"])
    pf.close()


@pytest.mark.gpu
def test_gpu_init_gaussian_rotated_covariance_vs_reference_sampler(oracle_built, dev):
    """a general rotation + three different sigmas, against PDFGaussian::sample of the compiled pdf_gaussian.cpp fed the
    SAME decomposition (a covariance that is diagonal in the rotated frame is passed pre-decomposed on both sides: the
    reference side through its diagonal path per axis, recombined here with the reference's own add order)"""
    import ctypes as C
    O, cabi = oracle_built, dev["cabi"]
    n = 20000
    cd = np.array([0.7, 0.05, 0.3])
    a = 0.6
    cr = np.array([[np.cos(a), -np.sin(a), 0.0], [np.sin(a), np.cos(a), 0.0], [0.0, 0.0, 1.0]])
    mean = np.array([1.5, -2.0, 0.4])
    # the draws r[j] = draw(cd[j]) in order, from the restated generator (pinned to pdf_gaussian.cpp's draw)
    rng = O.Rng(0)
    L = O.port_lib()
    L.orc_srand48(C.byref(rng), 77)
    r = np.array([[L.orc_gaussian_draw(C.byref(rng), cd[j]) for j in range(3)] for _ in range(n)])
    want = np.empty((n, 3))
    for i in range(3):   # v(i) = x[i]; v(i) += cr(i, j) * r[j], j = 0..2 (pdf_gaussian.cpp:65-69)
        t = np.full(n, mean[i])
        for j in range(3):
            t = t + cr[i, j] * r[:, j]
        want[:, i] = t
    pf = cabi.ParticleFilter(dev["ctx"], 100, n)
    pf.srand48(77)
    pf.init_gaussian(mean, cr, cd)
    assert pf.drand48_state == rng.x
    assert np.abs(pf.samples()[:, :3] - want).max() <= 1e-12
    pf.close()


# ------------------------------------------------------------------------------- the node itself, relinked
def test_reference_node_replayed_from_golden_inputs_is_the_golden_run(oracle_built):
    """run_node_2d / run_node_3d (the driver the relinked-node test below uses) on the all-reference library: bit-exact"""
    O = oracle_built
    if not O.have_ref():
        pytest.skip("oracle/_ref not built (needs the reference checkout)")
    for which, run in (("2d", nc.run_node_2d), ("3d", nc.run_node_3d), ("2d_mn", nc.run_node_2d_mn)):
        g = golden("node_%s.npz" % which)
        out = run(O, O.ref_lib(), g)
        for k, v in out.items():
            if k in g:
                assert np.array_equal(np.asarray(v), g[k], equal_nan=True), (which, k)
        assert int(out["init_state"]) == int(g["init_state"])


def _compare_relinked(out, g, pose_tol, same_particles):
    assert int(out["init_state"]) == int(g["init_state"])                       # initWithGaussian x 2: same draws accepted
    assert np.abs(out["init_samples"][:, :3] - g["init_samples"][:, :3]).max() <= 1e-12
    for k in range(nc.N_STEPS):
        s, w = out["s%d_samples" % k], g["s%d_samples" % k]
        assert out["s%d_resample_count" % k] == int(g["s%d_resample_count" % k])   # the node's update / resample schedule
        assert out["s%d_pose_count" % k] == int(g["s%d_pose_count" % k])           # ... and its publications
        assert s.shape == w.shape
        assert int(out["s%d_state" % k]) == int(g["s%d_state" % k])                # libc's drand48 stream stays in step
        if same_particles is not None:
            same = ((np.abs(s[:, :3] - w[:, :3]).max(axis=1) <= 1e-9) & np.isclose(s[:, 3], w[:, 3], rtol=2e-5, atol=1e-300)).mean()
            assert same >= same_particles, (k, same)
        if out["s%d_pose_count" % k]:
            assert np.abs(out["s%d_pose" % k][:2] - g["s%d_pose" % k][:2]).max() <= pose_tol
            if same_particles is not None:
                assert np.allclose(out["s%d_cov" % k], g["s%d_cov" % k], rtol=1e-2, atol=1e-4)


@pytest.mark.gpu
def test_gpu_the_reference_node_relinked_against_the_cuda_library_2d(oracle_built, dev):
    """The drop-in, executed: node/node.cpp + node_2d.cpp of the reference, unmodified, compiled against the mirror's
    headers and linked with the mirror + libbamcl_cuda.so (oracle/_ref/libamcl_node_gpu.so), fed the golden run's messages.
    Everything the node does itself (conversions, schedule) is identical; the filter underneath is the CUDA path with the
    reference's cache installed (OccupancyMap::setDistancesLUT of the mirror): same draws, same schedule, >= 99.9 % of the
    particles the reference's at every step (default fixed-point terms), published pose to a millimetre; and with the
    DEVICE-built cache (exact EDT, DESIGN 4 - the plain drop-in): same draws and schedule, the pose within centimetres."""
    O = oracle_built
    if not O.have_node_gpu():
        pytest.skip("oracle/_ref/libamcl_node_gpu.so not built (needs the reference checkout at build time)")
    g = golden("node_2d.npz")
    lib = O.node_gpu_lib()
    out = nc.run_node_2d(O, lib, g)
    assert lib.ref_node_last_device_error() == b""
    assert np.array_equal(out["free"], g["free"])
    for k in range(nc.N_STEPS):   # the node's own scan conversion (its tf2 / float arithmetic is the same object code path)
        assert np.array_equal(out["s%d_scan_ranges" % k], g["s%d_scan_ranges" % k], equal_nan=True)
        assert np.array_equal(out["s%d_scan_angles" % k], g["s%d_scan_angles" % k])
    _compare_relinked(out, g, pose_tol=0.05, same_particles=None)
    out = nc.run_node_2d(O, lib, g, reference_cache=True)
    assert lib.ref_node_last_device_error() == b""
    _compare_relinked(out, g, pose_tol=1e-3, same_particles=0.999)


@pytest.mark.gpu
def test_gpu_chain_2d_multinomial_with_recovery_vs_the_node(dev):
    """the C ABI driven in the node's order, random poses from the registered device pose source (exact terms: every
    resampled / injected particle the reference's, generator state exact)"""
    g = golden("node_2d_mn.npz")
    b = nc.CudaBackend2D(dev["cabi"], dev["ctx"], g, nc.PARAMS_2D_MN, exact_terms=True)
    bad = []
    nc.replay_2d(b, g, tolerant(bad, 1e-9, 1.0), nc.PARAMS_2D_MN)
    b.close()
    assert bad == []


@pytest.mark.gpu
def test_gpu_the_reference_node_relinked_multinomial_with_recovery(oracle_built, dev):
    """the relinked node with its DEFAULT resampler and recovery on: the mirror runs the branch loop of
    particle_filter.cpp:385-393 on the host (the node's pose function is an opaque std::function that draws from the same
    stream) and everything else on the device (bamcl_pf_resample_multinomial_drawn)"""
    O = oracle_built
    if not O.have_node_gpu():
        pytest.skip("oracle/_ref/libamcl_node_gpu.so not built (needs the reference checkout at build time)")
    g = golden("node_2d_mn.npz")
    lib = O.node_gpu_lib()
    out = nc.run_node_2d_mn(O, lib, g, reference_cache=True)
    assert lib.ref_node_last_device_error() == b""
    assert (out["s3_averages"] == 0.0).all() and int(out["s3_clusters"]) > 50   # random poses were injected here too
    _compare_relinked(out, g, pose_tol=2e-3, same_particles=0.995)


@pytest.mark.gpu
def test_gpu_the_reference_node_relinked_against_the_cuda_library_3d(oracle_built, dev):
    O = oracle_built
    if not O.have_node_gpu():
        pytest.skip("oracle/_ref/libamcl_node_gpu.so not built (needs the reference checkout at build time)")
    g = golden("node_3d.npz")
    lib = O.node_gpu_lib()
    out = nc.run_node_3d(O, lib, g)
    assert lib.ref_node_last_device_error() == b""
    assert out["free_count"] == (int(g["mx"][0]) - int(g["mn"][0])) * (int(g["mx"][1]) - int(g["mn"][1]))
    for k in range(nc.N_STEPS):
        assert np.array_equal(out["s%d_points" % k], g["s%d_points" % k])
    _compare_relinked(out, g, pose_tol=0.05, same_particles=None)
    out = nc.run_node_3d(O, lib, g, reference_cache=True)
    assert lib.ref_node_last_device_error() == b""
    _compare_relinked(out, g, pose_tol=1e-3, same_particles=0.999)


@pytest.mark.gpu
def test_gpu_relinked_node_events_against_the_reference_node_live(oracle_built, dev):
    """both node builds in one process on the same inputs: likelihood_field_prob + beam skipping, the global_localization
    service (initWithPoseFn through the node's generator), an /initialpose message (initWithGaussian) and a dynamic
    reconfigure (new filter, new model). Same schedule, same set sizes, libc's drand48 stream in step after every scan,
    nearly all particles identical, poses to millimetres."""
    O = oracle_built
    if not (O.have_node_gpu() and O.have_ref()):
        pytest.skip("needs oracle/_ref/libamcl_ref.so and libamcl_node_gpu.so (built where the reference checkout exists)")
    g = golden("node_2d.npz")
    want = nc.run_events_2d(O, O.ref_lib(), g)
    lib = O.node_gpu_lib()
    got = nc.run_events_2d(O, lib, g, cache=g["brushfire"])
    assert lib.ref_node_last_device_error() == b""
    assert [r["label"] for r in got] == [r["label"] for r in want]
    for a, b in zip(got, want):
        what = a["label"]
        assert (a["resample_count"], a["pose_count"], a["gl_active"]) == (b["resample_count"], b["pose_count"], b["gl_active"]), what
        assert a["samples"].shape == b["samples"].shape, what
        assert a["state"] == b["state"], what
        same = (np.abs(a["samples"][:, :3] - b["samples"][:, :3]).max(axis=1) <= 1e-9).mean()
        assert same >= 0.99, (what, same)
        assert np.abs(a["pose"][:2] - b["pose"][:2]).max() <= 5e-3, what
        assert a["converged"] == b["converged"], what


@pytest.mark.gpu
def test_gpu_relinked_node_3d_events_against_the_reference_node_live(oracle_built, dev):
    """3-D twin: OctoMap cropped to the occupancy map's bounds (setMapBounds), the free-space rectangle, clouds, the
    global_localization service, clouds again - both node builds live on the same inputs"""
    O = oracle_built
    if not (O.have_node_gpu() and O.have_ref()):
        pytest.skip("needs oracle/_ref/libamcl_ref.so and libamcl_node_gpu.so (built where the reference checkout exists)")
    g = golden("node_3d.npz")
    want = nc.run_events_3d(O, O.ref_lib(), g)
    lib = O.node_gpu_lib()
    got = nc.run_events_3d(O, lib, g, cache=want[1]["field"])
    assert lib.ref_node_last_device_error() == b""
    assert (got[0]["free"], got[1]["free"]) == (want[0]["free"], want[1]["free"])
    for a, b in zip(got[2:], want[2:]):
        what = a["label"]
        assert (a["resample_count"], a["pose_count"], a["gl_active"]) == (b["resample_count"], b["pose_count"], b["gl_active"]), what
        assert a["samples"].shape == b["samples"].shape, what
        assert a["state"] == b["state"], what
        same = (np.abs(a["samples"][:, :3] - b["samples"][:, :3]).max(axis=1) <= 1e-9).mean()
        assert same >= 0.99, (what, same)
        assert np.abs(a["pose"][:2] - b["pose"][:2]).max() <= 5e-3, what
