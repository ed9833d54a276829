"""The helpers either side of the hot path (global localisation, scan / map ingestion, pose read-out) live
in the reference's ROS node sources. Those translation units (node/node.cpp, node_2d.cpp, node_3d.cpp) are compiled
UNMODIFIED into oracle/_ref against header-only look-alikes of roscpp / tf2_ros / dynamic_reconfigure / message_filters /
yaml-cpp (oracle/shim) and driven in process here: parameters from a table, the map / scan / cloud messages delivered to
the node's own callbacks, tf from a table. What the node computes is then the reference's own object code, and the
restatements the CUDA path is tested against (oracle.py / amcl_oracle.c) are pinned to it bit for bit. tf2's own
arithmetic (quaternion <-> yaw) is the shim's restatement of the published tf2 code and stays "parity unpinned"."""
import ctypes as C

import numpy as np
import pytest


@pytest.fixture(scope="module")
def O(oracle_built):
    if not oracle_built.have_ref():
        pytest.skip("oracle/_ref not built (needs the reference checkout)")
    return oracle_built


def grid_world(h=120, w=160, seed=0):
    d = np.zeros((h, w), np.int8)
    d[0, :] = d[-1, :] = 100
    d[:, 0] = d[:, -1] = 100
    d[40:60, 50:55] = 100
    d[20:25, 120:140] = 100
    d[80:, 100:103] = -1
    rng = np.random.default_rng(seed)
    d[rng.integers(1, h - 1, 40), rng.integers(1, w - 1, 40)] = 100
    d[rng.integers(1, h - 1, 10), rng.integers(1, w - 1, 10)] = 37  # anything but 0 / 100 is unknown
    return d


def octo_world():
    keys = [(i, j, 0) for i in range(-20, 60) for j in range(-10, 50)]
    keys += [(i, j, k) for i in range(10, 14) for j in range(5, 30) for k in range(1, 8)]
    return np.array(keys, np.int32)


BASE = dict(min_particles=300, max_particles=1000, laser_max_beams=30, resample_model_type="systematic",
            recovery_alpha_slow=0.001, recovery_alpha_fast=0.1, laser_likelihood_max_dist=2.0, odom_integrator_enabled=0)


def drand_state(O):
    s = (C.c_ushort * 3)()
    O.ref_lib().ref_get_seed48(s)
    return s[0] | (s[1] << 16) | (s[2] << 32)


@pytest.mark.parametrize("scale", [1, 2, 3])
def test_node_convert_map(O, scale):
    """Node2D::mapMsgReceived -> convertMap (node_2d.cpp:202-295): cells, resolution (a FLOAT division) and origin"""
    data = grid_world(37, 53, seed=scale)
    n = O.RefNode(dict(BASE, map_scale_up_factor=scale))
    assert n.publish_map(data, 0.05, (-7.3, 2.15)) == 1
    m = n.map2d()
    cells, res, org = O.convert_map(data, 0.05, (-7.3, 2.15), scale)
    assert (m.size_y, m.size_x) == cells.shape
    assert np.array_equal(m.cells, cells)
    assert m.resolution == res and m.resolution != 0.05 / scale
    assert m.origin == (float(org[0]), float(org[1]))
    n.close()


@pytest.mark.parametrize("radius", [0.0, 0.3])
def test_node_free_space_list(O, radius):
    """Node2D::updateFreeSpaceIndices (node_2d.cpp:318-338) after the model built the distance cache"""
    data = grid_world()
    n = O.RefNode(dict(BASE, laser_scanner_non_free_space_radius=radius))
    n.publish_map(data, 0.05, (-4.0, -3.0))
    cells, res, org = O.convert_map(data, 0.05, (-4.0, -3.0))
    pm = O.PortMap2D(cells, res, org)
    assert np.array_equal(pm.update_distances(2.0), n.map2d().get_distances())
    fs = pm.free_space(radius)
    assert len(fs) > 1000 and np.array_equal(fs, n.free_space())
    n.close()


@pytest.mark.parametrize("thr,mult,after_scan", [(0.0, 0.0, False), (3.0, 0.5, False), (3.0, 0.5, True), (0.9, 0.5, True)])
def test_node_uniform_pose_generator(O, thr, mult, after_scan):
    """Node::uniformPoseGenerator / randomFreeSpacePose (node.cpp:823-868). With a scan present Node2D::scorePose scores a
    COPY of its fake sample and still returns 1.0 (node_2d.cpp:298-316): the retry count is fixed by the two parameters."""
    data = grid_world()
    n = O.RefNode(dict(BASE, uniform_pose_starting_weight_threshold=thr, uniform_pose_deweight_multiplier=mult))
    n.publish_map(data, 0.05, (-4.0, -3.0))
    if after_scan:
        n.set_odom(0.5, 0.2, 0.3)
        n.set_tf("base_link", "laser", (0.2, 0.0, 0.1), O.yaw_quat(0.1))
        r = np.random.default_rng(3).uniform(0.5, 8.0, 181).astype(np.float32)
        assert n.publish_scan(r, -1.5707964, 0.017453292, 0.1, 30.0, stamp=1.0) == 1
        assert n.scan_data() is not None
    cells, res, org = O.convert_map(data, 0.05, (-4.0, -3.0))
    pm = O.PortMap2D(cells, res, org)
    pm.update_distances(2.0)
    fs = pm.free_space(0.0)
    O.ref_lib().ref_srand48(99)
    got = n.uniform_poses(500)
    rng = O.Rng(0)
    O.port_lib().orc_srand48(C.byref(rng), 99)
    want = pm.uniform_poses(fs, rng, 500, thr, mult)
    assert np.array_equal(got, want)
    assert drand_state(O) == rng.x
    n.close()


def test_node_global_localization_service(O):
    """global_localization service -> ParticleFilter::initWithPoseFn(uniformPoseGenerator) (node.cpp:870-883)"""
    data = grid_world()
    n = O.RefNode(dict(BASE, global_localization_alpha_slow=0.002, global_localization_alpha_fast=0.2))
    n.publish_map(data, 0.05, (-4.0, -3.0))
    O.ref_lib().ref_srand48(5)
    assert n.global_localization() == 1
    st = n.state()
    assert st["global_localization_active"] and not st["odom_init"]
    cells, res, org = O.convert_map(data, 0.05, (-4.0, -3.0))
    pm = O.PortMap2D(cells, res, org)
    pm.update_distances(2.0)
    fs = pm.free_space(0.0)
    rng = O.Rng(0)
    O.port_lib().orc_srand48(C.byref(rng), 5)
    port = O.PortPF(300, 1000)
    port.init_with_poses(pm.uniform_poses(fs, rng, 1000))
    ref = n.pf()
    assert np.array_equal(ref.samples(), port.samples())
    assert ref.leaf_count() == port.leaf_count() and ref.cluster_count() == port.cluster_count()
    assert np.array_equal(ref.sample_clusters(), port.sample_clusters())
    assert drand_state(O) == rng.x
    n.close()


@pytest.mark.parametrize("smin,smax,mount_yaw", [(-1.0, -1.0, 0.1), (0.3, 12.0, 0.1), (0.05, 100.0, 3.0), (0.3, 12.0, -2.0)])
def test_node_scan_ingestion(O, smin, smax, mount_yaw):
    """Node2D::scanReceived -> updateScanner -> getAngleStats + updateLatestScanData (node_2d.cpp:340-392, 497-560)"""
    data = grid_world()
    n = O.RefNode(dict(BASE, laser_min_range=smin, laser_max_range=smax))
    n.publish_map(data, 0.05, (-4.0, -3.0))
    n.set_odom(0.5, 0.2, 0.3)
    n.set_tf("base_link", "laser", (0.2, 0.0, 0.1), O.yaw_quat(mount_yaw))
    r = np.random.default_rng(3).uniform(0.0, 35.0, 181).astype(np.float32)
    r[::50] = np.float32(0.1)
    r[7], r[9] = np.nan, np.inf
    assert n.publish_scan(r, -1.5707964, 0.017453292, 0.1, 30.0, stamp=1.0) == 1
    got = n.scan_data()
    stats = n.angle_stats(-1.5707964, 0.017453292)
    # the scan's bearings in the base frame: the mount yaw is added by a quaternion product (tf2 shim), increment kept
    assert abs(O.port_lib().orc_normalize_angle(stats[0] - (-1.5707964 + mount_yaw))) < 1e-6 and abs(stats[1] - 0.017453292) < 1e-6
    want = O.scan_from_laserscan(r, 0.1, 30.0, stats[0], stats[1], smin, smax)
    assert np.array_equal(got[0], want[0], equal_nan=True)
    assert np.array_equal(got[1], want[1])
    assert got[2] == want[2]
    n.close()


@pytest.mark.parametrize("count,max_beams", [(1000, 64), (64, 64), (65, 64), (3, 256), (40000, 30)])
def test_node_cloud_ingestion(O, count, max_beams):
    """Node3D::scanReceived -> updateScanner -> makePointCloudFromScan + updateLatestScanData (node_3d.cpp:320-366, 459-480)"""
    n = O.RefNode(dict(BASE, laser_max_beams=max_beams, wait_for_occupancy_map=0, laser_likelihood_max_dist=0.5), map_type=3)
    assert n.publish_octomap(0.1, octo_world()) == 1
    n.set_odom(0.5, 0.2, 0.3)
    n.set_tf("base_link", "lidar", (0.2, 0.0, 0.6), O.yaw_quat(0.1))
    xyz = np.random.default_rng(count).normal(0.0, 2.0, (count, 3)).astype(np.float32)
    assert n.publish_cloud(xyz, stamp=1.0) == 1
    assert np.array_equal(n.cloud_data(), O.cloud_subsample(xyz, max_beams))
    n.close()


@pytest.mark.parametrize("thr,mult", [(0.0, 0.0), (5.0, 0.4)])
def test_node_3d_free_space_and_generator(O, thr, mult):
    """Node3D::updateFreeSpaceIndices (node_3d.cpp:306-318) = the OctoMap's cropped cell rectangle, i outer; poses through
    OctoMap::convertMapToWorld (octomap.cpp:83-95)"""
    n = O.RefNode(dict(BASE, wait_for_occupancy_map=0, laser_likelihood_max_dist=0.5,
                       uniform_pose_starting_weight_threshold=thr, uniform_pose_deweight_multiplier=mult), map_type=3)
    n.publish_octomap(0.1, octo_world())
    mn, mx = n.map3d().bounds()
    fs = n.free_space()
    want = np.array([(i, j) for i in range(mn[0], mx[0]) for j in range(mn[1], mx[1])], np.int32)
    assert np.array_equal(fs, want)
    O.ref_lib().ref_srand48(17)
    got = n.uniform_poses(400)
    rng = O.Rng(0)
    O.port_lib().orc_srand48(C.byref(rng), 17)
    assert np.array_equal(got, O.uniform_poses_rect(mn, mx, 0.1, rng, 400, thr, mult))
    assert drand_state(O) == rng.x
    n.close()


def test_node_pose_readout(O):
    """Node2D::resamplePose -> getMaxWeightPose -> Node::updatePose (node_2d.cpp:572-617, node.cpp:359-433): the published
    amcl_pose carries the best cluster's mean and the SET covariance in the 6x6 layout"""
    data = grid_world()
    n = O.RefNode(dict(BASE, resample_interval=1))
    n.publish_map(data, 0.05, (-4.0, -3.0))
    n.set_odom(0.5, 0.2, 0.3)
    n.set_tf("base_link", "laser", (0.2, 0.0, 0.1), O.yaw_quat(0.1))
    r = np.random.default_rng(3).uniform(0.5, 8.0, 181).astype(np.float32)
    O.ref_lib().ref_srand48(3)
    n.publish_scan(r, -1.5707964, 0.017453292, 0.1, 30.0, stamp=1.0)
    pf = n.pf()
    w, pose, idx = pf.max_weight_pose()                    # the restatement over the reference's cluster table
    got = n.max_weight_pose()                              # Node2D::getMaxWeightPose itself
    assert got[0] == w and np.array_equal(got[1], pose)
    msg = n.last_pose()
    assert msg["count"] == 1 and msg["stamp"] == 1.0
    assert msg["position"][0] == pose[0] and msg["position"][1] == pose[1] and msg["position"][2] == 0.0
    assert np.array_equal(msg["covariance"], pf.pose_covariance())
    yaw = 2.0 * np.arctan2(msg["orientation"][2], msg["orientation"][3])
    assert abs(yaw - pose[2]) < 1e-12
    # the particle cloud message repeats the set
    cloud = n.last_particlecloud()
    s = pf.samples()
    assert len(cloud) == len(s) and np.array_equal(cloud[:, 0], s[:, 0]) and np.array_equal(cloud[:, 1], s[:, 1])
    n.close()


def test_gaussian_pdf_is_reference_code_on_diagonal_covariances(O):
    """pf/pdf_gaussian.cpp compiles against the Eigen::EigenSolver look-alike; for the diagonal covariances the node passes
    (node.cpp:281-287, 688-701) the decomposition is (diagonal, identity) in any correct solver, so sample() is the
    reference's: mean + sigma_i * draw, draws in order x, y, theta"""
    L = O.ref_lib()
    mean, cov = np.array([1.0, -2.0, 0.3]), np.diag([0.25, 0.09, (np.pi / 12.0) ** 2])
    out = np.empty((50, 3))
    L.ref_srand48(8)
    L.ref_gaussian_sample(O._ptr(mean), O._ptr(np.ascontiguousarray(cov)), 50, O._ptr(out))
    L.ref_srand48(8)
    sig = np.sqrt(np.diag(cov))
    for k in range(50):
        for a in range(3):
            assert out[k, a] == mean[a] + L.ref_gaussian_draw(sig[a])


def test_gaussian_pdf_shim_decomposition_is_a_decomposition(O):
    """off-diagonal covariance: the shim's Jacobi solver is not Eigen's algorithm (order / signs unpinned), but it must be a
    valid one: samples reproduce the covariance"""
    L = O.ref_lib()
    a = np.array([[0.5, 0.2, 0.05], [0.2, 0.3, -0.1], [0.05, -0.1, 0.2]])
    mean = np.zeros(3)
    out = np.empty((200000, 3))
    L.ref_srand48(2)
    L.ref_gaussian_sample(O._ptr(mean), O._ptr(np.ascontiguousarray(a)), len(out), O._ptr(out))
    assert np.allclose(np.cov(out.T), a, atol=6e-3)


def test_reference_gaussian_kat_on_the_compiled_translation_unit(O):
    """the reference's own known answers (test/test_badger_amcl.cpp:29-51: identity covariance, the process's unseeded
    drand48 stream = glibc's all-zero initial state) reproduced by ITS pdf_gaussian.cpp compiled against the EigenSolver
    look-alike"""
    L = O.ref_lib()
    L.ref_seed48((C.c_ushort * 3)(0, 0, 0))
    out = np.empty((2, 3))
    mean, cov = np.ones(3), np.eye(3)
    L.ref_gaussian_sample(O._ptr(mean), O._ptr(cov), 2, O._ptr(out))
    want = [[0.26562654174915334, 0.97172090090793528, -1.5856194295513539],
            [1.6262083813236745, 1.1142314205031041, 0.37407538872488655]]
    assert np.allclose(out, want, rtol=4e-16, atol=0.0)   # EXPECT_DOUBLE_EQ = 4 ulps


def test_node_draws_its_first_set_twice(O):
    """Node::initFromNewMap (node.cpp:670-708): initWithGaussian with the initial pose, then setInitialPose ->
    applyInitialPose -> initWithGaussian again; the particle set a fresh node holds is the SECOND batch of the stream"""
    n = O.RefNode(dict(BASE))
    O.ref_lib().ref_srand48(21)
    n.publish_map(grid_world(), 0.05, (-4.0, -3.0))
    got = n.pf().samples()
    state = drand_state(O)
    pf = O.RefPF(300, 1000)
    pf.srand48(21)
    cov = np.diag([0.5 * 0.5, 0.5 * 0.5, (np.pi / 12.0) * (np.pi / 12.0)])
    pf.init_gaussian(np.zeros(3), cov)
    assert not np.array_equal(pf.samples(), got)
    pf.init_gaussian(np.zeros(3), cov)
    assert np.array_equal(pf.samples(), got) and drand_state(O) == state
    n.close()


def test_tf2_look_alike_against_independent_formulas(O):
    """the tf2 arithmetic inside oracle/shim is a restatement of published third-party code (parity unpinned); what CAN be
    checked is that it is the mathematics it claims to be: yaw recovered through quaternion -> matrix -> getEulerYPR
    (Node::getOdomPose), quaternion product + getYaw (Node2D::getAngleStats), and Transform inverse / compose
    (Node::updatePose: latest_tf_ = (odom->base * (map->base)^-1), so latest_tf_^-1 * odom->base == map->base)"""
    n = O.RefNode(dict(BASE, resample_interval=1))
    n.publish_map(grid_world(), 0.05, (-4.0, -3.0))
    for yaw in (-3.1, -1.5707963267948966, -0.3, 0.0, 1e-9, 0.7, 2.9, 3.14159):
        n.set_odom(1.25, -0.5, yaw)
        got = n.odom_pose()
        assert abs(got[0] - 1.25) < 1e-15 and abs(got[1] + 0.5) < 1e-15 and abs(got[2] - yaw) < 1e-12
    for mount in (-2.0, 0.1, 3.0):
        n.set_tf("base_link", "laser", (0.2, 0.0, 0.1), O.yaw_quat(mount))
        a0, inc = n.angle_stats(-1.2, 0.01)
        want = np.arctan2(np.sin(-1.2 + mount), np.cos(-1.2 + mount))
        assert abs(a0 - want) < 1e-7 and abs(inc - 0.01) < 1e-7      # message fields are float32
    # one scan -> updatePose: map->base is the published pose; odom->base was set above
    n.set_odom(0.5, 0.2, 0.3)
    n.set_tf("base_link", "laser", (0.2, 0.0, 0.1), O.yaw_quat(0.1))
    r = np.random.default_rng(3).uniform(0.5, 8.0, 181).astype(np.float32)
    O.ref_lib().ref_srand48(3)
    assert n.publish_scan(r, -1.5707964, 0.017453292, 0.1, 30.0, stamp=1.0) == 1
    msg, tf = n.last_pose(), n.latest_tf()

    def mat(x, y, yaw):
        c, s = np.cos(yaw), np.sin(yaw)
        return np.array([[c, -s, x], [s, c, y], [0.0, 0.0, 1.0]])
    map_base = mat(msg["position"][0], msg["position"][1], 2.0 * np.arctan2(msg["orientation"][2], msg["orientation"][3]))
    odom_base = mat(0.5, 0.2, 0.3)
    latest = mat(tf[0], tf[1], 2.0 * np.arctan2(tf[5], tf[6]))
    # node.cpp:404-427: latest_tf_ = T(odom<-base) applied to (map<-base)^-1, i.e. odom<-map
    assert np.allclose(latest, odom_base @ np.linalg.inv(map_base), atol=1e-12)
    n.close()
