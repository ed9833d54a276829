"""One localisation run THROUGH THE REFERENCE'S OWN ROS NODE (node/node.cpp + node_2d.cpp / node_3d.cpp compiled unmodified
into oracle/_ref, see tests/test_node_pins.py), recorded step by step, and the same run replayed through the restated chain
(CPU) or the CUDA library (GPU). This pins the CALLER SEQUENCE around the hot path - map message -> model + cache ->
Gaussian init -> [tf odometry -> Odom::updateAction -> scan conversion -> updateSensor -> every resample_interval-th scan
updateResample -> max-weight pose + covariance] - not just the individual functions.

record_2d / record_3d need the compiled reference; their output is committed as tests/golden/node_2d.npz / node_3d.npz
(tests/golden/make_golden.py), so the replays run on the GPU box where /root/reference does not exist."""
import ctypes as C

import numpy as np

RES_MSG = 0.05          # nav_msgs/OccupancyGrid info.resolution (a float32 in the message)
ORIGIN = (-4.0, -3.0)
MOUNT = (0.2, 0.0, 0.1, 0.05)   # base_link -> laser: x, y, z, yaw
SCAN = dict(angle_min=-1.5707964, angle_increment=0.017453292, range_min=0.1, range_max=30.0, beams=181)
PARAMS_2D = dict(min_particles=300, max_particles=1500, laser_max_beams=30, resample_model_type="systematic",
                 recovery_alpha_slow=0.001, recovery_alpha_fast=0.1, laser_likelihood_max_dist=2.0,
                 odom_integrator_enabled=0, laser_min_range=0.3, laser_max_range=12.0, resample_interval=2,
                 update_min_d=0.2, update_min_a=0.5235987755982988, odom_model_type="diff", odom_alpha1=0.2,
                 odom_alpha2=0.2, odom_alpha3=0.2, odom_alpha4=0.2, odom_alpha5=0.2, kld_err=0.01, kld_z=0.99,
                 laser_z_hit=0.95, laser_z_rand=0.05, laser_sigma_hit=0.2, laser_model_type="likelihood_field")
N_STEPS = 6
CLOUD_MOUNT = (0.2, 0.0, 0.6, 0.1)
PARAMS_3D = dict(min_particles=200, max_particles=800, laser_max_beams=64, resample_model_type="systematic",
                 recovery_alpha_slow=0.001, recovery_alpha_fast=0.1, laser_likelihood_max_dist=0.5,
                 odom_integrator_enabled=0, resample_interval=2, update_min_d=0.2, update_min_a=0.5235987755982988,
                 odom_model_type="diff", odom_alpha1=0.2, odom_alpha2=0.2, odom_alpha3=0.2, odom_alpha4=0.2, odom_alpha5=0.2,
                 kld_err=0.01, kld_z=0.99, laser_z_hit=0.95, laser_z_rand=0.05, laser_sigma_hit=0.2,
                 laser_model_type="likelihood_field_gompertz", laser_gompertz_a=1.0, laser_gompertz_b=2.0,
                 laser_gompertz_c=6.0, laser_gompertz_input_shift=-0.2, laser_gompertz_input_scale=1.3,
                 laser_gompertz_output_shift=0.05, wait_for_occupancy_map=0)


def grid_world(h=120, w=160):
    d = np.zeros((h, w), np.int8)
    d[0, :] = d[-1, :] = 100
    d[:, 0] = d[:, -1] = 100
    d[40:60, 50:55] = 100
    d[20:25, 120:140] = 100
    d[85:, 100:103] = -1
    rng = np.random.default_rng(0)
    d[rng.integers(1, h - 1, 30), rng.integers(1, w - 1, 30)] = 100
    d[55:65, 75:90] = 0  # keep the start free
    return d


def octo_world():
    keys = [(i, j, 0) for i in range(-30, 70) for j in range(-30, 50)]
    keys += [(i, j, k) for i in range(15, 19) for j in range(-10, 30) for k in range(1, 12)]
    keys += [(i, 35, k) for i in range(-20, 60) for k in range(1, 10)]
    keys += [(-12, j, k) for j in range(-20, 30) for k in range(1, 8)]
    return np.array(keys, np.int32)


def truth_pose(step):
    """the robot's path in the odom (= map) frame"""
    return 0.25 * step, 0.05 * step, 0.12 * step


def normalize_angle(a):
    """angles::normalize_angle in the form oracle/shim/angles/angles.h and oracle/amcl_oracle.c restate it (ros/angles is a
    third-party dependency of the reference, not vendored)"""
    r = np.fmod(a + np.pi, 2.0 * np.pi)
    return r + np.pi if r <= 0.0 else r - np.pi


def drand_state(O):
    s = (C.c_ushort * 3)()
    O.ref_lib().ref_get_seed48(s)
    return s[0] | (s[1] << 16) | (s[2] << 32)


def _record_step(n, g, k):
    pf = n.pf()
    s = pf.samples()
    g["s%d_samples" % k] = s
    g["s%d_state" % k] = np.uint64(0)
    msg = n.last_pose()
    g["s%d_pose_count" % k] = msg["count"] if msg else 0
    if msg:
        g["s%d_pose" % k] = np.concatenate([msg["position"], msg["orientation"]])
        g["s%d_cov" % k] = msg["covariance"]
    st = n.state()
    g["s%d_resample_count" % k] = st["resample_count"]
    g["s%d_averages" % k] = np.array(pf.averages())
    g["s%d_leaf" % k] = pf.leaf_count()
    g["s%d_clusters" % k] = pf.cluster_count()
    g["s%d_labels" % k] = pf.sample_clusters()
    g["s%d_converged" % k] = int(pf.is_converged())
    mw = n.max_weight_pose()
    g["s%d_max_weight" % k] = np.concatenate([[mw[0]], mw[1]]) if mw else np.zeros(0)


# the node's default resampler (multinomial) with augmented-MCL recovery switched on, and a robot that is carried away after
# step 2 (the scans come from somewhere else than the odometry says): w_fast drops below w_slow and updateResample injects
# random poses through the node's own pose function, interleaved with the resampler's draws (particle_filter.cpp:385-393)
PARAMS_2D_MN = dict(PARAMS_2D, resample_model_type="multinomial", recovery_alpha_slow=0.05, recovery_alpha_fast=0.9,
                    min_particles=200, max_particles=1000)
KIDNAP = (1.6, -1.2, 0.9)   # added to the true pose the scans are taken from, steps 3..5


def record_2d(O, seed=21, params=None, kidnap_from=None):
    """run the 2-D scenario through the compiled node; returns the golden dict"""
    L = O.ref_lib()
    data = grid_world()
    g = dict(data=data)
    L.ref_srand48(seed)
    n = O.RefNode(params or PARAMS_2D, map_type=2)
    assert n.publish_map(data, RES_MSG, ORIGIN) == 1
    m = n.map2d()
    g["brushfire"] = m.get_distances()
    g["resolution"], g["origin"] = m.resolution, np.array(m.origin)
    g["free"] = n.free_space()
    pf = n.pf()
    g["init_samples"] = pf.samples()
    g["init_state"] = np.uint64(drand_state(O))
    g["init_leaf"], g["init_labels"] = pf.leaf_count(), pf.sample_clusters()
    n.set_tf("base_link", "laser", MOUNT[:3], O.yaw_quat(MOUNT[3]))
    noise = np.random.default_rng(4)
    for k in range(N_STEPS):
        x, y, yaw = truth_pose(k)
        n.set_odom(x, y, yaw)
        # what a laser at the true pose would see in this map, float32 like the message, a little noise, some drop-outs
        if kidnap_from is not None and k >= kidnap_from:
            x, y, yaw = x + KIDNAP[0], y + KIDNAP[1], yaw + KIDNAP[2]
        sx = x + MOUNT[0] * np.cos(yaw) - MOUNT[1] * np.sin(yaw)
        sy = y + MOUNT[0] * np.sin(yaw) + MOUNT[1] * np.cos(yaw)
        b = yaw + MOUNT[3] + SCAN["angle_min"] + SCAN["angle_increment"] * np.arange(SCAN["beams"])
        r = m.calc_ranges(np.full(len(b), sx), np.full(len(b), sy), b, SCAN["range_max"])
        r = (r + noise.normal(0.0, 0.02, len(r))).astype(np.float32)
        r[noise.random(len(r)) < 0.03] = np.float32(0.05)
        g["s%d_ranges" % k] = r
        g["s%d_odom" % k] = n.odom_pose(1.0 + k)  # yaw as the node recovers it from the quaternion (tf2 arithmetic)
        assert n.publish_scan(r, SCAN["angle_min"], SCAN["angle_increment"], SCAN["range_min"], SCAN["range_max"],
                              stamp=1.0 + k) == 1
        sd = n.scan_data()
        g["s%d_scan_ranges" % k], g["s%d_scan_angles" % k], g["s%d_scan_max" % k] = sd
        g["s%d_angle_stats" % k] = np.array(n.angle_stats(SCAN["angle_min"], SCAN["angle_increment"]))
        _record_step(n, g, k)
        g["s%d_state" % k] = np.uint64(drand_state(O))
    n.close()
    return g


def record_3d(O, seed=33):
    L = O.ref_lib()
    keys = octo_world()
    g = dict(keys=keys)
    L.ref_srand48(seed)
    n = O.RefNode(PARAMS_3D, map_type=3)
    assert n.publish_octomap(0.1, keys) == 1
    m = n.map3d()
    g["field"] = m.dense()
    mn, mx = m.bounds()
    g["mn"], g["mx"] = np.array(mn), np.array(mx)
    pf = n.pf()
    g["init_samples"] = pf.samples()
    g["init_state"] = np.uint64(drand_state(O))
    n.set_tf("base_link", "lidar", CLOUD_MOUNT[:3], O.yaw_quat(CLOUD_MOUNT[3]))
    rng = np.random.default_rng(6)
    for k in range(N_STEPS):
        x, y, yaw = truth_pose(k)
        n.set_odom(x, y, yaw)
        # points on the walls / floor around the robot, expressed in the lidar frame
        world = np.concatenate([
            np.stack([rng.uniform(1.5, 1.9, 300), rng.uniform(-1.0, 3.0, 300), rng.uniform(0.1, 1.2, 300)], 1),
            np.stack([rng.uniform(-2.0, 6.0, 300), np.full(300, 3.5), rng.uniform(0.1, 1.0, 300)], 1),
            np.stack([rng.uniform(-3.0, 7.0, 300), rng.uniform(-3.0, 5.0, 300), np.full(300, 0.05)], 1)])
        lx = x + CLOUD_MOUNT[0] * np.cos(yaw) - CLOUD_MOUNT[1] * np.sin(yaw)
        ly = y + CLOUD_MOUNT[0] * np.sin(yaw) + CLOUD_MOUNT[1] * np.cos(yaw)
        a = yaw + CLOUD_MOUNT[3]
        dx, dy = world[:, 0] - lx, world[:, 1] - ly
        pts = np.stack([np.cos(a) * dx + np.sin(a) * dy, -np.sin(a) * dx + np.cos(a) * dy, world[:, 2] - CLOUD_MOUNT[2]], 1)
        pts = (pts + rng.normal(0.0, 0.01, pts.shape)).astype(np.float32)
        g["s%d_cloud" % k] = pts
        g["s%d_odom" % k] = n.odom_pose(1.0 + k)
        assert n.publish_cloud(pts, stamp=1.0 + k) == 1
        g["s%d_points" % k] = n.cloud_data()
        _record_step(n, g, k)
        g["s%d_state" % k] = np.uint64(drand_state(O))
    n.close()
    return g


def odom_delta(pose, last):
    """Node::computeDelta (node.cpp:1019-1025)"""
    return np.array([pose[0] - last[0], pose[1] - last[1], normalize_angle(pose[2] - last[2])])


def should_update(delta, p):
    """Node::setScannersUpdateFlags without the integrator (node.cpp:1041-1046)"""
    return bool(abs(delta[0]) > p["update_min_d"] or abs(delta[1]) > p["update_min_d"] or abs(delta[2]) > p["update_min_a"])


class PortBackend2D:
    """the restated chain (oracle/amcl_oracle.c through oracle.py): bit-exact against the node"""

    def __init__(self, O, g, p=PARAMS_2D):
        self.O = O
        cells, res, org = O.convert_map(g["data"], RES_MSG, ORIGIN, 1)
        self.map = O.PortMap2D(cells, res, org)
        self.map.update_distances(p["laser_likelihood_max_dist"])
        self.geometry = (res, (float(org[0]), float(org[1])))
        self.free = self.map.free_space(0.0)
        self.pf = O.PortPF(p["min_particles"], p["max_particles"], p["recovery_alpha_slow"], p["recovery_alpha_fast"], 95.0,
                           systematic=p["resample_model_type"] == "systematic", pop_err=p["kld_err"], pop_z=p["kld_z"])
        self.pf.set_uniform_pose_source(self.map, self.free)
        self.model = O.planar_model(model_type=O.MODEL_LF, max_beams=p["laser_max_beams"], z_hit=p["laser_z_hit"],
                                    z_rand=p["laser_z_rand"], sigma_hit=p["laser_sigma_hit"],
                                    scanner_pose=(MOUNT[0], MOUNT[1], 0.0))  # mounting angle handled in the bearings

    def distances(self):
        return self.map.dist

    def init(self, poses, state):
        self.pf.init_with_poses(poses)
        self.pf.rng.x = int(state)

    def convert_scan(self, r, stats, p):
        return self.O.scan_from_laserscan(r, SCAN["range_min"], SCAN["range_max"], stats[0], stats[1], p["laser_min_range"],
                                          p["laser_max_range"])

    def action(self, alpha, pose, delta):
        self.pf.update_action(0, alpha, pose, delta, delta)

    def sensor(self, ranges, angles, range_max):
        self.pf.update_sensor(self.map, self.model, ranges, angles, range_max)

    def resample(self):
        self.pf.update_resample()

    def samples(self):
        return self.pf.samples()

    def state(self):
        return self.pf.rng.x

    def readout(self):
        w, pose, idx = self.pf.max_weight_pose()
        return w, pose, self.pf.pose_covariance()

    def labels(self):
        return self.pf.leaf_count(), self.pf.cluster_count(), self.pf.sample_clusters()


def replay_2d(backend, g, check, p=PARAMS_2D):
    """drive `backend` through the node's sequence (node_2d.cpp:340-392, node.cpp:300-328, 1019-1112); `check(name, got,
    want)` is called for every recorded quantity"""
    alpha = [p["odom_alpha%d" % i] for i in range(1, 6)]
    backend.init(g["init_samples"][:, :3], g["init_state"])
    check("init_samples", backend.samples(), g["init_samples"])
    odom_init, update, resample_count, last = False, False, 0, None
    for k in range(N_STEPS):
        pose = g["s%d_odom" % k]
        force_publication = resampled = False
        if odom_init:                                   # Node::updatePf
            delta = odom_delta(pose, last)
            update = update or should_update(delta, p)
            if update:
                backend.action(alpha, pose, delta)      # Node::updateOdom, absolute_motion = delta without the integrator
                last = pose
        else:                                           # Node::initOdom
            last, odom_init, update, force_publication, resample_count = pose, True, True, True, 0
        if update:                                      # Node2D::updateScanner
            ranges, angles, range_max = backend.convert_scan(g["s%d_ranges" % k], g["s%d_angle_stats" % k], p)
            check("s%d_scan_ranges" % k, ranges, g["s%d_scan_ranges" % k])
            check("s%d_scan_angles" % k, angles, g["s%d_scan_angles" % k])
            check("s%d_scan_max" % k, range_max, g["s%d_scan_max" % k])
            backend.sensor(ranges, angles, range_max)
            update = False
            resample_count += 1
            if resample_count % p["resample_interval"] == 0:
                backend.resample()
                resampled = True
        check("s%d_resample_count" % k, resample_count, g["s%d_resample_count" % k])
        check("s%d_samples" % k, backend.samples(), g["s%d_samples" % k])
        check("s%d_state" % k, backend.state(), int(g["s%d_state" % k]))
        if force_publication or resampled:              # Node2D::resamplePose -> Node::updatePose
            w, mean, cov = backend.readout()
            check("s%d_max_weight" % k, np.concatenate([[w], mean]), g["s%d_max_weight" % k])
            check("s%d_pose_xy" % k, mean[:2], g["s%d_pose" % k][:2])
            check("s%d_cov" % k, cov, g["s%d_cov" % k])
        leaf, clusters, labels = backend.labels()
        check("s%d_leaf" % k, leaf, g["s%d_leaf" % k])
        check("s%d_clusters" % k, clusters, g["s%d_clusters" % k])
        if resampled:  # between resamplings the particles move but the tree is not rebuilt: a per-particle label is only
            check("s%d_labels" % k, labels, g["s%d_labels" % k])  # defined right after computeClusterStats ran


class PortBackend3D:
    """the restated 3-D chain: OctoMap cache, getMapCloud + calcPointCloudModel(Gompertz), updateSensor, resample"""

    def __init__(self, O, g, p=PARAMS_3D):
        self.O = O
        self.map = O.PortMap3D(0.1, g["keys"], p["laser_likelihood_max_dist"])
        self.map.update_distances()
        self.pf = O.PortPF(p["min_particles"], p["max_particles"], p["recovery_alpha_slow"], p["recovery_alpha_fast"], 95.0,
                           systematic=p["resample_model_type"] == "systematic", pop_err=p["kld_err"], pop_z=p["kld_z"])
        self.model = O.cloud_model(model_type=1, max_beams=p["laser_max_beams"], z_hit=p["laser_z_hit"],
                                   z_rand=p["laser_z_rand"], sigma_hit=p["laser_sigma_hit"],
                                   gompertz=(p["laser_gompertz_a"], p["laser_gompertz_b"], p["laser_gompertz_c"],
                                             p["laser_gompertz_input_shift"], p["laser_gompertz_input_scale"],
                                             p["laser_gompertz_output_shift"]),
                                   mount_translation=CLOUD_MOUNT[:3], mount_rotation=O.yaw_quat(CLOUD_MOUNT[3]))

    def field(self):
        return self.map.dense()

    def bounds(self):
        return self.map.bounds()

    def init(self, poses, state):
        self.pf.init_with_poses(poses)
        self.pf.rng.x = int(state)

    def subsample(self, xyz, p):
        return self.O.cloud_subsample(xyz, p["laser_max_beams"])

    def action(self, alpha, pose, delta):
        self.pf.update_action(0, alpha, pose, delta, delta)

    def sensor(self, points):
        s = self.pf.samples()
        total, w = self.map.cloud_apply(self.model, points, s)
        s[:, 3] = w
        self.pf.set_samples(s)
        self.O.port_lib().orc_pf_normalise(self.pf.h, total)  # ParticleFilter::updateSensor, particle_filter.cpp:223-266

    def resample(self):
        self.pf.update_resample()

    def samples(self):
        return self.pf.samples()

    def state(self):
        return self.pf.rng.x

    def readout(self):
        w, pose, idx = self.pf.max_weight_pose()
        return w, pose, self.pf.pose_covariance()

    def labels(self):
        return self.pf.leaf_count(), self.pf.cluster_count(), self.pf.sample_clusters()


def replay_3d(backend, g, check, p=PARAMS_3D):
    """Node3D's sequence (node_3d.cpp:320-366): as replay_2d with the cloud sub-sampled instead of the scan converted"""
    alpha = [p["odom_alpha%d" % i] for i in range(1, 6)]
    backend.init(g["init_samples"][:, :3], g["init_state"])
    check("init_samples", backend.samples(), g["init_samples"])
    odom_init, update, resample_count, last = False, False, 0, None
    for k in range(N_STEPS):
        pose = g["s%d_odom" % k]
        force_publication = resampled = False
        if odom_init:
            delta = odom_delta(pose, last)
            update = update or should_update(delta, p)
            if update:
                backend.action(alpha, pose, delta)
                last = pose
        else:
            last, odom_init, update, force_publication, resample_count = pose, True, True, True, 0
        if update:
            points = backend.subsample(g["s%d_cloud" % k], p)
            check("s%d_points" % k, points, g["s%d_points" % k])
            backend.sensor(points)
            update = False
            resample_count += 1
            if resample_count % p["resample_interval"] == 0:
                backend.resample()
                resampled = True
        check("s%d_resample_count" % k, resample_count, g["s%d_resample_count" % k])
        check("s%d_samples" % k, backend.samples(), g["s%d_samples" % k])
        check("s%d_state" % k, backend.state(), int(g["s%d_state" % k]))
        if force_publication or resampled:
            w, mean, cov = backend.readout()
            check("s%d_max_weight" % k, np.concatenate([[w], mean]), g["s%d_max_weight" % k])
            check("s%d_pose_xy" % k, mean[:2], g["s%d_pose" % k][:2])
            check("s%d_cov" % k, cov, g["s%d_cov" % k])
        leaf, clusters, labels = backend.labels()
        check("s%d_leaf" % k, leaf, g["s%d_leaf" % k])
        check("s%d_clusters" % k, clusters, g["s%d_clusters" % k])
        if resampled:  # between resamplings the particles move but the tree is not rebuilt: a per-particle label is only
            check("s%d_labels" % k, labels, g["s%d_labels" % k])  # defined right after computeClusterStats ran


class CudaBackend2D:
    """the CUDA library through the C ABI (badger_amcl_b200.cabi), driven in the node's order"""

    def __init__(self, cabi, ctx, g, p=PARAMS_2D, reference_cache=True, exact_terms=True):
        self.cabi = cabi
        self.map = cabi.Map2D.from_occupancy_grid(ctx, g["data"], RES_MSG, ORIGIN, 1)
        if reference_cache:   # the reference's brushfire, so that weights compare at scorer precision
            self.map.upload_distances(g["brushfire"], p["laser_likelihood_max_dist"])
        else:
            self.map.build_distances(p["laser_likelihood_max_dist"])
        self.map.update_free_space(0.0)
        self.pf = cabi.ParticleFilter(ctx, p["min_particles"], p["max_particles"], p["recovery_alpha_slow"],
                                      p["recovery_alpha_fast"], 95.0, systematic=p["resample_model_type"] == "systematic", pop_err=p["kld_err"], pop_z=p["kld_z"])
        self.pf.set_random_pose_source(self.map)
        self.pf.set_exact_terms(exact_terms)
        self.model = cabi.planar_model(model_type=1, max_beams=p["laser_max_beams"], z_hit=p["laser_z_hit"],
                                       z_rand=p["laser_z_rand"], sigma_hit=p["laser_sigma_hit"],
                                       scanner_pose=(MOUNT[0], MOUNT[1], 0.0))

    def close(self):
        self.pf.close()
        self.map.close()

    def init(self, poses, state):
        self.pf.init_with_poses(poses)
        self.pf.drand48_state = int(state)

    def convert_scan(self, r, stats, p):
        return self.cabi.scan_from_laserscan(r, SCAN["range_min"], SCAN["range_max"], stats[0], stats[1],
                                             p["laser_min_range"], p["laser_max_range"])

    def action(self, alpha, pose, delta):
        self.pf.update_action(0, alpha, pose, delta, delta)

    def sensor(self, ranges, angles, range_max):
        self.pf.update_sensor(self.map, self.model, ranges, angles, range_max)

    def resample(self):
        self.pf.update_resample()

    def samples(self):
        return self.pf.samples()

    def state(self):
        return self.pf.drand48_state

    def readout(self):
        w, pose, idx = self.pf.max_weight_pose()
        return w, pose, self.pf.pose_covariance()

    def labels(self):
        return self.pf.last.leaf_count, self.pf.last.cluster_count, self.pf.sample_clusters()


class CudaBackend3D(CudaBackend2D):
    def __init__(self, cabi, ctx, g, p=PARAMS_3D, reference_cache=True):
        self.cabi = cabi
        self.map = cabi.Map3D(ctx, 0.1, g["keys"], p["laser_likelihood_max_dist"])
        if reference_cache:
            self.map.upload_distances(g["field"])
        else:
            self.map.build_distances()
        self.pf = cabi.ParticleFilter(ctx, p["min_particles"], p["max_particles"], p["recovery_alpha_slow"],
                                      p["recovery_alpha_fast"], 95.0, systematic=p["resample_model_type"] == "systematic", pop_err=p["kld_err"], pop_z=p["kld_z"])
        self.pf.set_random_pose_source(self.map)
        q = (0.0, 0.0, float(np.sin(0.5 * CLOUD_MOUNT[3])), float(np.cos(0.5 * CLOUD_MOUNT[3])))
        self.model = cabi.cloud_model(model_type=1, max_beams=p["laser_max_beams"], z_hit=p["laser_z_hit"],
                                      z_rand=p["laser_z_rand"], sigma_hit=p["laser_sigma_hit"],
                                      gompertz=(p["laser_gompertz_a"], p["laser_gompertz_b"], p["laser_gompertz_c"],
                                                p["laser_gompertz_input_shift"], p["laser_gompertz_input_scale"],
                                                p["laser_gompertz_output_shift"]),
                                      mount_translation=CLOUD_MOUNT[:3], mount_rotation=q)

    def subsample(self, xyz, p):
        return self.cabi.cloud_subsample(xyz, p["laser_max_beams"])

    def sensor(self, points):
        self.pf.update_sensor_cloud(self.map, self.model, points)


# ------------------------------------------------------------------------------------------------------------------------
# The same scenario through a NODE (either library): oracle/_ref/libamcl_ref.so = the reference's node on the reference's
# classes, oracle/_ref/libamcl_node_gpu.so = the reference's node on the host mirror + CUDA library. Inputs come from the
# golden run, so nothing here needs the reference checkout.
def run_node_2d(O, lib, g, seed=21, params=PARAMS_2D, reference_cache=False):
    out = {}
    lib.ref_srand48(seed)
    n = O.RefNode(params, map_type=2, lib=lib)
    assert n.publish_map(g["data"], RES_MSG, ORIGIN) == 1
    if reference_cache:
        n.install_cache_2d(g["brushfire"], params["laser_likelihood_max_dist"])
    out["free"] = n.free_space()
    out["init_samples"], out["init_state"] = n.samples(), n.drand48_state()
    n.set_tf("base_link", "laser", MOUNT[:3], O.yaw_quat(MOUNT[3]))
    for k in range(N_STEPS):
        x, y, yaw = truth_pose(k)
        n.set_odom(x, y, yaw)
        assert n.publish_scan(g["s%d_ranges" % k], SCAN["angle_min"], SCAN["angle_increment"], SCAN["range_min"],
                              SCAN["range_max"], stamp=1.0 + k) == 1
        _read_step(n, out, k)
        sd = n.scan_data()
        out["s%d_scan_ranges" % k], out["s%d_scan_angles" % k], out["s%d_scan_max" % k] = sd
    n.close()
    return out


def run_node_2d_mn(O, lib, g, reference_cache=False):
    return run_node_2d(O, lib, g, seed=44, params=PARAMS_2D_MN, reference_cache=reference_cache)


def run_node_3d(O, lib, g, seed=33, params=PARAMS_3D, reference_cache=False):
    out = {}
    lib.ref_srand48(seed)
    n = O.RefNode(params, map_type=3, lib=lib)
    assert n.publish_octomap(0.1, g["keys"]) == 1
    if reference_cache:
        n.install_cache_3d(g["field"])
    out["free_count"] = len(n.free_space())
    out["init_samples"], out["init_state"] = n.samples(), n.drand48_state()
    n.set_tf("base_link", "lidar", CLOUD_MOUNT[:3], O.yaw_quat(CLOUD_MOUNT[3]))
    for k in range(N_STEPS):
        x, y, yaw = truth_pose(k)
        n.set_odom(x, y, yaw)
        assert n.publish_cloud(g["s%d_cloud" % k], stamp=1.0 + k) == 1
        _read_step(n, out, k)
        out["s%d_points" % k] = n.cloud_data()
    n.close()
    return out


def _read_step(n, out, k):
    out["s%d_samples" % k] = n.samples()
    out["s%d_state" % k] = n.drand48_state()
    msg = n.last_pose()
    out["s%d_pose_count" % k] = msg["count"] if msg else 0
    if msg:
        out["s%d_pose" % k] = np.concatenate([msg["position"], msg["orientation"]])
        out["s%d_cov" % k] = msg["covariance"]
    out["s%d_resample_count" % k] = n.state()["resample_count"]
    info = n.filter_info()
    out["s%d_leaf" % k], out["s%d_clusters" % k] = info["leaf_count"], info["cluster_count"]
    out["s%d_averages" % k] = np.array([info["w_slow"], info["w_fast"]])
    out["s%d_converged" % k] = int(info["converged"])


# ------------------------------------------------------------------------------------------------------------------------
# A second scenario for the two node builds, exercising the node's other entry points live (no golden: on the GPU box both
# libraries are present and run the same inputs in the same process): likelihood_field_prob with beam skipping, the
# global_localization service, an /initialpose message and a dynamic reconfigure that swaps the filter and the model.
PARAMS_EVENTS = dict(PARAMS_2D, laser_model_type="likelihood_field_prob", do_beamskip=1, beam_skip_distance=0.5,
                     beam_skip_threshold=0.3, min_particles=200, max_particles=1200, resample_interval=1,
                     recovery_alpha_slow=0.001, recovery_alpha_fast=0.1)


def run_events_2d(O, lib, g, seed=77, cache=None):
    """g: the node_2d golden (map + scans are reused as inputs); returns a list of per-scan records"""
    rec = []
    lib.ref_srand48(seed)
    n = O.RefNode(PARAMS_EVENTS, map_type=2, lib=lib)
    assert n.publish_map(g["data"], RES_MSG, ORIGIN) == 1
    if cache is not None:
        n.install_cache_2d(cache, PARAMS_EVENTS["laser_likelihood_max_dist"])
    n.set_tf("base_link", "laser", MOUNT[:3], O.yaw_quat(MOUNT[3]))
    stamp = [1.0]

    def scan(k, label):
        x, y, yaw = truth_pose(k)
        n.set_odom(x, y, yaw)
        assert n.publish_scan(g["s%d_ranges" % k], SCAN["angle_min"], SCAN["angle_increment"], SCAN["range_min"],
                              SCAN["range_max"], stamp=stamp[0]) == 1
        stamp[0] += 1.0
        msg, st, info = n.last_pose(), n.state(), n.filter_info()
        rec.append(dict(label=label, samples=n.samples(), state=n.drand48_state(), pose_count=msg["count"] if msg else 0,
                        pose=np.concatenate([msg["position"], msg["orientation"]]) if msg else None,
                        resample_count=st["resample_count"], gl_active=st["global_localization_active"],
                        converged=info["converged"], clusters=info["cluster_count"]))

    scan(0, "start")
    scan(1, "tracking")
    assert n.global_localization() == 1                       # Node::globalLocalizationCallback -> initWithPoseFn
    scan(2, "global 1")
    scan(3, "global 2")
    cov = np.zeros(36)
    cov[0], cov[7], cov[35] = 0.04, 0.09, 0.01
    tx, ty, tyaw = truth_pose(3)
    assert n.publish_initial_pose(tx + 0.05, ty - 0.05, tyaw + 0.02, cov, stamp=stamp[0]) == 1   # -> initWithGaussian
    scan(4, "initialpose 1")
    scan(5, "initialpose 2")
    assert n.reconfigure(dict(laser_model_type="likelihood_field", laser_max_beams=40, max_particles=900,
                              update_min_d=0.1)) == 0       # Node::reconfigureCB: new filter, new model
    if cache is not None:
        pass                                                  # the map object (and its cache) survives a reconfigure
    scan(4, "reconfigured 1")
    scan(5, "reconfigured 2")
    n.close()
    return rec


PARAMS_EVENTS_3D = dict(PARAMS_3D, wait_for_occupancy_map=1, laser_model_type="likelihood_field", resample_interval=1,
                        min_particles=150, max_particles=600)


def run_events_3d(O, lib, g, seed=88, cache=None):
    """3-D: the OctoMap waits for the occupancy map's bounds (Node3D::occupancyMapMsgReceived -> setMapBounds, node_3d.cpp:
    178-197), clouds, the global_localization service over the cropped cell rectangle, clouds again. g: the node_3d golden."""
    rec = []
    lib.ref_srand48(seed)
    n = O.RefNode(PARAMS_EVENTS_3D, map_type=3, lib=lib)
    assert n.publish_octomap(0.1, g["keys"]) == 1
    rec.append(dict(label="octomap only", free=len(n.free_space())))
    assert n.publish_map(np.zeros((50, 60), np.int8), 0.1, (0.0, 0.0)) >= 1     # 6 m x 5 m bounds
    rec.append(dict(label="bounds", free=len(n.free_space())))
    if lib is O.ref_lib():
        rec[-1]["field"] = n.map3d().dense()      # the reference's cache over the cropped volume
    if cache is not None:
        n.install_cache_3d(cache)
    n.set_tf("base_link", "lidar", CLOUD_MOUNT[:3], O.yaw_quat(CLOUD_MOUNT[3]))
    stamp = [1.0]

    def cloud(k, label):
        x, y, yaw = truth_pose(k)
        n.set_odom(x, y, yaw)
        assert n.publish_cloud(g["s%d_cloud" % k], stamp=stamp[0]) == 1
        stamp[0] += 1.0
        msg, st = n.last_pose(), n.state()
        rec.append(dict(label=label, samples=n.samples(), state=n.drand48_state(), pose_count=msg["count"] if msg else 0,
                        pose=np.concatenate([msg["position"], msg["orientation"]]) if msg else None,
                        resample_count=st["resample_count"], gl_active=st["global_localization_active"]))

    cloud(0, "start")
    cloud(1, "tracking")
    assert n.global_localization() == 1
    cloud(2, "global 1")
    cloud(3, "global 2")
    n.close()
    return rec
