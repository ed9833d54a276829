"""oracle/oracle.py -- TEST INFRASTRUCTURE ONLY.

ctypes bindings for the two CPU checkers:

* ``port``  -- oracle/liboracle.so, the plain-C restatement (amcl_oracle.c)
* ``ref``   -- oracle/_ref/libamcl_ref.so, the UNMODIFIED reference translation units behind
               ref_harness.cpp (only exists where `make -C oracle ref` ran with /root/reference present;
               the built file travels to the GPU box)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libamcl_ref.so")

c_double_p = C.POINTER(C.c_double)
c_float_p = C.POINTER(C.c_float)
c_int_p = C.POINTER(C.c_int)


def build(ref=True):
    """Compile the restatement and, when the reference checkout is present, the reference itself."""
    subprocess.check_call(["make", "-s", "-C", HERE, "port"])
    if ref and os.path.isdir("/root/reference/src/amcl"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])
        # the reference's node on top of the host mirror + CUDA library (links libbamcl_cuda.so: built once that exists)
        if os.path.exists(os.path.join(HERE, "..", "badger_amcl_b200", "libbamcl_cuda.so")):
            subprocess.check_call(["make", "-s", "-C", HERE, "node_gpu"])


def _ptr(a, ty=None):
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)


class Map2DStruct(C.Structure):
    _fields_ = [("size_x", C.c_int), ("size_y", C.c_int), ("resolution", C.c_double), ("origin_x", C.c_float),
                ("origin_y", C.c_float), ("cells", C.c_void_p), ("distances", C.c_void_p), ("max_distance", C.c_double)]


class PlanarModel(C.Structure):
    _fields_ = [("model_type", C.c_int), ("max_beams", C.c_int), ("z_hit", C.c_double), ("z_short", C.c_double),
                ("z_max", C.c_double), ("z_rand", C.c_double), ("sigma_hit", C.c_double), ("lambda_short", C.c_double),
                ("gompertz_a", C.c_double), ("gompertz_b", C.c_double), ("gompertz_c", C.c_double),
                ("input_shift", C.c_double), ("input_scale", C.c_double), ("output_shift", C.c_double),
                ("do_beamskip", C.c_int), ("beam_skip_distance", C.c_double), ("beam_skip_threshold", C.c_double),
                ("beam_skip_error_threshold", C.c_double), ("off_map_factor", C.c_double),
                ("non_free_space_factor", C.c_double), ("non_free_space_radius", C.c_double),
                ("scanner_pose", C.c_double * 3)]


class CloudModel(C.Structure):
    _fields_ = [("model_type", C.c_int), ("max_beams", C.c_int), ("z_hit", C.c_double), ("z_rand", C.c_double),
                ("sigma_hit", C.c_double), ("gompertz_a", C.c_double), ("gompertz_b", C.c_double),
                ("gompertz_c", C.c_double), ("input_shift", C.c_double), ("input_scale", C.c_double),
                ("output_shift", C.c_double), ("off_map_factor", C.c_double), ("mount_translation", C.c_double * 3),
                ("mount_rotation", C.c_double * 4)]


class Rng(C.Structure):
    _fields_ = [("x", C.c_uint64)]


MODEL_BEAM, MODEL_LF, MODEL_LF_PROB, MODEL_LF_GOMPERTZ = 0, 1, 2, 3


def planar_model(model_type=MODEL_LF, max_beams=60, z_hit=0.95, z_short=0.1, z_max=0.05, z_rand=0.05, sigma_hit=0.2,
                 lambda_short=0.1, gompertz=(1.0, 1.0, 1.0, 0.0, 1.0, 0.0), do_beamskip=False, beam_skip_distance=0.5,
                 beam_skip_threshold=0.3, beam_skip_error_threshold=0.9, off_map_factor=1.0, non_free_space_factor=1.0,
                 non_free_space_radius=0.0, scanner_pose=(0.0, 0.0, 0.0)):
    """Defaults are the node defaults (node_2d.cpp:49-76)."""
    p = PlanarModel()
    p.model_type, p.max_beams = model_type, max_beams
    p.z_hit, p.z_short, p.z_max, p.z_rand, p.sigma_hit, p.lambda_short = z_hit, z_short, z_max, z_rand, sigma_hit, lambda_short
    (p.gompertz_a, p.gompertz_b, p.gompertz_c, p.input_shift, p.input_scale, p.output_shift) = gompertz
    p.do_beamskip = int(do_beamskip)
    p.beam_skip_distance, p.beam_skip_threshold, p.beam_skip_error_threshold = (
        beam_skip_distance, beam_skip_threshold, beam_skip_error_threshold)
    p.off_map_factor, p.non_free_space_factor, p.non_free_space_radius = (
        off_map_factor, non_free_space_factor, non_free_space_radius)
    p.scanner_pose[:] = scanner_pose
    return p


_port = None
_ref = None


def port_lib():
    global _port
    if _port is None:
        if not os.path.exists(PORT_SO):
            build(ref=False)
        L = C.CDLL(PORT_SO)
        D, I, L_, V, P = C.c_double, C.c_int, C.c_long, C.c_void_p, C.c_void_p
        sig = {
            "orc_pf_update_action": (None, [P, P, I, P, P, P, P]),
            "orc_map2d_free_space": (L_, [P, D, P]), "orc_uniform_poses": (None, [P, P, L_, P, D, D, L_, P]),
            "orc_pf_set_uniform_pose_source": (None, [P, P, P, L_, D, D]),
            "orc_srand48": (None, [P, L_]), "orc_drand48": (D, [P]), "orc_gaussian_draw": (D, [P, D]),
            "orc_map2d_world_to_map": (None, [P, D, D, P, P]), "orc_map2d_map_to_world": (None, [P, I, I, P, P]),
            "orc_map2d_is_valid": (I, [P, I, I]), "orc_map2d_distance": (C.c_float, [P, I, I]),
            "orc_map2d_calc_range": (D, [P, D, D, D, D, P]), "orc_map2d_update_distances": (None, [P, D, P]),
            "orc_map2d_exact_edt": (None, [P, D, P]),
            "orc_planar_state_create": (V, []), "orc_planar_state_destroy": (None, [P]),
            "orc_normalize_angle": (D, [D]), "orc_coord_add": (None, [P, P, P]), "orc_apply_gompertz": (D, [P, D]),
            "orc_planar_apply": (D, [P, P, P, P, P, I, D, P, L_, I]), "orc_planar_recalc_weight": (D, [P, P, P, L_]),
            "orc_kdtree_create": (V, []), "orc_kdtree_destroy": (None, [P]), "orc_kdtree_clear": (None, [P]),
            "orc_kdtree_insert": (None, [P, P, D]), "orc_kdtree_cluster": (None, [P]),
            "orc_kdtree_get_cluster": (I, [P, P]), "orc_kdtree_leaf_count": (I, [P]),
            "orc_kdtree_node_count": (L_, [P]), "orc_kdtree_get_nodes": (None, [P, P, P, P, P]),
            "orc_pose_key": (None, [P, P]),
            "orc_pf_create": (V, [I, I, D, D, D]), "orc_pf_destroy": (None, [P]),
            "orc_pf_set_resample_model": (None, [P, I]), "orc_pf_set_population_params": (None, [P, D, D]),
            "orc_pf_set_decay_rates": (None, [P, D, D]), "orc_pf_set_random_poses": (None, [P, P, L_]),
            "orc_pf_random_poses_used": (L_, [P]), "orc_pf_init_with_poses": (None, [P, P, L_]),
            "orc_pf_sample_count": (I, [P]), "orc_pf_get_samples": (None, [P, P]),
            "orc_pf_set_samples": (None, [P, P, L_]), "orc_pf_normalise": (None, [P, D]),
            "orc_pf_update_sensor_planar": (I, [P, P, P, P, P, P, I, D]), "orc_pf_update_resample": (None, [P, P]),
            "orc_pf_get_resample_sources": (None, [P, P]), "orc_pf_resample_limit": (I, [P, I]),
            "orc_pf_get_averages": (None, [P, P, P]), "orc_pf_set_averages": (None, [P, D, D]),
            "orc_pf_leaf_count": (I, [P]), "orc_pf_node_count": (L_, [P]), "orc_pf_is_converged": (I, [P]),
            "orc_pf_cluster_count": (I, [P]), "orc_pf_get_sample_clusters": (None, [P, P]),
            "orc_pf_get_cluster_stats": (None, [P, P]), "orc_pf_get_set_stats": (None, [P, P]),
            "orc_pf_cdf": (None, [P, P]),
            "orc_map3d_create": (V, [D, P, L_, D]), "orc_map3d_destroy": (None, [P]),
            "orc_map3d_set_bounds": (None, [P, P, P]), "orc_map3d_update_distances": (None, [P]),
            "orc_map3d_get_bounds": (None, [P, P, P]), "orc_map3d_distance": (D, [P, I, I, I]),
            "orc_map3d_get_dense": (None, [P, P]), "orc_map3d_set_dense": (None, [P, P]),
            "orc_map3d_exact_edt": (None, [P, P]),
            "orc_map3d_world_to_map": (None, [P, D, P]), "orc_map3d_is_pose_valid": (I, [P, I, I]),
            "orc_cloud_transform": (None, [P, P, P, L_, P]), "orc_cloud_apply": (D, [P, P, P, L_, P, L_]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _port = L
    return _port


def have_ref():
    return os.path.exists(REF_SO)


NODE_GPU_SO = os.path.join(HERE, "_ref", "libamcl_node_gpu.so")
_node_gpu = None


def _node_sigs(wrappers):
    """the node-driving surface of oracle/node_harness.inc (both libraries); wrappers: the reference-only handle getters"""
    D, I, L_, V, P, F = C.c_double, C.c_int, C.c_long, C.c_void_p, C.c_void_p, C.c_float
    sig = {
            "ref_ros_reset": (None, []), "ref_ros_set_param_num": (None, [C.c_char_p, D]),
            "ref_ros_set_param_str": (None, [C.c_char_p, C.c_char_p]), "ref_ros_set_time": (None, [D]),
            "ref_ros_set_tf": (None, [C.c_char_p, C.c_char_p, P, P]), "ref_ros_clear_tf": (None, [C.c_char_p, C.c_char_p]),
            "ref_ros_message_count": (L_, [C.c_char_p]), "ref_ros_call_service": (I, [C.c_char_p]),
            "ref_ros_fire_timers": (I, []), "ref_ros_reconfigure": (I, []),
            "ref_node_create": (V, []), "ref_node_destroy": (None, [P]),
            "ref_ros_publish_map": (I, [P, I, I, F, D, D]), "ref_ros_publish_octomap": (I, [D, P, L_, I]),
            "ref_ros_publish_scan": (I, [C.c_char_p, D, F, F, F, F, F, P, I]),
            "ref_ros_publish_cloud": (I, [C.c_char_p, D, P, L_]), "ref_ros_publish_odom": (I, [D, D, D, D]),
            "ref_ros_publish_initial_pose": (I, [C.c_char_p, D, D, D, D, P]),
            "ref_node_pf": (V, [P]), "ref_node_map2d": (V, [P]), "ref_node_map3d": (V, [P]),
            "ref_map2d_geometry": (None, [P, P, P, P]),
            "ref_node_free_space_count": (L_, [P]), "ref_node_free_space": (None, [P, P]),
            "ref_node_uniform_poses": (None, [P, L_, P]), "ref_node_scan_data": (I, [P, P, P, P, I]),
            "ref_node_angle_stats": (I, [P, C.c_char_p, D, F, F, P]), "ref_node_cloud_data": (L_, [P, P, L_]),
            "ref_node_max_weight_pose": (I, [P, P, P]), "ref_ros_last_pose": (L_, [P]),
            "ref_ros_last_particlecloud": (L_, [P, L_]), "ref_node_latest_tf": (I, [P, P]),
            "ref_node_state": (None, [P, P]), "ref_node_odom_pose": (I, [P, D, P]),
        "ref_node_sample_count": (I, [P]), "ref_node_samples": (None, [P, P]), "ref_node_filter_info": (None, [P, P]),
        "ref_node_sample_clusters": (None, [P, P]), "ref_node_get_seed48": (None, [P]),
    }
    if not wrappers:
        for k in ("ref_node_pf", "ref_node_map2d", "ref_node_map3d", "ref_map2d_geometry"):
            del sig[k]
    return sig


def have_node_gpu():
    return os.path.exists(NODE_GPU_SO)


def node_gpu_lib():
    """oracle/_ref/libamcl_node_gpu.so: the reference's node sources on top of the host mirror + CUDA library"""
    global _node_gpu
    if _node_gpu is None:
        if not have_node_gpu():
            raise RuntimeError("oracle/_ref/libamcl_node_gpu.so missing: run `make -C oracle node_gpu` where /root/reference exists")
        L = C.CDLL(NODE_GPU_SO)
        sig = _node_sigs(wrappers=False)
        sig["ref_srand48"] = (None, [C.c_long])
        sig["ref_node_last_device_error"] = (C.c_char_p, [])
        sig["ref_node_install_cache_2d"] = (C.c_int, [C.c_void_p, C.c_void_p, C.c_long, C.c_double])
        sig["ref_node_install_cache_3d"] = (C.c_int, [C.c_void_p, C.c_void_p, C.c_long])
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _node_gpu = L
    return _node_gpu


def ref_lib():
    global _ref
    if _ref is None:
        if not have_ref():
            raise RuntimeError("oracle/_ref/libamcl_ref.so missing: run `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(REF_SO)
        D, I, L_, V, P, F = C.c_double, C.c_int, C.c_long, C.c_void_p, C.c_void_p, C.c_float
        sig = {
            "ref_srand48": (None, [L_]), "ref_drand48": (D, []), "ref_seed48": (None, [P]), "ref_get_seed48": (None, [P]),
            "ref_gaussian_draw": (D, [D]),
            "ref_pf_set_uniform_pose_source": (None, [P, P, P, L_, D, D]),
            "ref_pf_update_action": (None, [P, I, P, P, P, P]),
            "ref_map2d_create": (V, [D, I, I, F, F]), "ref_map2d_destroy": (None, [P]),
            "ref_map2d_set_cells": (None, [P, P, L_]), "ref_map2d_get_cell": (I, [P, I, I]),
            "ref_map2d_update_distances": (None, [P, D]), "ref_map2d_get_distances": (None, [P, P, L_]),
            "ref_map2d_set_distances": (None, [P, P, L_, D]), "ref_map2d_distance": (F, [P, I, I]),
            "ref_map2d_calc_range": (D, [P, D, D, D, D]), "ref_map2d_calc_ranges": (None, [P, P, P, P, D, L_, P]),
            "ref_map2d_world_to_map": (None, [P, D, D, P]), "ref_map2d_map_to_world": (None, [P, I, I, P]),
            "ref_map2d_is_valid": (I, [P, I, I]),
            "ref_map3d_create": (V, [D, P, L_, D]), "ref_map3d_destroy": (None, [P]),
            "ref_map3d_set_bounds": (None, [P, P, P]), "ref_map3d_update_distances": (None, [P]),
            "ref_map3d_get_bounds": (None, [P, P, P]), "ref_map3d_distance": (D, [P, I, I, I]),
            "ref_map3d_get_dense": (None, [P, P]), "ref_map3d_world_to_map": (None, [P, P, P]),
            "ref_map3d_map_to_world": (None, [P, P, P]),
            "ref_scanner3d_create": (V, [P, I]), "ref_scanner3d_destroy": (None, [P]),
            "ref_scanner3d_model": (None, [P, I, D, D, D, D, D, D, D, D, D]),
            "ref_scanner3d_set_map_factors": (None, [P, D, D, D]), "ref_scanner3d_set_mount": (None, [P, P, P]),
            "ref_scanner3d_apply": (D, [P, P, L_, P, L_]),
            "ref_kdtree_create": (V, []), "ref_kdtree_destroy": (None, [P]), "ref_kdtree_clear": (None, [P]),
            "ref_kdtree_insert": (None, [P, D, D, D, D]), "ref_kdtree_insert_many": (None, [P, P, L_, D, P]),
            "ref_kdtree_cluster": (None, [P]), "ref_kdtree_get_cluster": (I, [P, D, D, D]),
            "ref_kdtree_leaf_count": (I, [P]), "ref_kdtree_node_count": (L_, [P]),
            "ref_kdtree_get_nodes": (None, [P, P, P, P, P]),
            "ref_scanner2d_create": (V, [P, I]), "ref_scanner2d_destroy": (None, [P]),
            "ref_scanner2d_set_pose": (None, [P, D, D, D]), "ref_scanner2d_set_map_factors": (None, [P, D, D, D]),
            "ref_scanner2d_model_beam": (None, [P, D, D, D, D, D, D]),
            "ref_scanner2d_model_lf": (None, [P, D, D, D, D]),
            "ref_scanner2d_model_lf_prob": (None, [P, D, D, D, D, I, D, D, D]),
            "ref_scanner2d_model_lf_gompertz": (None, [P, D, D, D, D, D, D, D, D, D, D]),
            "ref_scanner2d_model_lf_nocache": (None, [P, D, D, D]),
            "ref_scanner2d_model_lf_prob_nocache": (None, [P, D, D, D, I, D, D, D]),
            "ref_scanner2d_model_lf_gompertz_nocache": (None, [P, D, D, D, D, D, D, D, D, D]),
            "ref_scanner2d_apply": (D, [P, P, P, I, D, P, L_, I]),
            "ref_pf_create": (V, [I, I, D, D, D]), "ref_pf_destroy": (None, [P]),
            "ref_pf_set_resample_model": (None, [P, I]), "ref_pf_set_population_params": (None, [P, D, D]),
            "ref_pf_set_decay_rates": (None, [P, D, D]), "ref_pf_set_random_poses": (None, [P, P, L_]),
            "ref_pf_random_poses_used": (L_, [P]), "ref_pf_init_with_poses": (None, [P, P, L_]),
            "ref_pf_sample_count": (I, [P]), "ref_pf_get_samples": (None, [P, P]),
            "ref_pf_set_samples": (None, [P, P, L_]),
            "ref_pf_update_sensor_planar": (I, [P, P, P, P, I, D]), "ref_pf_update_resample": (None, [P]),
            "ref_pf_get_averages": (None, [P, P, P]), "ref_pf_set_averages": (None, [P, D, D]),
            "ref_pf_leaf_count": (I, [P]), "ref_pf_node_count": (L_, [P]), "ref_pf_is_converged": (I, [P]),
            "ref_pf_cluster_count": (I, [P]), "ref_pf_resample_limit": (I, [P, I]),
            "ref_pf_get_sample_clusters": (None, [P, P]), "ref_pf_get_cluster_stats": (None, [P, P]),
            "ref_pf_get_set_stats": (None, [P, P]),
            "ref_gaussian_sample": (None, [P, P, L_, P]), "ref_pf_init_gaussian": (None, [P, P, P]),
        }
        sig.update(_node_sigs(wrappers=True))
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _ref = L
    return _ref


# ----------------------------------------------------------------------------------------------
# thin object wrappers (same method names on both back-ends so tests can be parametrised)
# ----------------------------------------------------------------------------------------------
class PortMap2D:
    """orc_map2d: caller-owned arrays kept alive on the Python side."""

    def __init__(self, cells, resolution, origin=(0.0, 0.0)):
        self.L = port_lib()
        self.cells = np.ascontiguousarray(cells, dtype=np.int8)  # shape (size_y, size_x): index = i + j*size_x
        self.size_y, self.size_x = self.cells.shape
        self.resolution = float(resolution)
        self.dist = None
        self.s = Map2DStruct(self.size_x, self.size_y, self.resolution, origin[0], origin[1], _ptr(self.cells), None, 0.0)

    def update_distances(self, max_dist):
        self.dist = np.empty((self.size_y, self.size_x), np.float32)
        self.L.orc_map2d_update_distances(C.byref(self.s), max_dist, _ptr(self.dist))
        self.s.distances, self.s.max_distance = _ptr(self.dist), max_dist
        return self.dist

    def free_space(self, radius):
        """Node2D::updateFreeSpaceIndices (node_2d.cpp:320-338) -> (n, 2) int32 of (i, j)"""
        n = self.L.orc_map2d_free_space(C.byref(self.s), radius, None)
        out = np.empty((n, 2), np.int32)
        self.L.orc_map2d_free_space(C.byref(self.s), radius, _ptr(out))
        return out

    def uniform_poses(self, free_ij, rng, n, threshold=0.0, multiplier=0.0):
        """n calls of Node::uniformPoseGenerator (node.cpp:847-868) drawing from rng (an oracle.Rng)"""
        free_ij = np.ascontiguousarray(free_ij, np.int32)
        out = np.empty((n, 3))
        self.L.orc_uniform_poses(C.byref(self.s), _ptr(free_ij), len(free_ij), C.byref(rng), threshold, multiplier, n,
                                 _ptr(out))
        return out

    def exact_edt(self, max_dist):
        out = np.empty((self.size_y, self.size_x), np.float32)
        self.L.orc_map2d_exact_edt(C.byref(self.s), max_dist, _ptr(out))
        return out

    def set_distances(self, dist, max_dist):
        self.dist = np.ascontiguousarray(dist, dtype=np.float32)
        self.s.distances, self.s.max_distance = _ptr(self.dist), max_dist

    def calc_range(self, ox, oy, oa, max_range, with_steps=False):
        steps = C.c_long(0)
        r = self.L.orc_map2d_calc_range(C.byref(self.s), ox, oy, oa, max_range, C.byref(steps))
        return (r, steps.value) if with_steps else r

    def calc_ranges(self, ox, oy, oa, max_range):
        out = np.empty(len(ox))
        steps = 0
        st = C.c_long(0)
        for k in range(len(ox)):
            out[k] = self.L.orc_map2d_calc_range(C.byref(self.s), ox[k], oy[k], oa[k], max_range, C.byref(st))
            steps += st.value
        self.last_steps = steps
        return out

    def world_to_map(self, x, y):
        i, j = C.c_int(), C.c_int()
        self.L.orc_map2d_world_to_map(C.byref(self.s), x, y, C.byref(i), C.byref(j))
        return i.value, j.value

    def map_to_world(self, i, j):
        x, y = C.c_double(), C.c_double()
        self.L.orc_map2d_map_to_world(C.byref(self.s), i, j, C.byref(x), C.byref(y))
        return x.value, y.value

    def is_valid(self, i, j):
        return bool(self.L.orc_map2d_is_valid(C.byref(self.s), i, j))

    def apply(self, model, ranges, angles, range_max, samples, converged=0, state=None):
        """orc_planar_apply on a copy of `samples` (n x 4 AoS); returns (total, weights)."""
        s = np.array(samples, dtype=np.float64, order="C")
        ranges = np.ascontiguousarray(ranges, np.float64)
        angles = np.ascontiguousarray(angles, np.float64)
        own = state is None
        st = self.L.orc_planar_state_create() if own else state
        total = self.L.orc_planar_apply(C.byref(self.s), C.byref(model), st, _ptr(ranges), _ptr(angles), len(ranges),
                                        range_max, _ptr(s), len(s), converged)
        if own:
            self.L.orc_planar_state_destroy(st)
        return total, s[:, 3].copy()


class RefMap2D:
    def __init__(self, cells, resolution, origin=(0.0, 0.0)):
        self.L = ref_lib()
        self.cells = np.ascontiguousarray(cells, dtype=np.int8)
        self.size_y, self.size_x = self.cells.shape
        self.resolution = float(resolution)
        self.h = self.L.ref_map2d_create(resolution, self.size_x, self.size_y, origin[0], origin[1])
        self.L.ref_map2d_set_cells(self.h, _ptr(self.cells), self.cells.size)

    def __del__(self):
        try:
            self.L.ref_map2d_destroy(self.h)
        except Exception:
            pass

    def update_distances(self, max_dist):
        self.L.ref_map2d_update_distances(self.h, max_dist)
        return self.get_distances()

    def get_distances(self):
        out = np.empty((self.size_y, self.size_x), np.float32)
        self.L.ref_map2d_get_distances(self.h, _ptr(out), out.size)
        return out

    def set_distances(self, dist, max_dist):
        d = np.ascontiguousarray(dist, np.float32)
        self.L.ref_map2d_set_distances(self.h, _ptr(d), d.size, max_dist)

    def calc_range(self, ox, oy, oa, max_range):
        return self.L.ref_map2d_calc_range(self.h, ox, oy, oa, max_range)

    def calc_ranges(self, ox, oy, oa, max_range):
        ox, oy, oa = (np.ascontiguousarray(a, np.float64) for a in (ox, oy, oa))
        out = np.empty(len(ox))
        self.L.ref_map2d_calc_ranges(self.h, _ptr(ox), _ptr(oy), _ptr(oa), max_range, len(ox), _ptr(out))
        return out

    def world_to_map(self, x, y):
        ij = (C.c_int * 2)()
        self.L.ref_map2d_world_to_map(self.h, x, y, ij)
        return ij[0], ij[1]

    def map_to_world(self, i, j):
        xy = (C.c_double * 2)()
        self.L.ref_map2d_map_to_world(self.h, i, j, xy)
        return xy[0], xy[1]

    def is_valid(self, i, j):
        return bool(self.L.ref_map2d_is_valid(self.h, i, j))


class RefScanner2D:
    """PlanarScanner of the reference configured from the same PlanarModel struct the port uses."""

    def __init__(self, ref_map, model):
        self.L = ref_lib()
        self.map = ref_map
        self.h = self.L.ref_scanner2d_create(ref_map.h, model.max_beams)
        self.configure(model)

    def __del__(self):
        try:
            self.L.ref_scanner2d_destroy(self.h)
        except Exception:
            pass

    def configure(self, m):
        L, h = self.L, self.h
        L.ref_scanner2d_set_pose(h, *m.scanner_pose)
        L.ref_scanner2d_set_map_factors(h, m.off_map_factor, m.non_free_space_factor, m.non_free_space_radius)
        if m.model_type == MODEL_BEAM:
            L.ref_scanner2d_model_beam(h, m.z_hit, m.z_short, m.z_max, m.z_rand, m.sigma_hit, m.lambda_short)
        elif m.model_type == MODEL_LF:
            L.ref_scanner2d_model_lf_nocache(h, m.z_hit, m.z_rand, m.sigma_hit)
        elif m.model_type == MODEL_LF_PROB:
            L.ref_scanner2d_model_lf_prob_nocache(h, m.z_hit, m.z_rand, m.sigma_hit, m.do_beamskip, m.beam_skip_distance,
                                                  m.beam_skip_threshold, m.beam_skip_error_threshold)
        else:
            L.ref_scanner2d_model_lf_gompertz_nocache(h, m.z_hit, m.z_rand, m.sigma_hit, m.gompertz_a, m.gompertz_b,
                                                      m.gompertz_c, m.input_shift, m.input_scale, m.output_shift)

    def apply(self, ranges, angles, range_max, samples, converged=0):
        s = np.array(samples, dtype=np.float64, order="C")
        ranges = np.ascontiguousarray(ranges, np.float64)
        angles = np.ascontiguousarray(angles, np.float64)
        total = self.L.ref_scanner2d_apply(self.h, _ptr(ranges), _ptr(angles), len(ranges), range_max, _ptr(s), len(s),
                                           converged)
        return total, s[:, 3].copy()


class _PFBase:
    def samples(self):
        out = np.empty((self.sample_count(), 4))
        self._get_samples(out)
        return out

    def sample_clusters(self):
        out = np.empty(self.sample_count(), np.int32)
        self._get_sample_clusters(out)
        return out

    def cluster_stats(self):
        out = np.empty((self.cluster_count(), 10))
        self._get_cluster_stats(out)
        return out

    def set_stats(self):
        out = np.empty(8)
        self._get_set_stats(out)
        return out

    def max_weight_pose(self):
        """Node2D::getMaxWeightPose (node_2d.cpp:588-617) restated over getClusterStats: the first cluster whose weight
        is strictly greater than every earlier one -> (weight, mean pose, index); index -1 where the reference would
        read hyps[-1]. Pinned against Node2D::getMaxWeightPose of the compiled node (tests/test_node_pins.py)."""
        st = self.cluster_stats()
        best, idx = 0.0, -1
        for c in range(len(st)):
            if st[c, 1] > best:
                best, idx = st[c, 1], c
        return best, (st[idx, 2:5].copy() if idx >= 0 else None), idx

    def pose_covariance(self):
        """Node::updatePose's packing of set->cov into the 6x6 message covariance (node.cpp:381-395)"""
        s = self.set_stats()
        cov = np.zeros((6, 6))
        cov[0, 0], cov[0, 1], cov[1, 0], cov[1, 1], cov[5, 5] = s[3], s[4], s[5], s[6], s[7]
        return cov


class PortPF(_PFBase):
    def __init__(self, min_samples, max_samples, alpha_slow=0.001, alpha_fast=0.1, conv_thresh=95.0, systematic=True,
                 pop_err=0.01, pop_z=0.99):
        self.L = port_lib()
        self.max_samples = max_samples
        self.h = self.L.orc_pf_create(min_samples, max_samples, alpha_slow, alpha_fast, conv_thresh)
        self.L.orc_pf_set_resample_model(self.h, int(systematic))
        self.L.orc_pf_set_population_params(self.h, pop_err, pop_z)
        self.state = self.L.orc_planar_state_create()
        self.rng = Rng(0)

    def __del__(self):
        try:
            self.L.orc_planar_state_destroy(self.state)
            self.L.orc_pf_destroy(self.h)
        except Exception:
            pass

    def srand48(self, seed):
        self.L.orc_srand48(C.byref(self.rng), seed)

    def set_random_poses(self, poses):
        p = np.ascontiguousarray(poses, np.float64)
        self.L.orc_pf_set_random_poses(self.h, _ptr(p), len(p))

    def init_with_poses(self, poses):
        p = np.ascontiguousarray(poses, np.float64)
        self.L.orc_pf_init_with_poses(self.h, _ptr(p), len(p))

    def set_uniform_pose_source(self, port_map, free_ij, threshold=0.0, multiplier=0.0):
        """random_pose_fn_ = Node::uniformPoseGenerator over the map's free-space list (node.cpp:262-264, 847-868)"""
        self._uni = (port_map, np.ascontiguousarray(free_ij, np.int32))  # kept alive: the C side stores the pointers
        self.L.orc_pf_set_uniform_pose_source(self.h, C.byref(port_map.s), _ptr(self._uni[1]), len(self._uni[1]),
                                              threshold, multiplier)

    def sample_count(self):
        return self.L.orc_pf_sample_count(self.h)

    def _get_samples(self, out):
        self.L.orc_pf_get_samples(self.h, _ptr(out))

    def set_samples(self, aos):
        a = np.ascontiguousarray(aos, np.float64)
        self.L.orc_pf_set_samples(self.h, _ptr(a), len(a))

    def update_sensor(self, port_map, model, ranges, angles, range_max):
        ranges = np.ascontiguousarray(ranges, np.float64)
        angles = np.ascontiguousarray(angles, np.float64)
        return self.L.orc_pf_update_sensor_planar(self.h, C.byref(port_map.s), C.byref(model), self.state, _ptr(ranges),
                                                  _ptr(angles), len(ranges), range_max)

    def update_resample(self):
        self.L.orc_pf_update_resample(self.h, C.byref(self.rng))

    def update_action(self, model_type, alpha, pose, delta, absolute_motion=(0.0, 0.0, 0.0)):
        """Odom::updateAction (odom.cpp:74-311) on the current set, drawing from this filter's drand48 stream"""
        v = [np.ascontiguousarray(x, np.float64) for x in (alpha, pose, delta, absolute_motion)]
        self.L.orc_pf_update_action(self.h, C.byref(self.rng), int(model_type), *[_ptr(x) for x in v])

    def resample_sources(self):
        out = np.empty(self.sample_count(), np.int32)
        self.L.orc_pf_get_resample_sources(self.h, _ptr(out))
        return out

    def resample_limit(self, k):
        return self.L.orc_pf_resample_limit(self.h, k)

    def averages(self):
        a, b = C.c_double(), C.c_double()
        self.L.orc_pf_get_averages(self.h, C.byref(a), C.byref(b))
        return a.value, b.value

    def set_averages(self, w_slow, w_fast):
        self.L.orc_pf_set_averages(self.h, w_slow, w_fast)

    def leaf_count(self):
        return self.L.orc_pf_leaf_count(self.h)

    def node_count(self):
        return self.L.orc_pf_node_count(self.h)

    def is_converged(self):
        return bool(self.L.orc_pf_is_converged(self.h))

    def cluster_count(self):
        return self.L.orc_pf_cluster_count(self.h)

    def _get_sample_clusters(self, out):
        self.L.orc_pf_get_sample_clusters(self.h, _ptr(out))

    def _get_cluster_stats(self, out):
        self.L.orc_pf_get_cluster_stats(self.h, _ptr(out))

    def _get_set_stats(self, out):
        self.L.orc_pf_get_set_stats(self.h, _ptr(out))

    def cdf(self):
        out = np.empty(self.sample_count() + 1)
        self.L.orc_pf_cdf(self.h, _ptr(out))
        return out


class RefPF(_PFBase):
    def __init__(self, min_samples, max_samples, alpha_slow=0.001, alpha_fast=0.1, conv_thresh=95.0, systematic=True,
                 pop_err=0.01, pop_z=0.99):
        self.L = ref_lib()
        self.max_samples = max_samples
        self.h = self.L.ref_pf_create(min_samples, max_samples, alpha_slow, alpha_fast, conv_thresh)
        self.L.ref_pf_set_resample_model(self.h, int(systematic))
        self.L.ref_pf_set_population_params(self.h, pop_err, pop_z)

    def __del__(self):
        try:
            self.L.ref_pf_destroy(self.h)
        except Exception:
            pass

    def srand48(self, seed):
        self.L.ref_srand48(seed)  # process-global glibc stream

    def set_random_poses(self, poses):
        p = np.ascontiguousarray(poses, np.float64)
        self.L.ref_pf_set_random_poses(self.h, _ptr(p), len(p))

    def init_with_poses(self, poses):
        p = np.ascontiguousarray(poses, np.float64)
        self.L.ref_pf_init_with_poses(self.h, _ptr(p), len(p))

    def init_gaussian(self, mean, cov):
        """ParticleFilter::initWithGaussian of the compiled reference (draws from the process-global drand48 stream)"""
        m, c = np.ascontiguousarray(mean, np.float64), np.ascontiguousarray(cov, np.float64)
        self.L.ref_pf_init_gaussian(self.h, _ptr(m), _ptr(c))

    def set_uniform_pose_source(self, ref_map, free_ij, threshold=0.0, multiplier=0.0):
        f = np.ascontiguousarray(free_ij, np.int32)
        self._uni = ref_map
        self.L.ref_pf_set_uniform_pose_source(self.h, ref_map.h, _ptr(f), len(f), threshold, multiplier)

    def sample_count(self):
        return self.L.ref_pf_sample_count(self.h)

    def _get_samples(self, out):
        self.L.ref_pf_get_samples(self.h, _ptr(out))

    def set_samples(self, aos):
        a = np.ascontiguousarray(aos, np.float64)
        self.L.ref_pf_set_samples(self.h, _ptr(a), len(a))

    def update_sensor(self, scanner, ranges, angles, range_max):
        ranges = np.ascontiguousarray(ranges, np.float64)
        angles = np.ascontiguousarray(angles, np.float64)
        return self.L.ref_pf_update_sensor_planar(self.h, scanner.h, _ptr(ranges), _ptr(angles), len(ranges), range_max)

    def update_resample(self):
        self.L.ref_pf_update_resample(self.h)

    def update_action(self, model_type, alpha, pose, delta, absolute_motion=(0.0, 0.0, 0.0)):
        v = [np.ascontiguousarray(x, np.float64) for x in (alpha, pose, delta, absolute_motion)]
        self.L.ref_pf_update_action(self.h, int(model_type), *[_ptr(x) for x in v])

    def resample_limit(self, k):
        return self.L.ref_pf_resample_limit(self.h, k)

    def averages(self):
        a, b = C.c_double(), C.c_double()
        self.L.ref_pf_get_averages(self.h, C.byref(a), C.byref(b))
        return a.value, b.value

    def set_averages(self, w_slow, w_fast):
        self.L.ref_pf_set_averages(self.h, w_slow, w_fast)

    def leaf_count(self):
        return self.L.ref_pf_leaf_count(self.h)

    def node_count(self):
        return self.L.ref_pf_node_count(self.h)

    def is_converged(self):
        return bool(self.L.ref_pf_is_converged(self.h))

    def cluster_count(self):
        return self.L.ref_pf_cluster_count(self.h)

    def _get_sample_clusters(self, out):
        self.L.ref_pf_get_sample_clusters(self.h, _ptr(out))

    def _get_cluster_stats(self, out):
        self.L.ref_pf_get_cluster_stats(self.h, _ptr(out))

    def _get_set_stats(self, out):
        self.L.ref_pf_get_set_stats(self.h, _ptr(out))


class PortKDTree:
    def __init__(self):
        self.L = port_lib()
        self.h = self.L.orc_kdtree_create()

    def __del__(self):
        try:
            self.L.orc_kdtree_destroy(self.h)
        except Exception:
            pass

    def clear(self):
        self.L.orc_kdtree_clear(self.h)

    def insert(self, pose, value=0.0):
        p = (C.c_double * 3)(*pose)
        self.L.orc_kdtree_insert(self.h, p, value)

    def insert_many(self, poses, value=0.0, leaf_counts=False):
        poses = np.ascontiguousarray(poses, np.float64)
        lc = np.empty(len(poses), np.int32) if leaf_counts else None
        for k in range(len(poses)):
            self.L.orc_kdtree_insert(self.h, _ptr(poses[k]), value)
            if leaf_counts:
                lc[k] = self.L.orc_kdtree_leaf_count(self.h)
        return lc

    def cluster(self):
        self.L.orc_kdtree_cluster(self.h)

    def get_cluster(self, pose):
        p = (C.c_double * 3)(*pose)
        return self.L.orc_kdtree_get_cluster(self.h, p)

    def leaf_count(self):
        return self.L.orc_kdtree_leaf_count(self.h)

    def node_count(self):
        return self.L.orc_kdtree_node_count(self.h)

    def nodes(self):
        n = self.node_count()
        keys, cl, pv, val = np.empty((n, 3), np.int32), np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n)
        self.L.orc_kdtree_get_nodes(self.h, _ptr(keys), _ptr(cl), _ptr(pv), _ptr(val))
        return keys, cl, pv, val


class RefKDTree:
    def __init__(self):
        self.L = ref_lib()
        self.h = self.L.ref_kdtree_create()

    def __del__(self):
        try:
            self.L.ref_kdtree_destroy(self.h)
        except Exception:
            pass

    def clear(self):
        self.L.ref_kdtree_clear(self.h)

    def insert(self, pose, value=0.0):
        self.L.ref_kdtree_insert(self.h, pose[0], pose[1], pose[2], value)

    def insert_many(self, poses, value=0.0, leaf_counts=False):
        poses = np.ascontiguousarray(poses, np.float64)
        lc = np.empty(len(poses), np.int32) if leaf_counts else None
        self.L.ref_kdtree_insert_many(self.h, _ptr(poses), len(poses), value, _ptr(lc))
        return lc

    def cluster(self):
        self.L.ref_kdtree_cluster(self.h)

    def get_cluster(self, pose):
        return self.L.ref_kdtree_get_cluster(self.h, pose[0], pose[1], pose[2])

    def leaf_count(self):
        return self.L.ref_kdtree_leaf_count(self.h)

    def node_count(self):
        return self.L.ref_kdtree_node_count(self.h)

    def nodes(self):
        n = self.node_count()
        keys, cl, pv, val = np.empty((n, 3), np.int32), np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n)
        self.L.ref_kdtree_get_nodes(self.h, _ptr(keys), _ptr(cl), _ptr(pv), _ptr(val))
        return keys, cl, pv, val


class PortMap3D:
    def __init__(self, resolution, keys, max_dist):
        self.L = port_lib()
        self.keys = np.ascontiguousarray(keys, np.int32)
        self.resolution, self.max_dist = resolution, max_dist
        self.h = self.L.orc_map3d_create(resolution, _ptr(self.keys), len(self.keys), max_dist)

    def __del__(self):
        try:
            self.L.orc_map3d_destroy(self.h)
        except Exception:
            pass

    def bounds(self):
        mn, mx = (C.c_int * 3)(), (C.c_int * 3)()
        self.L.orc_map3d_get_bounds(self.h, mn, mx)
        return np.array(mn[:]), np.array(mx[:])

    def shape(self):
        mn, mx = self.bounds()
        return tuple((mx - mn + 1)[::-1])  # (D, H, W)

    def set_bounds(self, mn, mx):
        a, b = (C.c_double * 2)(*mn), (C.c_double * 2)(*mx)
        self.L.orc_map3d_set_bounds(self.h, a, b)

    def update_distances(self):
        self.L.orc_map3d_update_distances(self.h)

    def dense(self):
        out = np.empty(self.shape(), np.uint8)
        self.L.orc_map3d_get_dense(self.h, _ptr(out))
        return out

    def set_dense(self, field):
        f = np.ascontiguousarray(field, np.uint8)
        assert f.shape == self.shape()
        self.L.orc_map3d_set_dense(self.h, _ptr(f))

    def exact_edt(self):
        out = np.empty(self.shape(), np.uint8)
        self.L.orc_map3d_exact_edt(self.h, _ptr(out))
        return out

    def distance(self, i, j, k):
        return self.L.orc_map3d_distance(self.h, i, j, k)

    def cloud_apply(self, model, points, samples):
        s = np.array(samples, dtype=np.float64, order="C")
        pts = np.ascontiguousarray(points, np.float32)
        total = self.L.orc_cloud_apply(self.h, C.byref(model), _ptr(pts), len(pts), _ptr(s), len(s))
        return total, s[:, 3].copy()


class RefMap3D:
    def __init__(self, resolution, keys, max_dist):
        self.L = ref_lib()
        self.keys = np.ascontiguousarray(keys, np.int32)
        self.h = self.L.ref_map3d_create(resolution, _ptr(self.keys), len(self.keys), max_dist)

    def __del__(self):
        try:
            self.L.ref_map3d_destroy(self.h)
        except Exception:
            pass

    def bounds(self):
        mn, mx = (C.c_int * 3)(), (C.c_int * 3)()
        self.L.ref_map3d_get_bounds(self.h, mn, mx)
        return np.array(mn[:]), np.array(mx[:])

    def shape(self):
        mn, mx = self.bounds()
        return tuple((mx - mn + 1)[::-1])

    def set_bounds(self, mn, mx):
        a, b = (C.c_double * 2)(*mn), (C.c_double * 2)(*mx)
        self.L.ref_map3d_set_bounds(self.h, a, b)

    def update_distances(self):
        self.L.ref_map3d_update_distances(self.h)

    def dense(self):
        out = np.empty(self.shape(), np.uint8)
        self.L.ref_map3d_get_dense(self.h, _ptr(out))
        return out

    def distance(self, i, j, k):
        return self.L.ref_map3d_distance(self.h, i, j, k)

    def cloud_apply(self, model, points, samples):
        """PointCloudScanner::applyModelToSampleSet of the UNMODIFIED reference translation unit
        (sensors/point_cloud_scanner.cpp) over this map, configured from the CloudModel struct the port uses; its tf2 / pcl
        dependencies are the look-alikes under oracle/shim/"""
        h = self.L.ref_scanner3d_create(self.h, model.max_beams)
        try:
            self.L.ref_scanner3d_model(h, int(model.model_type != 0), model.z_hit, model.z_rand, model.sigma_hit,
                                       model.gompertz_a, model.gompertz_b, model.gompertz_c, model.input_shift,
                                       model.input_scale, model.output_shift)
            self.L.ref_scanner3d_set_map_factors(h, model.off_map_factor, 1.0, 0.0)
            t = (C.c_double * 3)(*model.mount_translation)
            q = (C.c_double * 4)(*model.mount_rotation)
            self.L.ref_scanner3d_set_mount(h, t, q)
            s = np.array(samples, dtype=np.float64, order="C")
            pts = np.ascontiguousarray(points, np.float32)
            total = self.L.ref_scanner3d_apply(h, _ptr(pts), len(pts), _ptr(s), len(s))
            return total, s[:, 3].copy()
        finally:
            self.L.ref_scanner3d_destroy(h)


def cloud_model(model_type=0, max_beams=4096, z_hit=0.95, z_rand=0.05, sigma_hit=0.2,
                gompertz=(1.0, 1.0, 1.0, 0.0, 1.0, 0.0), off_map_factor=1.0, mount_translation=(0.0, 0.0, 0.0),
                mount_rotation=(0.0, 0.0, 0.0, 1.0)):
    p = CloudModel()
    p.model_type, p.max_beams, p.z_hit, p.z_rand, p.sigma_hit = model_type, max_beams, z_hit, z_rand, sigma_hit
    (p.gompertz_a, p.gompertz_b, p.gompertz_c, p.input_shift, p.input_scale, p.output_shift) = gompertz
    p.off_map_factor = off_map_factor
    p.mount_translation[:] = mount_translation
    p.mount_rotation[:] = mount_rotation
    return p


# ---------------------------------------------------------------- ingestion (node_2d.cpp / node_3d.cpp)
# numpy restatements following the cited lines statement by statement; pinned bit for bit against the node translation
# units compiled into oracle/_ref and driven in process (RefNode below, tests/test_node_pins.py).
def convert_map(data, resolution, origin_xy, scale_up=1):
    """Node2D::convertMap (node_2d.cpp:265-295) -> (cells (size_y, size_x) int8, resolution, (origin_x, origin_y) f32).
    info.resolution is a float32 message field and `resolution / map_scale_up_factor_` (:267) is float / int = a float
    division; pinned against the compiled node (tests/test_oracle_port_vs_ref.py::test_node_convert_map)."""
    data = np.asarray(data, np.int8)
    height, width = data.shape
    res = float(np.float32(resolution) / np.float32(scale_up))
    size_x, size_y = width * scale_up, height * scale_up
    ox = np.float32(origin_xy[0] + (size_x // 2) * res)
    oy = np.float32(origin_xy[1] + (size_y // 2) * res)
    cells = np.empty((size_y, size_x), np.int8)
    for y in range(size_y):
        row = data[y // scale_up]
        v = row[np.arange(size_x) // scale_up]
        cells[y] = np.where(v == 0, -1, np.where(v == 100, 1, 0))
    return cells, res, (ox, oy)


def scan_from_laserscan(msg_ranges, msg_range_min, msg_range_max, angle_min, angle_increment, sensor_min_range=-1.0,
                        sensor_max_range=-1.0):
    """Node2D::updateLatestScanData (node_2d.cpp:534-560); message fields are float32"""
    r = np.asarray(msg_ranges, np.float32)
    rmin_msg, rmax_msg = np.float32(msg_range_min), np.float32(msg_range_max)
    range_max = float(min(rmax_msg, np.float32(sensor_max_range))) if sensor_max_range > 0.0 else float(rmax_msg)
    range_min = float(max(rmin_msg, np.float32(sensor_min_range))) if sensor_min_range > 0.0 else float(rmin_msg)
    ranges, angles = np.empty(len(r)), np.empty(len(r))
    for i in range(len(r)):
        ranges[i] = range_max if float(r[i]) <= range_min else float(r[i])
        angles[i] = angle_min + (i * angle_increment)
    return ranges, angles, range_max


def cloud_subsample(xyz, max_beams):
    """Node3D::updateLatestScanData (node_3d.cpp:467-480)"""
    xyz = np.asarray(xyz, np.float32)
    count = len(xyz)
    step = max(int((count - 1) / (max_beams - 1)), 1)  # C integer division truncates toward zero
    return xyz[0:count:step].copy()


def uniform_poses_rect(min_ij, max_ij, resolution, rng, n, threshold=0.0, multiplier=0.0):
    """n calls of Node::uniformPoseGenerator (node.cpp:847-868, scorePose == 1.0) with map_type 3: the free list of
    Node3D::updateFreeSpaceIndices (node_3d.cpp:306-318) is every (i, j) of [min, max), i outer; OctoMap::convertMapToWorld
    (octomap.cpp:83-95) puts cell (i, j) at (i * resolution, j * resolution). Draws from rng (an oracle.Rng)."""
    L = port_lib()
    ny = int(max_ij[1]) - int(min_ij[1])
    size = (int(max_ij[0]) - int(min_ij[0])) * ny
    out = np.empty((n, 3))

    def generate():
        k = int(L.orc_drand48(C.byref(rng)) * size)  # node.cpp:831
        th = L.orc_drand48(C.byref(rng)) * 2 * np.pi - np.pi  # :837
        return (int(min_ij[0]) + k // ny) * resolution, (int(min_ij[1]) + k % ny) * resolution, th

    for s in range(n):
        good = threshold
        p = generate()
        if good > 0.0 and 0.0 <= multiplier < 1.0:
            while 1.0 < good:
                p = generate()
                good *= multiplier
        out[s] = p
    return out


# ----------------------------------------------------------------------------------------------
# the reference's ROS node (node/node.cpp, node_2d.cpp, node_3d.cpp compiled against oracle/shim)
# ----------------------------------------------------------------------------------------------
def yaw_quat(yaw):
    """quaternion (x, y, z, w) of a rotation about z, for the tf table"""
    return (0.0, 0.0, float(np.sin(0.5 * yaw)), float(np.cos(0.5 * yaw)))


class RefNode:
    """badger_amcl::Node of the reference, driven in process: `params` fill the parameter table the node reads in its
    constructor (private names without the "~"), tf comes from set_tf, messages are delivered synchronously to the node's
    own callbacks. One node at a time (the table is process-global)."""

    def __init__(self, params=None, map_type=2, lib=None):
        self.L = lib if lib is not None else ref_lib()   # lib=node_gpu_lib(): the same node on the mirror + CUDA library
        self.L.ref_ros_reset()
        self.set_params(dict(map_type=map_type, save_pose=False))
        self.set_params(params or {})
        self.h = self.L.ref_node_create()

    def close(self):
        if self.h:
            self.L.ref_node_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_params(self, params):
        for k, v in params.items():
            if isinstance(v, str):
                self.L.ref_ros_set_param_str(k.encode(), v.encode())
            else:
                self.L.ref_ros_set_param_num(k.encode(), float(v))

    def reconfigure(self, params=None):
        """dynamic_reconfigure callback with the table's current values (Node::reconfigureCB, node.cpp:188-293)"""
        self.set_params(params or {})
        return self.L.ref_ros_reconfigure()

    def set_time(self, t):
        self.L.ref_ros_set_time(float(t))

    def set_tf(self, target, source, xyz=(0.0, 0.0, 0.0), quat=(0.0, 0.0, 0.0, 1.0)):
        a, q = np.array(xyz, np.float64), np.array(quat, np.float64)
        self.L.ref_ros_set_tf(target.encode(), source.encode(), _ptr(a), _ptr(q))

    def set_odom(self, x, y, yaw, odom="odom", base="base_link"):
        """the odom -> base_link transform Node::getOdomPose looks up (node.cpp:795-821)"""
        self.set_tf(odom, base, (x, y, 0.0), yaw_quat(yaw))

    def publish_map(self, data, resolution, origin_xy=(0.0, 0.0)):
        d = np.ascontiguousarray(data, np.int8)
        return self.L.ref_ros_publish_map(_ptr(d), d.shape[1], d.shape[0], resolution, origin_xy[0], origin_xy[1])

    def publish_octomap(self, resolution, keys, binary=True):
        k = np.ascontiguousarray(keys, np.int32)
        return self.L.ref_ros_publish_octomap(resolution, _ptr(k), len(k), int(binary))

    def publish_scan(self, ranges, angle_min, angle_increment, range_min, range_max, stamp=0.0, frame="laser",
                     angle_max=None):
        r = np.ascontiguousarray(ranges, np.float32)
        if angle_max is None:
            angle_max = angle_min + angle_increment * (len(r) - 1)
        return self.L.ref_ros_publish_scan(frame.encode(), stamp, angle_min, angle_max, angle_increment, range_min,
                                           range_max, _ptr(r), len(r))

    def publish_cloud(self, xyz, stamp=0.0, frame="lidar"):
        p = np.ascontiguousarray(xyz, np.float32)
        return self.L.ref_ros_publish_cloud(frame.encode(), stamp, _ptr(p), len(p))

    def publish_odom(self, x, y, yaw, stamp=0.0):
        return self.L.ref_ros_publish_odom(stamp, x, y, yaw)

    def publish_initial_pose(self, x, y, yaw, cov36=None, frame="map", stamp=0.0):
        c = np.zeros(36) if cov36 is None else np.ascontiguousarray(cov36, np.float64).reshape(36)
        return self.L.ref_ros_publish_initial_pose(frame.encode(), stamp, x, y, yaw, _ptr(c))

    def global_localization(self):
        return self.L.ref_ros_call_service(b"global_localization")

    def fire_timers(self):
        return self.L.ref_ros_fire_timers()

    def message_count(self, topic):
        return self.L.ref_ros_message_count(topic.encode())

    # ---- what the node holds / computed
    def pf(self):
        h = self.L.ref_node_pf(self.h)
        if not h:
            return None
        p = RefPF.__new__(RefPF)
        p.L, p.h, p.max_samples = self.L, h, None
        return p

    def map2d(self):
        """(cells (size_y, size_x) int8, resolution, (origin_x, origin_y), handle-backed RefMap2D) of Node2D::map_"""
        h = self.L.ref_node_map2d(self.h)
        if not h:
            return None
        size, res, org = (C.c_int * 2)(), C.c_double(), (C.c_double * 2)()
        self.L.ref_map2d_geometry(h, size, C.byref(res), org)
        m = RefMap2D.__new__(RefMap2D)
        m.L, m.h, m.size_x, m.size_y, m.resolution = self.L, h, size[0], size[1], res.value
        m.origin = (org[0], org[1])
        cells = np.empty((m.size_y, m.size_x), np.int8)
        for j in range(m.size_y):
            for i in range(m.size_x):
                cells[j, i] = self.L.ref_map2d_get_cell(h, i, j)
        m.cells = cells
        return m

    def map3d(self):
        h = self.L.ref_node_map3d(self.h)
        if not h:
            return None
        m = RefMap3D.__new__(RefMap3D)
        m.L, m.h = self.L, h
        return m

    def free_space(self):
        n = self.L.ref_node_free_space_count(self.h)
        out = np.empty((n, 2), np.int32)
        self.L.ref_node_free_space(self.h, _ptr(out))
        return out

    def uniform_poses(self, n):
        out = np.empty((n, 3))
        self.L.ref_node_uniform_poses(self.h, n, _ptr(out))
        return out

    def scan_data(self, capacity=4096):
        r, a, m = np.empty(capacity), np.empty(capacity), C.c_double()
        n = self.L.ref_node_scan_data(self.h, _ptr(r), _ptr(a), C.byref(m), capacity)
        return None if n < 0 else (r[:n].copy(), a[:n].copy(), m.value)

    def angle_stats(self, angle_min, angle_increment, frame="laser", stamp=0.0):
        out = np.empty(2)
        ok = self.L.ref_node_angle_stats(self.h, frame.encode(), stamp, angle_min, angle_increment, _ptr(out))
        return (out[0], out[1]) if ok == 1 else None

    def cloud_data(self, capacity=1 << 20):
        out = np.empty((capacity, 3), np.float32)
        n = self.L.ref_node_cloud_data(self.h, _ptr(out), capacity)
        return None if n < 0 else out[:n].copy()

    def max_weight_pose(self):
        w, p = C.c_double(), np.empty(3)
        ok = self.L.ref_node_max_weight_pose(self.h, C.byref(w), _ptr(p))
        return (w.value, p) if ok == 1 else None

    def last_pose(self):
        """last amcl_pose message: dict(count, position, orientation (x, y, z, w), stamp, covariance (6, 6))"""
        out = np.empty(44)
        n = self.L.ref_ros_last_pose(_ptr(out))
        if n == 0:
            return None
        return dict(count=n, position=out[0:3].copy(), orientation=out[3:7].copy(), stamp=out[7],
                    covariance=out[8:44].reshape(6, 6).copy())

    def last_particlecloud(self, capacity=1 << 20):
        out = np.empty((capacity, 7))
        n = self.L.ref_ros_last_particlecloud(_ptr(out), capacity)
        return None if n < 0 else out[:n].copy()

    def latest_tf(self):
        out = np.empty(7)
        return out if self.L.ref_node_latest_tf(self.h, _ptr(out)) else None

    def odom_pose(self, stamp=0.0):
        """Node::getOdomPose (node.cpp:795-821): (x, y, yaw) from the odom -> base_link transform, yaw through tf2"""
        out = np.empty(3)
        return out if self.L.ref_node_odom_pose(self.h, stamp, _ptr(out)) else None

    def install_cache_2d(self, dist, max_dist):
        """relinked node only: OccupancyMap::setDistancesLUT of the mirror (a cache computed by the reference)"""
        d = np.ascontiguousarray(dist, np.float32)
        assert self.L.ref_node_install_cache_2d(self.h, _ptr(d), d.size, max_dist) == 0

    def install_cache_3d(self, dense):
        d = np.ascontiguousarray(dense, np.uint8)
        assert self.L.ref_node_install_cache_3d(self.h, _ptr(d), d.size) == 0

    # ---- readers that work on both libraries (public API of the filter only)
    def samples(self):
        n = self.L.ref_node_sample_count(self.h)
        out = np.empty((max(n, 0), 4))
        self.L.ref_node_samples(self.h, _ptr(out))
        return out

    def filter_info(self):
        out = np.empty(5)
        self.L.ref_node_filter_info(self.h, _ptr(out))
        return dict(leaf_count=int(out[0]), cluster_count=int(out[1]), converged=bool(out[2]), w_slow=out[3], w_fast=out[4])

    def sample_clusters(self):
        out = np.empty(self.L.ref_node_sample_count(self.h), np.int32)
        self.L.ref_node_sample_clusters(self.h, _ptr(out))
        return out

    def drand48_state(self):
        s = (C.c_ushort * 3)()
        self.L.ref_node_get_seed48(s)
        return s[0] | (s[1] << 16) | (s[2] << 32)

    def state(self):
        out = np.empty(11)
        self.L.ref_node_state(self.h, _ptr(out))
        return dict(odom_init=bool(out[0]), global_localization_active=bool(out[1]), nd_global_localization_active=bool(out[2]),
                    resample_count=int(out[3]), scanners=int(out[4]), pf_odom_pose=out[5:8].copy(),
                    absolute_motion=out[8:11].copy())

