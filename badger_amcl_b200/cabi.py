"""ctypes binding of include/bamcl.h (libbamcl_cuda.so). Thin: numpy in, numpy out, errors raise.

This is plumbing for tests/ and bench.py; the product boundary is the C ABI itself. If the library has not
been built, or there is no CUDA device, construction fails loudly -- there is no CPU fallback.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libbamcl_cuda.so")

MODEL_BEAM, MODEL_LF, MODEL_LF_PROB, MODEL_LF_GOMPERTZ = 0, 1, 2, 3


class BamclError(RuntimeError):
    pass


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


def cloud_model(model_type=0, max_beams=4096, z_hit=0.95, z_rand=0.05, sigma_hit=0.2,
                gompertz=(1.0, 1.0, 1.0, 0.0, 1.0, 0.0), off_map_factor=1.0, mount_translation=(0.0, 0.0, 0.0),
                mount_rotation=(0.0, 0.0, 0.0, 1.0)):
    """Defaults are the reference node's (node_3d.cpp:58-94)."""
    p = CloudModel()
    p.model_type, p.max_beams, p.z_hit, p.z_rand, p.sigma_hit = model_type, max_beams, z_hit, z_rand, sigma_hit
    (p.gompertz_a, p.gompertz_b, p.gompertz_c, p.input_shift, p.input_scale, p.output_shift) = gompertz
    p.off_map_factor = off_map_factor
    p.mount_translation[:] = mount_translation
    p.mount_rotation[:] = mount_rotation
    return p


class ResampleParams(C.Structure):
    _fields_ = [("resample_systematic", C.c_int), ("min_samples", C.c_int), ("max_samples", C.c_int),
                ("pop_err", C.c_double), ("pop_z", C.c_double), ("w_diff", C.c_double), ("dist_threshold", C.c_double),
                ("convergence_threshold", C.c_double), ("drand48_state", C.c_uint64), ("random_poses", C.c_void_p),
                ("n_random_poses", C.c_long)]


class ResampleResult(C.Structure):
    _fields_ = [("sample_count", C.c_int), ("leaf_count", C.c_int), ("node_count", C.c_int), ("cluster_count", C.c_int),
                ("converged", C.c_int), ("random_poses_used", C.c_int), ("status_flags", C.c_int), ("reserved", C.c_int),
                ("drand48_state", C.c_uint64), ("mean", C.c_double * 3), ("cov", C.c_double * 5)]


ODOM_DIFF, ODOM_OMNI, ODOM_DIFF_CORRECTED, ODOM_OMNI_CORRECTED, ODOM_GAUSSIAN = 0, 1, 2, 3, 4  # odom.h:33-40


class OdomModel(C.Structure):
    _fields_ = [("model_type", C.c_int), ("alpha", C.c_double * 5)]


class OdomData(C.Structure):
    _fields_ = [("pose", C.c_double * 3), ("delta", C.c_double * 3), ("absolute_motion", C.c_double * 3)]


def scan_from_laserscan(msg_ranges, msg_range_min, msg_range_max, angle_min, angle_increment, sensor_min_range=-1.0,
                        sensor_max_range=-1.0):
    """Node2D::updateLatestScanData (node_2d.cpp:534-560) -> (ranges f64, angles f64, range_max)"""
    r = np.ascontiguousarray(msg_ranges, np.float32)
    ranges, angles, rmax = np.empty(len(r)), np.empty(len(r)), C.c_double()
    _check(lib().bamcl_scan_from_laserscan(_p(r), len(r), msg_range_min, msg_range_max, sensor_min_range, sensor_max_range,
                                           angle_min, angle_increment, _p(ranges), _p(angles), C.byref(rmax)))
    return ranges, angles, rmax.value


NOALLOC_MAX_SAMPLES = 1 << 24


def weights_normalise_noalloc(ctx, weights_dev_ptr, n, weights_out_dev_ptr, total_out_dev_ptr):
    """bamcl_weights_normalise_noalloc: DEVICE pointers as integers (e.g. tensor.data_ptr()); queued on the context's
    stream, not synchronised"""
    _check(lib().bamcl_weights_normalise_noalloc(ctx.h, C.c_void_p(weights_dev_ptr), n, C.c_void_p(weights_out_dev_ptr),
                                                 C.c_void_p(total_out_dev_ptr)))


def cloud_subsample(xyz, max_beams):
    """Node3D::updateLatestScanData (node_3d.cpp:467-480): stride sub-sampling of an (n, 3) float32 cloud"""
    xyz = np.ascontiguousarray(xyz, np.float32)
    out, n = np.empty_like(xyz), C.c_long()
    _check(lib().bamcl_cloud_subsample(_p(xyz), len(xyz), max_beams, _p(out), len(out), C.byref(n)))
    return out[:n.value].copy()


def planar_model(model_type=MODEL_LF, max_beams=60, z_hit=0.95, z_short=0.1, z_max=0.05, z_rand=0.05, sigma_hit=0.2,
                 lambda_short=0.1, gompertz=(1.0, 1.0, 1.0, 0.0, 1.0, 0.0), do_beamskip=False, beam_skip_distance=0.5,
                 beam_skip_threshold=0.3, beam_skip_error_threshold=0.9, off_map_factor=1.0, non_free_space_factor=1.0,
                 non_free_space_radius=0.0, scanner_pose=(0.0, 0.0, 0.0)):
    """Defaults are the reference node's (node_2d.cpp:49-76)."""
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


_lib = None

# every symbol include/bamcl.h declares: (restype, argtypes)
_D, _I, _L, _P, _F = C.c_double, C.c_int, C.c_long, C.c_void_p, C.c_float
SIGNATURES = {
    "bamcl_last_error": (C.c_char_p, []),
    "bamcl_version": (_I, []),
    "bamcl_ctx_create": (_I, [_I, _P]),
    "bamcl_ctx_destroy": (_I, [_P]),
    "bamcl_ctx_sync": (_I, [_P]),
    "bamcl_ctx_stream": (_P, [_P]),
    "bamcl_timer_start": (_I, [_P]),
    "bamcl_timer_stop_ms": (_I, [_P, _P]),
    "bamcl_ctx_launch_count": (_L, [_P]),
    "bamcl_ctx_flush_l2": (_I, [_P]),
    "bamcl_bench_gather": (_I, [_P, C.c_size_t, _L, _P]),
    "bamcl_host_alloc": (_I, [C.c_size_t, _P]),
    "bamcl_host_free": (_I, [_P]),
    "bamcl_host_register": (_I, [_P, C.c_size_t]),
    "bamcl_host_unregister": (_I, [_P]),
    "bamcl_pf_last_score_ms": (_I, [_P, _P]),
    "bamcl_pf_set_exact_terms": (_I, [_P, _I]),
    "bamcl_map2d_create": (_I, [_P, _I, _I, _D, _F, _F, _P]),
    "bamcl_map2d_destroy": (_I, [_P]),
    "bamcl_map2d_set_cells": (_I, [_P, _P, _L]),
    "bamcl_map2d_build_distances": (_I, [_P, _D]),
    "bamcl_map2d_upload_distances": (_I, [_P, _P, _L, _D]),
    "bamcl_map2d_download_distances": (_I, [_P, _P, _L]),
    "bamcl_map2d_calc_range": (_I, [_P, _P, _P, _P, _L, _D, _P, _P]),
    "bamcl_map3d_create": (_I, [_P, _D, _D, _P, _L, _P]),
    "bamcl_map3d_destroy": (_I, [_P]),
    "bamcl_map3d_get_bounds": (_I, [_P, _P, _P]),
    "bamcl_map3d_set_bounds": (_I, [_P, _P, _P]),
    "bamcl_map3d_build_distances": (_I, [_P]),
    "bamcl_map3d_upload_distances": (_I, [_P, _P, _L]),
    "bamcl_map3d_download_distances": (_I, [_P, _P, _L]),
    "bamcl_pf_score_cloud": (_I, [_P, _P, _P, _P, _L, _P]),
    "bamcl_pf_update_cloud": (_I, [_P, _P, _P, _P, _L, _P]),
    "bamcl_pf_stage_cloud": (_I, [_P, _P, _P, _P, _L]),
    "bamcl_pf_update_cloud_staged": (_I, [_P, _P, _P]),
    "bamcl_pf_create": (_I, [_P, _I, _I, _P]),
    "bamcl_pf_destroy": (_I, [_P]),
    "bamcl_pf_upload_samples": (_I, [_P, _P, _L]),
    "bamcl_pf_download_samples": (_I, [_P, _P, _L]),
    "bamcl_pf_sample_count": (_I, [_P, _P]),
    "bamcl_pf_upload_weights": (_I, [_P, _P, _L]),
    "bamcl_pf_download_weights": (_I, [_P, _P, _L]),
    "bamcl_pf_score_planar": (_I, [_P, _P, _P, _P, _P, _I, _D, _I, _P]),
    "bamcl_pf_normalise": (_I, [_P, _D]),
    "bamcl_pf_update_planar": (_I, [_P, _P, _P, _P, _P, _I, _D, _I, _P]),
    "bamcl_pf_stage_scan": (_I, [_P, _P, _P, _P, _P, _I, _D]),
    "bamcl_pf_update_planar_staged": (_I, [_P, _P, _I, _P]),
    "bamcl_pf_score_planar_staged": (_I, [_P, _P, _I, _P]),
    "bamcl_pf_update_planar_host": (_I, [_P, _P, _P, _P, _P, _I, _D, _I, _P, _L, _P]),
    "bamcl_pf_resample_host": (_I, [_P, _P, _P, _P, _L]),
    "bamcl_pf_cdf": (_I, [_P, _P, _L]),
    "bamcl_pf_resample": (_I, [_P, _P, _P]),
    "bamcl_pf_init_stats": (_I, [_P, _P]),
    "bamcl_pf_resample_sources": (_I, [_P, _P, _L]),
    "bamcl_pf_sample_clusters": (_I, [_P, _P, _L]),
    "bamcl_pf_cluster_stats": (_I, [_P, _P, _I]),
    "bamcl_resample_limit": (_I, [_I, _I, _I, _D, _D]),
    "bamcl_systematic_targets": (_I, [_D, _D, _L, _P]),
    "bamcl_pf_set_window": (_I, [_P, _L, _L]),
    "bamcl_systematic_plan": (_I, [_D, _D, _L, _P, _I, _P, _I, _P]),
    "bamcl_pf_leaf_count": (_I, [_P, _P]),
    "bamcl_comm_unique_id": (_I, [_P]),
    "bamcl_comm_create": (_I, [_P, _P, _I, _I, _P]),
    "bamcl_comm_destroy": (_I, [_P]),
    "bamcl_comm_rank": (_I, [_P, _P, _P]),
    "bamcl_comm_local_group": (_I, [_I, _P]),
    "bamcl_comm_local_group_free": (_I, [_P]),
    "bamcl_comm_create_local": (_I, [_P, _P, _I, _P]),
    "bamcl_comm_all_reduce_sum_i32": (_I, [_P, _P, _I]),
    "bamcl_comm_all_reduce_i32": (_I, [_P, _P, _I, _I]),
    "bamcl_comm_create_file": (_I, [_P, C.c_char_p, _I, _I, _I, _P]),
    "bamcl_pf_shard_attach": (_I, [_P, _P]),
    "bamcl_pf_shard_set_count": (_I, [_P, _L, _P, _P]),
    "bamcl_pf_shard_window": (_I, [_P, _P, _P]),
    "bamcl_pf_shard_init_stats": (_I, [_P, _P]),
    "bamcl_pf_shard_update_planar_staged": (_I, [_P, _P, _I, _P]),
    "bamcl_pf_shard_update_cloud_staged": (_I, [_P, _P, _P]),
    "bamcl_pf_shard_resample": (_I, [_P, _P, _P]),
    "bamcl_pf_shard_bounds": (_I, [_P, _P, _I]),
    "bamcl_shard_exchange_plan": (_I, [_P, _I, _L, _L, _I, _P, _I, _P]),
    "bamcl_pf_update_action": (_I, [_P, _P, _P, C.c_uint64, _P]),
    "bamcl_pf_max_weight_pose": (_I, [_P, _P, _P, _P]),
    "bamcl_pf_skip_counts": (_I, [_P, _P, _I, _P, _I, _P]),
    "bamcl_pf_set_skip_counts": (_I, [_P, _P, _I, _L]),
    "bamcl_pf_set_random_pose_source": (_I, [_P, _P, _D, _D]),
    "bamcl_map2d_from_occupancy_grid": (_I, [_P, _P, _I, _I, C.c_float, _D, _D, _I, _P]),
    "bamcl_map2d_download_cells": (_I, [_P, _P, _L]),
    "bamcl_map2d_geometry": (_I, [_P, _P, _P, _P]),
    "bamcl_scan_from_laserscan": (_I, [_P, _I, C.c_float, C.c_float, _D, _D, _D, _D, _P, _P, _P]),
    "bamcl_cloud_subsample": (_I, [_P, _L, _I, _P, _L, _P]),
    "bamcl_map2d_update_free_space": (_I, [_P, _D, _P]),
    "bamcl_map2d_download_free_space": (_I, [_P, _P, _L]),
    "bamcl_pf_init_uniform": (_I, [_P, _P, _D, _D, C.c_uint64, _P, _P]),
    "bamcl_pf_init_uniform_3d": (_I, [_P, _P, _D, _D, C.c_uint64, _P, _P]),
    "bamcl_pf_init_gaussian": (_I, [_P, _P, _P, _P, C.c_uint64, _P, _P]),
    "bamcl_pf_resample_multinomial_drawn": (_I, [_P, _P, _P, _P]),
    "bamcl_weights_normalise_noalloc": (_I, [_P, _P, _L, _P, _P]),
    "bamcl_pf_set_random_pose_source_3d": (_I, [_P, _P, _D, _D]),
    "bamcl_pose_covariance": (_I, [_P, _P]),
    "bamcl_pf_upload_window_samples": (_I, [_P, _P, _L, _L]),
    "bamcl_pf_download_window_samples": (_I, [_P, _P, _L, _L]),
}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise BamclError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(no CPU fallback exists)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError here = the library does not export what bamcl.h declares
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise BamclError("bamcl error %d: %s" % (rc, lib().bamcl_last_error().decode()))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Context:
    def __init__(self, device=0):
        self.L = lib()
        h = C.c_void_p()
        _check(self.L.bamcl_ctx_create(device, C.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            self.L.bamcl_ctx_destroy(self.h)
            self.h = None

    def sync(self):
        _check(self.L.bamcl_ctx_sync(self.h))

    def stream(self):
        return self.L.bamcl_ctx_stream(self.h)

    def timer_start(self):
        _check(self.L.bamcl_timer_start(self.h))

    def timer_stop_ms(self):
        ms = C.c_float()
        _check(self.L.bamcl_timer_stop_ms(self.h, C.byref(ms)))
        return ms.value

    def launch_count(self):
        return self.L.bamcl_ctx_launch_count(self.h)

    def flush_l2(self):
        _check(self.L.bamcl_ctx_flush_l2(self.h))

    def gather_rate(self, table_bytes, n_gathers=1 << 30):
        """random 2-byte gathers per second over a table of table_bytes (bamcl_bench_gather)"""
        ms = C.c_float()
        _check(self.L.bamcl_bench_gather(self.h, table_bytes, n_gathers, C.byref(ms)))
        return n_gathers / (ms.value * 1e-3)


def pinned_array(shape, dtype=np.float64):
    """numpy array over page-locked host memory from bamcl_host_alloc (kept alive by the array's base)."""
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = C.c_void_p()
    _check(lib().bamcl_host_alloc(max(n, 1), C.byref(p)))
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED.append((p, buf))
    return arr


_PINNED = []  # pinned blocks live for the life of the process (bench / tests only)


class Map2D:
    """OccupancyMap (include/amcl/map/occupancy_map.h:57-81) on the device."""

    def __init__(self, ctx, cells, resolution, origin=(0.0, 0.0)):
        self.ctx, self.L = ctx, ctx.L
        cells = np.ascontiguousarray(cells, np.int8)
        self.size_y, self.size_x = cells.shape
        self.resolution = float(resolution)
        h = C.c_void_p()
        _check(self.L.bamcl_map2d_create(ctx.h, self.size_x, self.size_y, resolution, origin[0], origin[1], C.byref(h)))
        self.h = h
        if cells.size:
            _check(self.L.bamcl_map2d_set_cells(self.h, _p(cells), cells.size))

    @classmethod
    def from_occupancy_grid(cls, ctx, data, resolution, origin_xy=(0.0, 0.0), scale_up=1):
        """Node2D::convertMap (node_2d.cpp:265-295): data = (height, width) int8 OccupancyGrid payload; resolution is the
        message's float32 field (narrowed here), so the map's resolution is float32(resolution) / scale_up in float"""
        data = np.ascontiguousarray(data, np.int8)
        self = cls.__new__(cls)
        self.L, self.ctx = lib(), ctx
        h = C.c_void_p()
        _check(self.L.bamcl_map2d_from_occupancy_grid(ctx.h, _p(data), data.shape[1], data.shape[0], resolution,
                                                      origin_xy[0], origin_xy[1], scale_up, C.byref(h)))
        self.h = h
        size, res, org = (C.c_int * 2)(), C.c_double(), (C.c_float * 2)()
        _check(self.L.bamcl_map2d_geometry(self.h, size, C.byref(res), org))
        self.size_x, self.size_y, self.resolution, self.origin = size[0], size[1], res.value, (org[0], org[1])
        return self

    def cells(self):
        out = np.empty((self.size_y, self.size_x), np.int8)
        _check(self.L.bamcl_map2d_download_cells(self.h, _p(out), out.size))
        return out

    def close(self):
        if self.h:
            self.L.bamcl_map2d_destroy(self.h)
            self.h = None

    def build_distances(self, max_distance):
        _check(self.L.bamcl_map2d_build_distances(self.h, max_distance))

    def upload_distances(self, dist, max_distance):
        d = np.ascontiguousarray(dist, np.float32)
        _check(self.L.bamcl_map2d_upload_distances(self.h, _p(d), d.size, max_distance))

    def update_free_space(self, radius):
        """Node2D::updateFreeSpaceIndices (node_2d.cpp:320-338) on the device; returns the number of free cells"""
        n = C.c_long()
        _check(self.L.bamcl_map2d_update_free_space(self.h, radius, C.byref(n)))
        self.free_count = n.value
        return n.value

    def free_space(self):
        out = np.empty((self.free_count, 2), np.int32)
        _check(self.L.bamcl_map2d_download_free_space(self.h, _p(out), len(out)))
        return out

    def download_distances(self):
        out = np.empty((self.size_y, self.size_x), np.float32)
        _check(self.L.bamcl_map2d_download_distances(self.h, _p(out), out.size))
        return out

    def calc_range(self, ox, oy, oa, max_range, with_steps=False):
        ox, oy, oa = (np.ascontiguousarray(np.broadcast_to(a, np.shape(oa)), np.float64) for a in (ox, oy, oa))
        out = np.empty(len(oa))
        steps = C.c_long(0)
        _check(self.L.bamcl_map2d_calc_range(self.h, _p(ox), _p(oy), _p(oa), len(oa), max_range, _p(out), C.byref(steps)))
        return (out, steps.value) if with_steps else out


class Map3D:
    """OctoMap (include/amcl/map/octomap.h:61-81) on the device; `keys` = occupied leaf keys (n, 3)."""

    def __init__(self, ctx, resolution, keys, max_distance):
        self.ctx, self.L = ctx, ctx.L
        keys = np.ascontiguousarray(keys, np.int32)
        h = C.c_void_p()
        _check(self.L.bamcl_map3d_create(ctx.h, resolution, max_distance, _p(keys), len(keys), C.byref(h)))
        self.h = h
        self.resolution, self.max_distance = resolution, max_distance

    def close(self):
        if self.h:
            self.L.bamcl_map3d_destroy(self.h)
            self.h = None

    def bounds(self):
        mn, mx = (C.c_int * 3)(), (C.c_int * 3)()
        _check(self.L.bamcl_map3d_get_bounds(self.h, mn, mx))
        return np.array(mn[:]), np.array(mx[:])

    def shape(self):
        mn, mx = self.bounds()
        return tuple(int(v) for v in (mx - mn + 1)[::-1])  # (D, H, W)

    def set_bounds(self, mn, mx):
        a, b = (C.c_double * 2)(*mn), (C.c_double * 2)(*mx)
        _check(self.L.bamcl_map3d_set_bounds(self.h, a, b))

    def build_distances(self):
        _check(self.L.bamcl_map3d_build_distances(self.h))

    def upload_distances(self, field):
        f = np.ascontiguousarray(field, np.uint8)
        assert f.shape == self.shape()
        _check(self.L.bamcl_map3d_upload_distances(self.h, _p(f), f.size))

    def download_distances(self):
        out = np.empty(self.shape(), np.uint8)
        _check(self.L.bamcl_map3d_download_distances(self.h, _p(out), out.size))
        return out


class ParticleFilter:
    """Device side of ParticleFilter (include/amcl/pf/particle_filter.h:96-130). The scalar bookkeeping the
    reference keeps in the class (w_slow_/w_fast_, current_set_, converged_) lives in the host C++ mirror
    (badger_amcl_b200/host); this wrapper re-states just enough of it for tests and the bench."""

    def __init__(self, ctx, min_samples, max_samples, alpha_slow=0.001, alpha_fast=0.1, conv_thresh=95.0,
                 systematic=True, pop_err=0.01, pop_z=0.99):
        self.ctx, self.L = ctx, ctx.L
        self.min_samples, self.max_samples = min_samples, max_samples
        self.alpha_slow, self.alpha_fast, self.conv_thresh = alpha_slow, alpha_fast, conv_thresh
        self.systematic, self.pop_err, self.pop_z = systematic, pop_err, pop_z
        self.w_slow = self.w_fast = 0.0
        self.converged = False
        self.drand48_state = 0
        self.random_poses = np.zeros((0, 3))
        self.random_next = 0
        self.last = None
        h = C.c_void_p()
        _check(self.L.bamcl_pf_create(ctx.h, min_samples, max_samples, C.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            self.L.bamcl_pf_destroy(self.h)
            self.h = None

    # -- sample exchange
    def sample_count(self):
        n = C.c_int()
        _check(self.L.bamcl_pf_sample_count(self.h, C.byref(n)))
        return n.value

    def upload_samples(self, aos):
        a = np.ascontiguousarray(aos, np.float64)
        assert a.ndim == 2 and a.shape[1] == 4
        _check(self.L.bamcl_pf_upload_samples(self.h, _p(a), len(a)))

    def samples(self):
        out = np.empty((self.sample_count(), 4))
        _check(self.L.bamcl_pf_download_samples(self.h, _p(out), len(out)))
        return out

    def upload_weights(self, w):
        w = np.ascontiguousarray(w, np.float64)
        _check(self.L.bamcl_pf_upload_weights(self.h, _p(w), len(w)))

    def weights(self):
        return self.weights_n(self.sample_count())

    def weights_n(self, n):
        """n = sample_count, or the window size of a distributed filter"""
        out = np.empty(n)
        _check(self.L.bamcl_pf_download_weights(self.h, _p(out), len(out)))
        return out

    def init_with_poses(self, poses):
        """initWithPoseFn (particle_filter.cpp:136-162): max_samples poses, weights 1/max_samples, kd-tree +
        cluster statistics, averages reset, not converged."""
        poses = np.asarray(poses, np.float64)
        n = self.max_samples
        aos = np.empty((n, 4))
        aos[:, :3] = poses[np.arange(n) % len(poses)]
        aos[:, 3] = 1.0 / n
        self.upload_samples(aos)
        self.w_slow = self.w_fast = 0.0
        self.converged = False
        res = ResampleResult()
        _check(self.L.bamcl_pf_init_stats(self.h, C.byref(res)))
        self.last = res
        return res

    def init_uniform(self, map2d, threshold=0.0, multiplier=0.0):
        """initWithPoseFn(uniformPoseGenerator) on the device (particle_filter.cpp:136-162, node.cpp:823-868); the map's
        free-space list must have been built (Map2D.update_free_space); a Map3D draws from its cropped cell rectangle
        (node_3d.cpp:306-318)"""
        res = ResampleResult()
        out = C.c_uint64()
        fn = self.L.bamcl_pf_init_uniform_3d if isinstance(map2d, Map3D) else self.L.bamcl_pf_init_uniform
        _check(fn(self.h, map2d.h, threshold, multiplier, self.drand48_state, C.byref(out), C.byref(res)))
        self.drand48_state = out.value
        self.w_slow = self.w_fast = 0.0
        self.converged = False
        self.last = res
        return res

    def init_gaussian(self, mean, cr, cd):
        """initWithGaussian on the device (particle_filter.cpp:105-133); cr (3, 3) rotation and cd (3,) standard deviations
        = the decomposition PDFGaussian's constructor computes (identity / sqrt(diagonal) for a diagonal covariance)"""
        m, r, d = (np.ascontiguousarray(v, np.float64) for v in (mean, cr, cd))
        assert m.shape == (3,) and r.shape == (3, 3) and d.shape == (3,)
        res = ResampleResult()
        out = C.c_uint64()
        _check(self.L.bamcl_pf_init_gaussian(self.h, _p(m), _p(r), _p(d), self.drand48_state, C.byref(out), C.byref(res)))
        self.drand48_state = out.value
        self.w_slow = self.w_fast = 0.0
        self.converged = False
        self.last = res
        return res

    def set_random_pose_source(self, map2d, threshold=0.0, multiplier=0.0):
        """random_pose_fn_ = Node::uniformPoseGenerator over map2d's free-space list (node.cpp:262-264, 847-868)"""
        fn = self.L.bamcl_pf_set_random_pose_source_3d if isinstance(map2d, Map3D) else self.L.bamcl_pf_set_random_pose_source
        _check(fn(self.h, map2d.h if map2d is not None else None, threshold, multiplier))
        self.device_pose_source = map2d is not None

    def init_with_current(self):
        """kd-tree + cluster statistics of the samples already on the device (bamcl_pf_init_stats)."""
        res = ResampleResult()
        _check(self.L.bamcl_pf_init_stats(self.h, C.byref(res)))
        self.converged = False
        self.last = res
        return res

    def download_samples_into(self, out):
        """current set -> caller's (n, 4) float64 buffer (e.g. a pinned_array slice)"""
        assert out.dtype == np.float64 and out.flags.c_contiguous and out.shape[1] == 4
        _check(self.L.bamcl_pf_download_samples(self.h, _p(out), len(out)))

    # -- sensor update
    def score_planar(self, map2d, model, ranges, angles, range_max):
        """PlanarScanner::applyModelToSampleSet: returns the total; weights stay un-normalised."""
        ranges = np.ascontiguousarray(ranges, np.float64)
        angles = np.ascontiguousarray(angles, np.float64)
        total = C.c_double()
        _check(self.L.bamcl_pf_score_planar(self.h, map2d.h, C.byref(model), _p(ranges), _p(angles), len(ranges),
                                            range_max, int(self.converged), C.byref(total)))
        return total.value

    def set_exact_terms(self, exact):
        _check(self.L.bamcl_pf_set_exact_terms(self.h, int(exact)))

    def last_score_ms(self):
        ms = C.c_float()
        _check(self.L.bamcl_pf_last_score_ms(self.h, C.byref(ms)))
        return ms.value

    def normalise(self, total):
        _check(self.L.bamcl_pf_normalise(self.h, total))

    def _averages(self, total):
        """particle_filter.cpp:237-257 (w_avg is the same sequential sum as the total, divided by the count)"""
        if total > 0.0:
            w_avg = total / self.sample_count()
            self.w_slow = w_avg if self.w_slow == 0.0 else self.w_slow + self.alpha_slow * (w_avg - self.w_slow)
            self.w_fast = w_avg if self.w_fast == 0.0 else self.w_fast + self.alpha_fast * (w_avg - self.w_fast)

    def update_sensor(self, map2d, model, ranges, angles, range_max):
        """PlanarScanner::updateSensor -> ParticleFilter::updateSensor, fused on the device."""
        if model.max_beams < 2:
            return False
        ranges = np.ascontiguousarray(ranges, np.float64)
        angles = np.ascontiguousarray(angles, np.float64)
        total = C.c_double()
        _check(self.L.bamcl_pf_update_planar(self.h, map2d.h, C.byref(model), _p(ranges), _p(angles), len(ranges),
                                             range_max, int(self.converged), C.byref(total)))
        self._averages(total.value)
        self.last_total = total.value
        return True

    def update_sensor_host(self, map2d, model, ranges, angles, range_max, aos):
        """updateSensor with the particles still in host memory (n x 4 AoS, ideally a pinned_array): the upload is
        pipelined with the scoring (bamcl_pf_update_planar_host)."""
        if model.max_beams < 2:
            return False
        ranges = np.ascontiguousarray(ranges, np.float64)
        angles = np.ascontiguousarray(angles, np.float64)
        assert aos.dtype == np.float64 and aos.flags.c_contiguous and aos.ndim == 2 and aos.shape[1] == 4
        total = C.c_double()
        _check(self.L.bamcl_pf_update_planar_host(self.h, map2d.h, C.byref(model), _p(ranges), _p(angles), len(ranges),
                                                  range_max, int(self.converged), _p(aos), len(aos), C.byref(total)))
        self._averages(total.value)
        self.last_total = total.value
        return True

    def stage_scan(self, map2d, model, ranges, angles, range_max):
        ranges = np.ascontiguousarray(ranges, np.float64)
        angles = np.ascontiguousarray(angles, np.float64)
        _check(self.L.bamcl_pf_stage_scan(self.h, map2d.h, C.byref(model), _p(ranges), _p(angles), len(ranges), range_max))

    def update_staged(self, map2d, want_total=True):
        total = C.c_double()
        _check(self.L.bamcl_pf_update_planar_staged(self.h, map2d.h, int(self.converged),
                                                    C.byref(total) if want_total else None))
        if want_total:
            self._averages(total.value)
            return total.value
        return None

    # -- 3-D sensor update
    def score_cloud(self, map3d, model, points):
        """PointCloudScanner::applyModelToSampleSet: returns the total; weights stay un-normalised."""
        pts = np.ascontiguousarray(points, np.float32)
        total = C.c_double()
        _check(self.L.bamcl_pf_score_cloud(self.h, map3d.h, C.byref(model), _p(pts), len(pts), C.byref(total)))
        return total.value

    def update_sensor_cloud(self, map3d, model, points):
        """PointCloudScanner::updateSensor -> ParticleFilter::updateSensor, fused on the device."""
        if model.max_beams < 2:
            return False
        pts = np.ascontiguousarray(points, np.float32)
        total = C.c_double()
        _check(self.L.bamcl_pf_update_cloud(self.h, map3d.h, C.byref(model), _p(pts), len(pts), C.byref(total)))
        self._averages(total.value)
        self.last_total = total.value
        return True

    def stage_cloud(self, map3d, model, points):
        pts = np.ascontiguousarray(points, np.float32)
        _check(self.L.bamcl_pf_stage_cloud(self.h, map3d.h, C.byref(model), _p(pts), len(pts)))

    def update_cloud_staged(self, map3d, want_total=True):
        total = C.c_double()
        _check(self.L.bamcl_pf_update_cloud_staged(self.h, map3d.h, C.byref(total) if want_total else None))
        if want_total:
            self._averages(total.value)
            return total.value
        return None

    # -- resampling
    def cdf(self):
        out = np.empty(self.sample_count() + 1)
        _check(self.L.bamcl_pf_cdf(self.h, _p(out), len(out)))
        return out

    def srand48(self, seed):
        self.drand48_state = ((seed & 0xFFFFFFFF) << 16) | 0x330E

    def set_random_poses(self, poses):
        self.random_poses = np.ascontiguousarray(poses, np.float64)
        self.random_next = 0

    def update_action(self, model_type, alpha, pose, delta, absolute_motion=(0.0, 0.0, 0.0)):
        """Odom::updateAction (odom.cpp:74-311) on the device-resident current set; the Gaussian draws continue this
        filter's drand48 stream (srand48 / the resampler's state)."""
        al = list(alpha) + [0.0] * (5 - len(alpha))
        m = OdomModel(int(model_type), (C.c_double * 5)(*al))
        d = OdomData((C.c_double * 3)(*pose), (C.c_double * 3)(*delta), (C.c_double * 3)(*absolute_motion))
        out = C.c_uint64()
        _check(self.L.bamcl_pf_update_action(self.h, C.byref(m), C.byref(d), self.drand48_state, C.byref(out)))
        self.drand48_state = out.value

    def resample_limit(self, k):
        return self.L.bamcl_resample_limit(k, self.min_samples, self.max_samples, self.pop_err, self.pop_z)

    def update_resample(self, w_diff=None, out=None):
        """ParticleFilter::updateResample (particle_filter.cpp:423-471). out = (cap x 4) float64 host buffer: the new
        set is also copied there, overlapped with the tree / cluster build (bamcl_pf_resample_host)."""
        if w_diff is None:
            with np.errstate(divide="ignore", invalid="ignore"):
                w_diff = float(1.0 - np.float64(self.w_fast) / np.float64(self.w_slow))
            if w_diff < 0.0:
                w_diff = 0.0
        rp = self.random_poses[self.random_next:]
        prm = ResampleParams(int(self.systematic), self.min_samples, self.max_samples, self.pop_err, self.pop_z, w_diff,
                             0.5, self.conv_thresh, self.drand48_state, _p(rp) if len(rp) else None, len(rp))
        res = ResampleResult()
        if out is None:
            _check(self.L.bamcl_pf_resample(self.h, C.byref(prm), C.byref(res)))
        else:
            assert out.dtype == np.float64 and out.flags.c_contiguous and out.ndim == 2 and out.shape[1] == 4
            _check(self.L.bamcl_pf_resample_host(self.h, C.byref(prm), C.byref(res), _p(out), len(out)))
        self.drand48_state = res.drand48_state
        self.random_next += res.random_poses_used
        if w_diff > 0.0:
            self.w_slow = self.w_fast = 0.0
        self.converged = bool(res.converged)
        self.last = res
        return res

    def update_resample_drawn(self, w_diff, draws, poses, states):
        """multinomial updateResample with the branch loop of particle_filter.cpp:385-393 run by the caller
        (bamcl_pf_resample_multinomial_drawn): draws[k] = second uniform of output k or -(1 + i) = poses[i]; states[k] = the
        caller's generator state before output k (states[max_samples] = after the last): the state after the kept outputs
        becomes this filter's drand48_state"""
        draws = np.ascontiguousarray(draws, np.float64)
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 3)
        assert len(draws) == self.max_samples and len(states) == self.max_samples + 1
        prm = ResampleParams(0, self.min_samples, self.max_samples, self.pop_err, self.pop_z, w_diff, 0.5, self.conv_thresh,
                             self.drand48_state, _p(poses) if len(poses) else None, len(poses))
        res = ResampleResult()
        _check(self.L.bamcl_pf_resample_multinomial_drawn(self.h, C.byref(prm), _p(draws), C.byref(res)))
        self.drand48_state = int(states[res.sample_count])
        if w_diff > 0.0:
            self.w_slow = self.w_fast = 0.0
        self.converged = bool(res.converged)
        self.last = res
        return res

    def resample_sources(self):
        return self.resample_sources_n(self.sample_count())

    def resample_sources_n(self, n):
        """n = sample_count, or the window size of a distributed filter"""
        out = np.empty(n, np.int32)
        _check(self.L.bamcl_pf_resample_sources(self.h, _p(out), len(out)))
        return out

    def sample_clusters(self):
        return self.sample_clusters_n(self.sample_count())

    def sample_clusters_n(self, n):
        out = np.empty(n, np.int32)
        _check(self.L.bamcl_pf_sample_clusters(self.h, _p(out), len(out)))
        return out

    def cluster_stats(self):
        n = self.last.cluster_count
        out = np.empty((n, 10))
        _check(self.L.bamcl_pf_cluster_stats(self.h, _p(out), n))
        return out

    def max_weight_pose(self):
        """Node2D::getMaxWeightPose (node_2d.cpp:588-617) -> (weight, pose[3], cluster index)"""
        w, idx = C.c_double(), C.c_int()
        pose = np.empty(3)
        _check(self.L.bamcl_pf_max_weight_pose(self.h, C.byref(w), _p(pose), C.byref(idx)))
        return w.value, pose, idx.value

    def pose_covariance(self):
        """the 6x6 covariance Node::updatePose publishes (node.cpp:381-395)"""
        out = np.empty(36)
        _check(self.L.bamcl_pose_covariance(C.byref(self.last), _p(out)))
        return out.reshape(6, 6)

    def set_stats(self):
        r = self.last
        return np.array(list(r.mean) + list(r.cov))
