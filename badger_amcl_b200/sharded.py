"""Distributed particle filter: ONE global sample set resampled exactly like the reference's single set
(particle_filter.cpp:223-267, 269-354, 423-471), its samples owned rank-major by G GPUs (one process per GPU).

All device work AND all rank-to-rank exchange live in the C library (include/bamcl.h, "distributed filter":
NCCL on the context's stream, csrc/comm.cu + csrc/resample.cu + csrc/kdtree.cu). This module is the thin Python
driver the tests and bench.py use: it creates the communicator (the 128-byte NCCL id travels over whatever host channel
the caller has -- torch.distributed here), keeps the filter's host-side scalars (w_slow / w_fast, the drand48 state),
and calls the collective entry points.
"""
import ctypes as C

import numpy as np

from . import cabi


def drand48_next(state):
    return (0x5DEECE66D * state + 0xB) & ((1 << 48) - 1)


def shard_bounds(n, world):
    """contiguous rank-major partition of n samples: rank r owns [b[r], b[r+1])"""
    return [(n * r) // world for r in range(world + 1)]


def systematic_plan(start, delta, n, bounds):
    """runs (k_begin, k_end, rank) of output slots by the rank whose CDF range holds the target (host arithmetic)"""
    b = np.ascontiguousarray(bounds, np.float64)
    g = len(b) - 1
    max_runs = 4 * g + 2048
    runs = np.zeros((max_runs, 3), np.int64)
    n_runs = C.c_int()
    cabi._check(cabi.lib().bamcl_systematic_plan(start, delta, n, cabi._p(b), g, cabi._p(runs), max_runs, C.byref(n_runs)))
    return [tuple(int(v) for v in runs[i]) for i in range(n_runs.value)]


def exchange_plan(runs, n_random, n, world):
    """transfers (source rank, destination rank, k_begin, k_end, position in the source's selected rows) of the pose
    exchange for a plan (host arithmetic, bamcl_shard_exchange_plan)"""
    r = np.ascontiguousarray(np.array(runs, np.int64).reshape(-1, 3))
    cap = (len(r) + 1) * (world + 1)
    moves = np.zeros((cap, 5), np.int64)
    n_moves = C.c_int()
    cabi._check(cabi.lib().bamcl_shard_exchange_plan(cabi._p(r), len(r), n_random, n, world, cabi._p(moves), cap,
                                                     C.byref(n_moves)))
    return [tuple(int(v) for v in moves[i]) for i in range(n_moves.value)]


class Comm:
    """bamcl_comm: NCCL (`Comm.nccl`) or the in-process test back end (`Comm.local`)."""

    def __init__(self, h, rank, world):
        self.h, self.rank, self.world, self.L = h, rank, world, cabi.lib()

    @classmethod
    def nccl(cls, ctx, rank, world, dist=None, id_bytes=None):
        """collective. The NCCL id is created on rank 0 and broadcast with torch.distributed (any backend) unless the
        caller already moved it (id_bytes)."""
        L = cabi.lib()
        if world > 1 and id_bytes is None:
            import torch
            buf = (C.c_ubyte * 128)()
            if rank == 0:
                cabi._check(L.bamcl_comm_unique_id(buf))
            t = torch.tensor(list(bytes(buf)), dtype=torch.uint8)
            if dist.get_backend() == "nccl":
                t = t.cuda()
            dist.broadcast(t, src=0)
            id_bytes = bytes(t.cpu().tolist())
        h = C.c_void_p()
        idb = (C.c_ubyte * 128)(*id_bytes) if id_bytes is not None else None
        cabi._check(L.bamcl_comm_create(ctx.h, idb, rank, world, C.byref(h)))
        return cls(h, rank, world)

    @staticmethod
    def local_group(world):
        g = C.c_void_p()
        cabi._check(cabi.lib().bamcl_comm_local_group(world, C.byref(g)))
        return g

    @staticmethod
    def free_local_group(g):
        cabi.lib().bamcl_comm_local_group_free(g)

    @classmethod
    def local(cls, ctx, group, rank, world):
        h = C.c_void_p()
        cabi._check(cabi.lib().bamcl_comm_create_local(ctx.h, group, rank, C.byref(h)))
        return cls(h, rank, world)

    def all_reduce_sum_i32(self, v):
        a = np.ascontiguousarray(v, np.int32).copy()
        cabi._check(self.L.bamcl_comm_all_reduce_sum_i32(self.h, cabi._p(a), len(a)))
        return a

    def close(self):
        if self.h:
            self.L.bamcl_comm_destroy(self.h)
            self.h = None


class ShardedFilter:
    """ParticleFilter::updateSensor / updateResample over a rank-major distributed set. `pf` is this rank's
    cabi.ParticleFilter (capacity = the GLOBAL max_samples), `comm` a Comm on the same context."""

    def __init__(self, pf, comm, map2d=None, alpha_slow=0.001, alpha_fast=0.1, conv_thresh=95.0, pop_err=0.01, pop_z=0.99):
        self.pf, self.comm, self.map, self.L = pf, comm, map2d, pf.L
        self.rank, self.world = comm.rank, comm.world
        self.min_samples, self.max_samples = pf.min_samples, pf.max_samples
        self.alpha_slow, self.alpha_fast, self.conv_thresh = alpha_slow, alpha_fast, conv_thresh
        self.pop_err, self.pop_z = pop_err, pop_z
        self.w_slow = self.w_fast = 0.0
        self.drand48_state = 0
        self.random_poses = np.zeros((0, 3))
        self.random_next = 0
        self.last = None
        self.converged = False
        self.cloud = None
        cabi._check(self.L.bamcl_pf_shard_attach(pf.h, comm.h))

    def srand48(self, seed):
        self.drand48_state = ((seed & 0xFFFFFFFF) << 16) | 0x330E

    def set_random_poses(self, poses):
        self.random_poses = np.ascontiguousarray(poses, np.float64)
        self.random_next = 0

    # -- the window this rank owns
    def window(self):
        off, cnt = C.c_long(), C.c_long()
        cabi._check(self.L.bamcl_pf_shard_window(self.pf.h, C.byref(off), C.byref(cnt)))
        return off.value, cnt.value

    def sample_count(self):
        return self.pf.sample_count()

    def set_global_samples(self, aos_global):
        """the global set (n x 4 AoS, the same array on every rank): every rank uploads the window it owns"""
        n = len(aos_global)
        off, cnt = C.c_long(), C.c_long()
        cabi._check(self.L.bamcl_pf_shard_set_count(self.pf.h, n, C.byref(off), C.byref(cnt)))
        self.upload_window(aos_global[off.value:off.value + cnt.value])

    def upload_window(self, aos_window):
        off, cnt = self.window()
        a = np.ascontiguousarray(aos_window, np.float64)
        assert len(a) == cnt
        if cnt:
            cabi._check(self.L.bamcl_pf_upload_window_samples(self.pf.h, cabi._p(a), off, cnt))

    def download_window(self, out=None):
        off, cnt = self.window()
        if out is None:
            out = np.empty((cnt, 4))
        if cnt:
            cabi._check(self.L.bamcl_pf_download_window_samples(self.pf.h, cabi._p(out), off, cnt))
        return out

    def init_stats(self):
        """initWithPoseFn's tail over the global set (kd-tree, clusters, statistics); collective"""
        res = cabi.ResampleResult()
        cabi._check(self.L.bamcl_pf_shard_init_stats(self.pf.h, C.byref(res)))
        self.w_slow = self.w_fast = 0.0
        self.converged = False
        self.pf.converged = False
        self.last = res
        return res

    def set_cloud(self, map3d):
        """score the staged 3-D point cloud (PointCloudScanner, point_cloud_scanner.cpp:111-129) instead of the planar scan"""
        self.cloud = map3d

    # -- ParticleFilter::updateSensor (particle_filter.cpp:223-267)
    def update_sensor(self):
        total = C.c_double()
        if self.cloud is not None:
            cabi._check(self.L.bamcl_pf_shard_update_cloud_staged(self.pf.h, self.cloud.h, C.byref(total)))
        else:
            # beam skipping (prob model, converged set): per-beam agreement counts summed over the ranks
            counts = np.zeros(4096, np.int32)
            n = C.c_int()
            cabi._check(self.L.bamcl_pf_skip_counts(self.pf.h, self.map.h, int(self.converged), cabi._p(counts), len(counts),
                                                    C.byref(n)))
            if n.value:
                summed = self.comm.all_reduce_sum_i32(counts[:n.value])
                cabi._check(self.L.bamcl_pf_set_skip_counts(self.pf.h, cabi._p(summed), len(summed), self.sample_count()))
            cabi._check(self.L.bamcl_pf_shard_update_planar_staged(self.pf.h, self.map.h, int(self.converged), C.byref(total)))
        total = total.value
        if total > 0.0:
            w_avg = total / self.sample_count()
            self.w_slow = w_avg if self.w_slow == 0.0 else self.w_slow + self.alpha_slow * (w_avg - self.w_slow)
            self.w_fast = w_avg if self.w_fast == 0.0 else self.w_fast + self.alpha_fast * (w_avg - self.w_fast)
        self.last_total = total
        return total

    # -- ParticleFilter::updateResample, systematic (particle_filter.cpp:423-471, 269-354)
    def update_resample(self, w_diff=None):
        if w_diff is None:
            with np.errstate(divide="ignore", invalid="ignore"):
                w_diff = float(1.0 - np.float64(self.w_fast) / np.float64(self.w_slow))
            if w_diff < 0.0:
                w_diff = 0.0
        if not w_diff >= 0.0:
            raise RuntimeError("updateResample before any successful updateSensor (w_diff is NaN, SURVEY Q11)")
        rp = self.random_poses[self.random_next:]
        prm = cabi.ResampleParams(1, self.min_samples, self.max_samples, self.pop_err, self.pop_z, w_diff, 0.5,
                                  self.conv_thresh, self.drand48_state, cabi._p(rp) if len(rp) else None, len(rp))
        res = cabi.ResampleResult()
        cabi._check(self.L.bamcl_pf_shard_resample(self.pf.h, C.byref(prm), C.byref(res)))
        self.drand48_state = res.drand48_state
        self.random_next += res.random_poses_used
        if w_diff > 0.0:
            self.w_slow = self.w_fast = 0.0
        self.converged = bool(res.converged)
        self.pf.converged = self.converged
        self.pf.last = res
        self.last = res
        return res

    def bounds(self):
        """exact CDF value at every rank boundary of the last resample"""
        out = np.empty(self.world + 1)
        cabi._check(self.L.bamcl_pf_shard_bounds(self.pf.h, cabi._p(out), len(out)))
        return out

    def window_sources(self):
        return self.pf.resample_sources_n(self.window()[1])

    def window_clusters(self):
        return self.pf.sample_clusters_n(self.window()[1])
