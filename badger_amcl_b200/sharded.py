"""Sharded particle filter: ONE global sample set resampled exactly like the reference's single set
(particle_filter.cpp:223-267, 269-354, 423-471), with the particles scored on G GPUs (one process per GPU).

Global order is rank-major, so the order-dependent reference semantics are the single-set ones:

  update    rank r scores its window [r*P/G, (r+1)*P/G); local totals are all-gathered and summed in rank order;
            every rank normalises its window by the global total (particle_filter.cpp:237-266)
  resample  the sequential CDF is chained over ranks (rank r starts at rank r-1's exact end value), the systematic
            target sequence is planned on the host (bamcl_systematic_plan: which rank's window holds the source of
            each output slot), each rank selects for its runs into a global (n x 3) buffer, the runs are broadcast,
            and every rank installs the new set and builds the global pose-bin tree / clusters / statistics
            (replicated: the KLD leaf count and the cluster numbering are properties of the global insertion order)

`ShardedFilter` holds the host logic and talks to torch.distributed (NCCL on the GPUs, gloo in the CPU tests); the
device phases live behind an *engine*: `CudaEngine` here (the C ABI); tests/ plug in a checker-based engine to
exercise the host logic without a GPU. No engine here computes on the CPU.
"""
import ctypes as C
import os
import time

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


class CudaEngine:
    """device phases of the sharded filter through the C ABI (include/bamcl.h, "sharded filter")"""

    def __init__(self, pf, map2d, device):
        import torch
        self.torch, self.pf, self.map, self.L = torch, pf, map2d, pf.L
        self.device = torch.device("cuda", device)
        self.buf = torch.empty((pf.max_samples, 3), dtype=torch.float64, device=self.device)
        self.src = None
        self.pack = self.pack_src = self.gathered = self.gathered_src = None

    def want_sources(self):
        self.src = self.torch.full((self.pf.max_samples,), -1, dtype=self.torch.int32, device=self.device)

    def scalar_tensor(self, v=0.0):
        return self.torch.tensor([v], dtype=self.torch.float64, device=self.device)

    def sample_count(self):
        return self.pf.sample_count()

    def leaf_count(self):
        k = C.c_int()
        cabi._check(self.L.bamcl_pf_leaf_count(self.pf.h, C.byref(k)))
        return k.value

    def set_window(self, off, cnt):
        cabi._check(self.L.bamcl_pf_set_window(self.pf.h, off, cnt))

    def set_cloud(self, map3d, model, points):
        """score 3-D point clouds (PointCloudScanner, point_cloud_scanner.cpp:111-129) instead of the staged planar scan"""
        self.cloud = (map3d, model, np.ascontiguousarray(points, np.float32))

    def score(self):
        if getattr(self, "cloud", None) is not None:
            map3d, model, pts = self.cloud
            return self.pf.score_cloud(map3d, model, pts)  # this rank's window, un-normalised
        total = C.c_double()
        cabi._check(self.L.bamcl_pf_score_planar_staged(self.pf.h, self.map.h, int(self.pf.converged), C.byref(total)))
        return total.value

    # -- update without host round trips: score -> all-gather of the totals -> global total + normalise, one stream
    def ext_stream(self):
        if getattr(self, "_ext", None) is None:
            self._ext = self.torch.cuda.ExternalStream(self.L.bamcl_ctx_stream(self.pf.ctx.h), device=self.device)
            self._t_local = self.torch.zeros(1, dtype=self.torch.float64, device=self.device)
            self._t_total = self.torch.zeros(1, dtype=self.torch.float64, device=self.device)
            self._t_flat = None
        return self._ext

    def update_dev(self, dist, world):
        """returns the global total (one device->host read at the end); None if this engine scores clouds"""
        if getattr(self, "cloud", None) is not None:
            return None
        t = self.torch
        with t.cuda.stream(self.ext_stream()):
            if self._t_flat is None or self._t_flat.numel() != world:
                self._t_flat = t.zeros(world, dtype=t.float64, device=self.device)
            cabi._check(self.L.bamcl_pf_score_planar_staged_dev(self.pf.h, self.map.h, int(self.pf.converged),
                                                                C.c_void_p(self._t_local.data_ptr())))
            if hasattr(dist, "all_gather_into_tensor") and dist.get_backend() == "nccl":
                dist.all_gather_into_tensor(self._t_flat, self._t_local)
            else:
                dist.all_gather([self._t_flat[r:r + 1] for r in range(world)], self._t_local)
            cabi._check(self.L.bamcl_pf_normalise_ranks_dev(self.pf.h, C.c_void_p(self._t_flat.data_ptr()), world,
                                                            C.c_void_p(self._t_total.data_ptr())))
            return float(self._t_total.item())

    def skip_counts(self):
        """this window's per-beam agreement counts (beam skipping), empty when the update does none"""
        if getattr(self, "cloud", None) is not None:
            return np.zeros(0, np.int32)
        out = np.zeros(4096, np.int32)
        n = C.c_int()
        cabi._check(self.L.bamcl_pf_skip_counts(self.pf.h, self.map.h, int(self.pf.converged), cabi._p(out), len(out),
                                                C.byref(n)))
        return out[:n.value].copy()

    def set_skip_counts(self, counts, total_n):
        c = np.ascontiguousarray(counts, np.int32)
        cabi._check(self.L.bamcl_pf_set_skip_counts(self.pf.h, cabi._p(c), len(c), total_n))

    def normalise(self, total):
        self.pf.normalise(total)

    def shard_cdf(self, c0):
        end = C.c_double()
        cabi._check(self.L.bamcl_pf_shard_cdf(self.pf.h, c0, C.byref(end)))
        return end.value

    def put_random(self, poses):
        if len(poses):
            self.buf[:len(poses)].copy_(self.torch.from_numpy(np.ascontiguousarray(poses, np.float64)))

    def select(self, start, delta, n_sys, row0, kb, ke, packed=False):
        """outputs [kb, ke) -> rows row0 + k of the global buffer, or of this rank's packed send buffer"""
        buf, src = (self.pack, self.pack_src) if packed else (self.buf, self.src)
        srcp = C.c_void_p(src.data_ptr()) if src is not None else None
        cabi._check(self.L.bamcl_pf_shard_select(self.pf.h, start, delta, n_sys, row0, kb, ke,
                                                 C.c_void_p(buf.data_ptr()), srcp))

    def reserve_packed(self, maxc, world):
        t = self.torch
        if self.pack is None or self.pack.shape[0] < maxc:
            self.pack = t.empty((maxc, 3), dtype=t.float64, device=self.device)
            self.pack_src = t.empty((maxc,), dtype=t.int32, device=self.device) if self.src is not None else None
        if self.gathered is None or self.gathered.shape[0] < world * maxc:
            self.gathered = t.empty((world * maxc, 3), dtype=t.float64, device=self.device)
            self.gathered_src = t.empty((world * maxc,), dtype=t.int32, device=self.device) if self.src is not None else None

    def gather_packed(self, dist, maxc, world):
        """every rank's packed rows -> gathered[r * maxc ...] (one all-gather instead of one broadcast per run)"""
        pairs = [(self.gathered, self.pack)] + ([(self.gathered_src, self.pack_src)] if self.src is not None else [])
        for out, inp in pairs:
            out, inp = out[:world * maxc], inp[:maxc]
            if getattr(dist, "get_backend", None) and dist.get_backend() == "nccl":
                dist.all_gather_into_tensor(out, inp)
            else:
                dist.all_gather([out[r * maxc:(r + 1) * maxc] for r in range(world)], inp)

    def unpack(self, r, maxc, off, row, cnt):
        self.buf[row:row + cnt].copy_(self.gathered[r * maxc + off:r * maxc + off + cnt])
        if self.src is not None:
            self.src[row:row + cnt].copy_(self.gathered_src[r * maxc + off:r * maxc + off + cnt])

    def before_exchange(self):
        self.pf.ctx.sync()  # selections were enqueued on the library's stream; NCCL runs on torch's

    def after_exchange(self):
        self.torch.cuda.current_stream(self.device).synchronize()

    def install(self, n, prm):
        res = cabi.ResampleResult()
        cabi._check(self.L.bamcl_pf_shard_install(self.pf.h, C.c_void_p(self.buf.data_ptr()), n, C.byref(prm), C.byref(res)))
        return res


class ShardedFilter:
    """ParticleFilter::updateSensor / updateResample over a rank-major sharded set. `dist` is torch.distributed (an
    initialised default group) or None for a single rank."""

    def __init__(self, engine, rank, world, dist, min_samples, max_samples, alpha_slow=0.001, alpha_fast=0.1,
                 conv_thresh=95.0, pop_err=0.01, pop_z=0.99):
        self.e, self.rank, self.world, self.dist = engine, rank, world, dist
        self.min_samples, self.max_samples = min_samples, max_samples
        self.alpha_slow, self.alpha_fast, self.conv_thresh = alpha_slow, alpha_fast, conv_thresh
        self.pop_err, self.pop_z = pop_err, pop_z
        self.w_slow = self.w_fast = 0.0
        self.drand48_state = 0
        self.random_poses = np.zeros((0, 3))
        self.random_next = 0
        self.last = None
        self.last_plan = None
        self.converged = False
        self.exchange = "allgather"  # or "broadcast": one broadcast per run of output slots
        self.last_exchange = None

    def srand48(self, seed):
        self.drand48_state = ((seed & 0xFFFFFFFF) << 16) | 0x330E

    def set_random_poses(self, poses):
        self.random_poses = np.ascontiguousarray(poses, np.float64)
        self.random_next = 0

    # -- windows
    def window(self):
        b = shard_bounds(self.e.sample_count(), self.world)
        return b[self.rank], b[self.rank + 1] - b[self.rank]

    def apply_window(self):
        off, cnt = self.window()
        self.e.set_window(off, cnt)

    # -- exchanges
    def _all_gather_scalar(self, v):
        if self.dist is None or self.world == 1:
            return [v]
        t = self.e.scalar_tensor(v)
        flat = self.e.scalar_tensor().repeat(self.world)
        self.dist.all_gather([flat[r:r + 1] for r in range(self.world)], t)
        return [float(x) for x in flat.tolist()]  # one device->host copy

    def _chain_cdf(self):
        """bounds[0..G]: rank r's CDF starts at the exact end value of rank r-1's (particle_filter.cpp:280-285)"""
        bounds = [0.0]
        for r in range(self.world):
            if r == self.rank:
                end = self.e.shard_cdf(bounds[r])
            else:
                end = 0.0
            if self.dist is not None and self.world > 1:
                t = self.e.scalar_tensor(end)
                self.dist.broadcast(t, src=r)
                end = float(t.item())
            bounds.append(end)
        return bounds

    # -- ParticleFilter::updateSensor (particle_filter.cpp:223-267) with PlanarScanner::applyModelToSampleSet
    def _all_reduce_counts(self, counts):
        if self.dist is None or self.world == 1:
            return counts
        import torch
        t = torch.from_numpy(np.ascontiguousarray(counts, np.int32)).to(self.e.scalar_tensor().device)
        parts = [torch.empty_like(t) for _ in range(self.world)]
        self.dist.all_gather(parts, t)
        return torch.stack(parts).sum(0).to(torch.int32).cpu().numpy()

    def update_sensor(self):
        self.apply_window()
        if hasattr(self.e, "skip_counts"):
            # beam skipping (prob model, converged set): per-beam agreement counts summed over the ranks
            counts = self.e.skip_counts()
            if len(counts):
                self.e.set_skip_counts(self._all_reduce_counts(counts), self.e.sample_count())
        total = None
        if hasattr(self.e, "update_dev") and self.dist is not None and self.world > 1:
            total = self.e.update_dev(self.dist, self.world)
        if total is None:
            local = self.e.score()
            totals = self._all_gather_scalar(local)
            total = 0.0
            for t in totals:  # rank order = sample order of the partial sums
                total += t
            self.e.normalise(total)
        if total > 0.0:
            w_avg = total / self.e.sample_count()
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
        tm = [time.perf_counter()] if os.environ.get("BAMCL_SHARD_TIMING") else None

        def mark():
            if tm is not None:
                self.e.before_exchange()
                tm.append(time.perf_counter())

        self.apply_window()
        bounds = self._chain_cdf()
        mark()
        # :291-306
        new_count = cabi.lib().bamcl_resample_limit(self.e.leaf_count(), self.min_samples, self.max_samples,
                                                   self.pop_err, self.pop_z)
        if w_diff > 0.0:
            new_count = int(new_count * (1.0 + w_diff))
            new_count = min(new_count, self.max_samples)
        n_rand = int(w_diff * new_count)
        n_sys = new_count - n_rand
        rp = self.random_poses[self.random_next:self.random_next + n_rand]
        if len(rp) < n_rand:
            raise RuntimeError("%d random poses needed, %d supplied" % (n_rand, len(rp)))
        self.e.put_random(rp)
        # :309-310
        state = drand48_next(self.drand48_state)
        start = float(np.ldexp(np.float64(state), -48))
        delta = 1.0 / n_sys if n_sys > 0 else 0.0
        runs = systematic_plan(start, delta, n_sys, bounds) if n_sys > 0 else []
        # exchange: every rank needs the whole new set (the pose-bin tree is global). Balanced plans use ONE all-gather
        # of packed per-rank rows; a plan where one rank sources most outputs (a converged filter) would pad that to
        # world x its size, so it falls back to one broadcast per run.
        counts = [0] * self.world
        for kb, ke, r in runs:
            counts[r] += ke - kb
        maxc = max(counts) if counts else 0
        gather = (self.dist is not None and self.world > 1 and self.exchange == "allgather" and maxc > 0 and
                  maxc <= 2 * (-(-n_sys // self.world)) + 1024)
        self.last_exchange = "allgather" if gather else "broadcast"
        if gather:
            self.e.reserve_packed(maxc, self.world)
            off = 0
            for kb, ke, r in runs:
                if r == self.rank:
                    self.e.select(start, delta, n_sys, off - kb, kb, ke, packed=True)
                    off += ke - kb
            self.e.before_exchange()
            mark()
            self.e.gather_packed(self.dist, maxc, self.world)
            offs = [0] * self.world
            for kb, ke, r in runs:
                self.e.unpack(r, maxc, offs[r], n_rand + kb, ke - kb)
                offs[r] += ke - kb
        else:
            for kb, ke, r in runs:
                if r == self.rank:
                    self.e.select(start, delta, n_sys, n_rand, kb, ke)
            self.e.before_exchange()
            mark()
            if self.dist is not None and self.world > 1:
                for kb, ke, r in runs:
                    self.dist.broadcast(self.e.buf[n_rand + kb:n_rand + ke], src=r)
                    if self.e.src is not None:
                        self.dist.broadcast(self.e.src[n_rand + kb:n_rand + ke], src=r)
        self.e.after_exchange()
        mark()
        prm = cabi.ResampleParams(1, self.min_samples, self.max_samples, self.pop_err, self.pop_z, w_diff, 0.5,
                                  self.conv_thresh, self.drand48_state, None, 0)
        res = self.e.install(new_count, prm)
        mark()
        if tm is not None and self.rank == 0:
            d = [1e3 * (b - a) for a, b in zip(tm, tm[1:])]
            print("[shard timing ms] cdf-chain %.3f plan+select %.3f exchange %.3f install+tree %.3f runs %d" %
                  (d[0], d[1], d[2], d[3], len(runs)), flush=True)
        res.random_poses_used = n_rand
        res.drand48_state = state
        self.drand48_state = state
        self.random_next += n_rand
        if w_diff > 0.0:
            self.w_slow = self.w_fast = 0.0
        self.converged = bool(res.converged)
        if hasattr(self.e, "pf"):
            self.e.pf.converged = self.converged
        self.last, self.last_plan, self.last_bounds = res, runs, bounds
        return res
