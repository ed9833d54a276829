"""Test-only helpers for the sharded filter: a thread-based stand-in for torch.distributed (several "ranks" in one
process, so one GPU can run the whole multi-rank path) and a checker-based engine (numpy + the oracle) that lets
the host logic of badger_amcl_b200/sharded.py run under gloo without a GPU. Nothing here ships."""
import threading

import numpy as np


class ThreadComm:
    def __init__(self, world):
        self.world = world
        self.bar = threading.Barrier(world)
        self.slots = [None] * world


class ThreadDist:
    """the two collectives ShardedFilter uses, over threads of one process"""

    def __init__(self, comm, rank):
        self.c, self.rank = comm, rank

    def all_gather(self, out, t):
        self.c.slots[self.rank] = t.clone()
        if t.is_cuda:  # every rank thread may be on its own stream: the clone must be complete before peers read it
            import torch
            torch.cuda.current_stream(t.device).synchronize()
        self.c.bar.wait()
        for i in range(self.c.world):
            out[i].copy_(self.c.slots[i])
        if t.is_cuda:
            import torch
            torch.cuda.current_stream(t.device).synchronize()
        self.c.bar.wait()

    def broadcast(self, t, src):
        if self.rank == src:
            self.c.slots[src] = t
        self.c.bar.wait()
        if self.rank != src:
            t.copy_(self.c.slots[src])
        self.c.bar.wait()


def run_ranks(world, fn):
    """fn(rank, dist) on `world` threads; returns the per-rank results, re-raising the first failure"""
    comm = ThreadComm(world)
    out, err = [None] * world, [None] * world

    def body(r):
        try:
            out[r] = fn(r, ThreadDist(comm, r))
        except BaseException as e:  # noqa: BLE001
            err[r] = e
            comm.bar.abort()

    ts = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    for e in err:
        if e is not None and not isinstance(e, threading.BrokenBarrierError):
            raise e
    for e in err:
        if e is not None:
            raise e
    return out


class CheckerEngine:
    """Device phases restated with numpy (sequential cumsum = the reference's CDF loop) and the oracle's kd-tree, for
    CPU tests of the exchange logic. `raw` = the un-normalised weights the scorer would produce for the global set."""

    def __init__(self, O, poses, raw, max_samples):
        import torch
        from badger_amcl_b200 import cabi
        self.O, self.torch, self.L = O, torch, cabi.lib()
        self.poses = np.array(poses, np.float64)
        self.raw = np.array(raw, np.float64)
        self.w = np.full(len(poses), 1.0 / len(poses))
        self.buf = torch.zeros((max_samples, 3), dtype=torch.float64)
        self.src = torch.full((max_samples,), -1, dtype=torch.int32)
        self.off, self.cnt = 0, len(poses)
        self._leaf = self._tree(self.poses)[0]
        self.c = None

    def _tree(self, poses):
        t = self.O.PortKDTree()
        t.insert_many(poses, 1.0)
        return t.leaf_count(), t.node_count()

    def scalar_tensor(self, v=0.0):
        return self.torch.tensor([v], dtype=self.torch.float64)

    def sample_count(self):
        return len(self.poses)

    def leaf_count(self):
        return self._leaf

    def set_window(self, off, cnt):
        self.off, self.cnt = off, cnt

    def score(self):
        sl = slice(self.off, self.off + self.cnt)
        self.w[sl] = self.w[sl] * self.raw[sl]
        return float(np.cumsum(self.w[sl])[-1]) if self.cnt else 0.0

    def normalise(self, total):
        sl = slice(self.off, self.off + self.cnt)
        self.w[sl] = self.w[sl] / total if total > 0.0 else 1.0 / len(self.poses)

    def shard_cdf(self, c0):
        sl = slice(self.off, self.off + self.cnt)
        self.c = np.cumsum(np.concatenate([[c0], self.w[sl]]))  # sequential, one rounding per element
        return float(self.c[-1])

    def put_random(self, poses):
        if len(poses):
            self.buf[:len(poses)] = self.torch.from_numpy(np.ascontiguousarray(poses))

    def select(self, start, delta, n_sys, row0, kb, ke, packed=False):
        from badger_amcl_b200 import cabi
        t = np.empty(n_sys)
        cabi._check(self.L.bamcl_systematic_targets(start, delta, n_sys, cabi._p(t)))
        j = np.searchsorted(self.c[1:], t[kb:ke], side="right")
        assert j.max() < self.cnt, "target beyond this rank's CDF"
        buf, src = (self.pack, self.pack_src) if packed else (self.buf, self.src)
        buf[row0 + kb:row0 + ke] = self.torch.from_numpy(self.poses[self.off + j])
        src[row0 + kb:row0 + ke] = self.torch.from_numpy((self.off + j).astype(np.int32))

    def reserve_packed(self, maxc, world):
        t = self.torch
        self.pack = t.zeros((maxc, 3), dtype=t.float64)
        self.pack_src = t.zeros((maxc,), dtype=t.int32)
        self.gathered = t.zeros((world * maxc, 3), dtype=t.float64)
        self.gathered_src = t.zeros((world * maxc,), dtype=t.int32)

    def gather_packed(self, dist, maxc, world):
        for out, inp in ((self.gathered, self.pack), (self.gathered_src, self.pack_src)):
            dist.all_gather([out[r * maxc:(r + 1) * maxc] for r in range(world)], inp[:maxc])

    def unpack(self, r, maxc, off, row, cnt):
        self.buf[row:row + cnt] = self.gathered[r * maxc + off:r * maxc + off + cnt]
        self.src[row:row + cnt] = self.gathered_src[r * maxc + off:r * maxc + off + cnt]

    def before_exchange(self):
        pass

    def after_exchange(self):
        pass

    def install(self, n, prm):
        from badger_amcl_b200 import cabi
        self.poses = self.buf[:n].numpy().copy()
        self.w = np.full(n, 1.0 / n)
        self.raw = np.ones(n)
        self.off, self.cnt = 0, n
        res = cabi.ResampleResult()
        res.sample_count = n
        res.leaf_count, res.node_count = self._tree(self.poses)
        self._leaf = res.leaf_count
        return res
