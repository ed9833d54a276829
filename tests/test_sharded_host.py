"""CPU checks of the multi-GPU host logic (the host arithmetic of the C library that plans the distributed resample):
the output-slot plan against a brute-force owner search, the transfer list of the pose exchange, and a world_size-2
gloo run that moves numpy rows exactly as that plan says and must reproduce the oracle's single-set resampler
(resampleSystematic, particle_filter.cpp:269-354). The device phases (CDF, select, pose bins) are GPU-tested in
tests/test_gpu_sharded.py; here numpy stands in for them so the exchange logic runs without a GPU."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from badger_amcl_b200 import cabi, sharded, synth  # noqa: E402


@pytest.fixture(scope="module")
def L():
    from badger_amcl_b200 import build
    build.build()
    return cabi.lib()


def _targets(L, start, delta, n):
    t = np.empty(n)
    cabi._check(L.bamcl_systematic_targets(start, delta, n, cabi._p(t)))
    return t


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("n", [1, 7, 1000, 65537])
@pytest.mark.parametrize("case", ["even", "skewed", "empty_ranks", "short_cdf"])
def test_plan_assigns_every_slot_to_the_rank_that_holds_its_target(L, world, n, case):
    rng = np.random.default_rng(world * 1000 + n)
    if case == "even":
        cuts = np.arange(1, world) / world
    elif case == "skewed":
        cuts = np.sort(rng.random(world - 1)) ** 3
    else:
        cuts = np.sort(rng.choice([0.0, 0.25, 0.25, 0.6, 1.0], world - 1))
    top = 1.0 if case != "short_cdf" else 1.0 - 1e-9  # sum of weights rounded below 1
    bounds = np.concatenate([[0.0], np.minimum(cuts, top), [top]])
    for start in (0.0, 0.37, 1.0 - 2.0 ** -40, rng.random()):
        delta = 1.0 / n
        runs = sharded.systematic_plan(start, delta, n, bounds)
        t = _targets(L, start, delta, n)
        owner = np.searchsorted(bounds[1:], t, side="right")
        last_mass = max(r for r in range(world) if bounds[r] < bounds[r + 1])
        owner[owner >= world] = last_mass
        got = np.full(n, -1)
        prev_end = 0
        for kb, ke, r in runs:
            assert kb == prev_end and ke > kb
            got[kb:ke] = r
            prev_end = ke
        assert prev_end == n
        assert np.array_equal(got, owner)
        assert len(runs) <= 2 * world + 2


@pytest.mark.parametrize("world", [1, 2, 5])
@pytest.mark.parametrize("n,n_random", [(1, 0), (1000, 0), (1000, 333), (65537, 5)])
def test_exchange_plan_moves_every_selected_row_to_the_owner_of_its_slot(L, world, n, n_random):
    rng = np.random.default_rng(world + n)
    cuts = np.sort(rng.random(world - 1))
    bounds = np.concatenate([[0.0], cuts, [1.0]])
    runs = sharded.systematic_plan(rng.random(), 1.0 / n, n, bounds)
    moves = sharded.exchange_plan(runs, n_random, n, world)
    ob = sharded.shard_bounds(n + n_random, world)
    covered = np.zeros(n, int)
    pos_expect = [0] * world
    run_pos = {}
    for kb, ke, r in runs:
        run_pos[(kb, ke)] = pos_expect[r]
        pos_expect[r] += ke - kb
    for src, dst, kb, ke, pos in moves:
        assert ke > kb
        covered[kb:ke] += 1
        assert ob[dst] <= n_random + kb and n_random + ke <= ob[dst + 1]      # the whole piece belongs to dst
        (rb, re_), = [(a, b) for a, b, r in runs if a <= kb and ke <= b and r == src]
        assert pos == run_pos[(rb, re_)] + (kb - rb)                           # where the source rank staged it
    assert np.all(covered == 1)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make_case(n, seed):
    rng = np.random.default_rng(seed)
    cells = synth.make_grid(400, 400, seed=7, wall_every=100)
    poses = synth.uniform_free_poses(cells, 0.05, n, seed=seed)
    w = np.exp(rng.normal(0.0, 1.5, n))
    return poses, w / w.sum()


def _gloo_worker(rank, world, port, n, n_new, seed, q):
    """one rank of the distributed resample with numpy in place of the device kernels: window CDF (sequential cumsum
    continued from the previous rank's exact end), the C library's plan + transfer list, rows moved with gloo"""
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    poses, w = _make_case(n, seed)
    b = sharded.shard_bounds(n, world)
    off, cnt = b[rank], b[rank + 1] - b[rank]
    # exact chained CDF (what the device resolves from the all-gathered tile summaries)
    bounds = [0.0]
    c = None
    for r in range(world):
        end = torch.zeros(1, dtype=torch.float64)
        if r == rank:
            c = np.cumsum(np.concatenate([[bounds[r]], w[off:off + cnt]]))
            end[0] = c[-1]
        dist.broadcast(end, src=r)
        bounds.append(float(end[0]))
    state = sharded.drand48_next(((4242 & 0xFFFFFFFF) << 16) | 0x330E)
    start, delta = float(np.ldexp(np.float64(state), -48)), 1.0 / n_new
    runs = sharded.systematic_plan(start, delta, n_new, bounds)
    moves = sharded.exchange_plan(runs, 0, n_new, world)
    t = np.empty(n_new)
    cabi._check(cabi.lib().bamcl_systematic_targets(start, delta, n_new, cabi._p(t)))
    rows = []
    for kb, ke, r in runs:
        if r == rank:
            j = np.searchsorted(c[1:], t[kb:ke], side="right")
            assert j.max() < cnt
            rows.append(np.concatenate([poses[off + j], (off + j)[:, None].astype(np.float64)], axis=1))
    staged = np.concatenate(rows) if rows else np.zeros((0, 4))
    ob = sharded.shard_bounds(n_new, world)
    mine = np.full((ob[rank + 1] - ob[rank], 4), np.nan)
    reqs = []
    for src, dst, kb, ke, pos in moves:
        if src == rank and dst == rank:
            mine[kb - ob[rank]:ke - ob[rank]] = staged[pos:pos + ke - kb]
        elif src == rank:
            reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(staged[pos:pos + ke - kb])), dst))
        elif dst == rank:
            buf = torch.empty((ke - kb, 4), dtype=torch.float64)
            dist.recv(buf, src)
            mine[kb - ob[rank]:ke - ob[rank]] = buf.numpy()
    for rq in reqs:
        rq.wait()
    q.put((rank, ob[rank], mine, len(runs)))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_exchange_reproduces_the_single_set_resample(L, oracle_built):
    import torch.multiprocessing as mp
    O = oracle_built
    n, seed, world = 6000, 3, 2
    poses, w = _make_case(n, seed)
    pf = O.PortPF(n // 10, n)
    pf.init_with_poses(poses)
    pf.srand48(4242)
    s = pf.samples()
    s[:, 3] = w
    pf.set_samples(s)
    pf.set_averages(1.0, 1.0)  # w_diff = 0
    pf.update_resample()
    n_new = pf.sample_count()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, world, port, n, n_new, seed, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = np.concatenate([pf.samples()[:, :3], pf.resample_sources()[:, None].astype(np.float64)], axis=1)
    for rank, o, mine, n_runs in got:
        assert np.array_equal(mine, want[o:o + len(mine)])
        assert n_runs <= 2 * world + 2
    assert sum(len(g[2]) for g in got) == n_new
