"""CPU checks of the multi-GPU host logic (badger_amcl_b200/sharded.py): the output-slot plan against a brute-force
owner search, and a world_size-2 gloo run of the sharded update + resample with the checker engine (numpy + the
oracle's kd-tree) against the oracle's single-set resampler (resampleSystematic, particle_filter.cpp:269-354)."""
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
    raw = [np.exp(rng.normal(0.0, 1.5, n)) for _ in range(2)]
    return poses, raw


def _gloo_worker(rank, world, port, n, seed, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import oracle as O
    from sharded_support import CheckerEngine
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    poses, raw = _make_case(n, seed)
    eng = CheckerEngine(O, poses, raw[0], n)
    f = sharded.ShardedFilter(eng, rank, world, dist, n // 10, n)
    f.srand48(4242)
    out = []
    for step in range(2):
        eng.raw = raw[step][:eng.sample_count()]
        total = f.update_sensor()
        res = f.update_resample()
        out.append((total, res.sample_count, res.leaf_count, eng.src[:res.sample_count].numpy().copy(), eng.poses.copy(),
                    list(f.last_plan)))
    if rank == 0:
        q.put(out)
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_sharded_resample_equals_the_single_set_resample(L, oracle_built):
    import torch.multiprocessing as mp
    O = oracle_built
    n, seed, world = 6000, 3, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, world, port, n, seed, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # the oracle's single set, fed the same un-normalised weights
    poses, raw = _make_case(n, seed)
    pf = O.PortPF(n // 10, n)
    pf.init_with_poses(poses)
    pf.srand48(4242)
    for step in range(2):
        m = pf.sample_count()
        s = pf.samples()
        w = s[:, 3] * raw[step][:m]
        total = float(np.cumsum(w)[-1])
        g_total, g_count, g_leaf, g_src, g_poses, plan = got[step]
        # per-window totals are summed rank by rank, the single set sums sample by sample: equal to rounding.
        # Normalising the single set by the sharded total makes the weights bit-identical, so everything after
        # (CDF, targets, indices, pose bins) must be bit-identical too.
        assert abs(g_total - total) <= 1e-12 * total
        s[:, 3] = w / g_total
        pf.set_samples(s)
        pf.set_averages(1.0, 1.0)  # w_diff = 0, as in the sharded run (first update sets slow = fast)
        pf.update_resample()
        assert g_count == pf.sample_count()
        assert np.array_equal(g_src, pf.resample_sources())
        assert np.array_equal(g_poses, pf.samples()[:, :3])
        assert g_leaf == pf.leaf_count()
        assert len(plan) <= 2 * world + 2
