"""GPU parity of the distributed filter: G ranks (threads of this process, one context + filter each, all
on cuda:0, joined by the library's in-process communicator) run the distributed update + resample of the C library;
the result must be the single-set result -- the oracle's (bit-identical indices, poses, leaf / node / cluster counts;
statistics to 1e-6) and the single-GPU device filter's. tests/test_gpu_nccl.py runs the same over real NCCL ranks."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from badger_amcl_b200 import synth  # noqa: E402

pytestmark = pytest.mark.gpu
RES = 0.05


@pytest.fixture(scope="module")
def env(oracle_built):
    from badger_amcl_b200 import cabi
    ctx = cabi.Context(0)
    cells = synth.make_grid(1000, 1000, seed=7)
    m = cabi.Map2D(ctx, cells, RES)
    m.build_distances(2.0)
    dist = m.download_distances()
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    yield dict(O=oracle_built, cabi=cabi, ctx=ctx, cells=cells, truth=truth, map=m, dist=dist)
    m.close()
    ctx.close()


def _rank_map(env, ctx):
    """every rank holds its own replica of the map (as one process per GPU would)"""
    cabi = env["cabi"]
    m = cabi.Map2D(ctx, env["cells"], RES)
    m.upload_distances(env["dist"], 2.0)
    return m


@pytest.mark.parametrize("slow_cap", [None, "1"])
@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("kind", ["uniform", "heavy", "zeros_first_rank", "dominant"])
def test_distributed_cdf_is_the_sequential_cdf(env, world, kind, slow_cap, monkeypatch):
    """particle_filter.cpp:280-285 across windows: concatenated window CDFs == the oracle's c[0..P], and the exact
    value at every rank boundary; slow_cap = 1 starts with one-tile summaries and forces the regrow path"""
    from badger_amcl_b200 import sharded
    from sharded_support import run_ranks
    O, cabi = env["O"], env["cabi"]
    if slow_cap:
        monkeypatch.setenv("BAMCL_CDF_SLOW_CAP", slow_cap)
    n = 50001
    rng = np.random.default_rng(world + len(kind))
    w = {"uniform": np.full(n, 1.0 / n), "heavy": np.exp(rng.normal(0.0, 6.0, n)), "zeros_first_rank": rng.random(n),
         "dominant": rng.random(n) * 1e-12}[kind]
    if kind == "zeros_first_rank":
        w[: n // world + 5] = 0.0
    if kind == "dominant":
        w[n // 2] = 1.0
    w = w / w.sum()
    aos = np.zeros((n, 4))
    aos[:, 3] = w
    port = O.PortPF(1, n)
    port.set_samples(aos)
    want = port.cdf()
    b = sharded.shard_bounds(n, world)

    def rank_fn(rank, group):
        ctx = cabi.Context(0)
        pf = cabi.ParticleFilter(ctx, 1, n)
        comm = sharded.Comm.local(ctx, group, rank, world)
        f = sharded.ShardedFilter(pf, comm)
        f.set_global_samples(aos)
        off, cnt = f.window()
        c = np.empty(cnt + 1)
        cabi._check(pf.L.bamcl_pf_cdf(pf.h, cabi._p(c), cnt + 1))
        bounds = f.bounds()
        pf.close()
        comm.close()
        ctx.close()
        return off, cnt, c, bounds

    got = run_ranks(world, rank_fn)
    for r, (off, cnt, c, bounds) in enumerate(got):
        assert (off, cnt) == (b[r], b[r + 1] - b[r])
        assert np.array_equal(c, want[off:off + cnt + 1]), "rank %d window CDF" % r
        assert np.array_equal(bounds, want[b]), "rank %d bounds" % r


def _run_distributed(env, world, poses, weights_per_step, seed, n_min):
    from badger_amcl_b200 import sharded
    from sharded_support import run_ranks
    cabi = env["cabi"]
    n = len(poses)
    aos0 = np.concatenate([poses, np.full((n, 1), 1.0 / n)], axis=1)

    def rank_fn(rank, group):
        ctx = cabi.Context(0)  # one context (stream) per rank, as in the one-process-per-GPU deployment
        pf = cabi.ParticleFilter(ctx, n_min, n)
        comm = sharded.Comm.local(ctx, group, rank, world)
        f = sharded.ShardedFilter(pf, comm)
        f.set_global_samples(aos0)
        init = f.init_stats()
        f.srand48(seed)
        out = [dict(res=init)]
        for w in weights_per_step:
            # staged parity: every rank installs the oracle's normalised weights of ITS window
            off, cnt = f.window()
            pf.upload_weights(w[off:off + cnt])
            f.w_slow = f.w_fast = 1.0
            res = f.update_resample()
            off, cnt = f.window()
            out.append(dict(res=res, off=off, cnt=cnt, src=f.window_sources(), samples=f.download_window(),
                            labels=f.window_clusters(), stats=np.array(list(res.mean) + list(res.cov))))
        pf.close()
        comm.close()
        ctx.close()
        return out

    return run_ranks(world, rank_fn)


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("n,cloud", [(5000, "global"), (100000, "tracking"), (100000, "global"), (7, "global")])
def test_distributed_resample_equals_the_oracles_single_set(env, world, n, cloud):
    O = env["O"]
    cells, truth = env["cells"], env["truth"]
    poses = (synth.uniform_free_poses(cells, RES, n, seed=1) if cloud == "global"
             else synth.gaussian_poses(truth, n, seed=2))
    rng = np.random.default_rng(n + world)
    port = O.PortPF(max(1, n // 10), n)
    port.init_with_poses(poses)
    init_want = (port.leaf_count(), port.node_count(), port.cluster_count())
    port.srand48(7)
    weights, want = [], []
    for _ in range(2):
        m = port.sample_count()
        w = np.exp(rng.normal(0.0, 2.0, m))
        w /= w.sum()
        s = port.samples()
        s[:, 3] = w
        port.set_samples(s)
        port.set_averages(1.0, 1.0)
        port.update_resample()
        weights.append(w)
        want.append(dict(count=port.sample_count(), src=port.resample_sources(), samples=port.samples(),
                         leaf=port.leaf_count(), nodes=port.node_count(), clusters=port.cluster_count(),
                         labels=port.sample_clusters(), stats=port.set_stats(), conv=port.is_converged()))
    got = _run_distributed(env, world, poses, weights, 7, max(1, n // 10))
    for rank in range(world):
        r0 = got[rank][0]["res"]
        assert (r0.leaf_count, r0.node_count, r0.cluster_count) == init_want
        for step in range(2):
            g, w_ = got[rank][step + 1], want[step]
            r = g["res"]
            sl = slice(g["off"], g["off"] + g["cnt"])
            assert r.sample_count == w_["count"]
            assert np.array_equal(g["src"], w_["src"][sl])
            assert np.array_equal(g["samples"][:, :3], w_["samples"][sl, :3])
            assert np.array_equal(g["samples"][:, 3], w_["samples"][sl, 3])
            assert (r.leaf_count, r.node_count, r.cluster_count) == (w_["leaf"], w_["nodes"], w_["clusters"])
            assert bool(r.converged) == w_["conv"]
            assert np.array_equal(g["labels"], w_["labels"][sl])
            assert np.allclose(g["stats"], w_["stats"], rtol=1e-6, atol=1e-9)
    # every sample is owned exactly once
    for step in range(2):
        assert sum(got[rank][step + 1]["cnt"] for rank in range(world)) == want[step]["count"]


def test_distributed_resample_with_random_pose_injection(env):
    """w_diff > 0 (particle_filter.cpp:297-324): the first n_rand outputs are the caller's random poses, wherever their
    owner sits"""
    O = env["O"]
    n, world = 20000, 3
    cells = env["cells"]
    poses = synth.uniform_free_poses(cells, RES, n, seed=1)
    rp = synth.uniform_free_poses(cells, RES, n, seed=9)
    rng = np.random.default_rng(3)
    w = np.exp(rng.normal(0.0, 2.0, n))
    w /= w.sum()
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    port.srand48(11)
    s = port.samples()
    s[:, 3] = w
    port.set_samples(s)
    port.set_averages(1.0, 0.7)  # w_slow, w_fast -> w_diff = 0.3
    port.set_random_poses(rp)
    port.update_resample()
    from badger_amcl_b200 import sharded
    from sharded_support import run_ranks
    cabi = env["cabi"]
    aos0 = np.concatenate([poses, np.full((n, 1), 1.0 / n)], axis=1)

    def rank_fn(rank, group):
        ctx = cabi.Context(0)
        pf = cabi.ParticleFilter(ctx, n // 10, n)
        comm = sharded.Comm.local(ctx, group, rank, world)
        f = sharded.ShardedFilter(pf, comm)
        f.set_global_samples(aos0)
        f.init_stats()
        f.srand48(11)
        f.set_random_poses(rp)
        off, cnt = f.window()
        pf.upload_weights(w[off:off + cnt])
        f.w_slow, f.w_fast = 1.0, 0.7
        res = f.update_resample()
        off, cnt = f.window()
        out = (res, off, cnt, f.window_sources(), f.download_window())
        pf.close()
        comm.close()
        ctx.close()
        return out

    got = run_ranks(world, rank_fn)
    want_s, want_src = port.samples(), port.resample_sources()
    for res, off, cnt, src, smp in got:
        assert res.sample_count == port.sample_count() and res.random_poses_used > 0
        assert np.array_equal(src, want_src[off:off + cnt])
        assert np.array_equal(smp, want_s[off:off + cnt])
        assert (res.leaf_count, res.node_count, res.cluster_count) == (port.leaf_count(), port.node_count(),
                                                                      port.cluster_count())


def test_distributed_update_matches_the_single_gpu_filter(env):
    """score + global total + normalise over 3 windows == one filter over the whole set (weights to 1e-12: only the
    order of the partial sums of the total differs)"""
    from badger_amcl_b200 import sharded
    from sharded_support import run_ranks
    cabi = env["cabi"]
    n, world = 30000, 3
    cells, truth = env["cells"], env["truth"]
    poses = synth.uniform_free_poses(cells, RES, n, seed=4)
    aos0 = np.concatenate([poses, np.full((n, 1), 1.0 / n)], axis=1)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=180)
    model = cabi.planar_model(model_type=cabi.MODEL_LF, max_beams=180, scanner_pose=(0.1, 0.0, 0.0))
    one = cabi.ParticleFilter(env["ctx"], n // 10, n)
    one.init_with_poses(poses)
    one.update_sensor(env["map"], model, ranges, angles, 30.0)
    want = one.weights()
    one.close()

    def rank_fn(rank, group):
        ctx = cabi.Context(0)
        m = _rank_map(env, ctx)
        pf = cabi.ParticleFilter(ctx, n // 10, n)
        comm = sharded.Comm.local(ctx, group, rank, world)
        f = sharded.ShardedFilter(pf, comm, m)
        f.set_global_samples(aos0)
        pf.stage_scan(m, model, ranges, angles, 30.0)
        total = f.update_sensor()
        off, cnt = f.window()
        w = pf.weights_n(cnt)
        pf.close()
        comm.close()
        m.close()
        ctx.close()
        return total, off, cnt, w

    got = run_ranks(world, rank_fn)
    assert len(set(g[0] for g in got)) == 1, "every rank must see the same global total"
    w = np.concatenate([g[3] for g in got])
    assert np.allclose(w, want, rtol=1e-12, atol=0.0)
    assert abs(w.sum() - 1.0) < 1e-9


@pytest.mark.parametrize("case", ["mask", "error_all_beams"])
@pytest.mark.parametrize("world", [2, 3])
def test_distributed_beam_skipping_matches_the_single_gpu_filter(env, world, case):
    """SURVEY 8e collective (2): per-beam agreement counts summed over the ranks' windows, the skip ratio against the
    whole set (planar_scanner.cpp:398-451, 483-494). The single-GPU device path is the one pinned to the oracle in
    tests/test_gpu_score.py; every rank's window must carry exactly its weights (same mask => same arithmetic)."""
    from badger_amcl_b200 import sharded
    from sharded_support import run_ranks
    cabi = env["cabi"]
    n, n_beams = 6000, 180
    cells, truth = env["cells"], env["truth"]
    kw = dict(model_type=cabi.MODEL_LF_PROB, do_beamskip=True, beam_skip_distance=0.5, beam_skip_threshold=0.3,
              beam_skip_error_threshold=0.9, scanner_pose=(0.1, 0.0, 0.0), max_beams=n_beams)
    if case == "error_all_beams":
        kw["beam_skip_threshold"] = 0.995
        poses = synth.uniform_free_poses(cells, RES, n, seed=3)
    else:
        # half of the cloud sits on the true pose, half is scattered: the counts differ from window to window
        poses = np.concatenate([synth.gaussian_poses(truth, n // 2, seed=3, sigma_xy=0.05, sigma_th=0.01),
                                synth.uniform_free_poses(cells, RES, n - n // 2, seed=4)])
    aos0 = np.concatenate([poses, np.full((n, 1), 1.0 / n)], axis=1)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=n_beams)
    rng = np.random.default_rng(9)
    for i in rng.choice(n_beams, 25, replace=False):
        ranges[i] = min(ranges[i], 29.0) * 0.45
    model = cabi.planar_model(**kw)
    one = cabi.ParticleFilter(env["ctx"], n // 10, n)
    one.init_with_poses(poses)
    one.set_exact_terms(True)
    one.converged = True
    one.update_sensor(env["map"], model, ranges, angles, 30.0)
    want, want_total = one.weights(), one.last_total
    plain = cabi.ParticleFilter(env["ctx"], n // 10, n)
    plain.init_with_poses(poses)
    plain.set_exact_terms(True)
    plain.update_sensor(env["map"], cabi.planar_model(**dict(kw, do_beamskip=False)), ranges, angles, 30.0)
    if case == "mask":
        assert plain.last_total != want_total, "the mask must really drop beams in this case"
    one.close()
    plain.close()

    def rank_fn(rank, group):
        ctx = cabi.Context(0)
        m = _rank_map(env, ctx)
        pf = cabi.ParticleFilter(ctx, n // 10, n)
        comm = sharded.Comm.local(ctx, group, rank, world)
        f = sharded.ShardedFilter(pf, comm, m)
        f.set_global_samples(aos0)
        pf.set_exact_terms(True)
        f.converged = True
        pf.stage_scan(m, model, ranges, angles, 30.0)
        total = f.update_sensor()
        off, cnt = f.window()
        w = pf.weights_n(cnt)
        pf.close()
        comm.close()
        m.close()
        ctx.close()
        return total, w

    got = run_ranks(world, rank_fn)
    assert len(set(g[0] for g in got)) == 1
    assert abs(got[0][0] - want_total) <= 1e-12 * abs(want_total)
    w = np.concatenate([g[1] for g in got])
    assert np.allclose(w, want, rtol=1e-12, atol=0.0)


def test_distributed_motion_update_matches_the_single_gpu_filter(env):
    """Odom::updateAction (odom.cpp:74-311) on a distributed filter: same drand48 stream position, every window moved
    exactly as the single-GPU filter moves those samples"""
    from badger_amcl_b200 import sharded
    from sharded_support import run_ranks
    cabi = env["cabi"]
    n, world = 20001, 3
    poses = synth.uniform_free_poses(env["cells"], RES, n, seed=6)
    aos0 = np.concatenate([poses, np.full((n, 1), 1.0 / n)], axis=1)
    odo = (cabi.ODOM_DIFF_CORRECTED, [0.2, 0.3, 0.25, 0.15, 0.1], [1.0, 2.0, 0.7], [0.3, 0.1, 0.2])
    one = cabi.ParticleFilter(env["ctx"], n // 10, n)
    one.upload_samples(aos0)
    one.srand48(77)
    one.update_action(*odo)
    want, want_state = one.samples(), one.drand48_state
    one.close()

    def rank_fn(rank, group):
        ctx = cabi.Context(0)
        pf = cabi.ParticleFilter(ctx, n // 10, n)
        comm = sharded.Comm.local(ctx, group, rank, world)
        f = sharded.ShardedFilter(pf, comm)
        f.set_global_samples(aos0)
        pf.srand48(77)
        pf.update_action(*odo)
        off, cnt = f.window()
        out = (off, cnt, f.download_window(), pf.drand48_state)
        pf.close()
        comm.close()
        ctx.close()
        return out

    for off, cnt, smp, state in run_ranks(world, rank_fn):
        assert state == want_state
        assert np.array_equal(smp, want[off:off + cnt])


def test_distributed_resample_falls_back_to_replicated_bins_in_a_huge_world(env, monkeypatch):
    """a world whose pose bins do not fit the dense grid (forced here with BAMCL_TREE_HASH): every rank receives the whole
    set and builds the hashed tree as a single GPU would -- same results"""
    monkeypatch.setenv("BAMCL_TREE_HASH", "1")
    O = env["O"]
    n, world = 30000, 3
    poses = synth.uniform_free_poses(env["cells"], RES, n, seed=1)
    rng = np.random.default_rng(17)
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    init_want = (port.leaf_count(), port.node_count(), port.cluster_count())
    port.srand48(7)
    w = np.exp(rng.normal(0.0, 2.0, n))
    w /= w.sum()
    s = port.samples()
    s[:, 3] = w
    port.set_samples(s)
    port.set_averages(1.0, 1.0)
    port.update_resample()
    got = _run_distributed(env, world, poses, [w], 7, n // 10)
    for rank in range(world):
        r0 = got[rank][0]["res"]
        assert (r0.leaf_count, r0.node_count, r0.cluster_count) == init_want
        g = got[rank][1]
        r = g["res"]
        sl = slice(g["off"], g["off"] + g["cnt"])
        assert r.sample_count == port.sample_count()
        assert np.array_equal(g["src"], port.resample_sources()[sl])
        assert np.array_equal(g["samples"], port.samples()[sl])
        assert (r.leaf_count, r.node_count, r.cluster_count) == (port.leaf_count(), port.node_count(), port.cluster_count())
        assert np.array_equal(g["labels"], port.sample_clusters()[sl])
        assert np.allclose(g["stats"], port.set_stats(), rtol=1e-6, atol=1e-9)
