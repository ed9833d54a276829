"""GPU parity of the sharded filter: G ranks (threads of this process, one filter object each, all on
cuda:0) run the sharded update + resample through the C ABI; the result must be the single-set result -- the
oracle's (bit-identical indices, poses, leaf / node / cluster counts; statistics to 1e-6) and the unsharded
device filter's."""
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
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    yield dict(O=oracle_built, cabi=cabi, ctx=ctx, cells=cells, truth=truth, map=m)
    m.close()
    ctx.close()


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("kind", ["uniform", "heavy", "zeros_first_rank", "dominant"])
def test_sharded_cdf_chain_is_the_sequential_cdf(env, world, kind):
    """particle_filter.cpp:280-285 across windows: concatenated local CDFs == the oracle's c[0..P]"""
    from badger_amcl_b200 import sharded
    O, cabi = env["O"], env["cabi"]
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
    pf = cabi.ParticleFilter(env["ctx"], 1, n)
    pf.upload_samples(aos)
    b = sharded.shard_bounds(n, world)
    c0 = 0.0
    for r in range(world):
        cabi._check(pf.L.bamcl_pf_set_window(pf.h, b[r], b[r + 1] - b[r]))
        end = cabi.C.c_double()
        cabi._check(pf.L.bamcl_pf_shard_cdf(pf.h, c0, cabi.C.byref(end)))
        assert end.value == want[b[r + 1]], "rank %d end value" % r
        c0 = end.value
    pf.close()


def _run_sharded(env, world, poses, weights_per_step, seed, n_min, exchange="allgather"):
    from badger_amcl_b200 import sharded
    from sharded_support import run_ranks
    cabi = env["cabi"]
    n = len(poses)

    def rank_fn(rank, dist):
        ctx = cabi.Context(0)  # one context (stream) per rank, as in the one-process-per-GPU deployment
        pf = cabi.ParticleFilter(ctx, n_min, n)
        pf.init_with_poses(poses)
        eng = sharded.CudaEngine(pf, env["map"], 0)
        eng.want_sources()
        f = sharded.ShardedFilter(eng, rank, world, dist, n_min, n)
        f.srand48(seed)
        f.exchange = exchange
        out = []
        for w in weights_per_step:
            m = pf.sample_count()
            # staged parity: every rank installs the oracle's normalised weights of ITS window
            off, cnt = f.window()
            full = np.zeros(m)
            full[off:off + cnt] = w[:m][off:off + cnt]
            pf.upload_weights(full)
            f.w_slow = f.w_fast = 1.0
            res = f.update_resample()
            out.append(dict(res=res, src=eng.src[:res.sample_count].cpu().numpy(), samples=pf.samples(),
                            labels=pf.sample_clusters(), stats=np.array(list(res.mean) + list(res.cov)),
                            runs=list(f.last_plan), exchange=f.last_exchange))
        pf.close()
        ctx.close()
        return out

    return run_ranks(world, rank_fn)


@pytest.mark.parametrize("exchange", ["allgather", "broadcast"])
@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("n,cloud", [(5000, "global"), (100000, "tracking"), (100000, "global")])
def test_sharded_resample_equals_the_oracles_single_set(env, world, n, cloud, exchange):
    O = env["O"]
    cells, truth = env["cells"], env["truth"]
    poses = (synth.uniform_free_poses(cells, RES, n, seed=1) if cloud == "global"
             else synth.gaussian_poses(truth, n, seed=2))
    rng = np.random.default_rng(n + world)
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
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
        weights.append(np.concatenate([w, np.zeros(n - m)]))
        want.append(dict(count=port.sample_count(), src=port.resample_sources(), samples=port.samples(),
                         leaf=port.leaf_count(), nodes=port.node_count(), clusters=port.cluster_count(),
                         labels=port.sample_clusters(), stats=port.set_stats(), conv=port.is_converged()))
    got = _run_sharded(env, world, poses, weights, 7, n // 10, exchange)
    assert got[0][0]["exchange"] == exchange  # balanced weights: the all-gather path is really taken when asked for
    for rank in range(world):
        for step in range(2):
            g, w_ = got[rank][step], want[step]
            r = g["res"]
            assert r.sample_count == w_["count"]
            assert np.array_equal(g["src"], w_["src"])
            assert np.array_equal(g["samples"][:, :3], w_["samples"][:, :3])
            assert np.array_equal(g["samples"][:, 3], w_["samples"][:, 3])
            assert (r.leaf_count, r.node_count, r.cluster_count) == (w_["leaf"], w_["nodes"], w_["clusters"])
            assert bool(r.converged) == w_["conv"]
            assert np.array_equal(g["labels"], w_["labels"])
            assert np.allclose(g["stats"], w_["stats"], rtol=1e-6, atol=1e-9)
            assert len(g["runs"]) <= 2 * world + 2


def test_sharded_update_matches_the_unsharded_device_filter(env):
    """score + global total + normalise over 3 windows == one filter over the whole set (weights to 1e-12: only the
    order of the partial sums of the total differs)"""
    from badger_amcl_b200 import sharded
    from sharded_support import run_ranks
    cabi = env["cabi"]
    n, world = 30000, 3
    cells, truth = env["cells"], env["truth"]
    poses = synth.uniform_free_poses(cells, RES, n, seed=4)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=180)
    model = cabi.planar_model(model_type=cabi.MODEL_LF, max_beams=180, scanner_pose=(0.1, 0.0, 0.0))
    one = cabi.ParticleFilter(env["ctx"], n // 10, n)
    one.init_with_poses(poses)
    one.update_sensor(env["map"], model, ranges, angles, 30.0)
    want = one.weights()
    one.close()

    def rank_fn(rank, dist):
        ctx = cabi.Context(0)
        pf = cabi.ParticleFilter(ctx, n // 10, n)
        pf.init_with_poses(poses)
        pf.stage_scan(env["map"], model, ranges, angles, 30.0)
        f = sharded.ShardedFilter(sharded.CudaEngine(pf, env["map"], 0), rank, world, dist, n // 10, n)
        total = f.update_sensor()
        off, cnt = f.window()
        w = pf.weights()[off:off + cnt]
        pf.close()
        ctx.close()
        return total, off, cnt, w

    got = run_ranks(world, rank_fn)
    assert len(set(g[0] for g in got)) == 1, "every rank must see the same global total"
    w = np.concatenate([g[3] for g in got])
    assert np.allclose(w, want, rtol=1e-12, atol=0.0)
    assert abs(w.sum() - 1.0) < 1e-9


@pytest.mark.parametrize("case", ["mask", "error_all_beams"])
@pytest.mark.parametrize("world", [2, 3])
def test_sharded_beam_skipping_matches_the_unsharded_device_filter(env, world, case):
    """SURVEY 8e collective (2): per-beam agreement counts summed over the ranks' windows, the skip ratio against the
    whole set (planar_scanner.cpp:398-451, 483-494). The unsharded device path is the one pinned to the oracle in
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

    def rank_fn(rank, dist):
        ctx = cabi.Context(0)
        pf = cabi.ParticleFilter(ctx, n // 10, n)
        pf.init_with_poses(poses)
        pf.set_exact_terms(True)
        pf.converged = True
        pf.stage_scan(env["map"], model, ranges, angles, 30.0)
        f = sharded.ShardedFilter(sharded.CudaEngine(pf, env["map"], 0), rank, world, dist, n // 10, n)
        total = f.update_sensor()
        off, cnt = f.window()
        w = pf.weights()[off:off + cnt]
        pf.close()
        ctx.close()
        return total, w

    got = run_ranks(world, rank_fn)
    assert len(set(g[0] for g in got)) == 1
    assert abs(got[0][0] - want_total) <= 1e-12 * abs(want_total)
    w = np.concatenate([g[1] for g in got])
    assert np.allclose(w, want, rtol=1e-12, atol=0.0)
