"""GPU parity: the OctoMap distance cache (K-e3) and the 3-D point-cloud scorer (K-c) vs the oracle.

Distance cache: the device builds the exact quantised transform; it must equal the oracle's exact transform bit
for bit, and be a lower bound of the reference's FIFO brushfire that agrees in > 99 % of the voxels.
Scorer: the oracle restatement follows tf2 / Eigen arithmetic that is not in the reference tree (PARITY UNPINNED,
SURVEY 8c); the device follows the same sequence, so weights agree to ~1e-12 except where a 1-ulp difference of
device sin/cos moves an f32 rotation entry -- bounded here by the 1e-5 log-weight tolerance of BASELINE.json."""
import numpy as np
import pytest

from badger_amcl_b200 import synth

pytestmark = pytest.mark.gpu
RES = 0.05
MAX_DIST = 0.36
LOGW_RTOL = 1e-5


@pytest.fixture(scope="module")
def env(oracle_built):
    from badger_amcl_b200 import cabi
    O = oracle_built
    ctx = cabi.Context(0)
    keys = synth.make_voxel_world(240, 200, 40, seed=3, wall_every=60)
    pm = O.PortMap3D(RES, keys, MAX_DIST)
    pm.update_distances()  # the reference's brushfire (restated; bit-exact vs the compiled reference in the CPU suite)
    m = cabi.Map3D(ctx, RES, keys, MAX_DIST)
    m.upload_distances(pm.dense())  # scorer parity runs against the SAME cache on both sides
    truth = (1.3, -0.7, 0.4)
    yield dict(O=O, cabi=cabi, ctx=ctx, keys=keys, pm=pm, m=m, truth=truth)
    m.close()
    ctx.close()


def test_bounds_match_oracle(env):
    mn, mx = env["m"].bounds()
    pmn, pmx = env["pm"].bounds()
    assert np.array_equal(mn, pmn) and np.array_equal(mx, pmx)


def test_exact_edt_matches_oracle_and_bounds_the_brushfire(env):
    cabi, O = env["cabi"], env["O"]
    keys = env["keys"]
    m = cabi.Map3D(env["ctx"], RES, keys, MAX_DIST)
    m.build_distances()
    got = m.download_distances()
    want = env["pm"].exact_edt()
    assert got.shape == want.shape
    assert np.array_equal(got, want)
    brush = env["pm"].dense()
    assert (got <= brush).all()
    assert (got == brush).mean() > 0.99
    m.close()


def test_set_bounds_crops_and_rebuilds(env):
    cabi, O = env["cabi"], env["O"]
    keys = env["keys"]
    pm = O.PortMap3D(RES, keys, MAX_DIST)
    pm.set_bounds((-2.0, -1.5), (3.0, 2.5))
    m = cabi.Map3D(env["ctx"], RES, keys, MAX_DIST)
    m.set_bounds((-2.0, -1.5), (3.0, 2.5))
    assert np.array_equal(m.bounds()[0], pm.bounds()[0]) and np.array_equal(m.bounds()[1], pm.bounds()[1])
    assert np.array_equal(m.download_distances(), pm.exact_edt())
    m.close()


def test_upload_download_round_trip(env):
    assert np.array_equal(env["m"].download_distances(), env["pm"].dense())


def _samples(env, n, seed=1):
    rng = np.random.default_rng(seed)
    mn, mx = env["pm"].bounds()
    glob = np.stack([rng.uniform(mn[0] * RES, mx[0] * RES, n // 2), rng.uniform(mn[1] * RES, mx[1] * RES, n // 2),
                     rng.uniform(-np.pi, np.pi, n // 2)], axis=1)
    track = synth.gaussian_poses(env["truth"], n - n // 2, seed=seed + 1)
    poses = np.concatenate([glob, track])
    poses[0, :2] = (500.0, 3.0)  # off the map
    poses[1, 2] = 3.1            # trace of the composed rotation <= 0: the other quaternion branches
    poses[2, 2] = -3.0
    w = rng.uniform(0.5, 1.5, n) / n
    return np.concatenate([poses, w[:, None]], axis=1)


MOUNTS = {
    "identity": dict(),
    "mounted": dict(mount_translation=(0.2, 0.0, 1.5), mount_rotation=synth.rpy_quaternion(0.0, 0.1, 0.0)),
    "tilted": dict(mount_translation=(0.1, -0.05, 1.2), mount_rotation=synth.rpy_quaternion(0.3, -0.2, 2.9)),
}


@pytest.mark.parametrize("mount", list(MOUNTS))
@pytest.mark.parametrize("model_type", [0, 1])
@pytest.mark.parametrize("n_points", [100, 4096])
def test_cloud_model_matches_oracle(env, mount, model_type, n_points):
    """applyModelToSampleSet (point_cloud_scanner.cpp:111-129): un-normalised weights and the returned total"""
    O, cabi = env["O"], env["cabi"]
    n = 2000 if n_points == 100 else 600
    kw = dict(model_type=model_type, max_beams=n_points, off_map_factor=0.3,
              gompertz=(1.0, 2.0, 6.0, -0.2, 1.3, 0.05))
    kw.update(MOUNTS[mount])
    points = synth.make_cloud(env["keys"], RES, env["truth"], n_points=n_points,
                              mount=kw.get("mount_translation", (0.0, 0.0, 1.0)))
    samples = _samples(env, n)
    want_total, want_w = env["pm"].cloud_apply(O.cloud_model(**kw), points, samples)
    pf = cabi.ParticleFilter(env["ctx"], n // 10, n)
    pf.upload_samples(samples)
    got_total = pf.score_cloud(env["m"], cabi.cloud_model(**kw), points)
    got_w = pf.weights()
    pos = want_w > 0
    assert np.array_equal(got_w > 0, pos)
    lg, lw = np.log(got_w[pos]), np.log(want_w[pos])
    assert (np.abs(lg - lw) <= LOGW_RTOL * np.abs(lw)).all()
    # nearly every particle is far tighter than the tolerance: same voxels, same table terms
    assert (np.abs(got_w - want_w) <= 1e-9 * want_w).mean() > 0.995
    assert abs(got_total - want_total) <= 1e-6 * abs(want_total)
    pf.close()


def test_fused_cloud_update_normalises(env):
    O, cabi = env["O"], env["cabi"]
    n = 1000
    kw = dict(model_type=0, max_beams=512, mount_translation=(0.2, 0.0, 1.5),
              mount_rotation=synth.rpy_quaternion(0.0, 0.1, 0.0))
    points = synth.make_cloud(env["keys"], RES, env["truth"], n_points=512)
    samples = _samples(env, n)
    want_total, want_w = env["pm"].cloud_apply(O.cloud_model(**kw), points, samples)
    pf = cabi.ParticleFilter(env["ctx"], n // 10, n)
    pf.upload_samples(samples)
    assert pf.update_sensor_cloud(env["m"], cabi.cloud_model(**kw), points)
    got = pf.weights()
    assert abs(got.sum() - 1.0) < 1e-12
    assert np.allclose(got, want_w / want_total, rtol=1e-6, atol=0.0)
    assert not pf.update_sensor_cloud(env["m"], cabi.cloud_model(max_beams=1), points)
    pf.close()


def test_cloud_scoring_needs_a_cache(env):
    cabi = env["cabi"]
    m = cabi.Map3D(env["ctx"], RES, env["keys"][:100], MAX_DIST)
    pf = cabi.ParticleFilter(env["ctx"], 10, 100)
    with pytest.raises(cabi.BamclError):
        pf.score_cloud(m, cabi.cloud_model(), np.zeros((8, 3), np.float32))
    pf.close()
    m.close()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_cloud_update_matches_the_unsharded_device_filter(env, world):
    """SURVEY 8e for the 3-D sensor: every rank scores its window of the one global set against the replicated OctoMap
    cache; totals all-gathered in rank order, one global normalisation"""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from badger_amcl_b200 import sharded
    from sharded_support import run_ranks
    cabi = env["cabi"]
    n = 3001
    kw = dict(model_type=0, max_beams=512, mount_translation=(0.2, 0.0, 1.5),
              mount_rotation=synth.rpy_quaternion(0.0, 0.1, 0.0))
    points = synth.make_cloud(env["keys"], RES, env["truth"], n_points=512)
    samples = _samples(env, n)
    model = cabi.cloud_model(**kw)
    one = cabi.ParticleFilter(env["ctx"], n // 10, n)
    one.upload_samples(samples)
    assert one.update_sensor_cloud(env["m"], model, points)
    want, want_total = one.weights(), one.last_total
    one.close()

    def rank_fn(rank, group):
        ctx = cabi.Context(0)
        m = cabi.Map3D(ctx, RES, env["keys"], MAX_DIST)  # the cache is replicated: one copy per rank
        m.upload_distances(env["pm"].dense())
        pf = cabi.ParticleFilter(ctx, n // 10, n)
        comm = sharded.Comm.local(ctx, group, rank, world)
        f = sharded.ShardedFilter(pf, comm)
        f.set_global_samples(samples)
        pf.stage_cloud(m, model, points)
        f.set_cloud(m)
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
    assert np.allclose(np.concatenate([g[1] for g in got]), want, rtol=1e-12, atol=0.0)
