"""GPU parity: K-a / K-b scoring kernels vs the oracle on the same seeded inputs.

Tolerance (BASELINE.json north_star): per-particle log-weights within 1e-5 relative. The kernels are far
tighter in practice (identical cells, host-built per-cell terms): in exact-term mode a 1e-9 relative bound on the
weights themselves is asserted as well (it would catch a flipped cell look-up); in the default fixed-point mode
the bound is the worst case of the 32-bit term quantisation, 2e-5 (typically 1e-7)."""
import numpy as np
import pytest

from badger_amcl_b200 import synth

pytestmark = pytest.mark.gpu
RES = 0.05
LOGW_RTOL = 1e-5


def assert_weights_close(got, want, rtol=1e-9):
    pos = want > 0
    assert np.array_equal(got > 0, pos)
    assert np.allclose(got, want, rtol=rtol, atol=0.0)
    lg, lw = np.log(got[pos]), np.log(want[pos])
    assert (np.abs(lg - lw) <= LOGW_RTOL * np.abs(lw) + 1e-300).all()


@pytest.fixture(scope="module")
def env(oracle_built):
    from badger_amcl_b200 import cabi
    O = oracle_built
    ctx = cabi.Context(0)
    cells = synth.make_grid(2000, 2000, seed=7)
    pm = O.PortMap2D(cells, RES)
    exact = pm.exact_edt(2.0)
    pm.set_distances(exact, 2.0)  # oracle and device score against the same cache
    m = cabi.Map2D(ctx, cells, RES)
    m.upload_distances(exact, 2.0)
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    yield dict(O=O, cabi=cabi, ctx=ctx, cells=cells, pm=pm, m=m, truth=truth)
    m.close()
    ctx.close()


def _samples(env, n, seed=1):
    glob = synth.uniform_free_poses(env["cells"], RES, n // 2, seed=seed)
    track = synth.gaussian_poses(env["truth"], n - n // 2, seed=seed + 1)
    poses = np.concatenate([glob, track])
    poses[0, :2] = (500.0, 3.0)    # off the map
    poses[1, :2] = (49.9, -49.9)   # near a corner: many end points off the map
    w = np.random.default_rng(seed).uniform(0.5, 1.5, n) / n
    return np.concatenate([poses, w[:, None]], axis=1)


MODELS = {
    "lf": dict(model_type=1),
    "lf_factors": dict(model_type=1, off_map_factor=0.2, non_free_space_factor=0.3, non_free_space_radius=0.4),
    "lf_subsample": dict(model_type=1, max_beams=30),
    "lf_mount_rotated": dict(model_type=1, scanner_pose=(0.1, 0.02, 0.3)),
    "gompertz": dict(model_type=3, gompertz=(1.0, 2.0, 6.0, -0.2, 1.3, 0.05)),
    "prob": dict(model_type=2),
    "beam": dict(model_type=0),
}


@pytest.mark.parametrize("terms", ["fixed", "exact"])
@pytest.mark.parametrize("name", list(MODELS))
@pytest.mark.parametrize("n_beams", [60, 1080])
def test_apply_model_matches_oracle(env, name, n_beams, terms):
    """applyModelToSampleSet: un-normalised weights and the returned total (C1 shape: 5k x 60; and 1080 beams)"""
    O, cabi = env["O"], env["cabi"]
    n = 5000 if n_beams == 60 else 1500
    kw = dict(max_beams=n_beams, scanner_pose=(0.1, 0.0, 0.0))
    kw.update(MODELS[name])
    if name == "lf_subsample":
        kw["max_beams"] = n_beams // 2
    ranges, angles = synth.make_scan(env["cells"], RES, env["truth"], n_beams=n_beams, dropouts=0.02)
    if name != "beam":
        ranges[3] = float("nan")
    samples = _samples(env, n)
    want_total, want_w = env["pm"].apply(O.planar_model(**kw), ranges, angles, 30.0, samples)
    pf = cabi.ParticleFilter(env["ctx"], n // 10, n)
    pf.set_exact_terms(terms == "exact")
    pf.upload_samples(samples)
    got_total = pf.score_planar(env["m"], cabi.planar_model(**kw), ranges, angles, 30.0)
    got_w = pf.weights()
    exact = terms == "exact" or name == "beam"
    assert_weights_close(got_w, want_w, rtol=1e-9 if exact else 2e-5)
    assert abs(got_total - want_total) <= (1e-12 if exact else 2e-5) * abs(want_total)
    pf.close()


def test_fused_update_matches_oracle_sequence(env):
    """PlanarScanner::updateSensor x2 through the fused path: normalised weights, w_slow/w_fast"""
    O, cabi = env["O"], env["cabi"]
    n = 5000
    kw = dict(model_type=1, max_beams=60, scanner_pose=(0.1, 0.0, 0.0))
    ranges, angles = synth.make_scan(env["cells"], RES, env["truth"], n_beams=60)
    poses = synth.uniform_free_poses(env["cells"], RES, n, seed=1)
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    pf = cabi.ParticleFilter(env["ctx"], n // 10, n)
    pf.set_exact_terms(True)
    aos = np.concatenate([poses, np.full((n, 1), 1.0 / n)], axis=1)
    pf.upload_samples(aos)
    for _ in range(2):
        port.update_sensor(env["pm"], O.planar_model(**kw), ranges, angles, 30.0)
        assert pf.update_sensor(env["m"], cabi.planar_model(**kw), ranges, angles, 30.0)
        assert_weights_close(pf.weights(), port.samples()[:, 3])
        ws, wf = port.averages()
        assert abs(pf.w_slow - ws) <= 1e-12 * ws and abs(pf.w_fast - wf) <= 1e-12 * wf
    pf.close()


def test_zero_total_resets_to_uniform_and_small_max_beams(env):
    """particle_filter.cpp:258-266 and planar_scanner.cpp:128,144"""
    cabi = env["cabi"]
    n = 1000
    ranges, angles = synth.make_scan(env["cells"], RES, env["truth"], n_beams=60)
    samples = _samples(env, n)
    samples[:, 3] = 0.0
    pf = cabi.ParticleFilter(env["ctx"], 100, n)
    pf.upload_samples(samples)
    assert pf.update_sensor(env["m"], cabi.planar_model(model_type=1, max_beams=60), ranges, angles, 30.0)
    assert np.array_equal(pf.weights(), np.full(n, 1.0 / n))
    before = pf.weights()
    assert not pf.update_sensor(env["m"], cabi.planar_model(model_type=1, max_beams=1), ranges, angles, 30.0)
    assert pf.score_planar(env["m"], cabi.planar_model(model_type=1, max_beams=1), ranges, angles, 30.0) == 0.0
    assert np.array_equal(pf.weights(), before)
    # empty scan and all-max-range scan: every particle keeps p = 1
    pf.upload_samples(_samples(env, n))
    w0 = pf.weights()
    pf.score_planar(env["m"], cabi.planar_model(model_type=1, max_beams=60), np.full(60, 30.0), angles, 30.0)
    assert np.array_equal(pf.weights(), w0 * 1.0)
    pf.close()


def test_scoring_needs_a_distance_cache(env):
    cabi = env["cabi"]
    m = cabi.Map2D(env["ctx"], np.full((50, 50), -1, np.int8), RES)
    pf = cabi.ParticleFilter(env["ctx"], 10, 100)
    with pytest.raises(cabi.BamclError):
        pf.score_planar(m, cabi.planar_model(), np.ones(60), np.linspace(-1, 1, 60), 30.0)
    pf.close()
    m.close()


# ---- likelihood_field_prob with beam skipping (planar_scanner.cpp:361-529)
def _skip_case(env, n, n_beams, cloud, dropouts=0.0, blockers=0):
    ranges, angles = synth.make_scan(env["cells"], RES, env["truth"], n_beams=n_beams, dropouts=dropouts)
    rng = np.random.default_rng(n_beams + blockers)
    for i in rng.choice(n_beams, blockers, replace=False):  # "a person in front of the robot": shortened readings
        ranges[i] = min(ranges[i], 29.0) * 0.45
    poses = (synth.gaussian_poses(env["truth"], n, seed=3, sigma_xy=0.05, sigma_th=0.01) if cloud == "tracking"
             else synth.uniform_free_poses(env["cells"], RES, n, seed=3))
    w = np.random.default_rng(1).uniform(0.5, 1.5, n) / n
    return ranges, angles, np.concatenate([poses, w[:, None]], axis=1)


SKIP_KW = dict(model_type=2, do_beamskip=True, beam_skip_distance=0.5, beam_skip_threshold=0.3,
               beam_skip_error_threshold=0.9, scanner_pose=(0.1, 0.0, 0.0))


@pytest.mark.parametrize("case", ["mask", "mask_dropouts", "error_all_beams", "not_converged", "subsampled"])
def test_beam_skipping_matches_oracle(env, case):
    O, cabi = env["O"], env["cabi"]
    n, n_beams = 3000, 180
    kw = dict(SKIP_KW, max_beams=n_beams)
    cloud, dropouts, blockers, converged = "tracking", 0.0, 25, 1
    if case == "mask_dropouts":
        dropouts = 0.05
    if case == "error_all_beams":
        # no beam reaches the agreement threshold: every beam is skipped -> all are integrated (:502-510)
        cloud, blockers, kw["beam_skip_threshold"] = "global", 0, 0.995
    if case == "not_converged":
        converged = 0
    if case == "subsampled":
        kw["max_beams"] = 60
    ranges, angles, samples = _skip_case(env, n, n_beams, cloud, dropouts, blockers)
    st = O.port_lib().orc_planar_state_create()
    pf = cabi.ParticleFilter(env["ctx"], n // 10, n)
    pf.set_exact_terms(True)
    for rep in range(2):  # the second call runs with temp_obs_ already allocated
        want_total, want_w = env["pm"].apply(O.planar_model(**kw), ranges, angles, 30.0, samples, converged=converged,
                                             state=st)
        pf.upload_samples(samples)
        pf.converged = bool(converged)
        got_total = pf.score_planar(env["m"], cabi.planar_model(**kw), ranges, angles, 30.0)
        assert_weights_close(pf.weights(), want_w, rtol=1e-9)
        assert abs(got_total - want_total) <= 1e-12 * abs(want_total)
    if case in ("mask", "mask_dropouts", "subsampled"):
        # the mask really dropped beams: the same scan without skipping weighs differently
        plain_total, _ = env["pm"].apply(O.planar_model(**dict(kw, do_beamskip=False)), ranges, angles, 30.0, samples)
        assert plain_total != want_total
    O.port_lib().orc_planar_state_destroy(st)
    pf.close()


def test_beam_skipping_never_written_slots_zero_every_weight(env):
    """in the integrate-everything case a slot no scan ever wrote holds 0.0 -> log 0 -> weight 0 -> total 0"""
    O, cabi = env["O"], env["cabi"]
    n, n_beams = 2000, 120
    kw = dict(SKIP_KW, max_beams=n_beams, beam_skip_threshold=0.995)
    ranges, angles, samples = _skip_case(env, n, n_beams, "global", dropouts=0.1)
    st = O.port_lib().orc_planar_state_create()
    want_total, want_w = env["pm"].apply(O.planar_model(**kw), ranges, angles, 30.0, samples, converged=1, state=st)
    O.port_lib().orc_planar_state_destroy(st)
    assert want_total == 0.0 and not want_w.any()
    pf = cabi.ParticleFilter(env["ctx"], n // 10, n)
    pf.upload_samples(samples)
    pf.converged = True
    assert pf.score_planar(env["m"], cabi.planar_model(**kw), ranges, angles, 30.0) == 0.0
    assert not pf.weights().any()
    # a second scan that would integrate the first scan's stale entries is refused, not approximated
    ranges2, angles2, _ = _skip_case(env, n, n_beams, "global", dropouts=0.0)
    pf.upload_samples(samples)
    pf.score_planar(env["m"], cabi.planar_model(**kw), ranges2, angles2, 30.0)  # writes every slot
    pf.upload_samples(samples)
    with pytest.raises(cabi.BamclError):
        pf.score_planar(env["m"], cabi.planar_model(**kw), ranges, angles, 30.0)
    pf.close()
