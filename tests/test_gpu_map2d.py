"""GPU parity: OccupancyMap device object vs the oracle (distance cache build/upload, calcRange)."""
import numpy as np
import pytest

from badger_amcl_b200 import synth

pytestmark = pytest.mark.gpu
RES = 0.05


@pytest.fixture(scope="module")
def env(oracle_built):
    from badger_amcl_b200 import cabi
    ctx = cabi.Context(0)
    yield dict(O=oracle_built, cabi=cabi, ctx=ctx)
    ctx.close()


@pytest.mark.parametrize("shape,max_dist", [((300, 400), 2.0), ((150, 100), 0.3), ((257, 263), 1.0)])
def test_edt_matches_exact_oracle_and_bounds_brushfire(env, shape, max_dist):
    """K-e2 builds the exact capped EDT with the reference's value formula: bit-identical to the oracle's
    exact EDT, never larger than the reference brushfire, equal to it in most cells."""
    O, cabi = env["O"], env["cabi"]
    cells = synth.make_grid(shape[1], shape[0], seed=3, wall_every=50)
    pm = O.PortMap2D(cells, RES)
    exact = pm.exact_edt(max_dist)
    brush = pm.update_distances(max_dist)
    m = cabi.Map2D(env["ctx"], cells, RES)
    m.build_distances(max_dist)
    got = m.download_distances()
    assert np.array_equal(got, exact)
    assert (got <= brush).all()
    frac = float((got == brush).mean())
    assert frac > 0.9
    m.close()


def test_upload_download_roundtrip_and_noncanonical(env):
    O, cabi = env["O"], env["cabi"]
    cells = synth.make_grid(200, 120, seed=4, wall_every=40)
    brush = O.PortMap2D(cells, RES).update_distances(2.0)
    m = cabi.Map2D(env["ctx"], cells, RES)
    m.upload_distances(brush, 2.0)
    assert np.array_equal(m.download_distances(), brush)
    bad = brush.copy()
    bad[5, 7] = 0.1234
    with pytest.raises(cabi.BamclError):
        m.upload_distances(bad, 2.0)
    m.close()


def test_empty_and_cacheless_errors(env):
    cabi = env["cabi"]
    m = cabi.Map2D(env["ctx"], np.full((10, 10), -1, np.int8), RES)
    with pytest.raises(cabi.BamclError):
        m.download_distances()
    m.build_distances(0.0)  # occupancy_map.cpp:141-145: a zero max distance builds nothing
    with pytest.raises(cabi.BamclError):
        m.download_distances()
    m.close()


def test_calc_range_matches_oracle(env):
    O, cabi = env["O"], env["cabi"]
    cells = synth.make_grid(400, 300, seed=7, wall_every=100)
    pm = O.PortMap2D(cells, RES)
    m = cabi.Map2D(env["ctx"], cells, RES)
    rng = np.random.default_rng(9)
    n = 20000
    ox, oy, oa = rng.uniform(-11, 11, n), rng.uniform(-8.5, 8.5, n), rng.uniform(-7, 7, n)
    for max_range in (0.0, 3.0, 30.0):
        want = pm.calc_ranges(ox, oy, oa, max_range)
        got, steps = m.calc_range(ox, oy, oa, max_range, with_steps=True)
        assert np.array_equal(got, want)
        assert steps == pm.last_steps
    # the reference's own known answers (test_badger_amcl.cpp:166-170)
    km = cabi.Map2D(env["ctx"], synth.reference_test_map(), RES, origin=(100 // 2 * RES, 150 // 2 * RES))
    r = km.calc_range(np.array([0.0, 0.05]), np.array([0.0, 0.0]), np.array([0.0, 1.5708]), 0.0)
    assert r[0] == 0.0
    r = km.calc_range(np.array([0.05]), np.array([0.0]), np.array([1.5708]), 0.5)
    assert abs(r[0] - 0.15) <= 4 * np.spacing(0.15)
    m.close()
    km.close()


def test_drop_in_path_without_cache_upload_against_the_reference(env):
    """The drop-in path of a user who never uploads a cache: updateDistancesLUT on the device (exact EDT) -> updateSensor,
    against the reference chain with ITS heap-order brushfire (occupancy_map.cpp:138-242). The exact EDT
    is <= the brushfire everywhere and equal in ~99 % of cells; the weights therefore differ: this test states by how much
    (the same figures at the bench maps are in profiles/r02_parity.json) instead of hiding the difference behind an
    uploaded cache. Scorer parity proper (1e-5 on log weights) is tested with the reference's cache uploaded."""
    O, cabi = env["O"], env["cabi"]
    cells = synth.make_grid(600, 500, seed=7, wall_every=100)
    pm = O.PortMap2D(cells, RES)
    brush = pm.update_distances(2.0)  # restatement of the reference brushfire, pinned bit for bit to the compiled one
    m = cabi.Map2D(env["ctx"], cells, RES)
    m.build_distances(2.0)
    dev = m.download_distances()
    assert (dev <= brush).all()
    assert (dev == brush).mean() > 0.98
    assert np.abs(dev - brush).max() <= 2.0 * RES * 1.5
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=360)
    n = 5000
    poses = synth.uniform_free_poses(cells, RES, n, seed=1)
    kw = dict(model_type=1, max_beams=360, scanner_pose=(0.1, 0.0, 0.0))
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    port.update_sensor(pm, O.planar_model(**kw), ranges, angles, 30.0)
    want = port.samples()[:, 3]
    pf = cabi.ParticleFilter(env["ctx"], n // 10, n)
    pf.init_with_poses(poses)
    pf.set_exact_terms(True)
    assert pf.update_sensor(m, cabi.planar_model(**kw), ranges, angles, 30.0)
    got = pf.weights()
    rel = np.abs(np.log(got) - np.log(want)) / np.abs(np.log(want))
    # a smaller distance never lowers a likelihood-field term: every un-normalised factor is >= the reference's
    assert rel.max() < 5e-2, "the exact cache moved a log weight by more than 5 %"
    assert np.median(rel) < 5e-3
    # ... and with the reference's cache uploaded the same call is inside the north_star tolerance
    m.upload_distances(brush, 2.0)
    pf.init_with_poses(poses)
    assert pf.update_sensor(m, cabi.planar_model(**kw), ranges, angles, 30.0)
    got = pf.weights()
    assert (np.abs(np.log(got) - np.log(want)) <= 1e-5 * np.abs(np.log(want))).all()
    pf.close()
    m.close()
