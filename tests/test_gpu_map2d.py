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
