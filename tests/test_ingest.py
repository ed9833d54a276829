"""Scan / map ingestion: Node2D::convertMap (node_2d.cpp:265-295), Node2D::updateLatestScanData
(node_2d.cpp:534-560), Node3D::updateLatestScanData (node_3d.cpp:467-480) against the oracle's numpy restatements
(pinned against the compiled node: tests/test_node_pins.py). Everything is integer / copy work or one exact expression
per element: bit-exact. The two scan functions are host arithmetic of the C ABI library and run without a GPU."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def cabi_built():
    from badger_amcl_b200 import build, cabi
    build.build()
    return cabi


@pytest.mark.parametrize("smin,smax", [(-1.0, -1.0), (0.5, 20.0), (0.05, 100.0)])
def test_laserscan_conversion(oracle_built, cabi_built, smin, smax):
    O, cabi = oracle_built, cabi_built
    rng = np.random.default_rng(0)
    r = rng.uniform(0.0, 35.0, 1081).astype(np.float32)
    r[::50] = np.float32(0.1)      # at the message's range_min
    r[7] = np.nan
    r[9] = np.inf
    got = cabi.scan_from_laserscan(r, 0.1, 30.0, -2.356194, 0.004363323, smin, smax)
    want = O.scan_from_laserscan(r, 0.1, 30.0, -2.356194, 0.004363323, smin, smax)
    assert np.array_equal(got[0], want[0], equal_nan=True)
    assert np.array_equal(got[1], want[1])
    assert got[2] == want[2]
    assert cabi.scan_from_laserscan(np.zeros(0, np.float32), 0.1, 30.0, 0.0, 0.1)[2] == float(np.float32(30.0))


@pytest.mark.parametrize("count,max_beams", [(0, 30), (1, 30), (29, 30), (30, 30), (31, 30), (1000, 30), (65536, 2), (7, 1000)])
def test_cloud_subsample(oracle_built, cabi_built, count, max_beams):
    O, cabi = oracle_built, cabi_built
    xyz = np.random.default_rng(count).normal(0.0, 5.0, (count, 3)).astype(np.float32)
    assert np.array_equal(cabi.cloud_subsample(xyz, max_beams), O.cloud_subsample(xyz, max_beams))


def test_cloud_subsample_refuses_the_reference_division_by_zero(cabi_built):
    with pytest.raises(cabi_built.BamclError):
        cabi_built.cloud_subsample(np.zeros((5, 3), np.float32), 1)


@pytest.mark.gpu
@pytest.mark.parametrize("scale", [1, 2, 3])
@pytest.mark.parametrize("shape", [(1, 1), (37, 53), (400, 300)])
def test_gpu_convert_map_vs_oracle(oracle_built, scale, shape):
    from badger_amcl_b200 import cabi
    O = oracle_built
    ctx = cabi.Context(0)
    rng = np.random.default_rng(scale + shape[0])
    data = rng.choice(np.array([0, 100, -1, 50, 99, 1], np.int8), size=shape, p=[0.6, 0.2, 0.1, 0.04, 0.03, 0.03])
    m = cabi.Map2D.from_occupancy_grid(ctx, data, 0.05, (-7.3, 2.15), scale)
    cells, res, org = O.convert_map(data, 0.05, (-7.3, 2.15), scale)
    assert (m.size_y, m.size_x) == cells.shape
    assert m.resolution == res
    assert m.origin == (float(org[0]), float(org[1]))
    assert np.array_equal(m.cells(), cells)
    # the converted map is a working map: the device cache build equals the one over the oracle's cells
    if shape[0] > 1:
        m.build_distances(0.5)
        m2 = cabi.Map2D(ctx, cells, res, (float(org[0]), float(org[1])))
        m2.build_distances(0.5)
        assert np.array_equal(m.download_distances(), m2.download_distances())
        m2.close()
    m.close()
    ctx.close()
