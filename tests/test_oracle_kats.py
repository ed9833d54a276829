"""Pin the oracle (C restatement AND compiled reference) to the reference's own known-answer tests,
test/test_badger_amcl.cpp. CPU only."""
import ctypes as C
import math

import numpy as np
import pytest

from badger_amcl_b200 import synth

BACKENDS = ["port", "ref"]


def _skip_if_no_ref(O, backend):
    if backend == "ref" and not O.have_ref():
        pytest.skip("compiled reference not available")


def test_drand48_posix_kat(oracle_built):
    """POSIX LCG; glibc srand48(12345) stream (checked against glibc itself through the ref harness)."""
    O = oracle_built
    P = O.port_lib()
    r = O.Rng(0)
    P.orc_srand48(C.byref(r), 12345)
    got = [P.orc_drand48(C.byref(r)) for _ in range(2)]
    assert got == [0.22532851279629895, 0.919183068533556]
    d = synth.Drand48(12345)
    assert [d.draw(), d.draw()] == got
    if O.have_ref():
        R = O.ref_lib()
        R.ref_srand48(12345)
        assert [R.ref_drand48(), R.ref_drand48()] == got


def test_pdf_gaussian_kat(oracle_built):
    """test_badger_amcl.cpp:29-49: unseeded drand48 stream through PDFGaussian::draw; identity covariance
    => sample = mean + draw(1.0). The second block samples the FIRST pdf again (:45)."""
    O = oracle_built
    P = O.port_lib()
    r = O.Rng(0)  # glibc's state before any srand48 is X = 0 (its drand48_data is zero-initialised)
    got = [1.0 + P.orc_gaussian_draw(C.byref(r), 1.0) for _ in range(6)]
    want = [0.26562654174915334, 0.97172090090793528, -1.5856194295513539, 1.6262083813236745, 1.1142314205031041,
            0.37407538872488655]
    for g, w in zip(got, want):  # EXPECT_DOUBLE_EQ = within 4 ulp
        assert abs(g - w) <= 4 * np.spacing(abs(w))


@pytest.mark.parametrize("backend", BACKENDS)
def test_pf_kdtree_kat(oracle_built, backend):
    """test_badger_amcl.cpp:51-82"""
    O = oracle_built
    _skip_if_no_ref(O, backend)
    t = O.PortKDTree() if backend == "port" else O.RefKDTree()
    assert t.leaf_count() == 0
    pose = (1, 1, 1)
    t.insert(pose, 0.0)
    assert t.leaf_count() == 1
    t.clear()
    assert t.leaf_count() == 0
    t.insert(pose, 0.0)
    assert t.get_cluster(pose) == -1
    t.cluster()
    assert t.get_cluster(pose) == 0
    pose2, pose3 = (0, 1, 1), (3, 0, 0)
    t.insert(pose2, 0.0)
    t.insert(pose3, 0.0)
    t.cluster()
    assert [t.get_cluster(p) for p in (pose, pose2, pose3)] == [0, 1, 2]
    assert t.leaf_count() == 2
    pose4 = (0.5, 1, 1)
    t.insert(pose4, 0.0)
    t.cluster()
    assert [t.get_cluster(p) for p in (pose, pose2, pose3, pose4)] == [0, 0, 1, 0]
    assert t.leaf_count() == 2


def test_octomap_conversions_kat(oracle_built):
    """test_badger_amcl.cpp:84-111 (res 0.05, incl. negative z)"""
    O = oracle_built
    P = O.port_lib()
    out = (C.c_int * 3)()
    P.orc_map3d_world_to_map((C.c_double * 3)(.15, .25, -.05), 0.05, out)
    assert out[:] == [3, 5, -1]
    P.orc_map3d_world_to_map((C.c_double * 3)(.05, .1, 0.0), 0.05, out)
    assert out[:2] == [1, 2]
    if O.have_ref():
        m = O.RefMap3D(0.05, np.array([[0, 0, 0]]), 0.3)
        ijk = (C.c_int * 3)()
        m.L.ref_map3d_world_to_map(m.h, (C.c_double * 3)(.15, .25, -.05), ijk)
        assert ijk[:] == [3, 5, -1]
        xyz = (C.c_double * 3)()
        m.L.ref_map3d_map_to_world(m.h, (C.c_int * 3)(3, 5, -1), xyz)
        for g, w in zip(xyz[:], (.15, .25, -.05)):
            assert abs(g - w) <= 4 * np.spacing(abs(w))


@pytest.mark.parametrize("backend", BACKENDS)
def test_occupancy_map_conversions_kat(oracle_built, backend):
    """test_badger_amcl.cpp:113-129: origin 0, size 0"""
    O = oracle_built
    _skip_if_no_ref(O, backend)
    cells = np.zeros((0, 0), np.int8)
    m = (O.PortMap2D if backend == "port" else O.RefMap2D)(cells, 0.05)
    x, y = m.map_to_world(1, 2)
    assert abs(x - .05) <= 4 * np.spacing(.05) and abs(y - .1) <= 4 * np.spacing(.1)
    assert m.world_to_map(.05, .1) == (1, 2)


@pytest.mark.parametrize("backend", BACKENDS)
def test_occupancy_map_distances_kat(oracle_built, backend):
    """test_badger_amcl.cpp:131-171"""
    O = oracle_built
    _skip_if_no_ref(O, backend)
    res = 0.05
    cells = synth.reference_test_map()
    origin = (100 // 2 * res, 150 // 2 * res)
    m = (O.PortMap2D if backend == "port" else O.RefMap2D)(cells, res, origin)
    assert m.is_valid(0, 0) and not m.is_valid(-1, 5) and m.is_valid(99, 149)
    assert not m.is_valid(100, 150) and not m.is_valid(149, 99)
    m.update_distances(0.3)
    r = m.calc_range(0, 0, 0, 0)
    assert r == 0.0
    r = m.calc_range(0.05, 0, 1.5708, 0.5)
    assert abs(r - 0.15) <= 4 * np.spacing(0.15)
