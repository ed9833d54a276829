"""Global localisation on the device: Node2D::updateFreeSpaceIndices (node_2d.cpp:320-338) and
initWithPoseFn(uniformPoseGenerator) (particle_filter.cpp:136-162, node.cpp:823-868).

The oracle restates the free-space list and the pose generator; both are pinned against the node sources compiled into
oracle/_ref (tests/test_node_pins.py); initWithPoseFn itself runs in the compiled
reference (oracle/_ref) fed with the oracle's poses. GPU: free list, poses, weights, drand48 state, pose-bin tree and
cluster labels bit-exact (integer arithmetic + one exact multiply-add per coordinate); statistics to 1e-6."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from badger_amcl_b200 import synth  # noqa: E402

RES = 0.05


def _world(O, origin=(0.0, 0.0)):
    cells = synth.make_grid(300, 200, seed=7, wall_every=60)
    pm = O.PortMap2D(cells, RES, origin)
    pm.update_distances(2.0)
    return cells, pm


def test_oracle_free_space_is_i_major_and_respects_the_radius(oracle_built):
    O = oracle_built
    cells, pm = _world(O)
    fs = pm.free_space(0.1)
    cond = (cells == -1) & (pm.dist.astype(np.float64) > 0.1)
    want = np.argwhere(cond.T)  # rows of cond.T are i: sorted by i, then j
    assert np.array_equal(fs, want.astype(np.int32))
    assert len(pm.free_space(0.0)) > len(fs) > len(pm.free_space(0.5))


def test_oracle_uniform_generator_consumes_two_draws_per_generation(oracle_built):
    O = oracle_built
    cells, pm = _world(O)
    fs = pm.free_space(0.0)
    L = O.port_lib()
    for thr, mult, per_pose in ((0.0, 0.0, 1), (0.9, 0.5, 1), (1.5, 0.5, 2), (5.0, 0.5, 4), (3.0, 1.0, 1), (3.0, 0.0, 2)):
        rng = O.Rng(0)
        L.orc_srand48(O.C.byref(rng), 11)
        poses = pm.uniform_poses(fs, rng, 50, thr, mult)
        chk = O.Rng(0)
        L.orc_srand48(O.C.byref(chk), 11)
        for s in range(50):
            for g in range(per_pose):
                u, v = L.orc_drand48(O.C.byref(chk)), L.orc_drand48(O.C.byref(chk))
            i, j = fs[int(u * len(fs))]
            assert poses[s, 0] == (i - 300 // 2) * RES and poses[s, 1] == (j - 200 // 2) * RES
            assert poses[s, 2] == v * 2 * np.pi - np.pi
        assert chk.x == rng.x


@pytest.fixture(scope="module")
def dev():
    from badger_amcl_b200 import cabi
    ctx = cabi.Context(0)
    yield dict(cabi=cabi, ctx=ctx)
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("radius", [0.0, 0.1, 0.5])
def test_gpu_free_space_list_vs_oracle(oracle_built, dev, radius):
    O, cabi = oracle_built, dev["cabi"]
    cells, pm = _world(O)
    m = cabi.Map2D(dev["ctx"], cells, RES)
    m.upload_distances(pm.dist, 2.0)
    n = m.update_free_space(radius)
    want = pm.free_space(radius)
    assert n == len(want)
    assert np.array_equal(m.free_space(), want)
    m.close()


@pytest.mark.gpu
@pytest.mark.parametrize("thr,mult", [(0.0, 0.0), (1.5, 0.5), (5.0, 0.5)])
@pytest.mark.parametrize("n", [1, 5000, 200000])
def test_gpu_init_uniform_vs_oracle(oracle_built, dev, n, thr, mult):
    O, cabi = oracle_built, dev["cabi"]
    cells, pm = _world(O, origin=(1.25, -0.75))
    m = cabi.Map2D(dev["ctx"], cells, RES, origin=(1.25, -0.75))
    m.upload_distances(pm.dist, 2.0)
    m.update_free_space(0.1)
    fs = pm.free_space(0.1)
    rng = O.Rng(0)
    O.port_lib().orc_srand48(O.C.byref(rng), 77)
    poses = pm.uniform_poses(fs, rng, n, thr, mult)
    ref = (O.RefPF if O.have_ref() else O.PortPF)(max(1, n // 10), n)  # initWithPoseFn: the compiled reference if here
    ref.init_with_poses(poses)
    pf = cabi.ParticleFilter(dev["ctx"], max(1, n // 10), n)
    pf.srand48(77)
    res = pf.init_uniform(m, thr, mult)
    got, want = pf.samples(), ref.samples()
    assert pf.drand48_state == rng.x
    assert np.array_equal(got, want)
    assert (res.sample_count, res.leaf_count, res.node_count, res.cluster_count) == (
        n, ref.leaf_count(), ref.node_count(), ref.cluster_count())
    assert np.array_equal(pf.sample_clusters(), ref.sample_clusters())
    assert np.allclose(pf.set_stats(), ref.set_stats(), rtol=1e-6, atol=1e-9)
    pf.close()
    m.close()


@pytest.mark.gpu
def test_gpu_init_uniform_needs_the_free_space_list(oracle_built, dev):
    O, cabi = oracle_built, dev["cabi"]
    cells, pm = _world(O)
    m = cabi.Map2D(dev["ctx"], cells, RES)
    m.upload_distances(pm.dist, 2.0)
    pf = cabi.ParticleFilter(dev["ctx"], 10, 100)
    with pytest.raises(cabi.BamclError):
        pf.init_uniform(m)
    m.update_free_space(100.0)  # nothing is that far from an obstacle: empty list
    with pytest.raises(cabi.BamclError):
        pf.init_uniform(m)
    pf.close()
    m.close()
