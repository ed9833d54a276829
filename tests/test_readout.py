"""Pose read-out: Node2D::getMaxWeightPose (node_2d.cpp:588-617) and the covariance packing of
Node::updatePose (node.cpp:381-395). The oracle restates both over the (compiled or restated) cluster statistics; the
device reduces its own cluster table. Cluster index bit-exact, weight / mean to 1e-6 (north_star's tolerance for means
and covariances)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from badger_amcl_b200 import synth  # noqa: E402


def test_covariance_packing_is_the_reference_layout(oracle_built):
    """host arithmetic of the C ABI library: runs without a GPU"""
    from badger_amcl_b200 import build, cabi
    build.build()
    L = cabi.lib()
    res = cabi.ResampleResult()
    res.cov = (cabi.C.c_double * 5)(0.11, 0.02, 0.03, 0.22, 0.33)
    out = np.full(36, 7.0)
    cabi._check(L.bamcl_pose_covariance(cabi.C.byref(res), cabi._p(out)))
    want = np.zeros(36)
    want[0], want[1], want[6], want[7], want[35] = 0.11, 0.02, 0.03, 0.22, 0.33  # COVARIANCE_AA = 6*5+5, node.h:67
    assert np.array_equal(out, want)


def test_oracle_max_weight_pose_takes_the_first_of_equal_maxima(oracle_built):
    O = oracle_built
    # three well separated blobs; the two big ones have the same count -> bit-equal weights -> the first one wins
    a = np.tile([0.1, 0.1, 0.0], (40, 1))
    b = np.tile([10.1, 0.1, 0.0], (20, 1))
    c = np.tile([20.1, 0.1, 0.0], (40, 1))
    poses = np.concatenate([b, a, c])
    pf = O.PortPF(10, len(poses))
    pf.init_with_poses(poses)
    w, pose, idx = pf.max_weight_pose()
    st = pf.cluster_stats()
    assert st[idx, 0] == 40 and idx == int(np.flatnonzero(st[:, 0] == 40)[0])
    assert abs(pose[0] - 0.1) < 1e-9


@pytest.fixture(scope="module")
def dev():
    from badger_amcl_b200 import cabi
    ctx = cabi.Context(0)
    yield dict(cabi=cabi, ctx=ctx)
    ctx.close()


def _compare(pf, port):
    w, pose, idx = pf.max_weight_pose()
    ww, wpose, widx = port.max_weight_pose()
    assert idx == widx
    assert abs(w - ww) <= 1e-6 * abs(ww)
    assert np.allclose(pose, wpose, rtol=1e-6, atol=1e-9)
    assert np.allclose(pf.pose_covariance(), port.pose_covariance(), rtol=1e-6, atol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("cloud", ["tracking", "global", "equal_blobs"])
def test_gpu_max_weight_pose_vs_oracle(oracle_built, dev, cloud):
    O, cabi = oracle_built, dev["cabi"]
    n = 20000
    if cloud == "tracking":
        poses = synth.gaussian_poses((2.0, 1.0, -0.3), n, seed=1)
    elif cloud == "global":
        cells = synth.make_grid(400, 400, seed=7, wall_every=100)
        poses = synth.uniform_free_poses(cells, 0.05, n, seed=2)
    else:
        blob = np.random.default_rng(0).normal(0.0, 0.05, (n // 4, 3))
        poses = np.concatenate([blob + [10 * k + 0.25, 0.25, 0.0] for k in range(4)])  # four clusters of equal count
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    port.srand48(3)
    pf = cabi.ParticleFilter(dev["ctx"], n // 10, n)
    pf.init_with_poses(poses)
    pf.srand48(3)
    _compare(pf, port)  # statistics of the initial set (initWithPoseFn, particle_filter.cpp:157)
    rng = np.random.default_rng(5)
    for _ in range(2):
        m = port.sample_count()
        s = port.samples()
        w = np.exp(rng.normal(0.0, 1.5, m))
        s[:, 3] = w / w.sum()
        port.set_samples(s)
        pf.upload_samples(s)
        port.set_averages(1.0, 1.0)
        pf.w_slow = pf.w_fast = 1.0
        port.update_resample()
        pf.update_resample()
        _compare(pf, port)
    pf.close()


@pytest.mark.gpu
def test_gpu_max_weight_pose_needs_statistics(dev):
    cabi = dev["cabi"]
    pf = cabi.ParticleFilter(dev["ctx"], 10, 100)
    with pytest.raises(cabi.BamclError):
        pf.max_weight_pose()
    pf.close()
