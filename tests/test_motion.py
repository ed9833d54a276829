"""Motion update: Odom::updateAction, odom.cpp:74-311, with PDFGaussian::draw, pdf_gaussian.cpp:77-97.

CPU: the C restatement against the golden fixture made from the compiled reference (tests/golden/motion.npz), against
oracle/_ref directly where it exists, and the parallel formulation of the rejection loop (attempt j uses generator
outputs 2j+1, 2j+2; the a-th accepted attempt is draw a) against the sequential draws.
GPU: bamcl_pf_update_action against the oracle -- the drand48 state after the update bit-exact (= the same attempts
were accepted), poses to 1e-12 (device log / sincos differ from glibc in the last ulp; tolerance stated here)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from badger_amcl_b200 import synth  # noqa: E402

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
POSE_TOL = 1e-12
CASES = [(mt, di) for mt in range(5) for di in range(3)]


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "motion.npz"))


def _port_run(O, g, mt, di, poses=None, seed=None):
    poses = g["poses"] if poses is None else poses
    pf = O.PortPF(40, len(poses))
    pf.init_with_poses(poses)
    pf.srand48(100 + 10 * mt + di if seed is None else seed)
    pf.update_action(mt, g["alpha"], g["pose"], g["deltas"][di], g["absolute_motion"])
    return pf


@pytest.mark.parametrize("mt,di", CASES)
def test_port_equals_golden(gold, oracle_built, mt, di):
    O = oracle_built
    pf = _port_run(O, gold, mt, di)
    assert np.array_equal(pf.samples()[:, :3], gold["m%d_d%d" % (mt, di)])
    assert O.port_lib().orc_drand48(O.C.byref(pf.rng)) == float(gold["m%d_d%d_next" % (mt, di)])


def test_port_equals_compiled_reference(gold, oracle_built):
    O = oracle_built
    if not O.have_ref():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    poses = synth.gaussian_poses((0.0, 0.0, 0.0), 3000, seed=3)
    for mt, di in CASES:
        a = _port_run(O, gold, mt, di, poses, seed=7)
        b = O.RefPF(40, len(poses))
        b.init_with_poses(poses)
        b.srand48(7)
        b.update_action(mt, gold["alpha"], gold["pose"], gold["deltas"][di], gold["absolute_motion"])
        assert np.array_equal(a.samples(), b.samples())
        assert O.port_lib().orc_drand48(O.C.byref(a.rng)) == O.ref_lib().ref_drand48()


def test_parallel_form_of_the_rejection_loop_gives_the_sequential_draws(oracle_built):
    """what motion.cu computes, in numpy: all attempts at once + rank of the accepted ones"""
    O = oracle_built
    rng = O.Rng(0)
    L = O.port_lib()
    L.orc_srand48(O.C.byref(rng), 99)
    x0 = rng.x
    n = 3000
    want = np.array([L.orc_gaussian_draw(O.C.byref(rng), 0.7) for _ in range(n)])
    m = 2 * n
    xs = np.empty(2 * m + 1, np.uint64)
    x = int(x0)
    for i in range(2 * m + 1):
        xs[i] = x
        x = (0x5DEECE66D * x + 0xB) & ((1 << 48) - 1)
    r = xs[1:].astype(np.float64) * 2.0 ** -48
    x1, x2 = 2.0 * r[0::2] - 1.0, 2.0 * r[1::2] - 1.0
    w = x1 * x1 + x2 * x2
    acc = ~((w > 1.0) | (w == 0.0))
    idx = np.flatnonzero(acc)[:n]
    got = 0.7 * x2[idx] * np.sqrt(-2.0 * np.log(w[idx]) / w[idx])
    assert np.allclose(got, want, rtol=1e-14, atol=0.0)
    assert int(xs[2 * (idx[-1] + 1)]) == rng.x  # generator state after the last accepted attempt


# ---------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def dev():
    from badger_amcl_b200 import cabi
    ctx = cabi.Context(0)
    yield dict(cabi=cabi, ctx=ctx)
    ctx.close()


def _gpu_vs_port(O, dev, gold, mt, di, poses, seed):
    cabi = dev["cabi"]
    port = _port_run(O, gold, mt, di, poses, seed)
    pf = cabi.ParticleFilter(dev["ctx"], 40, len(poses))
    pf.init_with_poses(poses)
    pf.srand48(seed)
    pf.update_action(mt, gold["alpha"], gold["pose"], gold["deltas"][di], gold["absolute_motion"])
    got, want = pf.samples(), port.samples()
    state = pf.drand48_state
    pf.close()
    assert state == port.rng.x, "the device consumed a different number of drand48 draws"
    assert np.array_equal(got[:, 3], want[:, 3])  # weights untouched
    err = np.abs(got[:, :3] - want[:, :3]).max()
    assert err <= POSE_TOL, err
    return err


@pytest.mark.gpu
@pytest.mark.parametrize("mt,di", CASES)
def test_gpu_update_action_vs_oracle(gold, oracle_built, dev, mt, di):
    _gpu_vs_port(oracle_built, dev, gold, mt, di, gold["poses"], 100 + 10 * mt + di)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 2, 3, 1000, 300000])
def test_gpu_update_action_sizes(gold, oracle_built, dev, n):
    """tiny sets (the +4096 attempt margin), and a set large enough for many scan tiles"""
    poses = synth.gaussian_poses((2.0, 1.0, -0.3), n, seed=n)
    for mt in (0, 3, 4):
        _gpu_vs_port(oracle_built, dev, gold, mt, 0, poses, 5 + n)


@pytest.mark.gpu
def test_gpu_update_action_vs_golden(gold, dev):
    cabi = dev["cabi"]
    for mt, di in CASES:
        pf = cabi.ParticleFilter(dev["ctx"], 40, len(gold["poses"]))
        pf.init_with_poses(gold["poses"])
        pf.srand48(100 + 10 * mt + di)
        pf.update_action(mt, gold["alpha"], gold["pose"], gold["deltas"][di], gold["absolute_motion"])
        got = pf.samples()[:, :3]
        pf.close()
        assert np.abs(got - gold["m%d_d%d" % (mt, di)]).max() <= POSE_TOL


@pytest.mark.gpu
def test_gpu_action_then_resample_keeps_one_drand48_stream(gold, oracle_built, dev):
    """updateAction and updateResample draw from the same stream (odom.cpp + particle_filter.cpp:309): run both on
    the device and in the oracle, weights staged from the oracle so the resample is comparable bit for bit"""
    O, cabi = oracle_built, dev["cabi"]
    n = 20000
    poses = synth.gaussian_poses((2.0, 1.0, -0.3), n, seed=1)
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    port.srand48(31)
    pf = cabi.ParticleFilter(dev["ctx"], n // 10, n)
    pf.init_with_poses(poses)
    pf.srand48(31)
    rng = np.random.default_rng(0)
    for step in range(2):
        args = (1 + step, gold["alpha"], gold["pose"], gold["deltas"][step], gold["absolute_motion"])
        port.update_action(*args)
        pf.update_action(*args)
        assert pf.drand48_state == port.rng.x
        m = port.sample_count()
        s = port.samples()
        assert np.abs(pf.samples()[:, :3] - s[:, :3]).max() <= POSE_TOL
        w = np.exp(rng.normal(0.0, 1.0, m))
        s[:, 3] = w / w.sum()
        port.set_samples(s)          # poses and weights staged from the oracle: the resample sees identical inputs
        pf.upload_samples(s)
        port.set_averages(1.0, 1.0)
        pf.w_slow = pf.w_fast = 1.0
        port.update_resample()
        res = pf.update_resample()
        assert res.sample_count == port.sample_count()
        assert np.array_equal(pf.resample_sources(), port.resample_sources())
        assert pf.drand48_state == port.rng.x
    pf.close()
