"""The C++ host mirror of the reference's class API (badger_amcl_b200/host): the reference's own unit tests
(test/test_badger_amcl.cpp:51-171) and a scan -> updateSensor -> updateResample flow, driven through
OccupancyMap / PlanarScanner / ParticleFilter exactly as Node2D drives them, compared with the oracle."""
import os
import struct
import subprocess

import numpy as np
import pytest

from badger_amcl_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RES = 0.05


@pytest.fixture(scope="module")
def exe():
    from badger_amcl_b200 import build
    build.build()
    build.build_host()
    return build.HOST_TEST


def test_host_library_builds_and_fails_loudly_without_a_gpu(exe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([exe, "kats"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 3
    assert "no CPU fallback" in p.stdout


@pytest.mark.gpu
def test_reference_unit_tests_pass_on_the_host_mirror(exe):
    p = subprocess.run([exe, "kats"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "0 failed" in p.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("cloud", ["global", "tracking"])
def test_flow_through_the_class_api_matches_oracle(exe, oracle_built, tmp_path, cloud):
    O = oracle_built
    n, n_beams, seed = 5000, 60, 4242
    cells = synth.make_grid(600, 500, seed=7, wall_every=100)
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=n_beams)
    poses = (synth.uniform_free_poses(cells, RES, n, seed=1) if cloud == "global"
             else synth.gaussian_poses(truth, n, seed=2))
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(fin, "wb") as f:
        f.write(struct.pack("<9i", 0x414D434C, cells.shape[1], cells.shape[0], n, n_beams, n // 10, 1, 1, seed))
        f.write(struct.pack("<3d", RES, 2.0, 30.0))
        f.write(np.ascontiguousarray(cells, np.int8).tobytes())
        f.write(np.ascontiguousarray(ranges, np.float64).tobytes())
        f.write(np.ascontiguousarray(angles, np.float64).tobytes())
        f.write(np.ascontiguousarray(poses, np.float64).tobytes())
    env = dict(os.environ, BAMCL_EXACT_TERMS="1")
    p = subprocess.run([exe, "flow", fin, fout], capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stdout + p.stderr
    got = np.fromfile(fout, np.float64)

    pm = O.PortMap2D(cells, RES)
    pm.set_distances(pm.exact_edt(2.0), 2.0)
    model = O.planar_model(model_type=1, max_beams=n_beams, scanner_pose=(0.1, 0.0, 0.0))
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    port.srand48(seed)
    pos = 0
    assert tuple(got[:3]) == (n, port.leaf_count(), port.cluster_count())
    pos = 3
    for rnd in range(2):
        for u in range(2):
            port.update_sensor(pm, model, ranges, angles, 30.0)
            if rnd == 0 and u == 0:
                w = got[pos:pos + n]
                pos += n
                assert np.allclose(w, port.samples()[:, 3], rtol=1e-9, atol=0.0)
        port.update_resample()
        m = port.sample_count()
        st = got[pos:pos + 12]
        pos += 12
        assert (int(st[0]), int(st[1]), int(st[2]), bool(st[3])) == (m, port.leaf_count(), port.cluster_count(),
                                                                     port.is_converged())
        assert np.allclose(st[4:12], port.set_stats(), rtol=1e-6, atol=1e-9)
        s = got[pos:pos + 4 * m].reshape(m, 4)
        pos += 4 * m
        assert np.array_equal(s, port.samples())  # same source indices => identical poses, weights 1/m
        best = got[pos:pos + 4]
        pos += 4
        cs = port.cluster_stats()
        k = int(np.argmax(cs[:, 1]))
        assert np.allclose(best, [cs[k, 1], cs[k, 2], cs[k, 3], cs[k, 4]], rtol=1e-6, atol=1e-9)
        moved = port.samples()
        moved[:, 0] += 0.01
        port.set_samples(moved)
    assert pos == len(got)


@pytest.mark.gpu
def test_odom_update_action_through_the_class_api_matches_oracle(exe, oracle_built, tmp_path):
    """the mirror's Odom::updateAction (host/src/odom.cpp -> bamcl_pf_update_action) against the oracle's restatement of
    odom.cpp:74-311: libc's drand48 state after every call bit-exact, poses to 1e-12 (device log / sincos)"""
    O = oracle_built
    n, seed = 4000, 99
    poses = synth.gaussian_poses((1.0, -2.0, 0.5), n, seed=3)
    updates = [(1, [0.2, 0.3, 0.25, 0.15, 0.1], [1.0, 2.0, 0.7], [0.3, 0.1, 0.2], [0.35, 0.05, 0.25]),
               (2, [0.1, 0.2, 0.3, 0.1, 0.0], [1.3, 2.1, 0.9], [0.001, 0.0, 1.5], [0.0, 0.0, 0.0]),
               (4, [0.2, 0.3, 0.25, 0.15, 0.1], [0.9, 2.3, -2.1], [-0.4, 0.2, -3.0], [0.45, 0.1, 3.0])]
    fin, fout = str(tmp_path / "odom_in.bin"), str(tmp_path / "odom_out.bin")
    with open(fin, "wb") as f:
        f.write(struct.pack("<3i", n, seed, len(updates)))
        f.write(np.ascontiguousarray(poses, np.float64).tobytes())
        for mt, alpha, pose, delta, am in updates:
            f.write(np.array([mt] + alpha + pose + delta + am, np.float64).tobytes())
    p = subprocess.run([exe, "odom", fin, fout], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    got = np.fromfile(fout, np.float64).reshape(len(updates), 4 * n + 1)
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    port.srand48(seed)
    for u, (mt, alpha, pose, delta, am) in enumerate(updates):
        port.update_action(mt, alpha, pose, delta, am)
        s = got[u, :4 * n].reshape(n, 4)
        assert int(got[u, 4 * n]) == port.rng.x
        assert np.abs(s[:, :3] - port.samples()[:, :3]).max() <= 1e-12
        assert np.array_equal(s[:, 3], port.samples()[:, 3])
        # continue from the device's poses so that rounding differences do not accumulate in the comparison
        port.set_samples(s)


@pytest.mark.gpu
def test_init_with_gaussian_through_the_class_api_matches_the_reference(exe, oracle_built, tmp_path):
    """the mirror's ParticleFilter::initWithGaussian (-> bamcl_pf_init_gaussian) against the compiled reference's
    (particle_filter.cpp:105-133 + pdf_gaussian.cpp): libc's drand48 state after the call bit-exact, poses to 1e-12"""
    O = oracle_built
    n, seed = 3000, 5
    mean, var = [1.5, -0.5, 0.3], [0.25, 0.09, 0.0685]
    fout = str(tmp_path / "gauss_out.bin")
    p = subprocess.run([exe, "gauss", str(seed), str(n)] + [repr(v) for v in mean + var] + [fout], capture_output=True,
                       text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    got = np.fromfile(fout, np.float64)
    s, tail = got[:4 * n].reshape(n, 4), got[4 * n:]
    ref = (O.RefPF if O.have_ref() else None)
    if ref is None:
        pytest.skip("oracle/_ref not built")
    pf = ref(n // 10, n)
    pf.srand48(seed)
    pf.init_gaussian(mean, np.diag(var))
    want = pf.samples()
    st = (O.C.c_ushort * 3)()
    O.ref_lib().ref_get_seed48(st)
    assert int(tail[0]) == st[0] | (st[1] << 16) | (st[2] << 32)
    assert np.abs(s[:, :3] - want[:, :3]).max() <= 1e-12
    assert np.array_equal(s[:, 3], want[:, 3])
    assert (int(tail[1]), int(tail[2])) == (pf.leaf_count(), pf.cluster_count())

