"""bamcl_weights_normalise_noalloc (include/bamcl.h): the allocation-free weight total + normalisation, through the C ABI,
against the existing entry points (bamcl_pf_score_planar's total, bamcl_pf_normalise's weights). Same tree shape and the same
division, so the comparison is bit-exact. Collected last on purpose (the file name)."""
import ctypes as C

import numpy as np
import pytest

from badger_amcl_b200 import synth

pytestmark = pytest.mark.gpu
RES = 0.05


@pytest.fixture(scope="module")
def env():
    import torch
    assert torch.cuda.is_available()
    from badger_amcl_b200 import build, cabi
    build.build()
    ctx = cabi.Context(0)
    cells = synth.make_grid(600, 500, seed=7)
    m = cabi.Map2D(ctx, cells, RES)
    m.build_distances(2.0)
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=60, dropouts=0.03)
    yield dict(torch=torch, cabi=cabi, ctx=ctx, cells=cells, m=m, truth=truth, ranges=ranges, angles=angles)
    m.close()
    ctx.close()


def existing_path(env, n, seed):
    """score n particles with the existing entry points -> (raw weights, total of the filter's own tree, normalised weights)"""
    cabi = env["cabi"]
    poses = np.concatenate([synth.uniform_free_poses(env["cells"], RES, n // 2, seed=seed),
                            synth.gaussian_poses(env["truth"], n - n // 2, seed=seed + 1)])
    w0 = np.random.default_rng(seed).uniform(0.5, 1.5, n) / n
    pf = cabi.ParticleFilter(env["ctx"], max(1, n // 10), n)
    pf.upload_samples(np.concatenate([poses, w0[:, None]], axis=1))
    model = cabi.planar_model(model_type=1, max_beams=60)
    total = pf.score_planar(env["m"], model, env["ranges"], env["angles"], 30.0)   # bamcl_pf_score_planar
    raw = pf.weights()
    pf.normalise(total)                                                            # bamcl_pf_normalise
    norm = pf.weights()
    pf.close()
    return raw, total, norm


def no_alloc_path(env, raw, in_place=False):
    torch, cabi = env["torch"], env["cabi"]
    dev = torch.device("cuda:0")
    w = torch.from_numpy(raw).to(dev)
    out = w if in_place else torch.full_like(w, -7.0)
    total = torch.full((1,), -7.0, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()   # the tensors were filled on torch's stream; the library works on its own
    cabi.weights_normalise_noalloc(env["ctx"], w.data_ptr(), len(raw), out.data_ptr(), total.data_ptr())
    env["ctx"].sync()
    return float(total.cpu()[0]), out.cpu().numpy()


def test_two_calls_in_a_row_match_the_existing_entry_points(env):
    torch = env["torch"]
    # the second call is the smaller one: the scratch still holds the first call's partials beyond what it uses
    cases = [existing_path(env, 100000, 1), existing_path(env, 5000, 11)]
    for k, (raw, total, norm) in enumerate(cases):
        assert total > 0.0 and not np.array_equal(raw, norm)
        free_before = torch.cuda.mem_get_info()[0]
        got_total, got = no_alloc_path(env, raw)
        if k == 1:   # (the first launch of a kernel may load its module; from then on device memory must not move)
            assert torch.cuda.mem_get_info()[0] == free_before
        assert got_total == total
        assert np.array_equal(got, norm)


def test_in_place_and_zero_total(env):
    raw, total, norm = existing_path(env, 5000, 3)
    got_total, got = no_alloc_path(env, raw.copy(), in_place=True)
    assert got_total == total and np.array_equal(got, norm)
    zeros = np.zeros(777)
    got_total, got = no_alloc_path(env, zeros)
    assert got_total == 0.0 and np.array_equal(got, np.full(777, 1.0 / 777))   # particle_filter.cpp:258-265


def test_sizes_beyond_the_static_scratch_are_refused(env):
    cabi = env["cabi"]
    torch = env["torch"]
    one = torch.zeros(1, dtype=torch.float64, device="cuda:0")
    with pytest.raises(cabi.BamclError, match="exceed"):
        cabi.weights_normalise_noalloc(env["ctx"], one.data_ptr(), cabi.NOALLOC_MAX_SAMPLES + 1, one.data_ptr(), one.data_ptr())
    with pytest.raises(cabi.BamclError):
        cabi.weights_normalise_noalloc(env["ctx"], 0, 10, one.data_ptr(), one.data_ptr())
    # n == 0 is a valid call: total 0, nothing else written
    cabi.weights_normalise_noalloc(env["ctx"], 0, 0, 0, one.data_ptr())
    env["ctx"].sync()
    assert float(one.cpu()[0]) == 0.0
