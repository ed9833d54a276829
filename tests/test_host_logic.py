"""CPU-only checks of the host-side logic that ships in libbamcl_cuda.so (no compute calls, no GPU):
the C ABI loads and exports every declared symbol, resampleLimit, and the closed-form systematic target
sequence against the reference's one-rounding-at-a-time loop (particle_filter.cpp:309,337-341)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from badger_amcl_b200 import cabi, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def L():
    from badger_amcl_b200 import build
    build.build()
    return cabi.lib()


def test_library_exports_every_declared_symbol(L):
    header = open(os.path.join(ROOT, "include", "bamcl.h")).read()
    declared = set(re.findall(r"\b(bamcl_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    missing = [name for name in sorted(declared) if not hasattr(L, name)]
    assert not missing, "declared in bamcl.h but not exported: %s" % missing
    assert declared == set(cabi.SIGNATURES), "cabi.SIGNATURES and bamcl.h disagree: %s" % (
        declared ^ set(cabi.SIGNATURES))


def test_no_device_fails_loudly(L):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(cabi.BamclError):
        cabi.Context(0)


def test_resample_limit_matches_oracle(L, oracle_built):
    O = oracle_built
    for (mn, mx) in ((100, 5000), (10, 100000), (1000, 1000000)):
        port = O.PortPF(mn, mx)
        for k in list(range(0, 50)) + [99, 1000, 18254, 57600, 10**6]:
            assert L.bamcl_resample_limit(k, mn, mx, 0.01, 0.99) == port.resample_limit(k)


def _sequential_targets(start, delta, n):
    out = np.empty(n)
    t = start
    for k in range(n):
        out[k] = t
        t = t + delta
        if t > 1.0:
            t -= 1.0
    return out


@pytest.mark.parametrize("n", [1, 2, 3, 7, 100, 4999, 5000, 65536, 100003])
def test_systematic_targets_bit_exact(L, n):
    rng = synth.Drand48(12345 + n)
    starts = [rng.draw() for _ in range(6)] + [0.0, 0.5, 1.0 - 2.0 ** -53, 2.0 ** -60, 0.75, 1e-300]
    delta = 1.0 / n
    for start in starts:
        got = np.empty(n)
        assert L.bamcl_systematic_targets(start, delta, n, got.ctypes.data_as(C.c_void_p)) == 0
        want = _sequential_targets(start, delta, n)
        assert np.array_equal(got, want), (n, start, int(np.argmax(got != want)))


def test_systematic_targets_tie_increments(L):
    """delta whose low bits make exact round-half-even ties in some binade"""
    for delta in (2.0 ** -20 + 2.0 ** -73, 2.0 ** -10 + 2.0 ** -63, 3 * 2.0 ** -12 + 2.0 ** -64):
        n = int(1.0 / delta) + 5
        for start in (0.3, 0.5 + 2.0 ** -53, 0.25 + 2.0 ** -54, 0.9999):
            got = np.empty(n)
            assert L.bamcl_systematic_targets(start, delta, n, got.ctypes.data_as(C.c_void_p)) == 0
            assert np.array_equal(got, _sequential_targets(start, delta, n))
