"""badger_amcl_b200: B200-native (sm_100a) implementation of badger_amcl's particle-filter
measurement-and-resample hot path.

Layout:
  csrc/        hand-written CUDA kernels + the C ABI (include/bamcl.h) -> libbamcl_cuda.so
  host/        C++ mirror of the reference's class API (OccupancyMap, PlanarScanner, ParticleFilter ...)
               implemented over the C ABI -> libbadger_amcl.so
  cabi.py      ctypes binding of the C ABI used by tests/ and bench.py
  synth.py     synthetic maps / scans / particle clouds (numpy)
There is no CPU fallback: every compute entry point needs the CUDA library and a GPU.
"""
from . import synth  # noqa: F401
