// drand48's generator (POSIX: X <- (0x5DEECE66D X + 0xB) mod 2^48, value X / 2^48) with skip-ahead, shared by the
// multinomial resampler (particle_filter.cpp:385,393) and the motion update's Gaussian draws (pdf_gaussian.cpp:85,90).
#pragma once

namespace bamcl
{
constexpr unsigned long long LCG_A = 0x5DEECE66DULL, LCG_C = 0xBULL, LCG_MASK = (1ULL << 48) - 1;

__host__ __device__ inline unsigned long long lcg_next(unsigned long long x) { return (LCG_A * x + LCG_C) & LCG_MASK; }

// state after j steps from x0: x -> A x + C with (A, C) built by repeated squaring of the one-step map
__host__ __device__ inline unsigned long long lcg_skip(unsigned long long x0, unsigned long long j)
{
  unsigned long long a = LCG_A, c = LCG_C, A = 1, C = 0;
  while (j)
  {
    if (j & 1)
    {
      A = (A * a) & LCG_MASK;
      C = (C * a + c) & LCG_MASK;
    }
    c = ((a + 1) * c) & LCG_MASK;
    a = (a * a) & LCG_MASK;
    j >>= 1;
  }
  return (A * x0 + C) & LCG_MASK;
}
}  // namespace bamcl
