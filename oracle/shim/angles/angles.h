// oracle/shim: TEST INFRASTRUCTURE ONLY. ros/angles is a third-party dependency of the reference
// (package.xml:42) that is not vendored; version unpinned. This restates its published
// normalize_angle in the fmod(a + pi, 2 pi) form => "parity unpinned" for the last ulp of theta.
#pragma once
#include <cmath>
namespace angles
{
static inline double normalize_angle_positive(double angle)
{
  const double result = std::fmod(angle, 2.0 * M_PI);
  if (result < 0)
    return result + 2.0 * M_PI;
  return result;
}
static inline double normalize_angle(double angle)
{
  const double result = std::fmod(angle + M_PI, 2.0 * M_PI);
  if (result <= 0.0)
    return result + M_PI;
  return result - M_PI;
}
static inline double shortest_angular_distance(double from, double to) { return normalize_angle(to - from); }
}  // namespace angles
