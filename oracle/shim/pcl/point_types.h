// oracle/shim: TEST INFRASTRUCTURE ONLY. pcl::PointXYZ / PointXYZI look-alikes (float members).
#pragma once
#include <atomic>
#include <Eigen/Dense>  // real PCL pulls Eigen in; octomap.h relies on that
namespace pcl
{
struct PointXYZ
{
  float x, y, z;
  PointXYZ() : x(0.f), y(0.f), z(0.f) {}
  PointXYZ(float a, float b, float c) : x(a), y(b), z(c) {}
};
struct PointXYZI
{
  float x, y, z, intensity;
  PointXYZI() : x(0.f), y(0.f), z(0.f), intensity(0.f) {}
};
}  // namespace pcl
