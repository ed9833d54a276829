// Header-only look-alikes of the third-party types that appear in the reference's hot-path signatures
// (SURVEY.md 8b): Eigen::Vector3d / Matrix3d, pcl::PointXYZ / PointCloud, geometry_msgs::Transform and
// octomap::OcTree. None of those libraries exists on the GPU box; only the members the reference's callers touch
// are provided, with the same names. Define BAMCL_HAVE_EIGEN / BAMCL_HAVE_PCL / BAMCL_HAVE_ROS / BAMCL_HAVE_OCTOMAP
// before including to use the real libraries instead.
#ifndef BAMCL_AMCL_COMPAT_H
#define BAMCL_AMCL_COMPAT_H
#include <cmath>
#include <cstddef>
#include <memory>
#include <string>
#include <vector>

#ifdef BAMCL_HAVE_EIGEN
#include <Eigen/Dense>
#else
namespace Eigen
{
struct Vector3d
{
  double v[3];
  Vector3d() {}
  Vector3d(double a, double b, double c) : v{a, b, c} {}
  double& operator[](int i) { return v[i]; }
  const double& operator[](int i) const { return v[i]; }
  double& operator()(int i) { return v[i]; }
  const double& operator()(int i) const { return v[i]; }
};
struct Vector3i
{
  int v[3];
  Vector3i() {}
  Vector3i(int a, int b, int c) : v{a, b, c} {}
  int& operator[](int i) { return v[i]; }
  const int& operator[](int i) const { return v[i]; }
};
struct Matrix3d
{
  double m[3][3];
  Matrix3d() {}
  double& operator()(int r, int c) { return m[r][c]; }
  const double& operator()(int r, int c) const { return m[r][c]; }
  static Matrix3d Zero()
  {
    Matrix3d z;
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++)
        z.m[r][c] = 0.0;
    return z;
  }
};
}  // namespace Eigen
#endif

#ifdef BAMCL_HAVE_PCL
#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#else
namespace pcl
{
struct PointXYZ
{
  float x, y, z;
  PointXYZ() : x(0.f), y(0.f), z(0.f) {}
  PointXYZ(float a, float b, float c) : x(a), y(b), z(c) {}
};
struct PCLHeader
{
  std::string frame_id;
};
template <typename PointT>
struct PointCloud
{
  typedef std::shared_ptr<PointCloud<PointT>> Ptr;
  PCLHeader header;
  unsigned width = 0, height = 1;
  std::vector<PointT> points;
  std::size_t size() const { return points.size(); }
  void push_back(const PointT& p)
  {
    points.push_back(p);
    width = (unsigned)points.size();
  }
};
}  // namespace pcl
#endif

#ifdef BAMCL_HAVE_ROS
#include <geometry_msgs/Transform.h>
#else
namespace geometry_msgs
{
struct Vector3
{
  double x = 0.0, y = 0.0, z = 0.0;
};
struct Quaternion
{
  double x = 0.0, y = 0.0, z = 0.0, w = 1.0;
};
struct Transform
{
  Vector3 translation;
  Quaternion rotation;
};
}  // namespace geometry_msgs
#endif

#ifdef BAMCL_HAVE_OCTOMAP
#include <octomap/OcTree.h>
#else
namespace octomap
{
// The occupied leaves of an octree as integer keys (leaf centre = (key + 0.5) * resolution, octomap's
// keyToCoord with the tree offset removed). This is all OctoMap::initFromOctree reads (octomap.cpp:53-74,
// 209-251: getMetricMin/Max, the leaf iterator, isNodeOccupied).
class OcTree
{
public:
  explicit OcTree(double resolution) : resolution_(resolution) {}
  double getResolution() const { return resolution_; }
  void addOccupiedKey(int kx, int ky, int kz)
  {
    keys_.push_back(kx);
    keys_.push_back(ky);
    keys_.push_back(kz);
  }
  // world coordinates of a point inside the leaf, as octomap::OcTree::updateNode(x, y, z, true) would key it
  void updateNode(double x, double y, double z, bool occupied)
  {
    if (occupied)
      addOccupiedKey((int)std::floor(x / resolution_), (int)std::floor(y / resolution_), (int)std::floor(z / resolution_));
  }
  void expand() {}
  const std::vector<int>& occupiedKeys() const { return keys_; }
  std::size_t getNumLeafNodes() const { return keys_.size() / 3; }

private:
  double resolution_;
  std::vector<int> keys_;
};
}  // namespace octomap
#endif


// The reference's hot path raises no exceptions: updateSensor returns false, a model returns 0.0, and
// nothing else happens. The mirror keeps that contract: a device failure on the path (updateSensor, updateResample,
// updateAction, the lazy sample synchronisation) is recorded here and the call returns like the reference's own failure
// paths, leaving the filter's weights untouched. Set-up calls (maps, constructors) still throw: without a CUDA device
// there is nothing to fall back to.
namespace badger_amcl
{
const char* lastDeviceError();  // "" while nothing failed since clearDeviceError()
void clearDeviceError();
bool deviceOk(int rc, const char* what);  // records bamcl_last_error() and returns false when rc != 0
void deviceFail(const char* what);
}  // namespace badger_amcl

#endif  // BAMCL_AMCL_COMPAT_H
