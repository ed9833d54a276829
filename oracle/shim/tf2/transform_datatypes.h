// oracle/shim: TEST INFRASTRUCTURE ONLY. tf2::Stamped<T>, ros::Time, the geometry_msgs / sensor_msgs structs the 3-D scanner
// touches (plain data, no arithmetic).
#pragma once
#include <string>
#include <vector>

#include <pcl/point_types.h>
namespace ros
{
struct Time
{
  static Time now() { return Time(); }
};
}  // namespace ros
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
struct TransformStamped
{
  std::string frame_id, child_frame_id;
  Transform transform;
};
}  // namespace geometry_msgs
namespace sensor_msgs
{
struct PointCloud2  // only the xyz payload the scanner moves around
{
  std::vector<pcl::PointXYZ> xyz;
};
}  // namespace sensor_msgs
namespace tf2
{
template <typename T>
class Stamped : public T
{
public:
  ros::Time stamp_;
  std::string frame_id_;
  Stamped() {}
  Stamped(const T& input, const ros::Time& timestamp, const std::string& frame_id) : T(input), stamp_(timestamp), frame_id_(frame_id) {}
};
}  // namespace tf2
