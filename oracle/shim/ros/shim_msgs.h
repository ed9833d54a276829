// oracle/shim: TEST INFRASTRUCTURE ONLY. Plain-data look-alikes of the ROS messages the reference's node sources touch
// (field names and C types as the .msg definitions give them: float32 -> float, float64 -> double, boost::array ->
// std::array, boost::shared_ptr -> std::shared_ptr). sensor_msgs/PointCloud2 and octomap_msgs/Octomap carry their
// decoded payload (xyz points / occupied leaf keys) instead of the serialised bytes: (de)serialisation is third-party
// byte shuffling, not arithmetic of the reference.
#pragma once
#include <array>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include <pcl/point_types.h>
#include <ros/shim_bus.h>
#define SHIM_MSG_PTRS(T)                  \
  typedef std::shared_ptr<T> T##Ptr;      \
  typedef std::shared_ptr<T const> T##ConstPtr;
namespace std_msgs
{
struct Header
{
  uint32_t seq = 0;
  ros::Time stamp;
  std::string frame_id;
};
}  // namespace std_msgs
namespace geometry_msgs
{
struct Point
{
  double x = 0.0, y = 0.0, z = 0.0;
};
struct Vector3
{
  double x = 0.0, y = 0.0, z = 0.0;
};
struct Quaternion
{
  double x = 0.0, y = 0.0, z = 0.0, w = 0.0;
};
struct Pose
{
  Point position;
  Quaternion orientation;
};
struct Pose2D
{
  double x = 0.0, y = 0.0, theta = 0.0;
};
struct PoseArray
{
  std_msgs::Header header;
  std::vector<Pose> poses;
};
struct PoseWithCovariance
{
  Pose pose;
  std::array<double, 36> covariance{};
};
struct PoseWithCovarianceStamped
{
  std_msgs::Header header;
  PoseWithCovariance pose;
};
SHIM_MSG_PTRS(PoseWithCovarianceStamped)
struct Transform
{
  Vector3 translation;
  Quaternion rotation;
};
struct TransformStamped
{
  std_msgs::Header header;
  std::string child_frame_id;
  Transform transform;
};
}  // namespace geometry_msgs
namespace nav_msgs
{
struct MapMetaData
{
  ros::Time map_load_time;
  float resolution = 0.f;
  uint32_t width = 0, height = 0;
  geometry_msgs::Pose origin;
};
struct OccupancyGrid
{
  std_msgs::Header header;
  MapMetaData info;
  std::vector<int8_t> data;
};
SHIM_MSG_PTRS(OccupancyGrid)
struct Odometry
{
  std_msgs::Header header;
  std::string child_frame_id;
  geometry_msgs::PoseWithCovariance pose;
};
SHIM_MSG_PTRS(Odometry)
}  // namespace nav_msgs
namespace sensor_msgs
{
struct LaserScan
{
  std_msgs::Header header;
  float angle_min = 0.f, angle_max = 0.f, angle_increment = 0.f, time_increment = 0.f, scan_time = 0.f;
  float range_min = 0.f, range_max = 0.f;
  std::vector<float> ranges, intensities;
};
SHIM_MSG_PTRS(LaserScan)
struct PointCloud2
{
  std_msgs::Header header;
  std::vector<pcl::PointXYZ> xyz;
};
SHIM_MSG_PTRS(PointCloud2)
}  // namespace sensor_msgs
namespace std_srvs
{
struct Empty
{
  struct Request
  {
  };
  struct Response
  {
  };
};
}  // namespace std_srvs
namespace octomap_msgs
{
struct Octomap
{
  std_msgs::Header header;
  bool binary = true;
  std::string id;
  double resolution = 0.0;
  std::vector<int> occupied_keys;  // decoded payload: 3 ints per occupied leaf (see shim/octomap/OcTree.h)
};
SHIM_MSG_PTRS(Octomap)
}  // namespace octomap_msgs
