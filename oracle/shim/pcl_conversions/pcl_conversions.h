// oracle/shim: TEST INFRASTRUCTURE ONLY. The (de)serialisations getMapCloud runs per particle
// (point_cloud_scanner.cpp:241-247) are byte copies in PCL / pcl_conversions; here they are plain vector copies.
#pragma once
#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#include <tf2/transform_datatypes.h>
namespace pcl
{
struct PCLPointCloud2
{
  std::vector<pcl::PointXYZ> xyz;
};
inline void toPCLPointCloud2(const pcl::PointCloud<pcl::PointXYZ>& cloud, pcl::PCLPointCloud2& msg) { msg.xyz = cloud.points; }
inline void fromPCLPointCloud2(const pcl::PCLPointCloud2& msg, pcl::PointCloud<pcl::PointXYZ>& cloud)
{
  cloud.points = msg.xyz;
  cloud.width = static_cast<uint32_t>(msg.xyz.size());
  cloud.height = 1;
}
}  // namespace pcl
namespace pcl_conversions
{
inline void fromPCL(const pcl::PCLPointCloud2& in, sensor_msgs::PointCloud2& out) { out.xyz = in.xyz; }
inline void toPCL(const sensor_msgs::PointCloud2& in, pcl::PCLPointCloud2& out) { out.xyz = in.xyz; }
}  // namespace pcl_conversions
