// oracle/shim: TEST INFRASTRUCTURE ONLY. pcl::toPCLPointCloud2 / fromPCLPointCloud2 are byte copies in PCL; here plain
// vector copies.
#pragma once
#include <pcl/PCLPointCloud2.h>
namespace pcl
{
inline void toPCLPointCloud2(const pcl::PointCloud<pcl::PointXYZ>& cloud, pcl::PCLPointCloud2& msg)
{
  msg.header = cloud.header;
  msg.xyz = cloud.points;
}
inline void fromPCLPointCloud2(const pcl::PCLPointCloud2& msg, pcl::PointCloud<pcl::PointXYZ>& cloud)
{
  cloud.header = msg.header;
  cloud.points = msg.xyz;
  cloud.width = static_cast<uint32_t>(msg.xyz.size());
  cloud.height = 1;
}
}  // namespace pcl
