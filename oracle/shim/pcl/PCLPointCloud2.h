// oracle/shim: TEST INFRASTRUCTURE ONLY. pcl::PCLPointCloud2 carrying the decoded xyz payload (see ros/shim_msgs.h).
#pragma once
#include <vector>

#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
namespace pcl
{
struct PCLPointCloud2
{
  ShimHeader header;
  std::vector<pcl::PointXYZ> xyz;
};
}  // namespace pcl
