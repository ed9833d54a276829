// oracle/shim: TEST INFRASTRUCTURE ONLY. The (de)serialisations getMapCloud runs per particle
// (point_cloud_scanner.cpp:241-247) and Node3D::makePointCloudFromScan (node_3d.cpp:459-465) are byte copies in PCL /
// pcl_conversions; here they are plain vector copies.
#pragma once
#include <pcl/conversions.h>
#include <tf2/transform_datatypes.h>
namespace pcl_conversions
{
inline void fromPCL(const pcl::PCLPointCloud2& in, sensor_msgs::PointCloud2& out)
{
  out.header.frame_id = in.header.frame_id;
  out.xyz = in.xyz;
}
inline void toPCL(const sensor_msgs::PointCloud2& in, pcl::PCLPointCloud2& out)
{
  out.header.frame_id = in.header.frame_id;
  out.xyz = in.xyz;
}
}  // namespace pcl_conversions
