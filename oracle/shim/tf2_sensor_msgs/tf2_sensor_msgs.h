// oracle/shim: TEST INFRASTRUCTURE ONLY. tf2::doTransform(sensor_msgs::PointCloud2), as ros/geometry2 tf2_sensor_msgs
// defines it: Eigen::Transform<float, 3, Affine> t = Translation3f(translation) * Quaternion<float>(rotation), then
// point' = t * point for every point, all in f32. Eigen's Quaternion::toRotationMatrix and its 3x3 * 3x1 product order are
// restated (no Eigen in this image; x86-64 build without FMA contraction): "parity unpinned" (SURVEY 8c).
#pragma once
#include <tf2/transform_datatypes.h>
namespace tf2
{
inline void doTransform(const sensor_msgs::PointCloud2& p_in, sensor_msgs::PointCloud2& p_out,
                        const geometry_msgs::TransformStamped& t_in)
{
  const float qx = static_cast<float>(t_in.transform.rotation.x), qy = static_cast<float>(t_in.transform.rotation.y);
  const float qz = static_cast<float>(t_in.transform.rotation.z), qw = static_cast<float>(t_in.transform.rotation.w);
  const float t[3] = {static_cast<float>(t_in.transform.translation.x), static_cast<float>(t_in.transform.translation.y),
                      static_cast<float>(t_in.transform.translation.z)};
  // Eigen::QuaternionBase::toRotationMatrix
  const float tx = 2.0f * qx, ty = 2.0f * qy, tz = 2.0f * qz;
  const float twx = tx * qw, twy = ty * qw, twz = tz * qw;
  const float txx = tx * qx, txy = ty * qx, txz = tz * qx;
  const float tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
  const float R[9] = {1.0f - (tyy + tzz), txy - twz, txz + twy, txy + twz, 1.0f - (txx + tzz), tyz - twx,
                      txz - twy, tyz + twx, 1.0f - (txx + tyy)};
  p_out.xyz.resize(p_in.xyz.size());
  for (size_t k = 0; k < p_in.xyz.size(); k++)
  {
    const float x = p_in.xyz[k].x, y = p_in.xyz[k].y, z = p_in.xyz[k].z;
    p_out.xyz[k].x = ((R[0] * x + R[1] * y) + R[2] * z) + t[0];
    p_out.xyz[k].y = ((R[3] * x + R[4] * y) + R[5] * z) + t[1];
    p_out.xyz[k].z = ((R[6] * x + R[7] * y) + R[8] * z) + t[2];
  }
}
}  // namespace tf2
