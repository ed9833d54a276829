// oracle/shim: TEST INFRASTRUCTURE ONLY. tf2::getYaw as ros/geometry2 tf2/utils.h + tf2/impl/utils.h define it
// (quaternion -> yaw with the two gimbal branches at |sarg| >= 0.99999).
#pragma once
#include <cmath>

#include <tf2_geometry_msgs/tf2_geometry_msgs.h>
namespace tf2
{
inline double getYaw(const tf2::Quaternion& q)
{
  double yaw;
  const double sqw = q.w() * q.w();
  const double sqx = q.x() * q.x();
  const double sqy = q.y() * q.y();
  const double sqz = q.z() * q.z();
  // normalization added from urdfom_headers
  const double sarg = -2 * (q.x() * q.z() - q.w() * q.y()) / (sqx + sqy + sqz + sqw);
  if (sarg <= -0.99999)
    yaw = -2 * std::atan2(q.y(), q.x());
  else if (sarg >= 0.99999)
    yaw = 2 * std::atan2(q.y(), q.x());
  else
    yaw = std::atan2(2 * (q.x() * q.y() + q.w() * q.z()), sqw + sqx - sqy - sqz);
  return yaw;
}
inline double getYaw(const geometry_msgs::Quaternion& m)
{
  tf2::Quaternion q;
  fromMsg(m, q);
  return getYaw(q);
}
}  // namespace tf2
