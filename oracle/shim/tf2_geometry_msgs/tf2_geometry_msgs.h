// oracle/shim: TEST INFRASTRUCTURE ONLY. tf2::fromMsg / tf2::toMsg between geometry_msgs and tf2 types, as
// ros/geometry2 tf2_geometry_msgs defines them (plain member copies; toMsg takes the rotation via Matrix3x3::getRotation).
#pragma once
#include <tf2/LinearMath/Transform.h>
#include <tf2/transform_datatypes.h>
namespace tf2
{
inline void fromMsg(const geometry_msgs::Transform& in, tf2::Transform& out)
{
  out = tf2::Transform(tf2::Quaternion(in.rotation.x, in.rotation.y, in.rotation.z, in.rotation.w),
                       tf2::Vector3(in.translation.x, in.translation.y, in.translation.z));
}
inline geometry_msgs::TransformStamped toMsg(const tf2::Stamped<tf2::Transform>& in)
{
  geometry_msgs::TransformStamped out;
  out.frame_id = in.frame_id_;
  out.transform.translation.x = in.getOrigin().x();
  out.transform.translation.y = in.getOrigin().y();
  out.transform.translation.z = in.getOrigin().z();
  const tf2::Quaternion q = in.getRotation();
  out.transform.rotation.x = q.x();
  out.transform.rotation.y = q.y();
  out.transform.rotation.z = q.z();
  out.transform.rotation.w = q.w();
  return out;
}
}  // namespace tf2
