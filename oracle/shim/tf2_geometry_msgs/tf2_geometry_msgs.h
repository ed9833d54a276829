// oracle/shim: TEST INFRASTRUCTURE ONLY. tf2::fromMsg / tf2::toMsg / tf2::doTransform between geometry_msgs and tf2 types,
// as ros/geometry2 tf2_geometry_msgs defines them: plain member copies; toMsg(Transform) takes the rotation via
// Matrix3x3::getRotation; doTransform composes tf2::Transforms (Pose, TransformStamped) or multiplies quaternions.
#pragma once
#include <tf2/LinearMath/Transform.h>
#include <tf2/transform_datatypes.h>
namespace tf2
{
inline geometry_msgs::Quaternion toMsg(const tf2::Quaternion& in)
{
  geometry_msgs::Quaternion out;
  out.w = in.w();
  out.x = in.x();
  out.y = in.y();
  out.z = in.z();
  return out;
}
inline void fromMsg(const geometry_msgs::Quaternion& in, tf2::Quaternion& out) { out = tf2::Quaternion(in.x, in.y, in.z, in.w); }
inline geometry_msgs::Vector3 toMsg(const tf2::Vector3& in)
{
  geometry_msgs::Vector3 out;
  out.x = in.x();
  out.y = in.y();
  out.z = in.z();
  return out;
}
inline geometry_msgs::Point& toMsg(const tf2::Vector3& in, geometry_msgs::Point& out)
{
  out.x = in.x();
  out.y = in.y();
  out.z = in.z();
  return out;
}
inline void fromMsg(const geometry_msgs::Point& in, tf2::Vector3& out) { out = tf2::Vector3(in.x, in.y, in.z); }
inline geometry_msgs::Transform toMsg(const tf2::Transform& in)
{
  geometry_msgs::Transform out;
  out.translation = toMsg(in.getOrigin());
  out.rotation = toMsg(in.getRotation());
  return out;
}
inline void fromMsg(const geometry_msgs::Transform& in, tf2::Transform& out)
{
  out = tf2::Transform(tf2::Quaternion(in.rotation.x, in.rotation.y, in.rotation.z, in.rotation.w),
                       tf2::Vector3(in.translation.x, in.translation.y, in.translation.z));
}
inline geometry_msgs::Pose& toMsg(const tf2::Transform& in, geometry_msgs::Pose& out)
{
  toMsg(in.getOrigin(), out.position);
  out.orientation = toMsg(in.getRotation());
  return out;
}
inline void fromMsg(const geometry_msgs::Pose& in, tf2::Transform& out)
{
  out.setOrigin(tf2::Vector3(in.position.x, in.position.y, in.position.z));
  out.setRotation(tf2::Quaternion(in.orientation.x, in.orientation.y, in.orientation.z, in.orientation.w));
}
inline geometry_msgs::TransformStamped toMsg(const tf2::Stamped<tf2::Transform>& in)
{
  geometry_msgs::TransformStamped out;
  out.header.stamp = in.stamp_;
  out.header.frame_id = in.frame_id_;
  out.transform = toMsg(static_cast<const tf2::Transform&>(in));
  return out;
}
inline void fromMsg(const geometry_msgs::TransformStamped& msg, tf2::Stamped<tf2::Transform>& out)
{
  out.stamp_ = msg.header.stamp;
  out.frame_id_ = msg.header.frame_id;
  tf2::Transform tmp;
  fromMsg(msg.transform, tmp);
  out.setData(tmp);
}
inline void doTransform(const geometry_msgs::Pose& t_in, geometry_msgs::Pose& t_out, const geometry_msgs::TransformStamped& transform)
{
  tf2::Vector3 v;
  fromMsg(t_in.position, v);
  tf2::Quaternion r;
  fromMsg(t_in.orientation, r);
  tf2::Transform t;
  fromMsg(transform.transform, t);
  tf2::Transform v_out = t * tf2::Transform(r, v);
  toMsg(v_out, t_out);
}
inline void doTransform(const geometry_msgs::Quaternion& t_in, geometry_msgs::Quaternion& t_out,
                        const geometry_msgs::TransformStamped& transform)
{
  tf2::Quaternion t, q_in;
  fromMsg(transform.transform.rotation, t);
  fromMsg(t_in, q_in);
  tf2::Quaternion q_out = t * q_in;
  t_out = toMsg(q_out);
}
inline void doTransform(const geometry_msgs::TransformStamped& t_in, geometry_msgs::TransformStamped& t_out,
                        const geometry_msgs::TransformStamped& transform)
{
  tf2::Transform input;
  fromMsg(t_in.transform, input);
  tf2::Transform t;
  fromMsg(transform.transform, t);
  tf2::Transform v_out = t * input;
  t_out.transform = toMsg(v_out);
  t_out.header.stamp = transform.header.stamp;
  t_out.header.frame_id = transform.header.frame_id;
}
}  // namespace tf2
