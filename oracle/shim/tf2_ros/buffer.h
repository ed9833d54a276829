// oracle/shim: TEST INFRASTRUCTURE ONLY. tf2_ros::Buffer as a table the harness fills: (target frame, source frame) ->
// transform, time-independent; a missing pair throws tf2::LookupException like the real buffer. Every Buffer instance
// sees the same table (the reference's Node and Node2D/3D each own one, all listening to the same /tf).
#pragma once
#include <map>
#include <string>
#include <utility>

#include <ros/shim_msgs.h>
#include <tf2/exceptions.h>
#include <tf2/transform_datatypes.h>
namespace tf2_ros
{
namespace shim
{
typedef std::map<std::pair<std::string, std::string>, geometry_msgs::Transform> TfTable;
inline TfTable& tf_table()
{
  static TfTable t;
  return t;
}
}  // namespace shim
class Buffer
{
public:
  Buffer() {}
  geometry_msgs::TransformStamped lookupTransform(const std::string& target_frame, const std::string& source_frame,
                                                  const ros::Time& time, const ros::Duration = ros::Duration(0.0)) const
  {
    shim::TfTable::const_iterator it = shim::tf_table().find(std::make_pair(target_frame, source_frame));
    if (it == shim::tf_table().end())
      throw tf2::LookupException("no transform from " + source_frame + " to " + target_frame);
    geometry_msgs::TransformStamped out;
    out.header.stamp = time;
    out.header.frame_id = target_frame;
    out.child_frame_id = source_frame;
    out.transform = it->second;
    return out;
  }
  // the "advanced" API (two times, a fixed frame): with a time-independent table both times see the same transforms
  geometry_msgs::TransformStamped lookupTransform(const std::string& target_frame, const ros::Time& target_time,
                                                  const std::string& source_frame, const ros::Time&,
                                                  const std::string&, const ros::Duration = ros::Duration(0.0)) const
  {
    return lookupTransform(target_frame, source_frame, target_time);
  }
};
}  // namespace tf2_ros
