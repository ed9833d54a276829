// oracle/shim: TEST INFRASTRUCTURE ONLY. sendTransform keeps the last transform on the bus topic "/tf".
#pragma once
#include <ros/shim_msgs.h>
namespace tf2_ros
{
class TransformBroadcaster
{
public:
  void sendTransform(const geometry_msgs::TransformStamped& t) { ros::Publisher("/tf").publish(t); }
};
}  // namespace tf2_ros
