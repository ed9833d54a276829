// oracle/shim: TEST INFRASTRUCTURE ONLY. message_filters::Subscriber only names the topic; tf2_ros::MessageFilter
// (shim) does the subscription.
#pragma once
#include <string>

#include <ros/shim_bus.h>
namespace message_filters
{
template <class M>
class Subscriber
{
public:
  Subscriber(ros::NodeHandle&, const std::string& topic, uint32_t) : topic_(topic) {}
  std::string getTopic() const { return topic_; }
private:
  std::string topic_;
};
}  // namespace message_filters
