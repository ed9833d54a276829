// oracle/shim: TEST INFRASTRUCTURE ONLY. tf2::Stamped<T> (plain data, no arithmetic) + the message structs the callers
// of this header expect to see with it.
#pragma once
#include <string>
#include <vector>

#include <ros/shim_msgs.h>
namespace tf2
{
template <typename T>
class Stamped : public T
{
public:
  ros::Time stamp_;
  std::string frame_id_;
  Stamped() {}
  Stamped(const T& input, const ros::Time& timestamp, const std::string& frame_id) : T(input), stamp_(timestamp), frame_id_(frame_id) {}
  void setData(const T& input) { *static_cast<T*>(this) = input; }
};
}  // namespace tf2
