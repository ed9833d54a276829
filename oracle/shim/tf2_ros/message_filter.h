// oracle/shim: TEST INFRASTRUCTURE ONLY. tf2_ros::MessageFilter passes every message straight to its callback (the real
// one holds a message until its frame can be transformed; the harness's tf table is complete before it publishes).
#pragma once
#include <functional>
#include <memory>
#include <string>

#include <message_filters/subscriber.h>
#include <tf2_ros/buffer.h>
namespace tf2_ros
{
template <class M>
class MessageFilter
{
public:
  template <class F>
  MessageFilter(F& f, Buffer&, const std::string&, uint32_t, const ros::NodeHandle&) : topic_(f.getTopic()), id_(0) {}
  ~MessageFilter()
  {
    if (id_)
      ros::shim::bus().unsubscribe(topic_, id_);
  }
  template <class C>
  void registerCallback(const C& callback)
  {
    std::function<void(const std::shared_ptr<M const>&)> cb = callback;
    id_ = ros::shim::bus().subscribe(topic_, [cb](const std::shared_ptr<const void>& m) { cb(std::static_pointer_cast<const M>(m)); });
  }
private:
  MessageFilter(const MessageFilter&);
  std::string topic_;
  long id_;
};
}  // namespace tf2_ros
