// oracle/shim: TEST INFRASTRUCTURE ONLY. dynamic_reconfigure::Server: setCallback calls the callback once with the
// start-up configuration (as the real server does) and leaves a hook on the bus with which the harness triggers a
// reconfigure from the current parameter table.
#pragma once
#include <cstdint>
#include <functional>

#include <ros/shim_bus.h>
namespace dynamic_reconfigure
{
template <class Config>
class Server
{
public:
  typedef std::function<void(Config&, uint32_t)> CallbackType;
  // the real constructor reads the configuration from the parameter server and writes it back complete, so parameters
  // the launch file leaves unset get the .cfg defaults BEFORE the node's own param(name, var, default) calls run
  explicit Server(const ros::NodeHandle&)
  {
    Config c;
    c.fromParameterTable();
    c.toParameterTable();
  }
  ~Server() { ros::shim::bus().reconfigure = nullptr; }
  void setCallback(const CallbackType& cb)
  {
    cb_ = cb;
    ros::shim::bus().reconfigure = [this]() {
      Config c;
      c.fromParameterTable();
      cb_(c, 0u);
    };
    Config c;
    c.fromParameterTable();
    cb_(c, ~0u);
  }
private:
  CallbackType cb_;
};
}  // namespace dynamic_reconfigure
