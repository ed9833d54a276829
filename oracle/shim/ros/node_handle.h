// oracle/shim: TEST INFRASTRUCTURE ONLY. No-op NodeHandle/Publisher (only used for a debug cloud).
#pragma once
#include <string>
#include <cstdint>
namespace ros
{
class Publisher
{
public:
  template <typename T>
  void publish(const T&) const {}
};
class NodeHandle
{
public:
  template <typename T>
  Publisher advertise(const std::string&, uint32_t, bool = false) { return Publisher(); }
};
}  // namespace ros
