// oracle/shim: TEST INFRASTRUCTURE ONLY. pcl::PointCloud<T> look-alike.
#pragma once
#include <memory>
#include <string>
#include <vector>
#include <cstdint>
namespace pcl
{
struct ShimHeader
{
  std::string frame_id;
};
template <typename T>
class PointCloud
{
public:
  typedef std::shared_ptr<PointCloud<T>> Ptr;
  typedef typename std::vector<T>::iterator iterator;
  ShimHeader header;
  uint32_t height = 0, width = 0;
  std::vector<T> points;
  iterator begin() { return points.begin(); }
  iterator end() { return points.end(); }
  std::size_t size() const { return points.size(); }
  const T& at(std::size_t n) const { return points.at(n); }
  T& at(std::size_t n) { return points.at(n); }
  void push_back(const T& pt)  // pcl::PointCloud::push_back
  {
    points.push_back(pt);
    width = static_cast<uint32_t>(points.size());
    height = 1;
  }
};
}  // namespace pcl
