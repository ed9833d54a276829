// Host mirror of include/amcl/map/map.h:35-57 of the reference (same namespace, class and method names).
#ifndef BAMCL_AMCL_MAP_MAP_H
#define BAMCL_AMCL_MAP_MAP_H
#include <atomic>
#include <vector>

#include "amcl/compat.h"

namespace badger_amcl
{
class Map
{
public:
  explicit Map(double resolution) : origin_(), resolution_(resolution), max_distance_to_object_(0.0), distances_lut_created_(false) {}
  virtual ~Map() = default;
  virtual void convertMapToWorld(const std::vector<int>& map_coords, std::vector<double>* world_coords) = 0;
  virtual void convertWorldToMap(const std::vector<double>& world_coords, std::vector<int>* map_coords) = 0;
  virtual bool isDistancesLUTCreated() { return distances_lut_created_; }  // map.cpp:33-36
  virtual pcl::PointXYZ getOrigin() { return origin_; }                    // map.cpp:38-41
  double getResolution() const { return resolution_; }

protected:
  pcl::PointXYZ origin_;
  double resolution_;
  double max_distance_to_object_;
  std::atomic<bool> distances_lut_created_;
};
}  // namespace badger_amcl
#endif
