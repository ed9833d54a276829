// Host mirror of include/amcl/map/octomap.h:38-124 of the reference. The distance cache lives on the device as
// a dense u8 field (csrc/map3d.cuh); getDistanceToObject reads a host copy fetched on first use.
#ifndef BAMCL_AMCL_MAP_OCTOMAP_H
#define BAMCL_AMCL_MAP_OCTOMAP_H
#include <cstdint>
#include <memory>
#include <vector>

#include "amcl/map/map.h"

struct bamcl_map3d;

namespace badger_amcl
{
class OctoMap : public Map
{
public:
  explicit OctoMap(double resolution);
  OctoMap(double resolution, bool publish_distances_lut);  // the debug cloud publisher is ROS glue: ignored
  virtual ~OctoMap();
  OctoMap(const OctoMap&) = delete;
  OctoMap& operator=(const OctoMap&) = delete;
  virtual void initFromOctree(std::shared_ptr<octomap::OcTree> octree, double max_distance_to_object);
  virtual void convertMapToWorld(const std::vector<int>& map_coords, std::vector<double>* world_coords);
  virtual void convertWorldToMap(const std::vector<double>& world_coords, std::vector<int>* map_coords);
  virtual bool isPoseValid(const int i, const int j);
  virtual bool isVoxelValid(const int i, const int j, const int k);
  virtual void getMinMaxCells(std::vector<int>* min_cells, std::vector<int>* max_cells);
  virtual void setMapBounds(const std::vector<double>& map_min, const std::vector<double>& map_max);
  virtual void updateDistancesLUT();
  virtual double getMaxDistanceToObject();
  double getDistanceToObject(int i, int j, int k);

  // ---- additions of the device port
  void setDistancesLUT(const std::vector<uint8_t>& dense);  // x fastest, then y, then z over getMinMaxCells
  bamcl_map3d* deviceHandle() { return dev_; }

private:
  void refreshBounds();
  bamcl_map3d* dev_;
  std::vector<int> cropped_min_cells_, cropped_max_cells_;
  std::vector<uint8_t> field_;  // host copy of the cache
  bool field_host_valid_;
  double max_distance_ratio_;
};
}  // namespace badger_amcl
#endif
