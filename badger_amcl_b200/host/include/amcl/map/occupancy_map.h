// Host mirror of include/amcl/map/occupancy_map.h:37-126 of the reference. The cell states live in a host
// vector (callers fill them one setCellState at a time, node_2d.cpp:265-295) and are pushed to the device when
// the distance cache is built; the distance cache and calcRange run on the device (libbamcl_cuda.so).
#ifndef BAMCL_AMCL_MAP_OCCUPANCY_MAP_H
#define BAMCL_AMCL_MAP_OCCUPANCY_MAP_H
#include <cstdint>
#include <vector>

#include "amcl/map/map.h"

struct bamcl_map2d;

namespace badger_amcl
{
enum MapCellState
{
  CELL_FREE = -1,
  CELL_UNKNOWN = 0,
  CELL_OCCUPIED = 1
};

class OccupancyMap : public Map
{
public:
  explicit OccupancyMap(double resolution);
  virtual ~OccupancyMap();
  OccupancyMap(const OccupancyMap&) = delete;
  OccupancyMap& operator=(const OccupancyMap&) = delete;
  virtual void convertMapToWorld(const std::vector<int>& map_coords, std::vector<double>* world_coords);
  virtual void convertWorldToMap(const std::vector<double>& world_coords, std::vector<int>* map_coords);
  virtual bool isValid(const std::vector<int>& coords);
  virtual void setOrigin(const pcl::PointXYZ& origin);
  virtual std::vector<int> getSize();
  virtual void setSize(std::vector<int> size_vec);
  virtual void updateDistancesLUT(double max_distance_to_object);
  virtual double calcRange(double ox, double oy, double oa, double max_range);
  virtual unsigned int computeCellIndex(int i, int j);
  virtual double getMaxDistanceToObject();
  virtual MapCellState getCellState(int i, int j);
  virtual void setCellState(int index, MapCellState state);
  float getDistanceToObject(int i, int j);

  // ---- additions of the device port
  // install a precomputed cache (e.g. one saved from the reference's own brushfire)
  void setDistancesLUT(const std::vector<float>& distances, double max_distance_to_object);
  // batch form of calcRange (one launch for n rays)
  void calcRanges(const std::vector<double>& ox, const std::vector<double>& oy, const std::vector<double>& oa,
                  double max_range, std::vector<double>* ranges);
  bamcl_map2d* deviceHandle();  // creates / refreshes the device object; used by PlanarScanner

protected:
  int size_x_, size_y_;
  std::vector<int8_t> cells_;
  std::vector<float> distances_lut_;  // host copy, fetched on the first getDistanceToObject after a build

private:
  void syncCells();
  bamcl_map2d* dev_;
  bool cells_dirty_, dist_host_valid_;
};
}  // namespace badger_amcl
#endif
