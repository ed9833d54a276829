// Host mirror of the reference's map classes over the C ABI of libbamcl_cuda.so.
// OccupancyMap: src/amcl/map/occupancy_map.cpp:30-135 (bookkeeping, conversions), :138-160 (updateDistancesLUT),
// :257-364 (calcRange). OctoMap: src/amcl/map/octomap.cpp:30-126, :128-150, :174-207, :336-350.
// Coordinate conversions are host arithmetic; everything that builds or walks a field runs on the device.
#include <cmath>
#include <cstdlib>
#include <stdexcept>
#include <string>

#include "amcl/compat.h"
#include "amcl/map/occupancy_map.h"
#include "amcl/map/octomap.h"
#include "amcl/pf/particle_filter.h"
#include "bamcl.h"

namespace badger_amcl
{
static void check(int rc, const char* what)
{
  if (rc != BAMCL_OK)
    throw std::runtime_error(std::string(what) + ": " + bamcl_last_error());
}

static thread_local std::string g_device_error;
const char* lastDeviceError() { return g_device_error.c_str(); }
void clearDeviceError() { g_device_error.clear(); }
void deviceFail(const char* what) { g_device_error = what; }
bool deviceOk(int rc, const char* what)
{
  if (rc == BAMCL_OK)
    return true;
  g_device_error = std::string(what) + ": " + bamcl_last_error();
  return false;
}

bamcl_ctx* deviceContext()
{
  static bamcl_ctx* ctx = nullptr;
  if (!ctx)
  {
    int device = 0;
    if (const char* e = std::getenv("BAMCL_DEVICE"))
      device = std::atoi(e);
    else if (const char* e2 = std::getenv("LOCAL_RANK"))
      device = std::atoi(e2);
    check(bamcl_ctx_create(device, &ctx), "bamcl_ctx_create");  // no CUDA device => throws: there is no CPU path
  }
  return ctx;
}

// ------------------------------------------------------------------------------------------ OccupancyMap
OccupancyMap::OccupancyMap(double resolution)
    : Map(resolution), size_x_(0), size_y_(0), dev_(nullptr), cells_dirty_(true), dist_host_valid_(false)
{
  max_distance_to_object_ = 0.0;
}

OccupancyMap::~OccupancyMap()
{
  if (dev_)
    bamcl_map2d_destroy(dev_);
}

void OccupancyMap::setOrigin(const pcl::PointXYZ& origin)
{
  origin_ = origin;
  if (dev_)  // the device object bakes the origin in
  {
    bamcl_map2d_destroy(dev_);
    dev_ = nullptr;
    cells_dirty_ = true;
    distances_lut_created_ = false;
  }
}

std::vector<int> OccupancyMap::getSize() { return {size_x_, size_y_}; }

void OccupancyMap::setSize(std::vector<int> size_vec)
{
  size_x_ = size_vec[0];
  size_y_ = size_vec[1];
  cells_.resize((size_t)size_vec[0] * size_vec[1]);  // value-initialised = CELL_UNKNOWN, as the reference's resize
  if (dev_)
  {
    bamcl_map2d_destroy(dev_);
    dev_ = nullptr;
    distances_lut_created_ = false;
  }
  cells_dirty_ = true;
}

double OccupancyMap::getMaxDistanceToObject() { return max_distance_to_object_; }

void OccupancyMap::convertMapToWorld(const std::vector<int>& map_coords, std::vector<double>* world_coords)
{
  (*world_coords)[0] = origin_.x + (map_coords[0] - size_x_ / 2) * resolution_;
  (*world_coords)[1] = origin_.y + (map_coords[1] - size_y_ / 2) * resolution_;
}

void OccupancyMap::convertWorldToMap(const std::vector<double>& world_coords, std::vector<int>* map_coords)
{
  (*map_coords)[0] = std::floor((world_coords[0] - origin_.x) / resolution_ + 0.5) + size_x_ / 2;
  (*map_coords)[1] = std::floor((world_coords[1] - origin_.y) / resolution_ + 0.5) + size_y_ / 2;
}

bool OccupancyMap::isValid(const std::vector<int>& coords)
{
  return (coords[0] >= 0) && (coords[0] < size_x_) && (coords[1] >= 0) && (coords[1] < size_y_);
}

unsigned int OccupancyMap::computeCellIndex(int i, int j) { return i + j * unsigned(size_x_); }

void OccupancyMap::setCellState(int index, MapCellState state)
{
  cells_[index] = (int8_t)state;
  cells_dirty_ = true;
}

MapCellState OccupancyMap::getCellState(int i, int j) { return (MapCellState)cells_[computeCellIndex(i, j)]; }

void OccupancyMap::syncCells()
{
  if (!dev_)
  {
    check(bamcl_map2d_create(deviceContext(), size_x_, size_y_, resolution_, origin_.x, origin_.y, &dev_), "bamcl_map2d_create");
    cells_dirty_ = true;
  }
  if (cells_dirty_ && !cells_.empty())
  {
    check(bamcl_map2d_set_cells(dev_, cells_.data(), (long)cells_.size()), "bamcl_map2d_set_cells");
    cells_dirty_ = false;
  }
}

bamcl_map2d* OccupancyMap::deviceHandle()
{
  syncCells();
  return dev_;
}

void OccupancyMap::updateDistancesLUT(double max_distance_to_object)
{
  max_distance_to_object_ = max_distance_to_object;
  if (max_distance_to_object_ == 0.0)  // occupancy_map.cpp:141-145
    return;
  syncCells();
  check(bamcl_map2d_build_distances(dev_, max_distance_to_object), "bamcl_map2d_build_distances");
  dist_host_valid_ = false;
  distances_lut_created_ = true;
}

void OccupancyMap::setDistancesLUT(const std::vector<float>& distances, double max_distance_to_object)
{
  max_distance_to_object_ = max_distance_to_object;
  syncCells();
  check(bamcl_map2d_upload_distances(dev_, distances.data(), (long)distances.size(), max_distance_to_object),
        "bamcl_map2d_upload_distances");
  distances_lut_ = distances;
  dist_host_valid_ = true;
  distances_lut_created_ = true;
}

float OccupancyMap::getDistanceToObject(int i, int j)
{
  if (i >= 0 && i < size_x_ && j >= 0 && j < size_y_)
  {
    if (!dist_host_valid_)
    {
      if (!distances_lut_created_)
        throw std::runtime_error("OccupancyMap::getDistanceToObject: no distance cache (the reference reads an empty vector here)");
      distances_lut_.resize((size_t)size_x_ * size_y_);
      check(bamcl_map2d_download_distances(dev_, distances_lut_.data(), (long)distances_lut_.size()),
            "bamcl_map2d_download_distances");
      dist_host_valid_ = true;
    }
    return distances_lut_[computeCellIndex(i, j)];
  }
  return max_distance_to_object_;
}

double OccupancyMap::calcRange(double ox, double oy, double oa, double max_range)
{
  syncCells();
  double r = 0.0;
  check(bamcl_map2d_calc_range(dev_, &ox, &oy, &oa, 1, max_range, &r, nullptr), "bamcl_map2d_calc_range");
  return r;
}

void OccupancyMap::calcRanges(const std::vector<double>& ox, const std::vector<double>& oy, const std::vector<double>& oa,
                              double max_range, std::vector<double>* ranges)
{
  syncCells();
  ranges->resize(oa.size());
  if (oa.empty())
    return;
  check(bamcl_map2d_calc_range(dev_, ox.data(), oy.data(), oa.data(), (long)oa.size(), max_range, ranges->data(), nullptr),
        "bamcl_map2d_calc_range");
}

// ------------------------------------------------------------------------------------------ OctoMap
OctoMap::OctoMap(double resolution) : OctoMap(resolution, false) {}

OctoMap::OctoMap(double resolution, bool)
    : Map(resolution), dev_(nullptr), cropped_min_cells_(3), cropped_max_cells_(3), field_host_valid_(false),
      max_distance_ratio_(0.0)
{
}

OctoMap::~OctoMap()
{
  if (dev_)
    bamcl_map3d_destroy(dev_);
}

void OctoMap::refreshBounds()
{
  check(bamcl_map3d_get_bounds(dev_, cropped_min_cells_.data(), cropped_max_cells_.data()), "bamcl_map3d_get_bounds");
}

void OctoMap::initFromOctree(std::shared_ptr<octomap::OcTree> octree, double max_distance_to_object)
{
  max_distance_to_object_ = max_distance_to_object;
  max_distance_ratio_ = max_distance_to_object_ / 255;
  if (dev_)
    bamcl_map3d_destroy(dev_);
  dev_ = nullptr;
#ifdef BAMCL_HAVE_OCTOMAP
  // a real octomap::OcTree: the occupied leaves through its public iterator (octomap.cpp:220-227 reads the same), leaf
  // centre (k + 0.5) * resolution -> integer key k
  std::vector<int> keys;
  octree->expand();
  for (octomap::OcTree::leaf_iterator it = octree->begin_leafs(), end = octree->end_leafs(); it != end; ++it)
    if (octree->isNodeOccupied(*it))
    {
      keys.push_back((int)std::floor(it.getX() / resolution_));
      keys.push_back((int)std::floor(it.getY() / resolution_));
      keys.push_back((int)std::floor(it.getZ() / resolution_));
    }
#else
  const std::vector<int>& keys = octree->occupiedKeys();
#endif
  check(bamcl_map3d_create(deviceContext(), resolution_, max_distance_to_object, keys.data(), (long)(keys.size() / 3), &dev_),
        "bamcl_map3d_create");
  refreshBounds();
  field_host_valid_ = false;
  distances_lut_created_ = false;
}

void OctoMap::getMinMaxCells(std::vector<int>* min_cells, std::vector<int>* max_cells)
{
  (*min_cells) = cropped_min_cells_;
  (*max_cells) = cropped_max_cells_;
}

void OctoMap::convertMapToWorld(const std::vector<int>& map_coords, std::vector<double>* world_coords)
{
  (*world_coords)[0] = map_coords[0] * resolution_;
  (*world_coords)[1] = map_coords[1] * resolution_;
  if (map_coords.size() > 2)
    (*world_coords)[2] = map_coords[2] * resolution_;
}

void OctoMap::convertWorldToMap(const std::vector<double>& world_coords, std::vector<int>* map_coords)
{
  (*map_coords)[0] = std::floor(world_coords[0] / resolution_ + 0.5);
  (*map_coords)[1] = std::floor(world_coords[1] / resolution_ + 0.5);
  if (world_coords.size() > 2)
    (*map_coords)[2] = std::floor(world_coords[2] / resolution_ + 0.5);
}

bool OctoMap::isPoseValid(const int i, const int j)
{
  return (i <= cropped_max_cells_[0] and i >= cropped_min_cells_[0] and j <= cropped_max_cells_[1] and
          j >= cropped_min_cells_[1]);
}

bool OctoMap::isVoxelValid(const int i, const int j, const int k)
{
  return isPoseValid(i, j) and k <= cropped_max_cells_[2] and k >= cropped_min_cells_[2];
}

double OctoMap::getMaxDistanceToObject() { return max_distance_to_object_; }

void OctoMap::setMapBounds(const std::vector<double>& map_min, const std::vector<double>& map_max)
{
  if (!dev_)
    throw std::runtime_error("OctoMap::setMapBounds before initFromOctree");
  check(bamcl_map3d_set_bounds(dev_, map_min.data(), map_max.data()), "bamcl_map3d_set_bounds");
  refreshBounds();
  field_host_valid_ = false;
  distances_lut_created_ = max_distance_to_object_ != 0.0;
}

void OctoMap::updateDistancesLUT()
{
  if (!dev_)
    throw std::runtime_error("OctoMap::updateDistancesLUT before initFromOctree");
  if (max_distance_to_object_ == 0.0)  // octomap.cpp:176-180
    return;
  check(bamcl_map3d_build_distances(dev_), "bamcl_map3d_build_distances");
  field_host_valid_ = false;
  distances_lut_created_ = true;
}

void OctoMap::setDistancesLUT(const std::vector<uint8_t>& dense)
{
  check(bamcl_map3d_upload_distances(dev_, dense.data(), (long)dense.size()), "bamcl_map3d_upload_distances");
  field_ = dense;
  field_host_valid_ = true;
  distances_lut_created_ = true;
}

double OctoMap::getDistanceToObject(int i, int j, int k)
{
  if (distances_lut_created_ and !isVoxelValid(i, j, k))  // octomap.cpp:338-341
    return max_distance_to_object_;
  const long W = cropped_max_cells_[0] - cropped_min_cells_[0] + 1, H = cropped_max_cells_[1] - cropped_min_cells_[1] + 1,
             D = cropped_max_cells_[2] - cropped_min_cells_[2] + 1;
  if (!field_host_valid_)
  {
    if (!distances_lut_created_)
      throw std::runtime_error("OctoMap::getDistanceToObject: no distance cache");
    field_.resize((size_t)(W * H * D));
    check(bamcl_map3d_download_distances(dev_, field_.data(), (long)field_.size()), "bamcl_map3d_download_distances");
    field_host_valid_ = true;
  }
  const long idx = ((long)(k - cropped_min_cells_[2]) * H + (j - cropped_min_cells_[1])) * W + (i - cropped_min_cells_[0]);
  return field_[(size_t)idx] * max_distance_ratio_;
}
}  // namespace badger_amcl
