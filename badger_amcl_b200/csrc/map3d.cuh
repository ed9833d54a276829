// OctoMap on the device (include/amcl/map/octomap.h:61-106): the distance cache as ONE dense u8 field over the
// map bounds, x fastest, then y, then z (the reference keeps lazily allocated z-columns behind a per-(x, y)
// offset table, octomap.h:96-106; on a 180 GB part the dense field is simply resident, one gather per look-up
// instead of two dependent ones). Value q = floor(min(d, max) / max * 255) (octomap.cpp:329-331); 255 also
// stands for "nothing within max_distance".
#pragma once
#include <vector>

#include "common.cuh"

struct bamcl_map3d
{
  bamcl_ctx* ctx = nullptr;
  double resolution = 0.0, max_distance = 0.0;
  int cmin[3] = {0, 0, 0}, cmax[3] = {0, 0, 0};  // cropped voxel bounds, inclusive (octomap.cpp:53-74, 128-150)
  int W = 0, H = 0, D = 0;
  std::vector<int> src;   // occupied voxels (i, j, k) = convertWorldToMap(leaf centre), octomap.cpp:220-227
  bamcl::DevBuf d_src;    // same, on the device
  bamcl::DevBuf field;    // u8[W*H*D]
  bool has_dist = false;
  uint64_t version = 0;
};
