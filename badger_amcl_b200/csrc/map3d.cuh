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
  // What the scorer gathers from (csrc/score_cloud.cu), derived from `field` whenever it changes: the volume in bricks of
  // 4 x 4 x 2 voxels (x fastest inside a brick, bricks x fastest). `bricks` holds the 32 distance bytes of a brick = one
  // 32-byte sector; `near_bits` one word per brick, bit = "this voxel is within max_distance of an obstacle" (value < 255).
  // The bit volume (1/8 of the field: 30 MB for 2001 x 2001 x 61 voxels) stays L2-resident and answers most look-ups.
  bamcl::DevBuf bricks, near_bits;
  int CW = 0, CH = 0, CD = 0;
  double near_fraction = 0.0;  // share of voxels within max_distance of an obstacle: the bit test pays below ~1/2
  bool has_dist = false;
  uint64_t version = 0;
};

namespace bamcl
{
int map3d_build_bricks(bamcl_map3d* m);  // field -> bricks + near_bits (map3d.cu)
}
