// OccupancyMap on the device: cell states, float distance cache, and the tiled u16 "squared cell distance"
// field the scoring kernels gather from.
#pragma once
#include "common.cuh"

struct bamcl_map2d
{
  bamcl_ctx* ctx = nullptr;
  int size_x = 0, size_y = 0;
  double resolution = 0.0;
  float origin_x = 0.f, origin_y = 0.f;
  // row-major, index = i + j*size_x (occupancy_map.cpp:107-110)
  bamcl::DevBuf cells;  // int8: -1 free, 0 unknown, +1 occupied
  bamcl::DevBuf dist;   // float distance cache (distances_lut_)
  // Tiled field for the particle x beam gathers: one u16 per cell holding the rank of k = di^2 + dj^2 of the
  // nearest obstacle (the cached distance is (float)(sqrt(k)*resolution), occupancy_map.cpp:125,229-232) or
  // rep_k.size() for cells beyond the radius. 8x8-cell tiles = one 128 B line; inside a tile
  // four 4x4 blocks = one 32 B sector each, so a scan's end points (which trace wall contours) share
  // sectors in every direction.
  bamcl::DevBuf kmap;
  bamcl::DevBuf rank_of_k;  // u16[radius^2 + 2]: k -> rank (0xFFFF = not a sum of two squares)
  std::vector<int> rep_k;   // rank -> k for the representable k <= radius^2 (host copy, for the value tables)
  int tiles_x = 0, tiles_y = 0;
  int cell_radius = 0;  // floor(max_distance / resolution), occupancy_map.cpp:125
  double max_distance = 0.0;
  bool has_cells = false, has_dist = false;
  // ray march (score_beam.cu): u8 per cell = floor(distance in cells to the nearest cell that stops a ray -- not free, or
  // off the grid), capped; the march jumps over the cells this proves free
  bamcl::DevBuf clear;
  // Node2D::updateFreeSpaceIndices (node_2d.cpp:320-338): (i, j) of the free cells far enough from obstacles, i-major
  bamcl::DevBuf free_ij;
  long free_count = -1;  // -1 = not built for the current cache
  uint64_t version = 0;  // bumped whenever the cache changes (scan staging keys on it)
  int layout = 0;        // tiled_offset layout (see below); fixed at creation
};

namespace bamcl
{
// u16 element offset of cell (i, j) in the tiled field.
//   layout 0: 8x8-cell tiles (one 128 B line), four 4x4 blocks (one 32 B sector each) inside a tile
//   layout 1: 8x8-cell tiles, row-major inside the tile (sector = 8x2 cells); cheaper address arithmetic
//   layout 2: 4x4-cell blocks (one sector), blocks row-major (line = 16x4 cells)
template <int LAYOUT>
__host__ __device__ __forceinline__ unsigned tiled_offset_t(int i, int j, int tiles_x)
{
  if (LAYOUT == 0)
  {
    const unsigned tile = (unsigned)(j >> 3) * (unsigned)tiles_x + (unsigned)(i >> 3);
    const unsigned low = (unsigned)(i & 3) | ((unsigned)(j & 3) << 2) | ((unsigned)(i & 4) << 2) | ((unsigned)(j & 4) << 3);
    return (tile << 6) | low;
  }
  else if (LAYOUT == 1)
  {
    // (j>>3)*pitch + (i>>3)*64 + (j&7)*8 + (i&7)  with pitch = tiles_x*64
    const unsigned fx = (unsigned)i + ((unsigned)i & ~7u) * 7u;
    const unsigned fy = (unsigned)j * 8u + ((unsigned)j >> 3) * ((unsigned)tiles_x * 64u - 64u);
    return fx + fy;
  }
  else
  {
    // ((j>>2)*bx + (i>>2))*16 + (j&3)*4 + (i&3)  with bx = tiles_x*2 blocks per row
    const unsigned fx = (unsigned)i + ((unsigned)i & ~3u) * 3u;
    const unsigned fy = (unsigned)j * 4u + ((unsigned)j >> 2) * ((unsigned)tiles_x * 32u - 16u);
    return fx + fy;
  }
}
__host__ __device__ __forceinline__ unsigned tiled_offset(int i, int j, int tiles_x, int layout)
{
  return layout == 0 ? tiled_offset_t<0>(i, j, tiles_x) : layout == 1 ? tiled_offset_t<1>(i, j, tiles_x) : tiled_offset_t<2>(i, j, tiles_x);
}
}  // namespace bamcl
