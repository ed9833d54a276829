// OccupancyMap device object: cell upload, distance-cache build (exact capped EDT, K-e2), cache
// upload/download, batched calcRange. Reference: src/amcl/map/occupancy_map.cpp.
#include <cmath>
#include <cstring>
#include <cstdlib>

#include "map2d.cuh"

using namespace bamcl;

namespace
{
constexpr int kFarG = 0xFFFF;  // "no obstacle within the radius along this column"

// ---- K-e2 phase 1: per column i, distance (in cells, along j) to the nearest occupied cell, capped.
// One thread per column; adjacent threads touch adjacent addresses of every row (coalesced).
__global__ void k_edt_columns(const int8_t* __restrict__ cells, uint16_t* __restrict__ g, int size_x, int size_y,
                              int radius)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= size_x)
    return;
  int d = kFarG;
  for (int j = 0; j < size_y; j++)
  {
    const size_t idx = (size_t)j * size_x + i;
    d = (cells[idx] == 1) ? 0 : (d >= radius ? kFarG : d + 1);
    g[idx] = (uint16_t)d;
  }
  d = kFarG;
  for (int j = size_y - 1; j >= 0; j--)
  {
    const size_t idx = (size_t)j * size_x + i;
    d = (cells[idx] == 1) ? 0 : (d >= radius ? kFarG : d + 1);
    if (d < g[idx])
      g[idx] = (uint16_t)d;
  }
}

// ---- K-e2 phase 2: per cell, k = min over |di| <= radius of di^2 + g(i+di, j)^2. A block owns a row
// segment; the g row (with halo) is staged in shared memory once and read 2*radius+1 times.
// Writes the float cache with the reference's value formula and the tiled u16 field.
template <int BLOCK>
__global__ void k_edt_rows(const uint16_t* __restrict__ g, float* __restrict__ dist, uint16_t* __restrict__ kmap,
                           const uint16_t* __restrict__ rank_of_k, int size_x, int size_y, int tiles_x, int layout, int radius,
                           double resolution, float far_value)
{
  extern __shared__ uint16_t s_g[];  // BLOCK + 2*radius
  const int j = blockIdx.y;
  const int i0 = blockIdx.x * BLOCK;
  const int span = BLOCK + 2 * radius;
  for (int t = threadIdx.x; t < span; t += BLOCK)
  {
    const int i = i0 - radius + t;
    s_g[t] = (i >= 0 && i < size_x) ? g[(size_t)j * size_x + i] : (uint16_t)kFarG;
  }
  __syncthreads();
  const int i = i0 + threadIdx.x;
  if (i >= size_x)
    return;
  const int r2 = radius * radius;
  int best = r2 + 1;
  for (int d = -radius; d <= radius; d++)
  {
    const int gv = s_g[threadIdx.x + radius + d];
    if (gv != kFarG)
    {
      const int k2 = d * d + gv * gv;
      best = min(best, k2);
    }
  }
  // occupancy_map.cpp:229-232: accepted iff sqrt(di^2+dj^2) <= cell_radius, stored as float(distance * res)
  float v = far_value;
  int k = r2 + 1;
  if (best <= r2)
  {
    v = (float)(sqrt((double)best) * resolution);
    k = best;
  }
  dist[(size_t)j * size_x + i] = v;
  kmap[tiled_offset(i, j, tiles_x, layout)] = rank_of_k[k];
}

// ---- uploaded cache -> tiled u16 field. A value is canonical if it is (float)(sqrt(k)*res) for an
// integer k <= radius^2, or equals the far value.
__global__ void k_index_uploaded(const float* __restrict__ dist, uint16_t* __restrict__ kmap,
                                 const uint16_t* __restrict__ rank_of_k, int size_x, int size_y, int tiles_x, int layout,
                                 int radius, double resolution, float far_value, int* __restrict__ bad)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = blockIdx.y;
  if (i >= size_x)
    return;
  const float d = dist[(size_t)j * size_x + i];
  const int r2 = radius * radius;
  int k = r2 + 1;
  if (d != far_value)
  {
    const double c = (double)d / resolution;
    const long kk = llrint(c * c);
    bool ok = false;
    for (long t = kk - 1; t <= kk + 1 && !ok; t++)
      if (t >= 0 && t <= r2 && (float)(sqrt((double)t) * resolution) == d)
      {
        k = (int)t;
        ok = true;
      }
    if (!ok || rank_of_k[k] == 0xFFFF)  // not a value the brushfire can produce (k must be a sum of two squares)
    {
      atomicAdd(bad, 1);
      k = r2 + 1;
    }
  }
  kmap[tiled_offset(i, j, tiles_x, layout)] = rank_of_k[k];
}

// ---- clearance field for the ray march (score_beam.cu): per cell, floor of the Euclidean distance in cells to the
// nearest cell a ray stops at -- any cell that is not CELL_FREE, or any cell off the grid (occupancy_map.cpp:315-358) --
// capped at CLEAR_CAP. 0 = the cell itself stops the ray. Exact (separable) below the cap, so it never overestimates.
constexpr int CLEAR_CAP = 64;
// pass 1, one thread per column: distance along the column to the nearest stopping cell (grid edge included)
__global__ void __launch_bounds__(128) k_clear_cols(const int8_t* __restrict__ cells, uint8_t* __restrict__ g, int size_x, int size_y)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= size_x)
    return;
  int run = 0;  // the cell at j = -1 stops rays
  for (int j = 0; j < size_y; j++)
  {
    run = cells[(size_t)j * size_x + i] != -1 ? 0 : min(run + 1, 255);
    g[(size_t)j * size_x + i] = (uint8_t)run;
  }
  run = 0;  // j = size_y
  for (int j = size_y - 1; j >= 0; j--)
  {
    const int here = g[(size_t)j * size_x + i];
    run = here == 0 ? 0 : min(run + 1, 255);
    g[(size_t)j * size_x + i] = (uint8_t)min(here, run);
  }
}
// pass 2, one thread per cell: min over the row window of dx^2 + g^2
__global__ void __launch_bounds__(256) k_clear_rows(const uint8_t* __restrict__ g, uint8_t* __restrict__ clear, int size_x, int size_y)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
  if (i >= size_x)
    return;
  const uint8_t* row = g + (size_t)j * size_x;
  int best = (int)row[i] * (int)row[i];
  const int edge = min(i + 1, size_x - i);  // columns -1 and size_x stop rays
  best = min(best, edge * edge);
  for (int d = 1; d <= CLEAR_CAP && d * d < best; d++)
  {
    if (i - d >= 0)
    {
      const int v = row[i - d];
      best = min(best, d * d + v * v);
    }
    if (i + d < size_x)
    {
      const int v = row[i + d];
      best = min(best, d * d + v * v);
    }
  }
  int r = (int)sqrtf((float)best);
  while (r * r > best)
    r--;
  while ((r + 1) * (r + 1) <= best)
    r++;
  clear[(size_t)j * size_x + i] = (uint8_t)min(r, CLEAR_CAP);
}

// ---- calcRange (occupancy_map.cpp:257-364), one thread per ray, the reference's own cell walk.
__global__ void k_calc_range(const int8_t* __restrict__ cells, int size_x, int size_y, double resolution, float ox_f,
                             float oy_f, const double* __restrict__ ox, const double* __restrict__ oy,
                             const double* __restrict__ oa, long n, double max_range, double* __restrict__ out,
                             unsigned long long* __restrict__ steps_total)
{
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n)
    return;
  int x0 = world_to_cell(ox[t], ox_f, resolution, size_x);
  int y0 = world_to_cell(oy[t], oy_f, resolution, size_y);
  int x1 = world_to_cell(ox[t] + max_range * cos(oa[t]), ox_f, resolution, size_x);
  int y1 = world_to_cell(oy[t] + max_range * sin(oa[t]), oy_f, resolution, size_y);
  unsigned long long steps = 0;
  double r = max_range;
  if (!(x0 == x1 && y0 == y1))
  {
    const bool steep = abs(y1 - y0) > abs(x1 - x0);
    if (steep)
    {
      int tmp = x0; x0 = y0; y0 = tmp;
      tmp = x1; x1 = y1; y1 = tmp;
    }
    const int deltax = abs(x1 - x0), deltay = abs(y1 - y0);
    const int xstep = (x0 < x1) ? 1 : -1, ystep = (y0 < y1) ? 1 : -1;
    int error = 0, x = x0, y = y0;
    for (;;)
    {
      const int cx = steep ? y : x, cy = steep ? x : y;
      steps++;
      const bool valid = cx >= 0 && cx < size_x && cy >= 0 && cy < size_y;
      if (!valid || cells[(size_t)cy * size_x + cx] != -1)
      {
        r = sqrt((double)((x - x0) * (x - x0) + (y - y0) * (y - y0))) * resolution;
        break;
      }
      if (x == x1 + xstep)
        break;
      x += xstep;
      error += deltay;
      if (2 * error >= deltax)
      {
        y += ystep;
        error -= deltax;
      }
    }
  }
  out[t] = r;
  if (steps_total)
    atomicAdd(steps_total, steps);
}
}  // namespace

// the reference loop tests the start cell, then steps until x == x1 + xstep *after* testing; the kernel
// above tests the cell at x1 + xstep too (the reference's while condition is checked before the step, so
// the last tested x is x1 + xstep) -- see the parity test against the oracle.

static int build_clearance(bamcl_map2d* m)  // the ray march's clearance field, from the cell states
{
  bamcl_ctx* ctx = m->ctx;
  const size_t n = (size_t)m->size_x * m->size_y;
  if (n == 0)
    return BAMCL_OK;
  bamcl::DevBuf g;
  BAMCL_TRY(g.reserve(n));
  BAMCL_TRY(m->clear.reserve(n));
  BAMCL_LAUNCH(ctx, k_clear_cols, (m->size_x + 127) / 128, 128, 0, m->cells.as<int8_t>(), g.as<uint8_t>(), m->size_x, m->size_y);
  dim3 grid((m->size_x + 255) / 256, m->size_y);
  BAMCL_LAUNCH(ctx, k_clear_rows, grid, 256, 0, g.as<uint8_t>(), m->clear.as<uint8_t>(), m->size_x, m->size_y);
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  g.release();
  return BAMCL_OK;
}

extern "C" {

int bamcl_map2d_create(bamcl_ctx* ctx, int size_x, int size_y, double resolution, float origin_x, float origin_y,
                       bamcl_map2d** out)
{
  if (!ctx || !out || size_x < 0 || size_y < 0 || !(resolution > 0.0))
    return set_error(BAMCL_E_ARG, "bamcl_map2d_create: bad arguments");
  bamcl_map2d* m = new bamcl_map2d();
  m->ctx = ctx;
  m->size_x = size_x;
  m->size_y = size_y;
  m->resolution = resolution;
  m->origin_x = origin_x;
  m->origin_y = origin_y;
  m->tiles_x = (size_x + 7) / 8;
  m->tiles_y = (size_y + 7) / 8;
  m->layout = 1;
  if (const char* e = getenv("BAMCL_LAYOUT"))  // layout experiments (profiles/), default 1
    m->layout = atoi(e) % 3;
  *out = m;
  return BAMCL_OK;
}

int bamcl_map2d_destroy(bamcl_map2d* m)
{
  if (!m)
    return BAMCL_OK;
  cudaStreamSynchronize(m->ctx->stream);
  m->cells.release();
  m->dist.release();
  m->kmap.release();
  m->rank_of_k.release();
  m->clear.release();
  m->free_ij.release();
  delete m;
  return BAMCL_OK;
}

int bamcl_map2d_set_cells(bamcl_map2d* m, const int8_t* cells, long n)
{
  if (!m || !cells || n != (long)m->size_x * m->size_y)
    return set_error(BAMCL_E_ARG, "bamcl_map2d_set_cells: need size_x*size_y cells");
  if (n == 0)
    return BAMCL_OK;
  BAMCL_TRY(m->cells.reserve((size_t)n));
  BAMCL_CUDA_OK(cudaMemcpyAsync(m->cells.p, cells, (size_t)n, cudaMemcpyHostToDevice, m->ctx->stream));
  m->has_cells = true;
  BAMCL_TRY(build_clearance(m));
  BAMCL_CUDA_OK(cudaStreamSynchronize(m->ctx->stream));
  return BAMCL_OK;
}

static int prepare_cache(bamcl_map2d* m, double max_distance)
{
  const size_t n = (size_t)m->size_x * m->size_y;
  m->max_distance = max_distance;
  m->cell_radius = (int)std::floor(max_distance / m->resolution);
  if (m->cell_radius > 250)
    return set_error(BAMCL_E_UNSUPPORTED, "max_distance/resolution = %d cells exceeds the u16 field (250 cells)",
                     m->cell_radius);
  // The field stores the RANK of k = di^2 + dj^2 among the values a brushfire can produce (sums of two squares
  // <= radius^2, about 0.28 * radius^2 of them), then one rank for "in the map but beyond the radius". The
  // scoring kernels index their per-scan value tables with it.
  {
    const int r2 = m->cell_radius * m->cell_radius;
    std::vector<uint16_t> rank((size_t)r2 + 2, (uint16_t)0xFFFF);
    std::vector<char> rep((size_t)r2 + 1, 0);
    for (int a = 0; a <= m->cell_radius; a++)
      for (int b = a; a * a + b * b <= r2; b++)
        rep[a * a + b * b] = 1;
    m->rep_k.clear();
    for (int k = 0; k <= r2; k++)
      if (rep[k])
      {
        rank[k] = (uint16_t)m->rep_k.size();
        m->rep_k.push_back(k);
      }
    rank[r2 + 1] = (uint16_t)m->rep_k.size();
    BAMCL_TRY(m->rank_of_k.reserve(sizeof(uint16_t) * rank.size()));
    BAMCL_CUDA_OK(cudaMemcpy(m->rank_of_k.p, rank.data(), sizeof(uint16_t) * rank.size(), cudaMemcpyHostToDevice));
  }
  BAMCL_TRY(m->dist.reserve(sizeof(float) * n));
  BAMCL_TRY(m->kmap.reserve(sizeof(uint16_t) * (size_t)m->tiles_x * m->tiles_y * 64));
  // cells of partial edge tiles that lie outside the map are never read; keep them defined anyway
  BAMCL_CUDA_OK(cudaMemsetAsync(m->kmap.p, 0xFF, sizeof(uint16_t) * (size_t)m->tiles_x * m->tiles_y * 64, m->ctx->stream));
  return BAMCL_OK;
}

int bamcl_map2d_build_distances(bamcl_map2d* m, double max_distance)
{
  if (!m)
    return set_error(BAMCL_E_ARG, "null map");
  if (!m->has_cells)
    return set_error(BAMCL_E_STATE, "bamcl_map2d_build_distances: set_cells first");
  m->max_distance = max_distance;
  if (max_distance == 0.0)  // occupancy_map.cpp:141-145: nothing is built
    return BAMCL_OK;
  BAMCL_TRY(prepare_cache(m, max_distance));
  bamcl_ctx* ctx = m->ctx;
  const size_t n = (size_t)m->size_x * m->size_y;
  DevBuf g;
  BAMCL_TRY(g.reserve(sizeof(uint16_t) * n));
  BAMCL_LAUNCH(ctx, k_edt_columns, (m->size_x + 127) / 128, 128, 0, m->cells.as<int8_t>(), g.as<uint16_t>(), m->size_x,
               m->size_y, m->cell_radius);
  constexpr int BLOCK = 256;
  dim3 grid((m->size_x + BLOCK - 1) / BLOCK, m->size_y);
  const size_t smem = sizeof(uint16_t) * (size_t)(BLOCK + 2 * m->cell_radius);
  BAMCL_LAUNCH(ctx, k_edt_rows<BLOCK>, grid, BLOCK, smem, g.as<uint16_t>(), m->dist.as<float>(), m->kmap.as<uint16_t>(),
               m->rank_of_k.as<uint16_t>(), m->size_x, m->size_y, m->tiles_x, m->layout, m->cell_radius, m->resolution, (float)max_distance);
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  g.release();
  m->has_dist = true;
  m->version++;
  m->free_count = -1;  // the free-space list belongs to the previous cache
  return BAMCL_OK;
}

int bamcl_map2d_upload_distances(bamcl_map2d* m, const float* distances, long n, double max_distance)
{
  if (!m || !distances || n != (long)m->size_x * m->size_y || !(max_distance > 0.0))
    return set_error(BAMCL_E_ARG, "bamcl_map2d_upload_distances: bad arguments");
  BAMCL_TRY(prepare_cache(m, max_distance));
  bamcl_ctx* ctx = m->ctx;
  BAMCL_CUDA_OK(cudaMemcpyAsync(m->dist.p, distances, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
  DevBuf bad;
  BAMCL_TRY(bad.reserve(sizeof(int)));
  BAMCL_CUDA_OK(cudaMemsetAsync(bad.p, 0, sizeof(int), ctx->stream));
  dim3 grid((m->size_x + 255) / 256, m->size_y);
  BAMCL_LAUNCH(ctx, k_index_uploaded, grid, 256, 0, m->dist.as<float>(), m->kmap.as<uint16_t>(), m->rank_of_k.as<uint16_t>(),
               m->size_x, m->size_y, m->tiles_x, m->layout, m->cell_radius, m->resolution, (float)max_distance, bad.as<int>());
  int h_bad = 0;
  BAMCL_CUDA_OK(cudaMemcpyAsync(&h_bad, bad.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  bad.release();
  if (h_bad != 0)
  {
    m->has_dist = false;
    return set_error(BAMCL_E_UNSUPPORTED,
                     "%d cached distances are not of the form float(sqrt(k)*resolution) or float(max_distance)", h_bad);
  }
  m->has_dist = true;
  m->version++;
  m->free_count = -1;  // the free-space list belongs to the previous cache
  return BAMCL_OK;
}

int bamcl_map2d_download_distances(bamcl_map2d* m, float* out, long n)
{
  if (!m || !out || n != (long)m->size_x * m->size_y)
    return set_error(BAMCL_E_ARG, "bamcl_map2d_download_distances: bad arguments");
  if (!m->has_dist)
    return set_error(BAMCL_E_STATE, "no distance cache has been built or uploaded");
  BAMCL_CUDA_OK(cudaMemcpyAsync(out, m->dist.p, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, m->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(m->ctx->stream));
  return BAMCL_OK;
}

int bamcl_map2d_calc_range(bamcl_map2d* m, const double* ox, const double* oy, const double* oa, long n,
                           double max_range, double* ranges_out, long* steps_total_or_null)
{
  if (!m || !ox || !oy || !oa || !ranges_out || n < 0)
    return set_error(BAMCL_E_ARG, "bamcl_map2d_calc_range: bad arguments");
  if (!m->has_cells)
    return set_error(BAMCL_E_STATE, "bamcl_map2d_calc_range: set_cells first");
  if (n == 0)
    return BAMCL_OK;
  bamcl_ctx* ctx = m->ctx;
  DevBuf buf;
  BAMCL_TRY(buf.reserve(sizeof(double) * 4 * (size_t)n + sizeof(unsigned long long)));
  double* d = buf.as<double>();
  unsigned long long* d_steps = reinterpret_cast<unsigned long long*>(d + 4 * n);
  BAMCL_CUDA_OK(cudaMemcpyAsync(d, ox, sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
  BAMCL_CUDA_OK(cudaMemcpyAsync(d + n, oy, sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
  BAMCL_CUDA_OK(cudaMemcpyAsync(d + 2 * n, oa, sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
  BAMCL_CUDA_OK(cudaMemsetAsync(d_steps, 0, sizeof(unsigned long long), ctx->stream));
  BAMCL_LAUNCH(ctx, k_calc_range, (unsigned)((n + 127) / 128), 128, 0, m->cells.as<int8_t>(), m->size_x, m->size_y,
               m->resolution, m->origin_x, m->origin_y, d, d + n, d + 2 * n, n, max_range, d + 3 * n, d_steps);
  unsigned long long h_steps = 0;
  BAMCL_CUDA_OK(cudaMemcpyAsync(ranges_out, d + 3 * n, sizeof(double) * n, cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaMemcpyAsync(&h_steps, d_steps, sizeof(h_steps), cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  buf.release();
  if (steps_total_or_null)
    *steps_total_or_null = (long)h_steps;
  return BAMCL_OK;
}

}  // extern "C"

// ---- Node2D::convertMap (src/amcl/node/node_2d.cpp:265-295): OccupancyGrid message -> cell states, with the integer
// up-scaling done on the device so that only width x height bytes cross PCIe
namespace
{
__global__ void __launch_bounds__(256) k_convert_grid(const signed char* __restrict__ data, int width, int scale, int size_x,
                                                      int size_y, signed char* __restrict__ cells)
{
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long)size_x * size_y)
    return;
  const int y = (int)(t / size_x), x = (int)(t % size_x);
  const signed char v = data[(size_t)(y / scale) * width + x / scale];
  cells[t] = v == 0 ? -1 : (v == 100 ? 1 : 0);  // CELL_FREE / CELL_OCCUPIED / CELL_UNKNOWN (map.h)
}
}  // namespace

extern "C" int bamcl_map2d_from_occupancy_grid(bamcl_ctx* ctx, const int8_t* data, int width, int height, float resolution,
                                               double origin_x, double origin_y, int scale_up, bamcl_map2d** out)
{
  if (!ctx || !data || !out || width <= 0 || height <= 0 || scale_up < 1 || !(resolution > 0.0f))
    return set_error(BAMCL_E_ARG, "bamcl_map2d_from_occupancy_grid: bad arguments");
  if ((long)width * scale_up > 0x7fffffffl || (long)height * scale_up > 0x7fffffffl)
    return set_error(BAMCL_E_ARG, "bamcl_map2d_from_occupancy_grid: scaled size overflows int");
  const double res = resolution / scale_up;  // :267, float / int evaluated in float, then widened
  const int size_x = width * scale_up, size_y = height * scale_up;
  const double x_origin = origin_x + (size_x / 2) * res;  // :275-276, narrowed to float by pcl::PointXYZ (:277)
  const double y_origin = origin_y + (size_y / 2) * res;
  bamcl_map2d* m = nullptr;
  BAMCL_TRY(bamcl_map2d_create(ctx, size_x, size_y, res, (float)x_origin, (float)y_origin, &m));
  auto body = [&]() -> int {
    const size_t n = (size_t)size_x * size_y, n_msg = (size_t)width * height;
    bamcl::DevBuf msg;
    BAMCL_TRY(msg.reserve(n_msg));
    int rc = m->cells.reserve(n);
    if (rc == BAMCL_OK && cudaMemcpyAsync(msg.p, data, n_msg, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess)
      rc = set_error(BAMCL_E_CUDA, "bamcl_map2d_from_occupancy_grid: upload failed");
    if (rc == BAMCL_OK)
    {
      k_convert_grid<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(msg.as<signed char>(), width, scale_up, size_x,
                                                                           size_y, m->cells.as<signed char>());
      ctx->launches++;
      m->has_cells = true;
      rc = build_clearance(m);
    }
    if (rc == BAMCL_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess)
      rc = set_error(BAMCL_E_CUDA, "bamcl_map2d_from_occupancy_grid: %s", cudaGetErrorString(cudaGetLastError()));
    msg.release();
    return rc;
  };
  const int rc = body();
  if (rc != BAMCL_OK)
  {
    bamcl_map2d_destroy(m);
    return rc;
  }
  *out = m;
  return BAMCL_OK;
}

extern "C" int bamcl_map2d_download_cells(bamcl_map2d* m, int8_t* cells, long n)
{
  if (!m || !cells || !m->has_cells || n != (long)m->size_x * m->size_y)
    return set_error(BAMCL_E_ARG, "bamcl_map2d_download_cells: need size_x*size_y cells of a map that has them");
  BAMCL_CUDA_OK(cudaMemcpyAsync(cells, m->cells.p, (size_t)n, cudaMemcpyDeviceToHost, m->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(m->ctx->stream));
  return BAMCL_OK;
}

extern "C" int bamcl_map2d_geometry(bamcl_map2d* m, int* size_xy, double* resolution, float* origin_xy)
{
  if (!m || !size_xy || !resolution || !origin_xy)
    return set_error(BAMCL_E_ARG, "bamcl_map2d_geometry: null argument");
  size_xy[0] = m->size_x;
  size_xy[1] = m->size_y;
  *resolution = m->resolution;
  origin_xy[0] = m->origin_x;
  origin_xy[1] = m->origin_y;
  return BAMCL_OK;
}
