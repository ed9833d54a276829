// K-b: beam model = per particle x beam ray march through the occupancy grid (calcRange) + the four-part
// mixture. Reference: src/amcl/sensors/planar_scanner.cpp:168-234, src/amcl/map/occupancy_map.cpp:257-364.
//
// A warp owns one particle at a time; lanes take adjacent beams (adjacent rays have similar lengths and
// touch neighbouring cells). Each lane walks the reference's own integer Bresenham sequence, so the
// first blocked cell -- and with it the returned range -- is the reference's.
#include <cmath>

#include "pf.cuh"

using namespace bamcl;

namespace
{
struct BeamArgs
{
  const double* x;
  const double* y;
  const double* th;
  double* w;
  int n;
  const double* ranges;   // strided original arrays (step applied in the kernel)
  const double4* beam;    // per used beam: cos b, sin b, short term, tail term
  int n_used;
  int step;
  const int8_t* cells;
  const float* dist;
  int size_x, size_y;
  double resolution;
  float origin_x, origin_y;
  double range_max;
  double scan_x, scan_y, cos_sa, sin_sa;
  double z_hit, z_hit_denom;
  double off_map_factor, non_free_space_factor, non_free_space_radius;
};

__device__ __forceinline__ double march(const BeamArgs& a, int x0, int y0, int x1, int y1)
{
  if (x0 == x1 && y0 == y1)
    return a.range_max;
  const bool steep = abs(y1 - y0) > abs(x1 - x0);
  if (steep)
  {
    int t = x0; x0 = y0; y0 = t;
    t = x1; x1 = y1; y1 = t;
  }
  const int deltax = abs(x1 - x0), deltay = abs(y1 - y0);
  const int xstep = (x0 < x1) ? 1 : -1, ystep = (y0 < y1) ? 1 : -1;
  const int xend = x1 + xstep;
  int error = 0, x = x0, y = y0;
  for (;;)
  {
    const int cx = steep ? y : x, cy = steep ? x : y;
    const bool valid = ((unsigned)cx < (unsigned)a.size_x) && ((unsigned)cy < (unsigned)a.size_y);
    if (!valid || a.cells[(size_t)cy * a.size_x + cx] != -1)
      return sqrt((double)((x - x0) * (x - x0) + (y - y0) * (y - y0))) * a.resolution;
    if (x == xend)
      return a.range_max;
    x += xstep;
    error += deltay;
    if (2 * error >= deltax)
    {
      y += ystep;
      error -= deltax;
    }
  }
}

__global__ void __launch_bounds__(256) k_score_beam(const BeamArgs a)
{
  const int lane = threadIdx.x & 31;
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int warps_total = (gridDim.x * blockDim.x) >> 5;
  for (int j = warp_global; j < a.n; j += warps_total)
  {
    const double px = a.x[j], py = a.y[j], pth = a.th[j];
    double s, c;
    sincos(pth, &s, &c);
    const double sx = px + a.scan_x * c - a.scan_y * s;  // coordAdd :693-701
    const double sy = py + a.scan_x * s + a.scan_y * c;
    const double c2 = c * a.cos_sa - s * a.sin_sa;
    const double s2 = s * a.cos_sa + c * a.sin_sa;
    const int x0 = world_to_cell(sx, a.origin_x, a.resolution, a.size_x);
    const int y0 = world_to_cell(sy, a.origin_y, a.resolution, a.size_y);
    double acc = 0.0;
    for (int u = lane; u < a.n_used; u += 32)
    {
      const double4 b = a.beam[u];
      const double obs_range = a.ranges[(size_t)u * a.step];
      const double ca = c2 * b.x - s2 * b.y;  // cos(theta + bearing)
      const double sa = s2 * b.x + c2 * b.y;
      const int x1 = world_to_cell(sx + a.range_max * ca, a.origin_x, a.resolution, a.size_x);
      const int y1 = world_to_cell(sy + a.range_max * sa, a.origin_y, a.resolution, a.size_y);
      const double map_range = march(a, x0, y0, x1, y1);
      const double z = obs_range - map_range;
      double pz = 0.0;
      pz += a.z_hit * exp(-(z * z) / a.z_hit_denom);  // :205
      if (z < 0)
        pz += b.z;  // :209
      pz += b.w;    // :212-217 (exactly one of the two applies unless the range exceeds range_max)
      acc += pz * pz * pz;
    }
    acc = warp_sum(acc);
    if (lane == 0)
    {
      double w = a.w[j] * (1.0 + acc);
      // recalcWeight :642-682
      const int ci = world_to_cell(px, a.origin_x, a.resolution, a.size_x);
      const int cj = world_to_cell(py, a.origin_y, a.resolution, a.size_y);
      double factor = 1.0;
      if (!(ci >= 0 && ci < a.size_x && cj >= 0 && cj < a.size_y))
        factor = a.off_map_factor;
      else if (a.cells[(size_t)cj * a.size_x + ci] != -1)
        factor = a.non_free_space_factor;
      else
      {
        const double distance = (double)a.dist[(size_t)cj * a.size_x + ci];
        if (distance < a.non_free_space_radius)
          factor = a.non_free_space_factor + (distance / a.non_free_space_radius) * (1.0 - a.non_free_space_factor);
      }
      a.w[j] = w * factor;
    }
  }
}
}  // namespace

namespace bamcl
{
int launch_beam_model(bamcl_pf* pf, bamcl_map2d* map)
{
  StagedScan& sc = pf->scan;
  SampleSet& set = pf->sets[pf->current];
  bamcl_ctx* ctx = pf->ctx;
  const bamcl_planar_model& m = sc.model;
  // per used beam constants, computed with the host libm in the reference's expression order
  std::vector<double> h_ranges(sc.range_count), h_angles(sc.range_count);
  BAMCL_CUDA_OK(cudaMemcpyAsync(h_ranges.data(), sc.ranges.p, sizeof(double) * sc.range_count, cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaMemcpyAsync(h_angles.data(), sc.angles.p, sizeof(double) * sc.range_count, cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  int n_used = 0;
  for (int i = 0; i < sc.range_count; i += sc.step)
    n_used++;
  BAMCL_TRY(pf->h_stage.reserve(sizeof(double) * 4 * (size_t)(n_used + 1)));
  double* hb = pf->h_stage.as<double>();
  for (int u = 0; u < n_used; u++)
  {
    const double obs_range = h_ranges[(size_t)u * sc.step], obs_bearing = h_angles[(size_t)u * sc.step];
    hb[4 * u] = std::cos(obs_bearing);
    hb[4 * u + 1] = std::sin(obs_bearing);
    hb[4 * u + 2] = m.z_short * m.lambda_short * std::exp(-m.lambda_short * obs_range);
    double tail = 0.0;
    if (obs_range == sc.range_max)
      tail = m.z_max * 1.0;
    if (obs_range < sc.range_max)
      tail = m.z_rand * 1.0 / sc.range_max;
    hb[4 * u + 3] = tail;
  }
  BAMCL_TRY(sc.beams.reserve(sizeof(double) * 4 * (size_t)(n_used + 1)));
  BAMCL_CUDA_OK(cudaMemcpyAsync(sc.beams.p, hb, sizeof(double) * 4 * n_used, cudaMemcpyHostToDevice, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  BeamArgs a;
  a.x = set.x.as<double>() + pf_win_off(pf);
  a.y = set.y.as<double>() + pf_win_off(pf);
  a.th = set.th.as<double>() + pf_win_off(pf);
  a.w = set.w.as<double>() + pf_win_off(pf);
  a.n = pf_win_cnt(pf);
  a.ranges = sc.ranges.as<double>();
  a.beam = sc.beams.as<double4>();
  a.n_used = n_used;
  a.step = sc.step;
  a.cells = map->cells.as<int8_t>();
  a.dist = map->dist.as<float>();
  a.size_x = map->size_x;
  a.size_y = map->size_y;
  a.resolution = map->resolution;
  a.origin_x = map->origin_x;
  a.origin_y = map->origin_y;
  a.range_max = sc.range_max;
  a.scan_x = m.scanner_pose[0];
  a.scan_y = m.scanner_pose[1];
  a.cos_sa = std::cos(m.scanner_pose[2]);
  a.sin_sa = std::sin(m.scanner_pose[2]);
  a.z_hit = m.z_hit;
  a.z_hit_denom = 2 * m.sigma_hit * m.sigma_hit;
  a.off_map_factor = m.off_map_factor;
  a.non_free_space_factor = m.non_free_space_factor;
  a.non_free_space_radius = m.non_free_space_radius;
  int blocks = (a.n + 7) / 8;
  if (blocks > kNumSMs * 8)
    blocks = kNumSMs * 8;
  BAMCL_CUDA_OK(cudaEventRecord(pf->score_ev0, ctx->stream));
  BAMCL_LAUNCH(ctx, k_score_beam, blocks, 256, 0, a);
  BAMCL_CUDA_OK(cudaEventRecord(pf->score_ev1, ctx->stream));
  pf->score_timed = true;
  return BAMCL_OK;
}
}  // namespace bamcl
