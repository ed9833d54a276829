// K-b: beam model = per particle x beam ray march through the occupancy grid (calcRange) + the four-part
// mixture. Reference: src/amcl/sensors/planar_scanner.cpp:168-234, src/amcl/map/occupancy_map.cpp:257-364.
//
// A warp owns one particle at a time; lanes take adjacent beams. Each lane follows the reference's own integer
// Bresenham sequence, so the first cell that stops the ray -- and with it the returned range -- is the reference's. It
// does not visit every cell of that sequence: the map's clearance field (map2d.cu, distance in cells to the nearest
// stopping cell) proves the next cells free, and the walk jumps over them, landing exactly on the sequence through its
// closed form y(t) = y0 + ystep * floor((t * 2dy + dx) / 2dx).
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
  const uint8_t* clear;  // clearance field: floor(distance in cells to the nearest stopping cell), capped
  int size_x, size_y;
  double resolution;
  float origin_x, origin_y;
  double range_max;
  double scan_x, scan_y, cos_sa, sin_sa;
  double z_hit, z_hit_denom;
  double off_map_factor, non_free_space_factor, non_free_space_radius;
};

// OccupancyMap::calcRange (occupancy_map.cpp:257-364) inside the scorer. The reference tests the cells of an integer
// Bresenham walk one by one: start cell, then `x += xstep; err += dy; if (2 err >= dx) { y += ystep; err -= dx; }` until
// x == x1 + xstep; the first cell that is off the grid or not CELL_FREE ends it. After t steps the walk is at
// x = x0 + t xstep, y = y0 + ystep * floor((t * 2dy + dx) / 2dx) (the step rule keeps 2 err in [-dx, dx)). From a cell
// with clearance d (true distance to the nearest stopping cell >= d) every cell within Euclidean distance < d is free;
// the cell s steps ahead lies within s / cos(phi) + 1 of the current one (phi = the ray's angle to its major axis, +1 for
// the rounding of y), so s < (d - 1) cos(phi) steps can be skipped untested.
// State of one lane's walk along its current ray.
struct RayWalk
{
  int x, y, x0, y0, t, dx, dy2, dx2, xstep, ystep;
  int left;        // cells still to test before the walk leaves the grid along x or has tested x1 + xstep
  int oob_at_end;  // running out of `left` means the next cell is off the grid (a hit) rather than "nothing hit"
  float fdir;      // 0.999 cos(phi)
  double inv_dx2;
  int steep;
  int u;  // beam index, -1 = no ray
};
__device__ __forceinline__ int walk_y(const RayWalk& r, int t)  // y after t steps
{
  const int num = t * r.dy2 + r.dx;
  int ny = (int)((double)num * r.inv_dx2);
  ny -= ny * r.dx2 > num ? 1 : 0;
  ny += (ny + 1) * r.dx2 <= num ? 1 : 0;
  return r.y0 + r.ystep * ny;
}

// Warp = one particle at a time; lanes = rays. Adjacent rays can differ 30x in length (one ends on the near wall, the
// next runs through a door), so a lane that finishes its ray scores it and takes the particle's next unassigned beam
// instead of idling until the longest ray of its group of 32 is done. Refills happen when at least REFILL_MIN lanes
// are idle (the set-up and the scoring are ~10x a walk step, so they are batched), or when nothing else is running.
#ifndef BEAM_REFILL_MIN
#define BEAM_REFILL_MIN 8
#endif
constexpr int REFILL_MIN = BEAM_REFILL_MIN;
// (swept on a B200, scripts/sweep_beam2.sh, 100k x 1080: chunk 2 / 4 / 8 / 16 jumps per refill check = 11.0 / 10.0 / 9.3 / 9.1
// ms; refill threshold 4 / 8 / 16 lanes = 9.7 / 9.3 / 9.3; 8 CTAs per SM (32 registers, spills) 8.75 against 9.3 at 4)
#ifndef BEAM_CHUNK
#define BEAM_CHUNK 16
#endif
#ifndef BEAM_MINB
#define BEAM_MINB 8
#endif

__global__ void __launch_bounds__(256, BEAM_MINB) k_score_beam(const BeamArgs a)
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
    const int px0 = world_to_cell(sx, a.origin_x, a.resolution, a.size_x);
    const int py0 = world_to_cell(sy, a.origin_y, a.resolution, a.size_y);
    double acc = 0.0;
    int next = 0;  // next unassigned beam (warp-uniform)
    RayWalk r;
    r.u = -1;
    bool walking = false;  // this lane is inside a walk
    double map_range = 0.0;
    bool have_result = false;  // a finished ray waits to be scored
    for (;;)
    {
      const unsigned busy = __ballot_sync(0xffffffffu, walking);
      const int idle = 32 - __popc(busy);
      if (next >= a.n_used && busy == 0)
      {
        // nothing left to hand out and nobody walks: score what is pending and stop
        if (have_result)
        {
          const double4 b = a.beam[r.u];
          const double z = a.ranges[(size_t)r.u * a.step] - map_range;
          double pz = 0.0;
          pz += a.z_hit * exp(-(z * z) / a.z_hit_denom);  // :205
          if (z < 0)
            pz += b.z;  // :209
          pz += b.w;    // :212-217
          acc += pz * pz * pz;
          have_result = false;
        }
        break;
      }
      if ((next < a.n_used && idle >= REFILL_MIN) || busy == 0)
      {
        if (!walking)
        {
          if (have_result)
          {
            const double4 b = a.beam[r.u];
            const double z = a.ranges[(size_t)r.u * a.step] - map_range;
            double pz = 0.0;
            pz += a.z_hit * exp(-(z * z) / a.z_hit_denom);
            if (z < 0)
              pz += b.z;
            pz += b.w;
            acc += pz * pz * pz;
            have_result = false;
          }
          // idle lanes take consecutive beams
          const unsigned idle_mask = ~busy;
          const int my = next + __popc(idle_mask & ((1u << lane) - 1));
          if (my < a.n_used)
          {
            const double4 b = a.beam[my];
            const double ca = c2 * b.x - s2 * b.y;  // cos(theta + bearing)
            const double sa = s2 * b.x + c2 * b.y;
            int x0 = px0, y0 = py0;
            int x1 = world_to_cell(sx + a.range_max * ca, a.origin_x, a.resolution, a.size_x);
            int y1 = world_to_cell(sy + a.range_max * sa, a.origin_y, a.resolution, a.size_y);
            r.u = my;
            if (x0 == x1 && y0 == y1)  // occupancy_map.cpp:279-280
            {
              map_range = a.range_max;
              have_result = true;
            }
            else
            {
              r.steep = abs(y1 - y0) > abs(x1 - x0);
              if (r.steep)
              {
                int t = x0; x0 = y0; y0 = t;
                t = x1; x1 = y1; y1 = t;
              }
              const int major_size = r.steep ? a.size_y : a.size_x, minor_size = r.steep ? a.size_x : a.size_y;
              r.x0 = x0;
              r.y0 = y0;
              r.dx = abs(x1 - x0);
              r.dy2 = 2 * abs(y1 - y0);
              r.dx2 = 2 * r.dx;
              r.xstep = (x0 < x1) ? 1 : -1;
              r.ystep = (y0 < y1) ? 1 : -1;
              r.t = 0;
              r.x = x0;
              r.y = y0;
              r.inv_dx2 = 1.0 / (double)r.dx2;
              r.fdir = 0.999f * (float)r.dx * rsqrtf((float)r.dx * (float)r.dx + 0.25f * (float)r.dy2 * (float)r.dy2);
              // cells x0 .. x1 + xstep are tested (dx + 2 of them); those with x off the grid count as hits
              const int to_test = r.dx + 2;
              const int in_grid = ((unsigned)x0 < (unsigned)major_size) ? (r.xstep > 0 ? major_size - x0 : x0 + 1) : 0;
              r.left = min(to_test, in_grid);
              r.oob_at_end = in_grid < to_test;
              if ((unsigned)y0 >= (unsigned)minor_size || r.left == 0)
              {
                map_range = 0.0;  // the start cell itself is off the grid: hit at distance 0 (:315-332)
                have_result = true;
              }
              else
                walking = true;
            }
          }
        }
        next += idle;  // warp-uniform: every idle lane was offered a beam (those past the end stay idle)
        if (next > a.n_used)
          next = a.n_used;
        continue;
      }
      // ---- one chunk of jumps for the lanes that have a ray
      if (walking)
      {
        const int minor_size = r.steep ? a.size_x : a.size_y;
        int x = r.x, y = r.y, t = r.t, left = r.left;
        // single-exit loop (no breaks): the lanes of a warp stay converged from one jump to the next
        int code = 0;  // 0 walking, 1 hit at (x, y), 2 nothing hit, 3 the next step leaves the grid along x
#pragma unroll 1
        for (int it = 0; it < BEAM_CHUNK && code == 0; it++)
        {
          const int d = __ldg(&a.clear[r.steep ? (size_t)x * a.size_x + y : (size_t)y * a.size_x + x]);
          if (d == 0)  // the cell stops the ray
            code = 1;
          else if (left == 1)  // that was the last testable cell
            code = r.oob_at_end ? 3 : 2;
          else
          {
            int s = (int)((float)(d - 1) * r.fdir);
            s = min(max(s, 1), left - 1);
            t += s;
            left -= s;
            x += s * r.xstep;
            y = walk_y(r, t);
            if ((unsigned)y >= (unsigned)minor_size)  // left the grid sideways: this cell is a hit
              code = 1;
          }
        }
        if (code == 3)
        {
          // the walk steps once more and lands off the grid: a hit there (:340-358)
          t++;
          x += r.xstep;
          y = walk_y(r, t);
          code = 1;
        }
        const bool done = code != 0, hit = code == 1;
        r.x = x;
        r.y = y;
        r.t = t;
        r.left = left;
        if (done)
        {
          const int ddx = x - r.x0, ddy = y - r.y0;
          map_range = hit ? sqrt((double)(ddx * ddx + ddy * ddy)) * a.resolution : a.range_max;
          walking = false;
          have_result = true;
        }
      }
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
  a.clear = map->clear.as<uint8_t>();
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
