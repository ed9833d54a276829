// Global localisation on the device: the free-space index list of Node2D::updateFreeSpaceIndices
// (src/amcl/node/node_2d.cpp:320-338) and ParticleFilter::initWithPoseFn (src/amcl/pf/particle_filter.cpp:136-162)
// fed by Node::uniformPoseGenerator / randomFreeSpacePose (src/amcl/node/node.cpp:823-868).
//
// The list is a stream compaction over the grid in the reference's order (i outer, j inner -- the transpose of the
// cell storage). Every randomFreeSpacePose call takes exactly two drand48 values, and Node2D::scorePose always returns
// 1.0 (it scores a copy of the fake sample, node_2d.cpp:310-313), so the de-weighting retry loop of
// uniformPoseGenerator runs a fixed number r of extra generations per pose: pose s is generation (1 + r) s + r of the
// stream and all poses are independent given the generator's skip-ahead (lcg.cuh).
#include <cmath>
#include <cstring>

#include "lcg.cuh"
#include "map3d.cuh"
#include "pf.cuh"
#include "scan.cuh"

using namespace bamcl;

namespace bamcl
{
int build_tree_and_stats(bamcl_pf* pf, int set_index, double dist_threshold, double convergence_threshold,
                         bamcl_resample_result* result);  // kdtree.cu
}

namespace
{
// t = i * size_y + j (the reference's loop order); cell storage is i + j * size_x
__global__ void __launch_bounds__(256) k_free_flags(const signed char* __restrict__ cells, const float* __restrict__ dist,
                                                    int size_x, int size_y, double radius, int* __restrict__ flag)
{
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long)size_x * size_y)
    return;
  const int i = (int)(t / size_y), j = (int)(t % size_y);
  const size_t c = (size_t)i + (size_t)j * size_x;
  flag[t] = (cells[c] == -1 && (double)dist[c] > radius) ? 1 : 0;  // CELL_FREE and getDistanceToObject > radius
}

__global__ void __launch_bounds__(256) k_free_scatter(const signed char* __restrict__ cells, const float* __restrict__ dist,
                                                      int size_x, int size_y, double radius, const int* __restrict__ pos,
                                                      int2* __restrict__ out)
{
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long)size_x * size_y)
    return;
  const int i = (int)(t / size_y), j = (int)(t % size_y);
  const size_t c = (size_t)i + (size_t)j * size_x;
  if (cells[c] == -1 && (double)dist[c] > radius)
    out[pos[t]] = make_int2(i, j);
}

struct UniformArgs
{
  unsigned long long state0;
  int n, per_pose;  // per_pose = 1 + r generations per returned pose; the last one is kept
  PoseSource src;
  double *x, *y, *th, *w;
  double weight;
};

__global__ void __launch_bounds__(256) k_uniform_poses(const UniformArgs a)
{
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= a.n)
    return;
  const unsigned long long gen = (unsigned long long)a.per_pose * (unsigned long long)s + (unsigned long long)(a.per_pose - 1);
  const unsigned long long u = lcg_skip(a.state0, 2ull * gen + 1ull), v = lcg_next(u);
  const double r1 = (double)u * 0x1p-48, r2 = (double)v * 0x1p-48;
  const unsigned int k = (unsigned int)(r1 * (double)(size_t)a.src.free_n);  // node.cpp:831
  a.src.world(k, &a.x[s], &a.y[s]);                                            // convertMapToWorld
  a.th[s] = r2 * 2 * M_PI - M_PI;  // node.cpp:837
  a.w[s] = a.weight;
}
// the same generator writing packed (x, y, theta) triples: the random poses of the systematic resampler
__global__ void __launch_bounds__(256) k_uniform_poses_aos(const UniformArgs a, double* __restrict__ out3)
{
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= a.n)
    return;
  const unsigned long long gen = (unsigned long long)a.per_pose * (unsigned long long)s + (unsigned long long)(a.per_pose - 1);
  const unsigned long long u = lcg_skip(a.state0, 2ull * gen + 1ull), v = lcg_next(u);
  const unsigned int k = (unsigned int)((double)u * 0x1p-48 * (double)(size_t)a.src.free_n);
  a.src.world(k, &out3[3 * s], &out3[3 * s + 1]);
  out3[3 * s + 2] = (double)v * 0x1p-48 * 2 * M_PI - M_PI;
}
}  // namespace

namespace bamcl
{
// n calls of the filter's registered pose function (bamcl_pf_set_random_pose_source) from generator state `state0`,
// written to out3_dev; *state_out = the state after them. Used by the systematic resampler (particle_filter.cpp:316-324).
static void source_of_map2d(const bamcl_map2d* map, PoseSource* s)
{
  *s = PoseSource();
  s->free_ij = map->free_ij.as<int2>();
  s->free_n = map->free_count;
  s->size_x = map->size_x;
  s->size_y = map->size_y;
  s->resolution = map->resolution;
  s->origin_x = (double)map->origin_x;
  s->origin_y = (double)map->origin_y;
}
// Node3D::updateFreeSpaceIndices (node_3d.cpp:306-318): for i in [min0, max0) for j in [min1, max1)
static bool source_of_map3d(const bamcl_map3d* map, PoseSource* s)
{
  *s = PoseSource();
  const long nx = (long)map->cmax[0] - map->cmin[0], ny = (long)map->cmax[1] - map->cmin[1];
  if (nx <= 0 || ny <= 0 || nx * ny > 0x7fffffffl)
    return false;
  s->free_n = nx * ny;
  s->resolution = map->resolution;
  s->min_i = map->cmin[0];
  s->min_j = map->cmin[1];
  s->ny = (int)ny;
  return true;
}
bool pf_pose_source(const bamcl_pf* pf, PoseSource* out)
{
  if (pf->rp_map && pf->rp_map->free_count > 0)
  {
    source_of_map2d(pf->rp_map, out);
    return true;
  }
  if (pf->rp_map3d)
    return source_of_map3d(pf->rp_map3d, out);
  return false;
}

int device_random_poses(bamcl_pf* pf, unsigned long long state0, int n, double* out3_dev, unsigned long long* state_out)
{
  long r = 0;
  if (pf->rp_threshold > 0.0 && pf->rp_multiplier < 1.0 && pf->rp_multiplier >= 0.0)
  {
    double good = pf->rp_threshold;
    while (1.0 < good)
    {
      good *= pf->rp_multiplier;
      if (++r > 1000000)
        return set_error(BAMCL_E_UNSUPPORTED, "resample: more than 1e6 de-weighting retries per random pose");
    }
  }
  UniformArgs a{};
  a.state0 = state0 & LCG_MASK;
  a.n = n;
  a.per_pose = (int)(1 + r);
  if (!pf_pose_source(pf, &a.src))
    return set_error(BAMCL_E_STATE, "resample: no pose source registered");
  if (n > 0)
    BAMCL_LAUNCH(pf->ctx, k_uniform_poses_aos, (unsigned)((n + 255) / 256), 256, 0, a, out3_dev);
  *state_out = lcg_skip(a.state0, 2ull * (unsigned long long)(1 + r) * (unsigned long long)n);
  return BAMCL_OK;
}
}  // namespace bamcl

extern "C" int bamcl_map2d_update_free_space(bamcl_map2d* m, double non_free_space_radius, long* count_out)
{
  if (!m || !count_out)
    return set_error(BAMCL_E_ARG, "bamcl_map2d_update_free_space: null argument");
  if (!m->has_cells || !m->has_dist)
    return set_error(BAMCL_E_STATE, "bamcl_map2d_update_free_space: needs the cells and the distance cache "
                                    "(node_2d.cpp:322-323)");
  bamcl_ctx* ctx = m->ctx;
  const long n = (long)m->size_x * m->size_y;
  DevBuf flags, scratch;
  int rc = flags.reserve(sizeof(int) * ((size_t)n + 2));
  if (rc == BAMCL_OK)
    rc = scratch.reserve(sizeof(int) * (scan_scratch_ints(n) + 8));
  if (rc != BAMCL_OK)
  {
    flags.release();
    scratch.release();
    return rc;
  }
  int* total = scratch.as<int>();
  const unsigned blocks = (unsigned)((n + 255) / 256);
  auto body = [&]() -> int {
    BAMCL_LAUNCH(ctx, k_free_flags, blocks, 256, 0, m->cells.as<signed char>(), m->dist.as<float>(), m->size_x, m->size_y,
                 non_free_space_radius, flags.as<int>());
    BAMCL_TRY(exclusive_scan_i32(ctx, flags.as<int>(), flags.as<int>(), n, total + 4, total));
    int h_total = 0;
    BAMCL_CUDA_OK(cudaMemcpyAsync(&h_total, total, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    BAMCL_TRY(m->free_ij.reserve(sizeof(int2) * (size_t)std::max(h_total, 1)));
    BAMCL_LAUNCH(ctx, k_free_scatter, blocks, 256, 0, m->cells.as<signed char>(), m->dist.as<float>(), m->size_x, m->size_y,
                 non_free_space_radius, flags.as<int>(), m->free_ij.as<int2>());
    BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    m->free_count = h_total;
    return BAMCL_OK;
  };
  rc = body();
  flags.release();
  scratch.release();
  if (rc != BAMCL_OK)
    return rc;
  *count_out = m->free_count;
  return BAMCL_OK;
}

extern "C" int bamcl_map2d_download_free_space(bamcl_map2d* m, int* ij_out, long n)
{
  if (!m || !ij_out || m->free_count < 0 || n != m->free_count)
    return set_error(BAMCL_E_ARG, "bamcl_map2d_download_free_space: need free_count pairs (update_free_space first)");
  if (n == 0)
    return BAMCL_OK;
  BAMCL_CUDA_OK(cudaMemcpyAsync(ij_out, m->free_ij.p, sizeof(int2) * (size_t)n, cudaMemcpyDeviceToHost, m->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(m->ctx->stream));
  return BAMCL_OK;
}

namespace
{
int init_uniform_from(bamcl_pf* pf, const PoseSource& src, double starting_weight_threshold, double deweight_multiplier,
                      uint64_t drand48_state, uint64_t* drand48_state_out, bamcl_resample_result* res)
{
  // node.cpp:849-866 with scorePose == 1.0: generations thrown away per returned pose
  long r = 0;
  if (starting_weight_threshold > 0.0 && deweight_multiplier < 1.0 && deweight_multiplier >= 0.0)
  {
    double good = starting_weight_threshold;
    while (1.0 < good)
    {
      good *= deweight_multiplier;
      if (++r > 1000000)
        return set_error(BAMCL_E_UNSUPPORTED, "bamcl_pf_init_uniform: more than 1e6 de-weighting retries per pose");
    }
  }
  bamcl_ctx* ctx = pf->ctx;
  pf->win_cnt = -1;
  SampleSet& set = pf->sets[pf->current];
  const int n = pf->max_samples;  // :147
  UniformArgs a{};
  a.state0 = drand48_state & LCG_MASK;
  a.n = n;
  a.per_pose = (int)(1 + r);
  a.src = src;
  a.x = set.x.as<double>();
  a.y = set.y.as<double>();
  a.th = set.th.as<double>();
  a.w = set.w.as<double>();
  a.weight = 1.0 / (double)pf->max_samples;  // :152
  BAMCL_LAUNCH(ctx, k_uniform_poses, (unsigned)((n + 255) / 256), 256, 0, a);
  set.sample_count = n;
  pf_apply_shard_window(pf);  // a distributed filter: every rank generated the whole set, and owns its window from here on
  memset(res, 0, sizeof(*res));
  BAMCL_TRY(build_tree_and_stats(pf, pf->current, 0.5, 1e300, res));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  res->sample_count = n;
  res->converged = 0;  // initConverged, :163-167
  *drand48_state_out = lcg_skip(a.state0, 2ull * (unsigned long long)(1 + r) * (unsigned long long)n);
  res->drand48_state = *drand48_state_out;
  return BAMCL_OK;
}
}  // namespace

extern "C" int bamcl_pf_init_uniform(bamcl_pf* pf, bamcl_map2d* map, double starting_weight_threshold,
                                     double deweight_multiplier, uint64_t drand48_state, uint64_t* drand48_state_out,
                                     bamcl_resample_result* res)
{
  if (!pf || !map || !drand48_state_out || !res)
    return set_error(BAMCL_E_ARG, "bamcl_pf_init_uniform: null argument");
  if (map->free_count <= 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_init_uniform: the free-space list is empty or not built (node.cpp:826-830 "
                                    "warns and returns an unset pose)");
  PoseSource src;
  source_of_map2d(map, &src);
  return init_uniform_from(pf, src, starting_weight_threshold, deweight_multiplier, drand48_state, drand48_state_out, res);
}

extern "C" int bamcl_pf_init_uniform_3d(bamcl_pf* pf, bamcl_map3d* map, double starting_weight_threshold,
                                        double deweight_multiplier, uint64_t drand48_state, uint64_t* drand48_state_out,
                                        bamcl_resample_result* res)
{
  if (!pf || !map || !drand48_state_out || !res)
    return set_error(BAMCL_E_ARG, "bamcl_pf_init_uniform_3d: null argument");
  PoseSource src;
  if (!source_of_map3d(map, &src))
    return set_error(BAMCL_E_STATE, "bamcl_pf_init_uniform_3d: the map's cropped cell rectangle is empty (node.cpp:826-830 "
                                    "warns and returns an unset pose)");
  return init_uniform_from(pf, src, starting_weight_threshold, deweight_multiplier, drand48_state, drand48_state_out, res);
}

extern "C" int bamcl_pf_set_random_pose_source_3d(bamcl_pf* pf, bamcl_map3d* map_or_null, double starting_weight_threshold,
                                                  double deweight_multiplier)
{
  if (!pf)
    return set_error(BAMCL_E_ARG, "bamcl_pf_set_random_pose_source_3d: null filter");
  pf->rp_map = nullptr;
  pf->rp_map3d = map_or_null;
  pf->rp_threshold = starting_weight_threshold;
  pf->rp_multiplier = deweight_multiplier;
  return BAMCL_OK;
}
