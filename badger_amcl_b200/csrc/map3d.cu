// K-e3 + OctoMap device object. Reference: src/amcl/map/octomap.cpp:53-74 (initFromOctree), 98-123
// (conversions / validity), 128-150 (setMapBounds), 174-333 (updateDistancesLUT + brushfire), 336-350
// (getDistanceToObject).
//
// The reference fills its cache with a FIFO brushfire that relaxes a voxel only when the new distance is better
// by more than one quantum (octomap.cpp:290-312); the result depends on queue order and differs from the exact
// transform in a few voxels per million. The device builds the EXACT quantised transform with
// the reference's value formula -- floor(min(sqrt(k2) * res, max) / max * 255) over the sources inside the
// (radius + 1) box -- by three windowed separable passes; tests report the mismatch against the reference and
// the scorer parity tests upload the reference's own cache instead.
#include <cmath>
#include <cstring>

#include "map3d.cuh"

using namespace bamcl;

namespace
{
constexpr int NONE8 = 255;
constexpr int NONE16 = 65535;

__global__ void k3_scatter_sources(const int* __restrict__ src, long n_src, uint8_t* __restrict__ occ, int i0, int j0, int k0,
                                   int W, int H, int D)
{
  const long s = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_src)
    return;
  const int i = src[3 * s] - i0, j = src[3 * s + 1] - j0, k = src[3 * s + 2] - k0;
  if ((unsigned)i < (unsigned)W && (unsigned)j < (unsigned)H && (unsigned)k < (unsigned)D)  // isVoxelValid, :118-123
    occ[((size_t)k * H + j) * W + i] = 1;
}

// pass 1: nearest source along x within the window
__global__ void k3_pass_x(const uint8_t* __restrict__ occ, uint8_t* __restrict__ g1, int W, size_t n, int R)
{
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n)
    return;
  const int i = (int)(t % W);
  const uint8_t* row = occ + (t - i);
  int best = NONE8;
  for (int d = 0; d <= R; d++)
  {
    if ((i - d >= 0 && row[i - d]) || (i + d < W && row[i + d]))
    {
      best = d;
      break;
    }
  }
  g1[t] = (uint8_t)best;
}

// pass 2: min over dy of dy^2 + g1^2
__global__ void k3_pass_y(const uint8_t* __restrict__ g1, uint16_t* __restrict__ g2, int W, int H, size_t n, int R)
{
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n)
    return;
  const int j = (int)((t / W) % H);
  int best = NONE16;
  for (int d = -R; d <= R; d++)
  {
    const int jj = j + d;
    if (jj < 0 || jj >= H)
      continue;
    const int g = g1[t + (long)d * W];
    if (g != NONE8)
      best = min(best, d * d + g * g);
  }
  g2[t] = (uint16_t)best;
}

// pass 3: min over dz, then the reference's quantisation (octomap.cpp:329-331)
__global__ void k3_pass_z(const uint16_t* __restrict__ g2, uint8_t* __restrict__ field, int W, int H, int D, size_t n, int R,
                          double resolution, double max_distance)
{
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n)
    return;
  const size_t plane = (size_t)W * H;
  const int k = (int)(t / plane);
  int best = -1;
  for (int d = -R; d <= R; d++)
  {
    const int kk = k + d;
    if (kk < 0 || kk >= D)
      continue;
    const int g = g2[t + (long)d * (long)plane];
    if (g != NONE16)
    {
      const int v = d * d + g;
      best = best < 0 ? v : min(best, v);
    }
  }
  int q = 255;
  if (best >= 0)
  {
    double dd = sqrt((double)best) * resolution;
    dd = dd < max_distance ? dd : max_distance;
    dd = dd / max_distance * 255;
    q = (int)floor(dd);
  }
  field[t] = (uint8_t)q;
}
}  // namespace

static void recount(bamcl_map3d* m)
{
  m->W = m->cmax[0] - m->cmin[0] + 1;
  m->H = m->cmax[1] - m->cmin[1] + 1;
  m->D = m->cmax[2] - m->cmin[2] + 1;
}

static inline int world_to_voxel(double v, double res) { return (int)std::floor(v / res + 0.5); }  // octomap.cpp:98-109

namespace
{
// one thread per brick of 4 x 4 x 2 voxels: its 32 distance bytes (255 beyond the bounds) and its "near" bits
__global__ void __launch_bounds__(256) k3_build_bricks(const uint8_t* __restrict__ field, int W, int H, int D, int CW, int CH,
                                                       long n_bricks, uint8_t* __restrict__ bricks, unsigned* __restrict__ bits,
                                                       unsigned long long* __restrict__ near_count)
{
  const long b = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_bricks)
    return;
  const int bx = (int)(b % CW), by = (int)((b / CW) % CH), bz = (int)(b / ((long)CW * CH));
  unsigned word = 0;
  unsigned packed[8];
#pragma unroll
  for (int q = 0; q < 8; q++)
    packed[q] = 0;
#pragma unroll
  for (int pos = 0; pos < 32; pos++)
  {
    const int i = bx * 4 + (pos & 3), j = by * 4 + ((pos >> 2) & 3), k = bz * 2 + (pos >> 4);
    unsigned v = 255;
    if (i < W && j < H && k < D)
      v = field[((size_t)k * H + j) * W + i];
    if (v < 255)
      word |= 1u << pos;
    packed[pos >> 2] |= v << (8 * (pos & 3));
  }
  uint4* dst = reinterpret_cast<uint4*>(bricks + (size_t)b * 32);
  dst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
  dst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
  bits[b] = word;
  const int cnt = warp_sum_i(__popc(word));
  if ((threadIdx.x & 31) == 0 && cnt)
    atomicAdd(near_count, (unsigned long long)cnt);
}
}  // namespace

namespace bamcl
{
int map3d_build_bricks(bamcl_map3d* m)
{
  m->CW = (m->W + 3) / 4;
  m->CH = (m->H + 3) / 4;
  m->CD = (m->D + 1) / 2;
  const long n_bricks = (long)m->CW * m->CH * m->CD;
  if (n_bricks * 32 >= (1l << 32))
    return set_error(BAMCL_E_UNSUPPORTED, "3-D map of %ld bricks exceeds the scorer's 32-bit brick addressing", n_bricks);
  BAMCL_TRY(m->bricks.reserve((size_t)n_bricks * 32));
  BAMCL_TRY(m->near_bits.reserve((size_t)n_bricks * 4 + 16));
  unsigned long long* counter = reinterpret_cast<unsigned long long*>(m->near_bits.as<unsigned>() + ((n_bricks + 1) & ~1l));
  BAMCL_CUDA_OK(cudaMemsetAsync(counter, 0, sizeof(unsigned long long), m->ctx->stream));
  BAMCL_LAUNCH(m->ctx, k3_build_bricks, (unsigned)((n_bricks + 255) / 256), 256, 0, m->field.as<uint8_t>(), m->W, m->H, m->D, m->CW,
               m->CH, n_bricks, m->bricks.as<uint8_t>(), m->near_bits.as<unsigned>(), counter);
  unsigned long long near = 0;
  BAMCL_CUDA_OK(cudaMemcpyAsync(&near, counter, sizeof(near), cudaMemcpyDeviceToHost, m->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(m->ctx->stream));
  m->near_fraction = (double)near / ((double)m->W * m->H * m->D);
  return BAMCL_OK;
}
}  // namespace bamcl

extern "C" {

int bamcl_map3d_create(bamcl_ctx* ctx, double resolution, double max_distance, const int* keys, long n_keys,
                       bamcl_map3d** out)
{
  if (!ctx || !out || !(resolution > 0.0) || n_keys < 0 || (n_keys > 0 && !keys))
    return set_error(BAMCL_E_ARG, "bamcl_map3d_create: bad arguments");
  bamcl_map3d* m = new bamcl_map3d();
  m->ctx = ctx;
  m->resolution = resolution;
  m->max_distance = max_distance;
  // initFromOctree (octomap.cpp:53-74): bounds = getMetricMin/Max over the leaves (octomap's calcMinMax: leaf
  // centre -/+ half a leaf), converted with convertWorldToMap; sources = convertWorldToMap(leaf centre) (:220-227)
  double mn[3], mx[3];
  for (int a = 0; a < 3; a++)
  {
    mn[a] = 1.7976931348623157e308;
    mx[a] = -1.7976931348623157e308;
  }
  m->src.resize((size_t)3 * n_keys);
  for (long k = 0; k < n_keys; k++)
    for (int a = 0; a < 3; a++)
    {
      const double c = (keys[3 * k + a] + 0.5) * resolution;
      const double lo = c - resolution / 2.0;
      if (lo < mn[a])
        mn[a] = lo;
      if (lo + resolution > mx[a])
        mx[a] = lo + resolution;
      m->src[3 * k + a] = world_to_voxel(c, resolution);
    }
  if (n_keys == 0)
    for (int a = 0; a < 3; a++)
      mn[a] = mx[a] = 0.0;
  for (int a = 0; a < 3; a++)
  {
    m->cmin[a] = world_to_voxel(mn[a], resolution);
    m->cmax[a] = world_to_voxel(mx[a], resolution);
  }
  recount(m);
  if (n_keys > 0)
  {
    BAMCL_TRY(m->d_src.reserve(sizeof(int) * 3 * (size_t)n_keys));
    BAMCL_CUDA_OK(cudaMemcpyAsync(m->d_src.p, m->src.data(), sizeof(int) * 3 * (size_t)n_keys, cudaMemcpyHostToDevice,
                                  ctx->stream));
    BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  }
  *out = m;
  return BAMCL_OK;
}

int bamcl_map3d_destroy(bamcl_map3d* m)
{
  if (!m)
    return BAMCL_OK;
  cudaStreamSynchronize(m->ctx->stream);
  m->d_src.release();
  m->field.release();
  m->bricks.release();
  m->near_bits.release();
  delete m;
  return BAMCL_OK;
}

int bamcl_map3d_get_bounds(bamcl_map3d* m, int mn[3], int mx[3])
{
  if (!m || !mn || !mx)
    return set_error(BAMCL_E_ARG, "bamcl_map3d_get_bounds: bad arguments");
  memcpy(mn, m->cmin, sizeof(int) * 3);
  memcpy(mx, m->cmax, sizeof(int) * 3);
  return BAMCL_OK;
}

int bamcl_map3d_build_distances(bamcl_map3d* m)
{
  if (!m)
    return set_error(BAMCL_E_ARG, "null map");
  if (m->max_distance == 0.0)  // octomap.cpp:176-180: nothing is built
    return BAMCL_OK;
  if (m->W <= 0 || m->H <= 0 || m->D <= 0)
    return set_error(BAMCL_E_ARG, "bamcl_map3d_build_distances: empty bounds");
  const int R = (int)std::floor(m->max_distance / m->resolution) + 1;
  if (R > 180)
    return set_error(BAMCL_E_UNSUPPORTED, "max_distance/resolution = %d voxels exceeds the u16 pass (180)", R);
  bamcl_ctx* ctx = m->ctx;
  const size_t n = (size_t)m->W * m->H * m->D;
  BAMCL_TRY(m->field.reserve(n));
  DevBuf occ, g1, g2;
  BAMCL_TRY(occ.reserve(n));
  BAMCL_TRY(g1.reserve(n));
  BAMCL_TRY(g2.reserve(sizeof(uint16_t) * n));
  BAMCL_CUDA_OK(cudaMemsetAsync(occ.p, 0, n, ctx->stream));
  const long n_src = (long)(m->src.size() / 3);
  if (n_src > 0)
    BAMCL_LAUNCH(ctx, k3_scatter_sources, (unsigned)((n_src + 255) / 256), 256, 0, m->d_src.as<int>(), n_src, occ.as<uint8_t>(),
                 m->cmin[0], m->cmin[1], m->cmin[2], m->W, m->H, m->D);
  const unsigned blocks = (unsigned)((n + 255) / 256);
  BAMCL_LAUNCH(ctx, k3_pass_x, blocks, 256, 0, occ.as<uint8_t>(), g1.as<uint8_t>(), m->W, n, R);
  BAMCL_LAUNCH(ctx, k3_pass_y, blocks, 256, 0, g1.as<uint8_t>(), g2.as<uint16_t>(), m->W, m->H, n, R);
  BAMCL_LAUNCH(ctx, k3_pass_z, blocks, 256, 0, g2.as<uint16_t>(), m->field.as<uint8_t>(), m->W, m->H, m->D, n, R,
               m->resolution, m->max_distance);
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  occ.release();
  g1.release();
  g2.release();
  BAMCL_TRY(bamcl::map3d_build_bricks(m));
  m->has_dist = true;
  m->version++;
  return BAMCL_OK;
}

int bamcl_map3d_set_bounds(bamcl_map3d* m, const double mn[2], const double mx[2])
{
  if (!m || !mn || !mx)
    return set_error(BAMCL_E_ARG, "bamcl_map3d_set_bounds: bad arguments");
  // setMapBounds (octomap.cpp:128-150): crop to the occupancy map's extent grown by max_distance, then rebuild
  for (int a = 0; a < 2; a++)
  {
    const int lo = (int)std::floor((mn[a] - m->max_distance) / m->resolution + 0.5);
    const int hi = (int)std::floor((mx[a] + m->max_distance) / m->resolution + 0.5);
    if (lo > m->cmin[a])
      m->cmin[a] = lo;
    if (hi < m->cmax[a])
      m->cmax[a] = hi;
  }
  recount(m);
  m->has_dist = false;
  return bamcl_map3d_build_distances(m);
}

int bamcl_map3d_upload_distances(bamcl_map3d* m, const uint8_t* dense, long n)
{
  if (!m || !dense || n != (long)m->W * m->H * m->D || n <= 0)
    return set_error(BAMCL_E_ARG, "bamcl_map3d_upload_distances: need W*H*D values (x fastest, then y, then z)");
  BAMCL_TRY(m->field.reserve((size_t)n));
  BAMCL_CUDA_OK(cudaMemcpyAsync(m->field.p, dense, (size_t)n, cudaMemcpyHostToDevice, m->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(m->ctx->stream));
  BAMCL_TRY(bamcl::map3d_build_bricks(m));
  m->has_dist = true;
  m->version++;
  return BAMCL_OK;
}

int bamcl_map3d_download_distances(bamcl_map3d* m, uint8_t* out, long n)
{
  if (!m || !out || n != (long)m->W * m->H * m->D)
    return set_error(BAMCL_E_ARG, "bamcl_map3d_download_distances: need W*H*D values");
  if (!m->has_dist)
    return set_error(BAMCL_E_STATE, "no 3-D distance cache has been built or uploaded");
  BAMCL_CUDA_OK(cudaMemcpyAsync(out, m->field.p, (size_t)n, cudaMemcpyDeviceToHost, m->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(m->ctx->stream));
  return BAMCL_OK;
}

}  // extern "C"
