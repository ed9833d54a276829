// K-c: 3-D point-cloud likelihood of every particle against the OctoMap distance field.
// Reference: src/amcl/sensors/point_cloud_scanner.cpp:92-129 (updateSensor / applyModelToSampleSet), 132-167
// (calcPointCloudModel), 169-203 (Gompertz), 205-229 (recalcWeight), 231-248 (getMapCloud), 250-260 (applyGompertz).
//
// getMapCloud delegates to tf2 / tf2_sensor_msgs / Eigen (not in the reference tree, versions unpinned); the
// arithmetic sequence restated in oracle/amcl_oracle.c (orc_cloud_transform) is followed step by step: f64
// compose of R_z(theta), (x, y, 0) with the mount transform, matrix -> quaternion (Shepperd branches), cast to
// f32, Eigen's quaternion -> matrix in f32, then p' = ((R0 x + R1 y) + R2 z) + t per point in f32 without FMA.
//
// Thread mapping as in K-a: a warp owns PB particles (their 12 f32 transform entries sit in shared memory) and
// walks the cloud in passes of 32 * NPL points; lane l keeps NPL points in registers, adjacent lanes = adjacent
// points of the scan = neighbouring voxels. The per-point term comes from a 257-entry f64 table (one entry per
// quantised distance + one for "outside the map"), built on the host with glibc's exp in the reference's
// expression order and held in shared memory in 16 interleaved copies (bank-conflict free).
//
// What is gathered (map3d.cuh): the volume in bricks of 4 x 4 x 2 voxels. A look-up first tests the voxel's bit in the
// brick's "near an obstacle" word (the bit volume is 1/8 of the field and L2-resident); only voxels within max_distance of
// an obstacle -- a minority of the evals of any cloud -- then read their distance byte from the brick's 32-byte sector.
// The staged cloud is ordered along a Morton curve, so the 32 points of a lane group fall into a handful of bricks.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <utility>
#include <vector>

#include "map3d.cuh"
#include "pf.cuh"

using namespace bamcl;

namespace
{
#ifndef CLOUD_PB
#define CLOUD_PB 16
#endif
#ifndef CLOUD_NPL
#define CLOUD_NPL 2
#endif
#ifndef CLOUD_GB
#define CLOUD_GB 8
#endif
#ifndef CLOUD_MINB
#define CLOUD_MINB 3
#endif
// (swept on a B200, scripts/sweep_cloud.sh, 1M x 4096: 16/4/4/2 CTAs 48.5 ms, 8/4/4/3 29.5, 16/2/8/3 28.1, 16/2/4/4 29.4,
// 16/1/16/4 32.2, 32/4/4/2 57.3: occupancy -- registers -- decides, three CTAs per SM is the sweet spot)
constexpr int PB = CLOUD_PB;    // particles per warp batch
constexpr int NPL = CLOUD_NPL;  // points per lane per pass
constexpr int GB = CLOUD_GB;    // particles per gather batch
constexpr int WARPS = 8;
constexpr int PASS_POINTS = 32 * NPL;
constexpr int TABLE_N = 257;

struct CloudArgs
{
  const double *x, *y, *th;
  double* w;
  int n;
  const float *px, *py, *pz;
  int n_points;
  const uint8_t* bricks;    // [n_bricks][32]
  const unsigned* near_bits;  // [n_bricks]
  int i0, j0, k0, W, H, D, CW, CH;
  const int* perm;  // scoring order (window-local particle indices), or null
  int n_dropped;  // points with a non-finite coordinate: outside every map, removed from the staged cloud
  double resolution, inv_resolution;
  float inv_resolution_f;
  const double* table;
  int model;
  double gompertz_a, gompertz_b, gompertz_c, input_shift, input_scale, output_shift;
  double off_map_factor;
  double mt[3];   // mount translation
  double R2[9];   // mount rotation as a basis (tf2::Matrix3x3::setRotation of the quaternion)
};

struct Xform  // Eigen::Transform<float,3,Affine> of one particle
{
  float R[9];
  float t[3];
};

// floor(v / res + 0.5) exactly as octomap.cpp:98-109 computes it. The product with 1/res is within 1e-10 of the
// true quotient for any |v / res| < 1e5, so only values that close to a rounding boundary pay for the division.
__device__ __forceinline__ int world_to_voxel(float vf, double res, double inv_res)
{
  const double v = (double)vf;
  const double f = v * inv_res + 0.5;
  const double fl = floor(f);
  if (f - fl < 1e-7 || fl + 1.0 - f < 1e-7)
    return (int)floor(v / res + 0.5);
  return (int)fl;
}

// The same index from one f32 FMA when the result is clear of a rounding boundary by more than the f32 error of that
// FMA (|f| * 1.2e-7: inv_res rounded to f32 + the FMA's own rounding); the reference's f64 expression decides the rest.
__device__ __forceinline__ int world_to_voxel_fast(float vf, float inv_res_f, double res, double inv_res)
{
  const float f = __fmaf_rn(vf, inv_res_f, 0.5f);
  const float fl = floorf(f);
  if (fabsf((f - fl) - 0.5f) > 0.5f - (4e-7f * fabsf(f) + 1e-6f))
    return world_to_voxel(vf, res, inv_res);
  return (int)fl;
}

__device__ void particle_transform(const CloudArgs& a, double px, double py, double pth, Xform* out)
{
  // tf2::Quaternion::setRPY(0, 0, yaw) -> (0, 0, sin(yaw/2), cos(yaw/2)); tf2::Matrix3x3::setRotation
  const double half = pth * 0.5;
  double sh, ch;
  sincos(half, &sh, &ch);
  const double q1[4] = {0.0, 0.0, sh, ch};
  double R1[9];
  {
    const double d = q1[0] * q1[0] + q1[1] * q1[1] + q1[2] * q1[2] + q1[3] * q1[3];
    const double s = 2.0 / d;
    const double xs = q1[0] * s, ys = q1[1] * s, zs = q1[2] * s;
    const double wx = q1[3] * xs, wy = q1[3] * ys, wz = q1[3] * zs;
    const double xx = q1[0] * xs, xy = q1[0] * ys, xz = q1[0] * zs;
    const double yy = q1[1] * ys, yz = q1[1] * zs, zz = q1[2] * zs;
    R1[0] = 1.0 - (yy + zz); R1[1] = xy - wz;         R1[2] = xz + wy;
    R1[3] = xy + wz;         R1[4] = 1.0 - (xx + zz); R1[5] = yz - wx;
    R1[6] = xz - wy;         R1[7] = yz + wx;         R1[8] = 1.0 - (xx + yy);
  }
  // footprint_to_map * scanner_to_footprint
  double R[9], o[3];
#pragma unroll
  for (int r = 0; r < 3; r++)
  {
#pragma unroll
    for (int c = 0; c < 3; c++)
      R[r * 3 + c] = R1[r * 3] * a.R2[c] + R1[r * 3 + 1] * a.R2[3 + c] + R1[r * 3 + 2] * a.R2[6 + c];
    const double o1 = r == 0 ? px : r == 1 ? py : 0.0;
    o[r] = R1[r * 3] * a.mt[0] + R1[r * 3 + 1] * a.mt[1] + R1[r * 3 + 2] * a.mt[2] + o1;
  }
  // tf2::Matrix3x3::getRotation
  double q[4];
  const double trace = R[0] + R[4] + R[8];
  if (trace > 0.0)
  {
    double s = sqrt(trace + 1.0);
    q[3] = s * 0.5;
    s = 0.5 / s;
    q[0] = (R[7] - R[5]) * s;
    q[1] = (R[2] - R[6]) * s;
    q[2] = (R[3] - R[1]) * s;
  }
  else
  {
    const int i = R[0] < R[4] ? (R[4] < R[8] ? 2 : 1) : (R[0] < R[8] ? 2 : 0);
    const int j = (i + 1) % 3, k = (i + 2) % 3;
    double s = sqrt(R[i * 3 + i] - R[j * 3 + j] - R[k * 3 + k] + 1.0);
    double qq[4];
    qq[i] = s * 0.5;
    s = 0.5 / s;
    qq[3] = (R[k * 3 + j] - R[j * 3 + k]) * s;
    qq[j] = (R[j * 3 + i] + R[i * 3 + j]) * s;
    qq[k] = (R[k * 3 + i] + R[i * 3 + k]) * s;
    q[0] = qq[0];
    q[1] = qq[1];
    q[2] = qq[2];
    q[3] = qq[3];
  }
  // tf2::doTransform(PointCloud2): Translation3f(t) * Quaternionf(q); Eigen::Quaternionf::toRotationMatrix
  const float qx = (float)q[0], qy = (float)q[1], qz = (float)q[2], qw = (float)q[3];
  const float tx = __fmul_rn(2.0f, qx), ty = __fmul_rn(2.0f, qy), tz = __fmul_rn(2.0f, qz);
  const float twx = __fmul_rn(tx, qw), twy = __fmul_rn(ty, qw), twz = __fmul_rn(tz, qw);
  const float txx = __fmul_rn(tx, qx), txy = __fmul_rn(ty, qx), txz = __fmul_rn(tz, qx);
  const float tyy = __fmul_rn(ty, qy), tyz = __fmul_rn(tz, qy), tzz = __fmul_rn(tz, qz);
  out->R[0] = __fsub_rn(1.0f, __fadd_rn(tyy, tzz));
  out->R[1] = __fsub_rn(txy, twz);
  out->R[2] = __fadd_rn(txz, twy);
  out->R[3] = __fadd_rn(txy, twz);
  out->R[4] = __fsub_rn(1.0f, __fadd_rn(txx, tzz));
  out->R[5] = __fsub_rn(tyz, twx);
  out->R[6] = __fsub_rn(txz, twy);
  out->R[7] = __fadd_rn(tyz, twx);
  out->R[8] = __fsub_rn(1.0f, __fadd_rn(txx, tyy));
  out->t[0] = (float)o[0];
  out->t[1] = (float)o[1];
  out->t[2] = (float)o[2];
}

template <bool TAIL, bool BITS>
__device__ __forceinline__ void cloud_pass(const CloudArgs& a, const double* __restrict__ tab, const Xform* stash, int p0,
                                           int lane, double (&acc)[PB])
{
  float x[NPL], y[NPL], z[NPL];
  bool live[NPL];
#pragma unroll
  for (int q = 0; q < NPL; q++)
  {
    const int t = p0 + q * 32 + lane;
    live[q] = !TAIL || t < a.n_points;
    x[q] = live[q] ? __ldg(&a.px[t]) : 0.f;
    y[q] = live[q] ? __ldg(&a.py[t]) : 0.f;
    z[q] = live[q] ? __ldg(&a.pz[t]) : 0.f;
  }
  const unsigned copy = (unsigned)(lane & 15);
  constexpr unsigned OUTSIDE = 0xFFFFFFFFu;
#pragma unroll
  for (int g0 = 0; g0 < PB; g0 += GB)
  {
    unsigned addr[GB][NPL], word[GB][NPL];
#pragma unroll
    for (int gg = 0; gg < GB; gg++)
    {
      const Xform& T = stash[g0 + gg];
#pragma unroll
      for (int q = 0; q < NPL; q++)
      {
        int v[3];
#pragma unroll
        for (int r = 0; r < 3; r++)
        {
          const float m = __fadd_rn(
              __fadd_rn(__fadd_rn(__fmul_rn(T.R[r * 3], x[q]), __fmul_rn(T.R[r * 3 + 1], y[q])), __fmul_rn(T.R[r * 3 + 2], z[q])),
              T.t[r]);
          v[r] = world_to_voxel_fast(m, a.inv_resolution_f, a.resolution, a.inv_resolution);
        }
        const int i = v[0] - a.i0, j = v[1] - a.j0, k = v[2] - a.k0;
        unsigned ad = OUTSIDE;  // outside the (cropped) bounds: getDistanceToObject returns max_distance, :338-341
        unsigned wd = 0;
        if ((unsigned)i < (unsigned)a.W && (unsigned)j < (unsigned)a.H && (unsigned)k < (unsigned)a.D)
        {
          const unsigned brick = (unsigned)(((k >> 1) * a.CH + (j >> 2)) * a.CW + (i >> 2));
          ad = (brick << 5) | (unsigned)(((k & 1) << 4) | ((j & 3) << 2) | (i & 3));
          wd = BITS ? __ldg(&a.near_bits[brick]) : 0xFFFFFFFFu;
        }
        addr[gg][q] = ad;
        word[gg][q] = wd;
      }
    }
#pragma unroll
    for (int gg = 0; gg < GB; gg++)
#pragma unroll
      for (int q = 0; q < NPL; q++)
      {
        const unsigned ad = addr[gg][q];
        unsigned qd = ad == OUTSIDE ? 256u : 255u;  // far from every obstacle: the quantised distance saturates at 255
        if ((word[gg][q] >> (ad & 31u)) & 1u)
          qd = __ldg(&a.bricks[ad]);
        addr[gg][q] = qd;
      }
#pragma unroll
    for (int gg = 0; gg < GB; gg++)
#pragma unroll
      for (int q = 0; q < NPL; q++)
      {
        const double v = tab[(addr[gg][q] << 4) | copy];
        if (!TAIL || live[q])
          acc[g0 + gg] += v;
      }
  }
}

// BITS: test the "near an obstacle" bit before touching the distance byte (pays when most of the volume is far from
// obstacles, the usual map); maps where most voxels are near something read the byte directly
template <bool BITS>
__global__ void __launch_bounds__(WARPS * 32, CLOUD_MINB) k_score_cloud(const CloudArgs a)
{
  extern __shared__ __align__(16) unsigned char s_raw[];
  double* s_tab = reinterpret_cast<double*>(s_raw);                                    // [TABLE_N][16]
  Xform* s_stash = reinterpret_cast<Xform*>(s_raw + sizeof(double) * TABLE_N * 16);    // [WARPS][PB]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int t = threadIdx.x; t < TABLE_N * 16; t += blockDim.x)
    s_tab[t] = a.table[t >> 4];
  __syncthreads();
  Xform* stash = s_stash + warp * PB;
  const int n_batches = (a.n + PB - 1) / PB;
  const int warps_total = gridDim.x * WARPS;
  const int n_full = a.n_points / PASS_POINTS;

  for (int batch = blockIdx.x * WARPS + warp; batch < n_batches; batch += warps_total)
  {
    const int slot = batch * PB + lane;
    const bool mine = lane < PB && slot < a.n;
    const int idx = a.perm ? a.perm[min(slot, a.n - 1)] : slot;
    double w = 0.0, factor = 1.0;
    __syncwarp();
    if (lane < PB)
    {
      const int src = a.perm ? idx : min(idx, a.n - 1);
      const double px = a.x[src], py = a.y[src], pth = a.th[src];
      w = a.w[src];
      particle_transform(a, px, py, pth, &stash[lane]);
      // recalcWeight, point_cloud_scanner.cpp:205-229: only the off-map factor, by the robot's (x, y) voxel
      const int vi = (int)floor(px / a.resolution + 0.5), vj = (int)floor(py / a.resolution + 0.5);
      if (!((unsigned)(vi - a.i0) < (unsigned)a.W && (unsigned)(vj - a.j0) < (unsigned)a.H))
        factor = a.off_map_factor;
    }
    __syncwarp();
    double acc[PB];
#pragma unroll
    for (int g = 0; g < PB; g++)
      acc[g] = 0.0;
    for (int pass = 0; pass < n_full; pass++)
      cloud_pass<false, BITS>(a, s_tab, stash, pass * PASS_POINTS, lane, acc);
    if (n_full * PASS_POINTS < a.n_points)
      cloud_pass<true, BITS>(a, s_tab, stash, n_full * PASS_POINTS, lane, acc);
    double sum = 0.0;
#pragma unroll
    for (int g = 0; g < PB; g++)
    {
      const double v = warp_sum(acc[g]);
      if (lane == g)
        sum = v;
    }
    if (a.n_dropped)
      sum += (double)a.n_dropped * s_tab[256 << 4];
    if (mine)
    {
      double p;
      if (a.model == BAMCL_CLOUD_PLAIN)
        p = 1.0 + sum;  // :145,:161
      else
      {
        p = sum / (a.n_points + a.n_dropped);  // :197 (no zero guard in the reference: an empty cloud gives NaN)
        p = p * a.input_scale + a.input_shift;  // applyGompertz :250-260
        p = a.gompertz_a * exp(-1.0 * a.gompertz_b * exp(-1.0 * a.gompertz_c * p));
        p += a.output_shift;
      }
      w *= p;
      w *= factor;
      a.w[idx] = w;
    }
  }
}

// SoA split of the packed xyz cloud
__global__ void k_split_points(const float* __restrict__ xyz, float* __restrict__ px, float* __restrict__ py,
                               float* __restrict__ pz, int n)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n)
    return;
  px[t] = xyz[3 * t];
  py[t] = xyz[3 * t + 1];
  pz[t] = xyz[3 * t + 2];
}
}  // namespace

// =============================================================================================
// host side
// =============================================================================================
namespace bamcl
{
static void quat_to_basis(const double q[4], double R[9])  // tf2::Matrix3x3::setRotation
{
  const double d = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  const double s = 2.0 / d;
  const double xs = q[0] * s, ys = q[1] * s, zs = q[2] * s;
  const double wx = q[3] * xs, wy = q[3] * ys, wz = q[3] * zs;
  const double xx = q[0] * xs, xy = q[0] * ys, xz = q[0] * zs;
  const double yy = q[1] * ys, yz = q[1] * zs, zz = q[2] * zs;
  R[0] = 1.0 - (yy + zz); R[1] = xy - wz;         R[2] = xz + wy;
  R[3] = xy + wz;         R[4] = 1.0 - (xx + zz); R[5] = yz - wx;
  R[6] = xz - wy;         R[7] = yz + wx;         R[8] = 1.0 - (xx + yy);
}

int stage_cloud(bamcl_pf* pf, bamcl_map3d* map, const bamcl_cloud_model* model, const float* points_xyz, long n_points)
{
  StagedCloud& sc = pf->cloud;
  sc.valid = false;
  if (!pf || !map || !model || n_points < 0 || (n_points > 0 && !points_xyz))
    return set_error(BAMCL_E_ARG, "stage_cloud: bad arguments");
  if (!map->has_dist)
    return set_error(BAMCL_E_STATE, "the 3-D map has no distance cache (build_distances / upload_distances first)");
  if (n_points > (1L << 30))
    return set_error(BAMCL_E_ARG, "stage_cloud: too many points");
  sc.model = *model;
  sc.n_points = (int)n_points;
  sc.map_version = map->version;
  bamcl_ctx* ctx = pf->ctx;
  // per-quantum value table, in the reference's expression order (point_cloud_scanner.cpp:138-160, 175-193)
  BAMCL_TRY(pf->h_stage.reserve(sizeof(double) * TABLE_N + 3 * sizeof(float) * (size_t)std::max<long>(n_points, 1)));
  double* ht = pf->h_stage.as<double>();
  const double z_hit_denom = 2 * model->sigma_hit * model->sigma_hit;
  const double z_rand_mult = 1.0 / map->max_distance;
  const double max_distance_ratio = map->max_distance / 255;  // octomap.cpp:57
  for (int q = 0; q < TABLE_N; q++)
  {
    const double z = q < 256 ? q * max_distance_ratio : map->max_distance;  // :348 / :338-341
    double pz = 0.0;
    pz += model->z_hit * std::exp(-(z * z) / z_hit_denom);
    if (model->model_type == BAMCL_CLOUD_PLAIN)
    {
      pz += model->z_rand * z_rand_mult;
      ht[q] = pz * pz * pz;
    }
    else
    {
      pz += model->z_rand;
      ht[q] = pz;
    }
  }
  BAMCL_TRY(sc.table.reserve(sizeof(double) * TABLE_N));
  BAMCL_CUDA_OK(cudaMemcpyAsync(sc.table.p, ht, sizeof(double) * TABLE_N, cudaMemcpyHostToDevice, ctx->stream));
  // The per-particle sum over the cloud is order-free up to f64 rounding (tests: 1e-6 on weights), so the staged cloud
  // is ordered along a Morton curve of its sensor-frame cells (0.2 m = one brick): adjacent lanes then look at adjacent
  // voxels under any rigid transform. Points with a non-finite coordinate lie outside every map ((int)NaN is INT_MIN in
  // the reference, octomap.cpp:98-109): they are dropped here and their constant term is added back by the kernel.
  std::vector<std::pair<unsigned long long, int>> order;
  order.reserve((size_t)n_points);
  sc.n_dropped = 0;
  for (long t = 0; t < n_points; t++)
  {
    const float* p = points_xyz + 3 * t;
    if (!std::isfinite(p[0]) || !std::isfinite(p[1]) || !std::isfinite(p[2]))
    {
      sc.n_dropped++;
      continue;
    }
    unsigned long long code = 0;
    unsigned c[3];
    for (int r = 0; r < 3; r++)
    {
      const double cell = std::floor((double)p[r] / 0.2) + 1048576.0;  // 21 bits per axis around the sensor
      c[r] = (unsigned)std::min(2097151.0, std::max(0.0, cell));
    }
    for (int bit = 0; bit < 21; bit++)
      for (int r = 0; r < 3; r++)
        code |= (unsigned long long)((c[r] >> bit) & 1u) << (3 * bit + r);
    order.emplace_back(code, (int)t);
  }
  std::stable_sort(order.begin(), order.end(), [](const auto& l, const auto& r) { return l.first < r.first; });
  sc.n_points = (int)order.size();
  if (sc.n_points > 0)
  {
    const size_t nb = sizeof(float) * (size_t)sc.n_points;
    float* hx = reinterpret_cast<float*>(ht + TABLE_N);
    for (size_t t = 0; t < order.size(); t++)
      memcpy(hx + 3 * t, points_xyz + 3 * (size_t)order[t].second, 3 * sizeof(float));
    BAMCL_TRY(sc.xyz.reserve(3 * nb));
    BAMCL_TRY(sc.px.reserve(nb));
    BAMCL_TRY(sc.py.reserve(nb));
    BAMCL_TRY(sc.pz.reserve(nb));
    BAMCL_CUDA_OK(cudaMemcpyAsync(sc.xyz.p, hx, 3 * nb, cudaMemcpyHostToDevice, ctx->stream));
    BAMCL_LAUNCH(ctx, k_split_points, (unsigned)((sc.n_points + 255) / 256), 256, 0, sc.xyz.as<float>(), sc.px.as<float>(),
                 sc.py.as<float>(), sc.pz.as<float>(), sc.n_points);
  }
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));  // the pinned table buffer and the caller's cloud are reusable
  sc.valid = true;
  return BAMCL_OK;
}

int launch_score_cloud(bamcl_pf* pf, bamcl_map3d* map)
{
  StagedCloud& sc = pf->cloud;
  if (!sc.valid)
    return set_error(BAMCL_E_STATE, "no cloud staged");
  if (sc.map_version != map->version)
    return set_error(BAMCL_E_STATE, "the 3-D map's distance cache changed after the cloud was staged");
  SampleSet& set = pf->sets[pf->current];
  if (set.sample_count == 0 || sc.model.max_beams < 2)
    return BAMCL_OK;
  bamcl_ctx* ctx = pf->ctx;
  CloudArgs a;
  a.x = set.x.as<double>() + pf_win_off(pf);
  a.y = set.y.as<double>() + pf_win_off(pf);
  a.th = set.th.as<double>() + pf_win_off(pf);
  a.w = set.w.as<double>() + pf_win_off(pf);
  a.n = pf_win_cnt(pf);
  a.px = sc.px.as<float>();
  a.py = sc.py.as<float>();
  a.pz = sc.pz.as<float>();
  a.n_points = sc.n_points;
  a.bricks = map->bricks.as<uint8_t>();
  a.near_bits = map->near_bits.as<unsigned>();
  a.CW = map->CW;
  a.CH = map->CH;
  a.n_dropped = sc.n_dropped;
  a.inv_resolution_f = (float)(1.0 / map->resolution);
  a.i0 = map->cmin[0];
  a.j0 = map->cmin[1];
  a.k0 = map->cmin[2];
  a.W = map->W;
  a.H = map->H;
  a.D = map->D;
  a.resolution = map->resolution;
  a.inv_resolution = 1.0 / map->resolution;
  a.table = sc.table.as<double>();
  a.model = sc.model.model_type;
  a.gompertz_a = sc.model.gompertz_a;
  a.gompertz_b = sc.model.gompertz_b;
  a.gompertz_c = sc.model.gompertz_c;
  a.input_shift = sc.model.input_shift;
  a.input_scale = sc.model.input_scale;
  a.output_shift = sc.model.output_shift;
  a.off_map_factor = sc.model.off_map_factor;
  for (int r = 0; r < 3; r++)
    a.mt[r] = sc.model.mount_translation[r];
  quat_to_basis(sc.model.mount_rotation, a.R2);
  const int n_batches = (a.n + PB - 1) / PB;
  const int blocks = std::max(1, std::min((n_batches + WARPS - 1) / WARPS, kNumSMs * 2 * CLOUD_MINB));
  const size_t smem = sizeof(double) * TABLE_N * 16 + sizeof(Xform) * WARPS * PB;
  // particles that sit close together look at the same part of the field: score them together (L2 reuse)
  a.perm = nullptr;
  static const bool no_order = getenv("BAMCL_NO_SCORE_ORDER") != nullptr;  // A/B switch
  if (a.n >= 65536 && !no_order)
  {
    const double x0 = map->cmin[0] * map->resolution, y0 = map->cmin[1] * map->resolution;
    const double extent = std::max(map->W, map->H) * map->resolution;
    BAMCL_TRY(pf_spatial_order(pf, x0, y0, std::max(extent / 64.0, 1e-3), &a.perm));
  }
  BAMCL_CUDA_OK(cudaEventRecord(pf->score_ev0, ctx->stream));
  static const char* force_bits = getenv("BAMCL_CLOUD_BITS");  // A/B switch: "0" / "1"
  const bool bits = force_bits ? force_bits[0] == '1' : map->near_fraction < 0.5;
  if (bits)
    BAMCL_LAUNCH(ctx, k_score_cloud<true>, blocks, WARPS * 32, smem, a);
  else
    BAMCL_LAUNCH(ctx, k_score_cloud<false>, blocks, WARPS * 32, smem, a);
  BAMCL_CUDA_OK(cudaEventRecord(pf->score_ev1, ctx->stream));
  pf->score_timed = true;
  return BAMCL_OK;
}
}  // namespace bamcl

namespace bamcl
{
int pf_shard_finish_update(bamcl_pf* pf, double* total_out);  // pf.cu
}

extern "C" {

int bamcl_pf_stage_cloud(bamcl_pf* pf, bamcl_map3d* map, const bamcl_cloud_model* model, const float* points_xyz,
                         long n_points)
{
  return stage_cloud(pf, map, model, points_xyz, n_points);
}

int bamcl_pf_score_cloud(bamcl_pf* pf, bamcl_map3d* map, const bamcl_cloud_model* model, const float* points_xyz,
                         long n_points, double* total_out)
{
  if (!total_out)
    return set_error(BAMCL_E_ARG, "bamcl_pf_score_cloud: total_out is required");
  BAMCL_TRY(stage_cloud(pf, map, model, points_xyz, n_points));
  if (model->max_beams < 2)  // point_cloud_scanner.cpp:114
  {
    *total_out = 0.0;
    return BAMCL_OK;
  }
  BAMCL_TRY(launch_score_cloud(pf, map));
  BAMCL_TRY(pf_reduce_total(pf, pf->scalars.as<double>()));
  return pf_read_scalar(pf, 0, total_out);
}

int bamcl_pf_update_cloud_staged(bamcl_pf* pf, bamcl_map3d* map, double* total_out_or_null)
{
  if (!pf || !map)
    return set_error(BAMCL_E_ARG, "bamcl_pf_update_cloud_staged: bad arguments");
  if (pf->cloud.valid && pf->cloud.model.max_beams < 2)  // updateSensor returns false (:95)
  {
    if (total_out_or_null)
      *total_out_or_null = 0.0;
    return BAMCL_OK;
  }
  BAMCL_TRY(launch_score_cloud(pf, map));
  BAMCL_TRY(pf_reduce_total(pf, pf->scalars.as<double>()));
  BAMCL_TRY(pf_normalise_by_device_total(pf));
  if (total_out_or_null)
    return pf_read_scalar(pf, 0, total_out_or_null);
  return BAMCL_OK;
}

int bamcl_pf_shard_update_cloud_staged(bamcl_pf* pf, bamcl_map3d* map, double* total_out)
{
  if (!pf || !map)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_update_cloud_staged: bad arguments");
  if (!pf->comm)
    return set_error(BAMCL_E_STATE, "bamcl_pf_shard_update_cloud_staged: no communicator attached");
  pf_apply_shard_window(pf);
  if (pf->cloud.valid && pf->cloud.model.max_beams < 2)  // updateSensor returns false (:95)
  {
    if (total_out)
      *total_out = 0.0;
    return BAMCL_OK;
  }
  BAMCL_TRY(launch_score_cloud(pf, map));
  BAMCL_TRY(pf_reduce_total(pf, pf->scalars.as<double>()));
  return pf_shard_finish_update(pf, total_out);
}

int bamcl_pf_update_cloud(bamcl_pf* pf, bamcl_map3d* map, const bamcl_cloud_model* model, const float* points_xyz,
                          long n_points, double* total_out)
{
  BAMCL_TRY(stage_cloud(pf, map, model, points_xyz, n_points));
  return bamcl_pf_update_cloud_staged(pf, map, total_out);
}

}  // extern "C"
