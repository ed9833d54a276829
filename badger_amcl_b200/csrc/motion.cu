// Motion (action) update on the device: Odom::updateAction, src/amcl/sensors/odom.cpp:74-311, with the Gaussian draws
// of src/amcl/pf/pdf_gaussian.cpp:77-97.
//
// The reference walks the samples in order and takes three PDFGaussian::draw values per sample from drand48. A draw
// is a rejection loop: attempts of two uniforms (x1, x2) in (-1, 1)^2 until w = x1^2 + x2^2 lies in (0, 1], then
// sigma * x2 * sqrt(-2 ln w / w). Attempt j therefore uses generator outputs 2j+1 and 2j+2 whatever happened before
// it (as long as no output is exactly 0, which the inner `while (r == 0.0)` loops would re-draw), and the a-th
// ACCEPTED attempt is draw a % 3 of sample a / 3. Parallel form:
//   1. k_attempt_flags : acceptance of M attempts at once (generator skip-ahead, lcg.cuh);
//   2. exclusive prefix sum of the flags = rank of every accepted attempt;
//   3. k_attempt_draws : accepted attempts of rank < 3P write sigma[rank % 3] * x2 * sqrt(-2 ln w / w);
//   4. k_apply_*       : one thread per sample applies the model's pose update.
// M is sized for the acceptance rate pi/4 plus a margin; a short batch is followed by another one.
// Loop invariants (sigmas, bearings) are computed on the host with the reference's expressions.
#include <cmath>
#include <cstring>

#include "lcg.cuh"
#include "pf.cuh"
#include "scan.cuh"

using namespace bamcl;

namespace
{
struct AttemptArgs
{
  unsigned long long state0;
  long long j0;  // first attempt of this batch
  int m;         // attempts in this batch
  int* flag;     // m (+1)
  int* meta;     // [0] accepted in this batch, [1] bit0: a generator output was 0, [2..3] attempt holding draw 3P-1
  long long base;  // draws accepted by earlier batches
  long long need;  // 3P
  double sigma[3];
  double* draws;  // 3P
};

// the two uniforms of attempt j -> (x1, x2, w); false if a generator output is exactly 0
__device__ __forceinline__ bool attempt(unsigned long long state0, long long j, double* x2, double* w)
{
  const unsigned long long a = lcg_skip(state0, 2ull * (unsigned long long)j + 1ull);
  const unsigned long long b = lcg_next(a);
  const double r1 = (double)a * 0x1p-48, r2 = (double)b * 0x1p-48;  // drand48: X / 2^48, exact
  const double x1 = 2.0 * r1 - 1.0;
  *x2 = 2.0 * r2 - 1.0;
  *w = x1 * x1 + *x2 * *x2;
  return a != 0ull && b != 0ull;
}

__global__ void __launch_bounds__(256) k_attempt_flags(const AttemptArgs a)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= a.m)
    return;
  double x2, w;
  if (!attempt(a.state0, a.j0 + t, &x2, &w))
    atomicOr(&a.meta[1], 1);
  a.flag[t] = (w > 1.0 || w == 0.0) ? 0 : 1;  // pdf_gaussian.cpp:94
}

__global__ void __launch_bounds__(256) k_attempt_draws(const AttemptArgs a, const int* __restrict__ rank)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= a.m)
    return;
  double x2, w;
  attempt(a.state0, a.j0 + t, &x2, &w);
  if (w > 1.0 || w == 0.0)
    return;
  const long long r = a.base + rank[t];
  if (r >= a.need)
    return;
  a.draws[r] = a.sigma[r % 3] * x2 * sqrt(-2.0 * log(w) / w);  // :96
  if (r == a.need - 1)
    *reinterpret_cast<long long*>(&a.meta[2]) = a.j0 + t;
}

// angles::normalize_angle in the fmod(a + pi, 2 pi) form (third party, see oracle/shim/angles); Odom::angleDiff(a, b)
// = normalize_angle(a - b), odom.cpp:318-321
__device__ __forceinline__ double angle_diff(double a, double b)
{
  const double r = fmod((a - b) + M_PI, 2.0 * M_PI);
  return r <= 0.0 ? r + M_PI : r - M_PI;
}

struct ApplyArgs
{
  int model, n;
  double *x, *y, *th;
  const double* g;  // 3 draws per sample, in the model's draw order
  double trans, rot;        // delta_trans, delta_rot
  double bearing0;          // angleDiff(atan2(dy, dx), old_theta)
  double rot1, rot2;        // diff models
  double half_rot;          // delta[2] / 2 (gaussian model)
};

__global__ void __launch_bounds__(256) k_apply_motion(const ApplyArgs a)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n)
    return;
  const double g0 = a.g[3 * i], g1 = a.g[3 * i + 1], g2 = a.g[3 * i + 2];
  double x = a.x[i], y = a.y[i], th = a.th[i];
  if (a.model == BAMCL_ODOM_OMNI || a.model == BAMCL_ODOM_OMNI_CORRECTED)
  {
    // odom.cpp:105-121 / :188-204, draws: trans, rot, strafe
    double sb, cb;
    sincos(a.bearing0 + th, &sb, &cb);
    const double trans_hat = a.trans + g0, rot_hat = a.rot + g1, strafe_hat = 0 + g2;
    x += (trans_hat * cb + strafe_hat * sb);
    y += (trans_hat * sb - strafe_hat * cb);
    th += rot_hat;
  }
  else if (a.model == BAMCL_ODOM_DIFF || a.model == BAMCL_ODOM_DIFF_CORRECTED)
  {
    // odom.cpp:150-167 / :231-253, draws: rot1, trans, rot2
    const double rot1_hat = angle_diff(a.rot1, g0);
    const double trans_hat = a.trans - g1;
    const double rot2_hat = angle_diff(a.rot2, g2);
    double sh, ch;
    sincos(th + rot1_hat, &sh, &ch);
    x += trans_hat * ch;
    y += trans_hat * sh;
    th += rot1_hat + rot2_hat;
  }
  else
  {
    // odom.cpp:280-308, draws: trans, strafe, rot
    double sh, ch, sb, cb;
    sincos(th + a.half_rot, &sh, &ch);
    sincos(a.bearing0 + th, &sb, &cb);
    x += (a.trans * cb);
    y += (a.trans * sb);
    th += a.rot;
    x += (g0 * ch + g1 * sh);
    y += (g0 * sh - g1 * ch);
    th += g2;
  }
  a.x[i] = x;
  a.y[i] = y;
  a.th[i] = th;
}

// host copies of the reference's loop invariants
double h_normalize(double a)
{
  const double r = fmod(a + M_PI, 2.0 * M_PI);
  return r <= 0.0 ? r + M_PI : r - M_PI;
}
double h_angle_diff(double a, double b) { return h_normalize(a - b); }
double h_rot_noise(double rot)  // odom.cpp:145-148
{
  const double z = fabs(h_angle_diff(rot, 0.0)), p = fabs(h_angle_diff(rot, M_PI));
  return p < z ? p : z;
}
}  // namespace

namespace
{
// `need` PDFGaussian::draw values from generator state `state0`, draw r with sigma[r % 3], left in pf->scratch (device);
// *state_out = the generator state after the last attempt consumed
int gaussian_draws(bamcl_pf* pf, uint64_t state0, const double sigma[3], long long need, const char* who, double** draws_out,
                   uint64_t* state_out)
{
  bamcl_ctx* ctx = pf->ctx;
  const long long first = (long long)std::ceil((double)need * 1.2732395447351628 * 1.01) + 4096;
  if (first > 0x7fffff00ll)
    return set_error(BAMCL_E_UNSUPPORTED, "%s: sample set too large for one attempt batch", who);
  BAMCL_TRY(pf->scratch.reserve(sizeof(double) * (size_t)need));
  BAMCL_TRY(pf->scratch2.reserve(sizeof(int) * ((size_t)first + 1 + scan_scratch_ints(first + 1) + 16)));
  AttemptArgs at;
  at.state0 = state0 & LCG_MASK;
  at.meta = pf->scratch2.as<int>();
  at.flag = at.meta + 8;
  int* scan_scratch = at.flag + first + 1;
  at.need = need;
  at.draws = pf->scratch.as<double>();
  for (int k = 0; k < 3; k++)
    at.sigma[k] = sigma[k];
  BAMCL_CUDA_OK(cudaMemsetAsync(at.meta, 0, sizeof(int) * 8, ctx->stream));
  long long base = 0, j0 = 0, last_attempt = -1;
  while (base < need)
  {
    const long long left = need - base;
    long long m = (long long)std::ceil((double)left * 1.2732395447351628 * 1.01) + 4096;
    if (m > first)
      m = first;
    at.j0 = j0;
    at.m = (int)m;
    at.base = base;
    const unsigned blocks = (unsigned)((m + 255) / 256);
    BAMCL_LAUNCH(ctx, k_attempt_flags, blocks, 256, 0, at);
    BAMCL_TRY(exclusive_scan_i32(ctx, at.flag, at.flag, m, scan_scratch, at.meta));
    BAMCL_LAUNCH(ctx, k_attempt_draws, blocks, 256, 0, at, at.flag);
    int h_meta[4];
    BAMCL_CUDA_OK(cudaMemcpyAsync(h_meta, at.meta, sizeof(h_meta), cudaMemcpyDeviceToHost, ctx->stream));
    BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    if (h_meta[1] & 1)
      return set_error(BAMCL_E_UNSUPPORTED,
                       "%s: drand48 produced exactly 0.0 inside the draw range "
                       "(pdf_gaussian.cpp:83-92 re-draws, which shifts the pairing of the stream)", who);
    base += h_meta[0] < left ? h_meta[0] : left;
    j0 += m;
    memcpy(&last_attempt, &h_meta[2], sizeof(long long));
  }
  *draws_out = at.draws;
  *state_out = lcg_skip(at.state0, 2ull * (unsigned long long)(last_attempt + 1));
  return BAMCL_OK;
}
}  // namespace

extern "C" int bamcl_pf_update_action(bamcl_pf* pf, const bamcl_odom_model* model, const bamcl_odom_data* data,
                                      uint64_t drand48_state, uint64_t* drand48_state_out)
{
  if (!pf || !model || !data || !drand48_state_out)
    return set_error(BAMCL_E_ARG, "bamcl_pf_update_action: null argument");
  if (model->model_type < BAMCL_ODOM_DIFF || model->model_type > BAMCL_ODOM_GAUSSIAN)
    return set_error(BAMCL_E_ARG, "bamcl_pf_update_action: unknown model_type (odom.h:33-40)");
  bamcl_ctx* ctx = pf->ctx;
  SampleSet& set = pf->sets[pf->current];
  const int P = set.sample_count;
  *drand48_state_out = drand48_state & LCG_MASK;
  if (P <= 0)
    return BAMCL_OK;
  const int mt = model->model_type;
  const double a1 = model->alpha[0], a2 = model->alpha[1], a3 = model->alpha[2], a4 = model->alpha[3],
               a5 = model->alpha[4];
  const double* d = data->delta;
  const double old_theta = data->pose[2] - d[2];  // odom.cpp:83-85
  const double trans = std::sqrt(d[0] * d[0] + d[1] * d[1]);
  const double rot = d[2];

  ApplyArgs ap;
  memset(&ap, 0, sizeof(ap));
  ap.model = mt;
  ap.n = P;
  ap.x = set.x.as<double>();
  ap.y = set.y.as<double>();
  ap.th = set.th.as<double>();
  ap.trans = trans;
  ap.rot = rot;
  ap.half_rot = d[2] / 2;
  double sigma[3];
  if (mt == BAMCL_ODOM_OMNI || mt == BAMCL_ODOM_OMNI_CORRECTED)
  {
    sigma[0] = a3 * (trans * trans) + a1 * (rot * rot);  // :97-102 (variances handed to draw as they are) / :180-185
    sigma[1] = a4 * (rot * rot) + a2 * (trans * trans);
    sigma[2] = a1 * (rot * rot) + a5 * (trans * trans);
    ap.bearing0 = h_angle_diff(std::atan2(d[1], d[0]), old_theta);
  }
  else if (mt == BAMCL_ODOM_DIFF || mt == BAMCL_ODOM_DIFF_CORRECTED)
  {
    ap.rot1 = trans < 0.01 ? 0.0 : h_angle_diff(std::atan2(d[1], d[0]), old_theta);  // :133-137
    ap.rot2 = h_angle_diff(d[2], ap.rot1);
    const double n1 = h_rot_noise(ap.rot1), n2 = h_rot_noise(ap.rot2);
    sigma[0] = a1 * n1 * n1 + a2 * trans * trans;  // :155-161
    sigma[1] = a3 * trans * trans + a4 * n1 * n1 + a4 * n2 * n2;
    sigma[2] = a1 * n2 * n2 + a2 * trans * trans;
  }
  else
  {
    const double* am = data->absolute_motion;
    const double at2 = am[0] * am[0], as2 = am[1] * am[1], ar2 = am[2] * am[2];
    sigma[0] = a3 * at2 + a4 * ar2;  // trans, strafe, rot in draw order (:276-278, :293-295)
    sigma[1] = a4 * ar2 + a5 * as2;
    sigma[2] = a1 * ar2 + a2 * at2;
    ap.bearing0 = h_angle_diff(std::atan2(d[1], d[0]), old_theta);
  }
  if (mt != BAMCL_ODOM_DIFF && mt != BAMCL_ODOM_OMNI)
    for (int k = 0; k < 3; k++)
      sigma[k] = std::sqrt(sigma[k]);

  double* draws = nullptr;
  uint64_t state_after = 0;
  BAMCL_TRY(gaussian_draws(pf, drand48_state, sigma, 3ll * P, "bamcl_pf_update_action", &draws, &state_after));
  ap.g = draws;
  if (pf_ranks(pf) > 1)
  {
    // a distributed filter: every rank replays the draw stream of the WHOLE set (the attempt flags above: the number of
    // draws a sample consumes depends on every sample before it) and moves the window it owns
    long off = 0, cnt = 0;
    pf_owned(pf, &off, &cnt);
    ap.x += off;
    ap.y += off;
    ap.th += off;
    ap.g += 3 * off;
    ap.n = (int)cnt;
  }
  if (ap.n > 0)
    BAMCL_LAUNCH(ctx, k_apply_motion, (unsigned)((ap.n + 255) / 256), 256, 0, ap);
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  *drand48_state_out = state_after;
  return BAMCL_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// ParticleFilter::initWithGaussian (particle_filter.cpp:105-133) on the device: max_samples poses, pose s =
// PDFGaussian::sample() (pdf_gaussian.cpp:52-71): r[j] = draw(cd[j]) for j = 0, 1, 2 in this order - accepted attempts
// 3s, 3s+1, 3s+2 of the stream, exactly the motion update's pairing - then v(i) = x[i]; v(i) += cr(i, j) * r[j] for j = 0..2.
namespace bamcl
{
int build_tree_and_stats(bamcl_pf* pf, int set_index, double dist_threshold, double convergence_threshold,
                         bamcl_resample_result* result);  // kdtree.cu
}
namespace
{
struct GaussInitArgs
{
  int n;
  double *x, *y, *th, *w;
  const double* g;
  double mean[3], cr[9];
  double weight;
};
__global__ void __launch_bounds__(256) k_gaussian_init(const GaussInitArgs a)
{
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= a.n)
    return;
  const double r0 = a.g[3 * s], r1 = a.g[3 * s + 1], r2 = a.g[3 * s + 2];
  double v[3];
#pragma unroll
  for (int i = 0; i < 3; i++)
  {
    double t = a.mean[i];  // pdf_gaussian.cpp:65-69 (no contraction: the library is built with -fmad=false)
    t += a.cr[3 * i] * r0;
    t += a.cr[3 * i + 1] * r1;
    t += a.cr[3 * i + 2] * r2;
    v[i] = t;
  }
  a.x[s] = v[0];
  a.y[s] = v[1];
  a.th[s] = v[2];
  a.w[s] = a.weight;
}
}  // namespace

extern "C" int bamcl_pf_init_gaussian(bamcl_pf* pf, const double* mean3, const double* cr9, const double* cd3,
                                      uint64_t drand48_state, uint64_t* drand48_state_out, bamcl_resample_result* res)
{
  if (!pf || !mean3 || !cr9 || !cd3 || !drand48_state_out || !res)
    return set_error(BAMCL_E_ARG, "bamcl_pf_init_gaussian: null argument");
  bamcl_ctx* ctx = pf->ctx;
  pf->win_cnt = -1;
  SampleSet& set = pf->sets[pf->current];
  const int n = pf->max_samples;  // :113
  double* draws = nullptr;
  uint64_t state_after = 0;
  BAMCL_TRY(gaussian_draws(pf, drand48_state, cd3, 3ll * n, "bamcl_pf_init_gaussian", &draws, &state_after));
  GaussInitArgs a;
  a.n = n;
  a.x = set.x.as<double>();
  a.y = set.y.as<double>();
  a.th = set.th.as<double>();
  a.w = set.w.as<double>();
  a.g = draws;
  for (int k = 0; k < 3; k++)
    a.mean[k] = mean3[k];
  for (int k = 0; k < 9; k++)
    a.cr[k] = cr9[k];
  a.weight = 1.0 / (double)pf->max_samples;  // :119
  BAMCL_LAUNCH(ctx, k_gaussian_init, (unsigned)((n + 255) / 256), 256, 0, a);
  set.sample_count = n;
  pf_apply_shard_window(pf);  // a distributed filter: every rank generated the whole set and owns its window from here on
  memset(res, 0, sizeof(*res));
  BAMCL_TRY(build_tree_and_stats(pf, pf->current, 0.5, 1e300, res));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  res->sample_count = n;
  res->converged = 0;  // initConverged, :132
  *drand48_state_out = state_after;
  res->drand48_state = state_after;
  return BAMCL_OK;
}
