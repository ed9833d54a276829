// K-d (multinomial): KLD-adaptive multinomial resampling. Reference: src/amcl/pf/particle_filter.cpp:356-420.
//
// The reference draws one output at a time -- drand48() < w_diff ? random pose : (r = drand48(); linear search
// c[i] <= r < c[i+1]) -- inserts it into set_b's pose-bin tree and stops as soon as the output count exceeds
// resampleLimit(leaf count so far) (:416). Parallel form, for w_diff == 0 (two draws per output, the first one
// unused because `drand48() < 0` is never true):
//   1. exact sequential CDF (resample.cu);
//   2. all max_samples candidate outputs at once: output k uses draw 2k+2 of the generator, reached by the LCG's
//      skip-ahead (X_j = A^j X_0 + C_j mod 2^48), and the same upper_bound as the systematic select;
//   3. the pose-bin tree of ALL candidates (general build, which keeps every node's first foreign key). A node is
//      a leaf from its creation until its first child is created, so after m insertions
//         leaf(m) = #{nodes created by a sample < m} - #{nodes whose first child was created by a sample < m}
//      (SURVEY H2): two +-1 events per node on the sample time line and one prefix sum;
//   4. the smallest m with m > resampleLimit(leaf(m)), else max_samples: the new set is the first m candidates;
//   5. tree, clusters and statistics of those m samples as for any new set.
// w_diff > 0 interleaves random_pose_fn_() with the resampler's draws one output at a time. The node's pose function is
// Node::uniformPoseGenerator (node.cpp:262-264), which itself takes a fixed number g = 2 (1 + retries) of drand48
// values per pose (init.cu). With bamcl_pf_set_random_pose_source the device replays that too: output k starts at
// stream position p_k, p_0 = 0, p_{k+1} = p_k + (draw p_k + 1 < w_diff ? 1 + g : 2). All p_k at once by binary lifting:
// next[p] for every position, level_{j+1} = level_j o level_j, and p_k = the composition of the levels of k's set bits
// (powers of one function commute). Without a registered source (an arbitrary host callback) the case is refused.
#include <cmath>
#include <cstring>

#include "lcg.cuh"
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
constexpr int NONE = INT_MAX;

struct MultiArgs
{
  const double* c;  // P + 1
  int P;
  const double *ax, *ay, *ath;
  double *bx, *by, *bth, *bw;
  int* src;
  int n;
  unsigned long long state0;
  int* flags;
  // random-pose injection (w_diff > 0): stream position of every output, the free-space list and the map geometry
  const int* pos;  // null: output k starts at position 2k
  double w_diff;
  int retries;  // generations thrown away per random pose (uniformPoseGenerator's de-weighting loop)
  PoseSource rp;
  // host-drawn mode (bamcl_pf_resample_multinomial_drawn): output k's second uniform, or -(1 + i) = the i-th handed-over pose
  const double* host_draws;
  const double* host_poses;
};

__global__ void __launch_bounds__(256) k_select_multinomial(const MultiArgs a)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= a.n)
    return;
  double r;
  if (a.host_draws)
  {
    r = a.host_draws[k];
    if (r < 0.0)
    {
      const long i = (long)(-r) - 1;  // random_pose_fn_() was called on the host for this output (:385-388)
      a.src[k] = -1;
      a.bx[k] = a.host_poses[3 * i];
      a.by[k] = a.host_poses[3 * i + 1];
      a.bth[k] = a.host_poses[3 * i + 2];
      a.bw[k] = 1.0;
      return;
    }
  }
  else
  {
  const unsigned long long p = a.pos ? (unsigned long long)a.pos[k] : 2ull * (unsigned long long)k;
  const unsigned long long x1 = lcg_skip(a.state0, p + 1ull);  // :385
  if (a.w_diff > 0.0 && (double)x1 * 0x1p-48 < a.w_diff)
  {
    // random_pose_fn_() = the last of (1 + retries) randomFreeSpacePose generations (node.cpp:823-868)
    const unsigned long long u = lcg_skip(a.state0, p + 2ull + 2ull * (unsigned long long)a.retries), v = lcg_next(u);
    const unsigned int cell = (unsigned int)((double)u * 0x1p-48 * (double)(size_t)a.rp.free_n);
    a.src[k] = -1;
    a.rp.world(cell, &a.bx[k], &a.by[k]);
    a.bth[k] = (double)v * 0x1p-48 * 2 * M_PI - M_PI;
    a.bw[k] = 1.0;
    return;
  }
  const unsigned long long x = lcg_next(x1);
  r = scalbn((double)x, -48);  // drand48: X / 2^48, exact (:393)
  }
  int L = 0, H = a.P;  // first i with c[i] <= r < c[i+1] == number of entries of c[1..P] that are <= r
  while (L < H)
  {
    const int mid = (L + H) >> 1;
    if (__ldg(&a.c[mid + 1]) <= r)
      L = mid + 1;
    else
      H = mid;
  }
  int j = L;
  if (j >= a.P)
  {
    // r >= c[P]: the reference's search runs off the end and copies samples[P] (:391-403); flagged, last particle
    // with weight taken
    atomicOr(a.flags, 1);
    const double top = a.c[a.P];
    int l2 = 0, h2 = a.P;
    while (l2 < h2)
    {
      const int mid = (l2 + h2) >> 1;
      if (a.c[mid + 1] < top)
        l2 = mid + 1;
      else
        h2 = mid;
    }
    j = min(l2, a.P - 1);
  }
  a.src[k] = j;
  a.bx[k] = a.ax[j];
  a.by[k] = a.ay[j];
  a.bth[k] = a.ath[j];
  a.bw[k] = 1.0;
}

// next[p] = stream position after the output that starts at p; positions >= n_pos collapse into the sentinel n_pos
__global__ void __launch_bounds__(256) k_mn_next(unsigned long long state0, double w_diff, int token_random, int n_pos,
                                                 int* __restrict__ next)
{
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p > n_pos)
    return;
  if (p == n_pos)
  {
    next[p] = n_pos;
    return;
  }
  const unsigned long long x = lcg_skip(state0, (unsigned long long)p + 1ull);
  const long long q = (long long)p + ((double)x * 0x1p-48 < w_diff ? token_random : 2);
  next[p] = q >= n_pos ? n_pos : (int)q;
}
__global__ void __launch_bounds__(256) k_mn_apply(int* __restrict__ cur, const int* __restrict__ level, int bit, int n)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n && ((k >> bit) & 1))
    cur[k] = level[cur[k]];
}
__global__ void __launch_bounds__(256) k_mn_square(const int* __restrict__ in, int* __restrict__ out, int n)
{
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < n)
    out[p] = in[in[p]];
}
__global__ void __launch_bounds__(256) k_count_random(const int* __restrict__ src, int n, int* __restrict__ out)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const bool r = k < n && src[k] < 0;
  const unsigned m = __ballot_sync(0xffffffffu, r);
  if ((threadIdx.x & 31) == 0 && m)
    atomicAdd(out, __popc(m));
}

// +1 at the sample that created node v, -1 at the sample that created v's first child
__global__ void k_leaf_events(const int* __restrict__ nfirst, const int* __restrict__ first_at, int K, int* __restrict__ ev)
{
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= K)
    return;
  atomicAdd(&ev[nfirst[v]], 1);
  const int f = first_at[v];
  if (f != NONE)
    atomicAdd(&ev[nfirst[f]], -1);
}

// particle_filter.cpp:475-502 on the device (same operations in the same order; the library is built without FMA
// contraction)
__device__ __forceinline__ int d_resample_limit(int k, int min_samples, int max_samples, double pop_err, double pop_z)
{
  if (k <= 1)
    return max_samples;
  const double kd = (double)k;
  const double a = 1;
  const double b = 2 / (9 * (kd - 1));
  const double c = sqrt(2 / (9 * (kd - 1))) * pop_z;
  const double x = a - b + c;
  const int n = (int)ceil((k - 1) / (2 * pop_err) * x * x * x);
  if (n < min_samples)
    return min_samples;
  if (n > max_samples)
    return max_samples;
  return n;
}

// leaf_before[m] = leaf count after m insertions; the first m with m > resampleLimit(leaf) stops the loop (:416)
__global__ void k_find_stop(const int* __restrict__ leaf_before, int n, int min_samples, int max_samples, double pop_err,
                            double pop_z, int* __restrict__ stop)
{
  const int m = blockIdx.x * blockDim.x + threadIdx.x + 1;
  if (m > n)
    return;
  if (m > d_resample_limit(leaf_before[m], min_samples, max_samples, pop_err, pop_z))
    atomicMin(stop, m);
}

__global__ void k_fill_f64(double* __restrict__ p, double v, int n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    p[i] = v;
}
}  // namespace

namespace bamcl
{
// host_draws / host_poses (host pointers, or null): the draws were made on the host by the reference's own loop shape
// (bamcl_pf_resample_multinomial_drawn); then nothing is drawn here and the generator state is returned unchanged
int resample_multinomial(bamcl_pf* pf, const bamcl_resample_params* prm, bamcl_resample_result* res, const double* host_draws,
                         const double* host_poses, long n_host_poses)
{
  const bool inject = prm->w_diff > 0.0 && !host_draws;
  PoseSource rp_src;
  if (inject && !pf_pose_source(pf, &rp_src))
    return set_error(BAMCL_E_UNSUPPORTED,
                     "multinomial resampling with w_diff > 0 interleaves random_pose_fn_() with the resampler's drand48 "
                     "draws one output at a time (particle_filter.cpp:385-387): register the node's pose function with "
                     "bamcl_pf_set_random_pose_source (map with a free-space list), or use the systematic model");
  long retries = 0;
  if (inject && pf->rp_threshold > 0.0 && pf->rp_multiplier < 1.0 && pf->rp_multiplier >= 0.0)
  {
    double good = pf->rp_threshold;  // node.cpp:849-866 with scorePose == 1.0 (init.cu)
    while (1.0 < good)
    {
      good *= pf->rp_multiplier;
      if (++retries > 1000000)
        return set_error(BAMCL_E_UNSUPPORTED, "resample: more than 1e6 de-weighting retries per random pose");
    }
  }
  bamcl_ctx* ctx = pf->ctx;
  SampleSet& set_a = pf->sets[pf->current];
  SampleSet& set_b = pf->sets[1 - pf->current];
  const int P = set_a.sample_count;
  const int n = prm->max_samples;
  if (P <= 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_resample: the current set is empty");
  BAMCL_TRY(pf_build_cdf(pf));
  BAMCL_TRY(pf->src.reserve(sizeof(int) * (size_t)pf->max_samples));
  MultiArgs a;
  a.c = pf->cdf.as<double>();
  a.P = P;
  a.ax = set_a.x.as<double>();
  a.ay = set_a.y.as<double>();
  a.ath = set_a.th.as<double>();
  a.bx = set_b.x.as<double>();
  a.by = set_b.y.as<double>();
  a.bth = set_b.th.as<double>();
  a.bw = set_b.w.as<double>();
  a.src = pf->src.as<int>();
  a.n = n;
  a.state0 = prm->drand48_state & LCG_MASK;
  a.flags = pf_flags(pf);
  a.pos = nullptr;
  a.w_diff = 0.0;
  a.retries = 0;
  a.rp = PoseSource();
  a.host_draws = a.host_poses = nullptr;
  if (host_draws)
  {
    const size_t nd = (size_t)n, np3 = 3 * (size_t)std::max(n_host_poses, 0l);
    BAMCL_TRY(pf->mn_buf.reserve(sizeof(double) * (nd + np3 + 1)));
    double* d = pf->mn_buf.as<double>();
    BAMCL_CUDA_OK(cudaMemcpyAsync(d, host_draws, sizeof(double) * nd, cudaMemcpyHostToDevice, ctx->stream));
    if (np3)
      BAMCL_CUDA_OK(cudaMemcpyAsync(d + nd, host_poses, sizeof(double) * np3, cudaMemcpyHostToDevice, ctx->stream));
    a.host_draws = d;
    a.host_poses = d + nd;
  }
  int* d_pos = nullptr;
  if (inject)
  {
    // stream position of every output (and of "output n" = the draws n outputs consume) by binary lifting
    const long long token = 1 + 2 * (1 + retries);
    const long long worst = (long long)n * token + 2;
    long long n_pos = (long long)std::ceil((double)n * (2.0 + prm->w_diff * (double)(token - 2)) * 1.05) + 4096 + token;
    if (n_pos > worst)
      n_pos = worst;
    for (;;)
    {
      if (n_pos > 0x7ffffff0ll)
        return set_error(BAMCL_E_UNSUPPORTED, "resample: drand48 stream positions exceed 2^31");
      BAMCL_TRY(pf->mn_buf.reserve(sizeof(int) * ((size_t)(n + 1) + 2 * (size_t)(n_pos + 1) + 8)));
      d_pos = pf->mn_buf.as<int>();
      int* level[2] = {d_pos + n + 1, d_pos + n + 1 + n_pos + 1};
      BAMCL_CUDA_OK(cudaMemsetAsync(d_pos, 0, sizeof(int) * ((size_t)n + 1), ctx->stream));
      const unsigned pb = (unsigned)((n_pos + 1 + 255) / 256), kb = (unsigned)((n + 1 + 255) / 256);
      BAMCL_LAUNCH(ctx, k_mn_next, pb, 256, 0, a.state0, prm->w_diff, (int)token, (int)n_pos, level[0]);
      int bits = 0;
      while ((1ll << bits) <= (long long)n)
        bits++;
      for (int j = 0; j < bits; j++)
      {
        BAMCL_LAUNCH(ctx, k_mn_apply, kb, 256, 0, d_pos, level[j & 1], j, n + 1);
        if (j + 1 < bits)
          BAMCL_LAUNCH(ctx, k_mn_square, pb, 256, 0, level[j & 1], level[(j + 1) & 1], (int)n_pos + 1);
      }
      int h_end = 0;
      BAMCL_CUDA_OK(cudaMemcpyAsync(&h_end, d_pos + n, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
      BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
      if (h_end < n_pos || n_pos == worst)
        break;
      n_pos = worst;  // more random poses than the margin allowed for: the worst case always fits
    }
    a.pos = d_pos;
    a.w_diff = prm->w_diff;
    a.retries = (int)retries;
    a.rp = rp_src;
  }
  BAMCL_LAUNCH(ctx, k_select_multinomial, (n + 255) / 256, 256, 0, a);

  // tree of all candidates with every node's first foreign key kept (general build)
  set_b.sample_count = n;
  pf->current = 1 - pf->current;
  pf->force_general_tree = true;
  bamcl_resample_result all;
  memset(&all, 0, sizeof(all));
  const int rc = build_tree_and_stats(pf, pf->current, prm->dist_threshold, prm->convergence_threshold, &all);
  pf->force_general_tree = false;
  if (rc != BAMCL_OK)
    return rc;
  const int K = all.node_count;
  // events on the sample time line, prefix sum, first stopping point
  BAMCL_TRY(pf->scratch.reserve(sizeof(int) * ((size_t)n + 8)));
  BAMCL_TRY(pf->scratch2.reserve(sizeof(int) * (scan_scratch_ints((long)n + 1) + 8)));
  int* ev = pf->scratch.as<int>();
  int* stop = pf->scratch2.as<int>();
  int* scan_scratch = stop + 4;
  const int sentinel = INT_MAX;
  BAMCL_CUDA_OK(cudaMemsetAsync(ev, 0, sizeof(int) * ((size_t)n + 1), ctx->stream));
  BAMCL_CUDA_OK(cudaMemcpyAsync(stop, &sentinel, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
  BAMCL_LAUNCH(ctx, k_leaf_events, (K + 255) / 256, 256, 0, pf->t_nfirst.as<int>(), pf->t_first_at.as<int>(), K, ev);
  BAMCL_TRY(exclusive_scan_i32(ctx, ev, ev, (long)n + 1, scan_scratch, stop + 1));
  BAMCL_LAUNCH(ctx, k_find_stop, (n + 255) / 256, 256, 0, ev, n, prm->min_samples, prm->max_samples, prm->pop_err,
               prm->pop_z, stop);
  int h_stop = 0;
  BAMCL_CUDA_OK(cudaMemcpyAsync(&h_stop, stop, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  const int count = h_stop < n ? h_stop : n;

  // the new set: the first `count` candidates, weights 1/count (:404-405 then :457-462)
  set_b.sample_count = count;
  BAMCL_LAUNCH(ctx, k_fill_f64, (count + 255) / 256, 256, 0, set_b.w.as<double>(), 1.0 / (double)count, count);
  BAMCL_TRY(build_tree_and_stats(pf, pf->current, prm->dist_threshold, prm->convergence_threshold, res));
  int flags = 0;
  BAMCL_CUDA_OK(cudaMemcpyAsync(&flags, pf_flags(pf), sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  if (flags & 2)
    return set_error(BAMCL_E_UNSUPPORTED, "resample: a CDF running sum left the supported range");
  res->sample_count = count;
  res->random_poses_used = 0;
  res->status_flags = flags & 1;
  unsigned long long consumed = host_draws ? 0ull : 2ull * (unsigned long long)count;
  if (host_draws)
  {
    int* d_cnt = stop + 2;
    int h = 0;
    BAMCL_CUDA_OK(cudaMemsetAsync(d_cnt, 0, sizeof(int), ctx->stream));
    BAMCL_LAUNCH(ctx, k_count_random, (count + 255) / 256, 256, 0, pf->src.as<int>(), count, d_cnt);
    BAMCL_CUDA_OK(cudaMemcpyAsync(&h, d_cnt, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    res->random_poses_used = h;
  }
  if (inject)
  {
    int* d_cnt = stop + 2;
    int h[2] = {0, 0};
    BAMCL_CUDA_OK(cudaMemsetAsync(d_cnt, 0, sizeof(int), ctx->stream));
    BAMCL_LAUNCH(ctx, k_count_random, (count + 255) / 256, 256, 0, pf->src.as<int>(), count, d_cnt);
    BAMCL_CUDA_OK(cudaMemcpyAsync(&h[0], d_cnt, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    BAMCL_CUDA_OK(cudaMemcpyAsync(&h[1], d_pos + count, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    res->random_poses_used = h[0];
    consumed = (unsigned long long)h[1];
  }
  res->drand48_state = lcg_skip(a.state0, consumed);
  return BAMCL_OK;
}
}  // namespace bamcl

extern "C" int bamcl_pf_set_random_pose_source(bamcl_pf* pf, bamcl_map2d* map_or_null, double starting_weight_threshold,
                                               double deweight_multiplier)
{
  if (!pf)
    return set_error(BAMCL_E_ARG, "bamcl_pf_set_random_pose_source: null filter");
  pf->rp_map = map_or_null;
  pf->rp_map3d = nullptr;
  pf->rp_threshold = starting_weight_threshold;
  pf->rp_multiplier = deweight_multiplier;
  return BAMCL_OK;
}
