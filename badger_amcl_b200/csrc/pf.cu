// ParticleFilter device object: sample-set storage, host<->device sample exchange, weight total and
// normalisation, and the update entry points. Reference: src/amcl/pf/particle_filter.cpp:38-98, 223-267.
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include "pf.cuh"

using namespace bamcl;

namespace bamcl
{
int stage_scan(bamcl_pf* pf, bamcl_map2d* map, const bamcl_planar_model* model, const double* ranges,
               const double* angles, int range_count, double range_max);
int launch_score(bamcl_pf* pf, bamcl_map2d* map);
}  // namespace bamcl

namespace
{
__global__ void k_aos_to_soa(const double4* __restrict__ aos, double* __restrict__ x, double* __restrict__ y,
                             double* __restrict__ th, double* __restrict__ w, int n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  const double4 s = aos[i];
  x[i] = s.x;
  y[i] = s.y;
  th[i] = s.z;
  w[i] = s.w;
}
__global__ void k_soa_to_aos(double4* __restrict__ aos, const double* __restrict__ x, const double* __restrict__ y,
                             const double* __restrict__ th, const double* __restrict__ w, int n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  aos[i] = make_double4(x[i], y[i], th[i], w[i]);
}
__global__ void k_fill(double* __restrict__ p, double v, int n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    p[i] = v;
}

// ---- scoring order (pf_spatial_order): counting sort of the window by coarse-cell Morton key, 64 x 64 cells
constexpr int ORDER_BITS = 6, ORDER_BINS = 1 << (2 * ORDER_BITS);
__device__ __forceinline__ unsigned order_key(double x, double y, double x0, double y0, double inv_cell)
{
  const int cx = min(max((int)floor((x - x0) * inv_cell), 0), (1 << ORDER_BITS) - 1);
  const int cy = min(max((int)floor((y - y0) * inv_cell), 0), (1 << ORDER_BITS) - 1);
  unsigned k = 0;
#pragma unroll
  for (int b = 0; b < ORDER_BITS; b++)
    k |= (((unsigned)cx >> b) & 1u) << (2 * b) | (((unsigned)cy >> b) & 1u) << (2 * b + 1);
  return k;
}
__global__ void __launch_bounds__(256) k_order_count(const double* __restrict__ x, const double* __restrict__ y, int n, double x0,
                                                     double y0, double inv_cell, int* __restrict__ hist)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    atomicAdd(&hist[order_key(x[i], y[i], x0, y0, inv_cell)], 1);
}
__global__ void __launch_bounds__(1024) k_order_scan(int* __restrict__ hist)  // exclusive scan of ORDER_BINS counts, one block
{
  __shared__ int s[1024];
  constexpr int PER = ORDER_BINS / 1024;
  int v[PER], sum = 0;
#pragma unroll
  for (int q = 0; q < PER; q++)
  {
    v[q] = hist[threadIdx.x * PER + q];
    sum += v[q];
  }
  s[threadIdx.x] = sum;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1)
  {
    const int t = threadIdx.x >= o ? s[threadIdx.x - o] : 0;
    __syncthreads();
    s[threadIdx.x] += t;
    __syncthreads();
  }
  int run = s[threadIdx.x] - sum;
#pragma unroll
  for (int q = 0; q < PER; q++)
  {
    hist[threadIdx.x * PER + q] = run;
    run += v[q];
  }
}
__global__ void __launch_bounds__(256) k_order_scatter(const double* __restrict__ x, const double* __restrict__ y, int n, double x0,
                                                       double y0, double inv_cell, int* __restrict__ cursor, int* __restrict__ perm)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    perm[atomicAdd(&cursor[order_key(x[i], y[i], x0, y0, inv_cell)], 1)] = i;
}

// ---- deterministic weight total: fixed-shape two-level tree (block partials, then one block).
constexpr int RED_BLOCK = 256;
constexpr int RED_ITEMS = 8;
__global__ void __launch_bounds__(RED_BLOCK) k_sum_partials(const double* __restrict__ w, int n, double* __restrict__ partials)
{
  __shared__ double s[RED_BLOCK / 32];
  const int base = blockIdx.x * RED_BLOCK * RED_ITEMS;
  double v = 0.0;
#pragma unroll
  for (int t = 0; t < RED_ITEMS; t++)
  {
    const int i = base + t * RED_BLOCK + threadIdx.x;
    if (i < n)
      v += w[i];
  }
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0)
    s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32)
  {
    double t = threadIdx.x < RED_BLOCK / 32 ? s[threadIdx.x] : 0.0;
    t = warp_sum(t);
    if (threadIdx.x == 0)
      partials[blockIdx.x] = t;
  }
}
__global__ void __launch_bounds__(1024) k_sum_final(const double* __restrict__ partials, int m, double* __restrict__ total)
{
  __shared__ double s[32];
  double v = 0.0;
  for (int i = threadIdx.x; i < m; i += 1024)
    v += partials[i];
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0)
    s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32)
  {
    double t = s[threadIdx.x];
    t = warp_sum(t);
    if (threadIdx.x == 0)
      *total = t;
  }
}

// ---- ParticleFilter::updateSensor, particle_filter.cpp:237-266
__global__ void k_normalise(double* __restrict__ w, int n, const double* __restrict__ total_p, int n_set)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  const double total = *total_p;
  if (total > 0.0)
    w[i] = w[i] / total;
  else
    w[i] = 1.0 / n_set;  // the whole set's count (a sharded filter normalises only its window)
}
// sharded update: the ranks' window totals (rank order = sample order of the partial sums) -> global total -> normalise
__global__ void k_normalise_ranks(double* __restrict__ w, int n, const double* __restrict__ totals, int n_ranks, int n_set,
                                  double* __restrict__ total_out)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double total = 0.0;
  for (int r = 0; r < n_ranks; r++)
    total += totals[r];
  if (i == 0)
    *total_out = total;
  if (i >= n)
    return;
  if (total > 0.0)
    w[i] = w[i] / total;
  else
    w[i] = 1.0 / n_set;
}
}  // namespace

namespace
{
struct Marks
{
  std::vector<std::pair<std::string, double>> rows;
  std::chrono::steady_clock::time_point last;
  bool started = false;
};
thread_local Marks g_marks;
bool marks_on()
{
  static const bool on = getenv("BAMCL_SHARD_TIMING") != nullptr;
  return on;
}
}  // namespace

namespace bamcl
{
void pf_mark(bamcl_pf* pf, const char* phase)
{
  if (!marks_on())
    return;
  cudaStreamSynchronize(pf->ctx->stream);
  const auto now = std::chrono::steady_clock::now();
  if (g_marks.started && phase)
    g_marks.rows.emplace_back(phase, std::chrono::duration<double, std::milli>(now - g_marks.last).count());
  g_marks.started = true;
  g_marks.last = std::chrono::steady_clock::now();
}
void pf_mark_report(bamcl_pf* pf)
{
  if (!marks_on())
    return;
  if (pf_rank(pf) == 0)
  {
    std::string line = "[shard timing ms]";
    char buf[96];
    for (auto& r : g_marks.rows)
    {
      snprintf(buf, sizeof(buf), " %s %.3f", r.first.c_str(), r.second);
      line += buf;
    }
    fprintf(stderr, "%s\n", line.c_str());
  }
  g_marks.rows.clear();
  g_marks.started = false;
}

int pf_ensure_streams(bamcl_pf* pf)
{
  if (pf->aux_stream)
    return BAMCL_OK;
  BAMCL_CUDA_OK(cudaStreamCreateWithPriority(&pf->aux_stream, cudaStreamNonBlocking, pf->ctx->stream_priority));
  BAMCL_CUDA_OK(cudaStreamCreateWithPriority(&pf->aux_stream_lo, cudaStreamNonBlocking, pf->ctx->low_priority));
  BAMCL_CUDA_OK(cudaEventCreateWithFlags(&pf->ev_fork, cudaEventDisableTiming));
  BAMCL_CUDA_OK(cudaEventCreateWithFlags(&pf->ev_join, cudaEventDisableTiming));
  BAMCL_CUDA_OK(cudaEventCreateWithFlags(&pf->ev_bucket_fork, cudaEventDisableTiming));
  for (int i = 0; i < 2; i++)
  {
    BAMCL_CUDA_OK(cudaStreamCreateWithPriority(&pf->bucket_stream[i], cudaStreamNonBlocking, pf->ctx->stream_priority));
    BAMCL_CUDA_OK(cudaEventCreateWithFlags(&pf->ev_bucket_join[i], cudaEventDisableTiming));
  }
  BAMCL_CUDA_OK(cudaStreamCreateWithFlags(&pf->copy_stream, cudaStreamNonBlocking));
  for (int i = 0; i < 8; i++)
    BAMCL_CUDA_OK(cudaEventCreateWithFlags(&pf->ev_copy[i], cudaEventDisableTiming));
  return BAMCL_OK;
}

// current set -> host AoS, enqueued on the copy stream behind everything the main stream holds now; the caller
// synchronises the copy stream before it lets the host read `aos`
int pf_download_async(bamcl_pf* pf, double* aos, long n)
{
  BAMCL_TRY(pf_ensure_streams(pf));
  SampleSet& s = pf->sets[pf->current];
  BAMCL_TRY(pf->d_stage.reserve(sizeof(double) * 4 * (size_t)n));
  BAMCL_CUDA_OK(cudaEventRecord(pf->ev_copy[0], pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamWaitEvent(pf->copy_stream, pf->ev_copy[0], 0));
  k_soa_to_aos<<<(unsigned)((n + 255) / 256), 256, 0, pf->copy_stream>>>(pf->d_stage.as<double4>(), s.x.as<double>(),
                                                                        s.y.as<double>(), s.th.as<double>(),
                                                                        s.w.as<double>(), (int)n);
  pf->ctx->launches++;
  BAMCL_CUDA_OK(cudaGetLastError());
  BAMCL_CUDA_OK(cudaMemcpyAsync(aos, pf->d_stage.p, sizeof(double) * 4 * (size_t)n, cudaMemcpyDeviceToHost, pf->copy_stream));
  return BAMCL_OK;
}

int pf_spatial_order(bamcl_pf* pf, double x0, double y0, double cell, const int** perm_out)
{
  SampleSet& set = pf->sets[pf->current];
  const int n = pf_win_cnt(pf);
  const long off = pf_win_off(pf);
  BAMCL_TRY(pf->order_perm.reserve(sizeof(int) * (size_t)std::max(n, 1)));
  BAMCL_TRY(pf->order_hist.reserve(sizeof(int) * ORDER_BINS));
  BAMCL_CUDA_OK(cudaMemsetAsync(pf->order_hist.p, 0, sizeof(int) * ORDER_BINS, pf->ctx->stream));
  if (n > 0)
  {
    const unsigned blocks = (unsigned)((n + 255) / 256);
    const double inv = 1.0 / cell;
    BAMCL_LAUNCH(pf->ctx, k_order_count, blocks, 256, 0, set.x.as<double>() + off, set.y.as<double>() + off, n, x0, y0, inv,
                 pf->order_hist.as<int>());
    BAMCL_LAUNCH(pf->ctx, k_order_scan, 1, 1024, 0, pf->order_hist.as<int>());
    BAMCL_LAUNCH(pf->ctx, k_order_scatter, blocks, 256, 0, set.x.as<double>() + off, set.y.as<double>() + off, n, x0, y0, inv,
                 pf->order_hist.as<int>(), pf->order_perm.as<int>());
  }
  *perm_out = pf->order_perm.as<int>();
  return BAMCL_OK;
}

int pf_reduce_total(bamcl_pf* pf, double* d_total)
{
  SampleSet& set = pf->sets[pf->current];
  const int n = pf_win_cnt(pf);
  const int blocks = (n + RED_BLOCK * RED_ITEMS - 1) / (RED_BLOCK * RED_ITEMS);
  BAMCL_TRY(pf->partials.reserve(sizeof(double) * (size_t)(blocks + 1)));
  if (blocks > 0)
    BAMCL_LAUNCH(pf->ctx, k_sum_partials, blocks, RED_BLOCK, 0, set.w.as<double>() + pf_win_off(pf), n,
                 pf->partials.as<double>());
  BAMCL_LAUNCH(pf->ctx, k_sum_final, 1, 1024, 0, pf->partials.as<double>(), blocks, d_total);
  return BAMCL_OK;
}
}  // namespace bamcl

namespace bamcl
{
int pf_normalise_by_device_total(bamcl_pf* pf)
{
  SampleSet& s = pf->sets[pf->current];
  if (pf_win_cnt(pf) > 0)
    BAMCL_LAUNCH(pf->ctx, k_normalise, (pf_win_cnt(pf) + 255) / 256, 256, 0, s.w.as<double>() + pf_win_off(pf),
                 pf_win_cnt(pf), pf->scalars.as<double>(), s.sample_count);
  return BAMCL_OK;
}
int pf_read_scalar(bamcl_pf* pf, int slot, double* out);
}  // namespace bamcl

static int read_scalar(bamcl_pf* pf, int slot, double* out) { return bamcl::pf_read_scalar(pf, slot, out); }

int bamcl::pf_read_scalar(bamcl_pf* pf, int slot, double* out)
{
  BAMCL_CUDA_OK(cudaMemcpyAsync(pf->h_scalars.as<double>() + slot, pf->scalars.as<double>() + slot, sizeof(double),
                                cudaMemcpyDeviceToHost, pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  *out = pf->h_scalars.as<double>()[slot];
  return BAMCL_OK;
}

namespace bamcl
{
int pf_shard_finish_update(bamcl_pf* pf, double* total_out);
}

// Distributed update, after the scoring kernel ran over this rank's window and its total sits in scalars[0]: the ranks'
// window totals are all-gathered and summed in rank order (= the sample order in which ParticleFilter::updateSensor adds
// up, particle_filter.cpp:246-252), every rank normalises its window by that global total. One host read at the end.
int bamcl::pf_shard_finish_update(bamcl_pf* pf, double* total_out)
{
  const int G = pf_ranks(pf), me = pf_rank(pf);
  if (G > 384)
    return set_error(BAMCL_E_UNSUPPORTED, "more than 384 ranks");
  SampleSet& s = pf->sets[pf->current];
  double* sc = pf->scalars.as<double>();
  double* totals = sc + 64;
  BAMCL_CUDA_OK(cudaMemcpyAsync(totals + me, sc, sizeof(double), cudaMemcpyDeviceToDevice, pf->ctx->stream));
  BAMCL_TRY(comm_all_gather(pf->comm, totals + me, totals, sizeof(double)));
  const int cnt = s.sample_count == 0 ? 0 : pf_win_cnt(pf);
  BAMCL_LAUNCH(pf->ctx, k_normalise_ranks, (std::max(cnt, 1) + 255) / 256, 256, 0, s.w.as<double>() + pf_win_off(pf), cnt, totals,
               G, s.sample_count, sc + 1);
  if (total_out)
    return read_scalar(pf, 1, total_out);
  return BAMCL_OK;
}

extern "C" {

int bamcl_pf_create(bamcl_ctx* ctx, int min_samples, int max_samples, bamcl_pf** out)
{
  if (!ctx || !out || max_samples <= 0 || min_samples < 0)
    return set_error(BAMCL_E_ARG, "bamcl_pf_create: bad arguments");
  bamcl_pf* pf = new bamcl_pf();
  pf->ctx = ctx;
  pf->min_samples = min_samples;
  pf->max_samples = max_samples;
  const size_t bytes = sizeof(double) * (size_t)max_samples;
  for (int j = 0; j < 2; j++)
  {
    SampleSet& s = pf->sets[j];
    BAMCL_TRY(s.x.reserve(bytes));
    BAMCL_TRY(s.y.reserve(bytes));
    BAMCL_TRY(s.th.reserve(bytes));
    BAMCL_TRY(s.w.reserve(bytes));
    BAMCL_TRY(s.label.reserve(sizeof(int) * (size_t)max_samples));
    // particle_filter.cpp:69-79: max_samples samples at the origin with weight 1/max_samples
    s.sample_count = max_samples;
    BAMCL_CUDA_OK(cudaMemsetAsync(s.x.p, 0, bytes, ctx->stream));
    BAMCL_CUDA_OK(cudaMemsetAsync(s.y.p, 0, bytes, ctx->stream));
    BAMCL_CUDA_OK(cudaMemsetAsync(s.th.p, 0, bytes, ctx->stream));
    BAMCL_LAUNCH(ctx, k_fill, (max_samples + 255) / 256, 256, 0, s.w.as<double>(), 1.0 / max_samples, max_samples);
  }
  BAMCL_CUDA_OK(cudaEventCreate(&pf->score_ev0));
  BAMCL_CUDA_OK(cudaEventCreate(&pf->score_ev1));
  BAMCL_TRY(pf->scalars.reserve(sizeof(double) * 512));
  BAMCL_TRY(pf->h_scalars.reserve(sizeof(double) * 512));
  BAMCL_CUDA_OK(cudaMemsetAsync(pf->scalars.p, 0, sizeof(double) * 512, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  *out = pf;
  return BAMCL_OK;
}

int bamcl_pf_destroy(bamcl_pf* pf)
{
  if (!pf)
    return BAMCL_OK;
  cudaStreamSynchronize(pf->ctx->stream);
  for (int j = 0; j < 2; j++)
  {
    SampleSet& s = pf->sets[j];
    s.x.release();
    s.y.release();
    s.th.release();
    s.w.release();
    s.label.release();
    s.cluster_stats.release();
  }
  pf->scan.beams.release();
  pf->scan.table.release();
  pf->scan.table_q.release();
  pf->scan.ranges.release();
  pf->scan.angles.release();
  for (bamcl::DevBuf* b : {&pf->cloud.xyz, &pf->cloud.px, &pf->cloud.py, &pf->cloud.pz, &pf->cloud.table})
    b->release();
  pf->scalars.release();
  pf->h_scalars.release();
  pf->h_stage.release();
  pf->d_stage.release();
  pf->partials.release();
  pf->cdf.release();
  pf->src.release();
  pf->scratch.release();
  pf->scratch2.release();
  pf->random_poses.release();
  pf->seg_buf.release();
  pf->mn_buf.release();
  for (bamcl::DevBuf* b : {&pf->t_hkeys, &pf->t_hfirst, &pf->t_hnode, &pf->t_slot, &pf->t_flag, &pf->t_nkey, &pf->t_cur,
                           &pf->t_child, &pf->t_first_at, &pf->t_side, &pf->t_parent, &pf->t_cid, &pf->t_nfirst, &pf->t_root, &pf->t_scan,
                           &pf->t_out, &pf->t_sums, &pf->t_fast})
    b->release();
  for (bamcl::DevBuf* b : {&pf->g_first, &pf->g_node, &pf->t_bits, &pf->t_bpre, &pf->t_setpart, &pf->t_ints, &pf->stage_x,
                           &pf->stage_y, &pf->scan.agree, &pf->scan.skip_count, &pf->scan.beams_sel, &pf->order_perm, &pf->order_hist})
    b->release();
  pf->t_hout.release();
  if (pf->aux_stream)
    cudaStreamDestroy(pf->aux_stream);
  if (pf->aux_stream_lo)
    cudaStreamDestroy(pf->aux_stream_lo);
  if (pf->ev_fork)
    cudaEventDestroy(pf->ev_fork);
  if (pf->ev_join)
    cudaEventDestroy(pf->ev_join);
  for (int i = 0; i < 2; i++)
  {
    if (pf->bucket_stream[i])
      cudaStreamDestroy(pf->bucket_stream[i]);
    if (pf->ev_bucket_join[i])
      cudaEventDestroy(pf->ev_bucket_join[i]);
  }
  if (pf->ev_bucket_fork)
    cudaEventDestroy(pf->ev_bucket_fork);
  if (pf->copy_stream)
    cudaStreamDestroy(pf->copy_stream);
  for (int i = 0; i < 8; i++)
    if (pf->ev_copy[i])
      cudaEventDestroy(pf->ev_copy[i]);
  if (pf->score_ev0)
    cudaEventDestroy(pf->score_ev0);
  if (pf->score_ev1)
    cudaEventDestroy(pf->score_ev1);
  delete pf;
  return BAMCL_OK;
}

int bamcl_pf_upload_samples(bamcl_pf* pf, const double* aos, long n)
{
  if (!pf || !aos || n < 0 || n > pf->max_samples)
    return set_error(BAMCL_E_ARG, "bamcl_pf_upload_samples: need 0 <= n <= max_samples");
  SampleSet& s = pf->sets[pf->current];
  if (pf_ranks(pf) > 1)
    return set_error(BAMCL_E_STATE, "bamcl_pf_upload_samples: a distributed filter uploads its window "
                                    "(bamcl_pf_shard_set_count + bamcl_pf_upload_window_samples)");
  s.sample_count = (int)n;
  pf->win_cnt = -1;  // a new set: back to the whole-set window
  if (n == 0)
    return BAMCL_OK;
  bamcl_ctx* ctx = pf->ctx;
  BAMCL_TRY(pf->d_stage.reserve(sizeof(double) * 4 * (size_t)n));
  BAMCL_CUDA_OK(cudaMemcpyAsync(pf->d_stage.p, aos, sizeof(double) * 4 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
  BAMCL_LAUNCH(ctx, k_aos_to_soa, (unsigned)((n + 255) / 256), 256, 0, pf->d_stage.as<double4>(), s.x.as<double>(),
               s.y.as<double>(), s.th.as<double>(), s.w.as<double>(), (int)n);
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  return BAMCL_OK;
}

int bamcl_pf_upload_window_samples(bamcl_pf* pf, const double* aos, long offset, long n)
{
  if (!pf || !aos || offset < 0 || n < 0 || offset + n > pf->sets[pf->current].sample_count)
    return set_error(BAMCL_E_ARG, "bamcl_pf_upload_window_samples: [offset, offset + n) must lie inside the current set");
  if (n == 0)
    return BAMCL_OK;
  SampleSet& s = pf->sets[pf->current];
  bamcl_ctx* ctx = pf->ctx;
  BAMCL_TRY(pf->d_stage.reserve(sizeof(double) * 4 * (size_t)n));
  BAMCL_CUDA_OK(cudaMemcpyAsync(pf->d_stage.p, aos, sizeof(double) * 4 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
  BAMCL_LAUNCH(ctx, k_aos_to_soa, (unsigned)((n + 255) / 256), 256, 0, pf->d_stage.as<double4>(), s.x.as<double>() + offset,
               s.y.as<double>() + offset, s.th.as<double>() + offset, s.w.as<double>() + offset, (int)n);
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  return BAMCL_OK;
}

int bamcl_pf_download_window_samples(bamcl_pf* pf, double* aos, long offset, long n)
{
  if (!pf || !aos || offset < 0 || n < 0 || offset + n > pf->sets[pf->current].sample_count)
    return set_error(BAMCL_E_ARG, "bamcl_pf_download_window_samples: [offset, offset + n) must lie inside the current set");
  if (n == 0)
    return BAMCL_OK;
  SampleSet& s = pf->sets[pf->current];
  bamcl_ctx* ctx = pf->ctx;
  BAMCL_TRY(pf->d_stage.reserve(sizeof(double) * 4 * (size_t)n));
  BAMCL_LAUNCH(ctx, k_soa_to_aos, (unsigned)((n + 255) / 256), 256, 0, pf->d_stage.as<double4>(), s.x.as<double>() + offset,
               s.y.as<double>() + offset, s.th.as<double>() + offset, s.w.as<double>() + offset, (int)n);
  BAMCL_CUDA_OK(cudaMemcpyAsync(aos, pf->d_stage.p, sizeof(double) * 4 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  return BAMCL_OK;
}

int bamcl_pf_download_samples(bamcl_pf* pf, double* aos, long n)
{
  if (!pf || !aos)
    return set_error(BAMCL_E_ARG, "bamcl_pf_download_samples: bad arguments");
  SampleSet& s = pf->sets[pf->current];
  if (n != s.sample_count)
    return set_error(BAMCL_E_ARG, "bamcl_pf_download_samples: n must equal sample_count (%d)", s.sample_count);
  if (n == 0)
    return BAMCL_OK;
  bamcl_ctx* ctx = pf->ctx;
  BAMCL_TRY(pf->d_stage.reserve(sizeof(double) * 4 * (size_t)n));
  BAMCL_LAUNCH(ctx, k_soa_to_aos, (unsigned)((n + 255) / 256), 256, 0, pf->d_stage.as<double4>(), s.x.as<double>(),
               s.y.as<double>(), s.th.as<double>(), s.w.as<double>(), (int)n);
  BAMCL_CUDA_OK(cudaMemcpyAsync(aos, pf->d_stage.p, sizeof(double) * 4 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  return BAMCL_OK;
}

int bamcl_pf_sample_count(bamcl_pf* pf, int* n)
{
  if (!pf || !n)
    return set_error(BAMCL_E_ARG, "bamcl_pf_sample_count: bad arguments");
  *n = pf->sets[pf->current].sample_count;
  return BAMCL_OK;
}

int bamcl_pf_upload_weights(bamcl_pf* pf, const double* w, long n)
{
  long off = 0, cnt = 0;
  if (pf)
    pf_owned(pf, &off, &cnt);
  if (!pf || !w || n != cnt)
    return set_error(BAMCL_E_ARG, "bamcl_pf_upload_weights: n must equal sample_count (a distributed filter: its window)");
  if (n == 0)
    return BAMCL_OK;
  BAMCL_CUDA_OK(cudaMemcpyAsync(pf->sets[pf->current].w.as<double>() + off, w, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice,
                                pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  return BAMCL_OK;
}

int bamcl_pf_download_weights(bamcl_pf* pf, double* w, long n)
{
  long off = 0, cnt = 0;
  if (pf)
    pf_owned(pf, &off, &cnt);
  if (!pf || !w || n != cnt)
    return set_error(BAMCL_E_ARG, "bamcl_pf_download_weights: n must equal sample_count (a distributed filter: its window)");
  if (n == 0)
    return BAMCL_OK;
  BAMCL_CUDA_OK(cudaMemcpyAsync(w, pf->sets[pf->current].w.as<double>() + off, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost,
                                pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  return BAMCL_OK;
}

int bamcl_pf_stage_scan(bamcl_pf* pf, bamcl_map2d* map, const bamcl_planar_model* model, const double* ranges,
                        const double* angles, int range_count, double range_max)
{
  return stage_scan(pf, map, model, ranges, angles, range_count, range_max);
}

int bamcl_pf_score_planar(bamcl_pf* pf, bamcl_map2d* map, const bamcl_planar_model* model, const double* ranges,
                          const double* angles, int range_count, double range_max, int converged, double* total_out)
{
  if (pf)
    pf->set_converged = converged;
  if (!total_out)
    return set_error(BAMCL_E_ARG, "bamcl_pf_score_planar: total_out is required");
  BAMCL_TRY(stage_scan(pf, map, model, ranges, angles, range_count, range_max));
  if (model->max_beams < 2)  // planar_scanner.cpp:144
  {
    *total_out = 0.0;
    return BAMCL_OK;
  }
  BAMCL_TRY(launch_score(pf, map));
  BAMCL_TRY(pf_reduce_total(pf, pf->scalars.as<double>()));
  return read_scalar(pf, 0, total_out);
}

int bamcl_pf_set_exact_terms(bamcl_pf* pf, int exact)
{
  if (!pf)
    return set_error(BAMCL_E_ARG, "null pf");
  pf->exact_terms = exact != 0;
  return BAMCL_OK;
}

int bamcl_pf_last_score_ms(bamcl_pf* pf, float* ms)
{
  if (!pf || !ms)
    return set_error(BAMCL_E_ARG, "bamcl_pf_last_score_ms: bad arguments");
  if (!pf->score_timed)
    return set_error(BAMCL_E_STATE, "no scoring kernel has been launched yet");
  BAMCL_CUDA_OK(cudaEventSynchronize(pf->score_ev1));
  BAMCL_CUDA_OK(cudaEventElapsedTime(ms, pf->score_ev0, pf->score_ev1));
  return BAMCL_OK;
}

int bamcl_pf_normalise(bamcl_pf* pf, double total)
{
  if (!pf)
    return set_error(BAMCL_E_ARG, "null pf");
  SampleSet& s = pf->sets[pf->current];
  if (s.sample_count == 0 || pf_win_cnt(pf) == 0)
    return BAMCL_OK;
  pf->h_scalars.as<double>()[1] = total;
  BAMCL_CUDA_OK(cudaMemcpyAsync(pf->scalars.as<double>() + 1, pf->h_scalars.as<double>() + 1, sizeof(double),
                                cudaMemcpyHostToDevice, pf->ctx->stream));
  BAMCL_LAUNCH(pf->ctx, k_normalise, (pf_win_cnt(pf) + 255) / 256, 256, 0, s.w.as<double>() + pf_win_off(pf),
               pf_win_cnt(pf), pf->scalars.as<double>() + 1, s.sample_count);
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  return BAMCL_OK;
}

int bamcl_pf_update_planar_staged(bamcl_pf* pf, bamcl_map2d* map, int converged, double* total_out_or_null)
{
  if (pf)
    pf->set_converged = converged;
  if (!pf || !map)
    return set_error(BAMCL_E_ARG, "bamcl_pf_update_planar_staged: bad arguments");
  SampleSet& s = pf->sets[pf->current];
  if (pf->scan.valid && pf->scan.model.max_beams < 2)  // updateSensor returns false, nothing happens (:128)
  {
    if (total_out_or_null)
      *total_out_or_null = 0.0;
    return BAMCL_OK;
  }
  BAMCL_TRY(launch_score(pf, map));
  BAMCL_TRY(pf_reduce_total(pf, pf->scalars.as<double>()));
  if (pf_win_cnt(pf) > 0)
    BAMCL_LAUNCH(pf->ctx, k_normalise, (pf_win_cnt(pf) + 255) / 256, 256, 0, s.w.as<double>() + pf_win_off(pf),
                 pf_win_cnt(pf), pf->scalars.as<double>(), s.sample_count);
  if (total_out_or_null)
    return read_scalar(pf, 0, total_out_or_null);
  return BAMCL_OK;
}

int bamcl_pf_score_planar_staged(bamcl_pf* pf, bamcl_map2d* map, int converged, double* total_out)
{
  if (pf)
    pf->set_converged = converged;
  if (!pf || !map || !total_out)
    return set_error(BAMCL_E_ARG, "bamcl_pf_score_planar_staged: bad arguments");
  if (pf->scan.valid && pf->scan.model.max_beams < 2)
  {
    *total_out = 0.0;
    return BAMCL_OK;
  }
  BAMCL_TRY(launch_score(pf, map));
  BAMCL_TRY(pf_reduce_total(pf, pf->scalars.as<double>()));
  return read_scalar(pf, 0, total_out);
}

int bamcl_pf_shard_update_planar_staged(bamcl_pf* pf, bamcl_map2d* map, int converged, double* total_out)
{
  if (!pf || !map)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_update_planar_staged: bad arguments");
  if (!pf->comm)
    return set_error(BAMCL_E_STATE, "bamcl_pf_shard_update_planar_staged: no communicator attached");
  pf->set_converged = converged;
  pf_apply_shard_window(pf);
  if (pf->scan.valid && pf->scan.model.max_beams < 2)  // updateSensor returns false on every rank, nothing happens (:128)
  {
    if (total_out)
      *total_out = 0.0;
    return BAMCL_OK;
  }
  BAMCL_TRY(launch_score(pf, map));
  BAMCL_TRY(pf_reduce_total(pf, pf->scalars.as<double>()));
  return pf_shard_finish_update(pf, total_out);
}

int bamcl_pf_update_planar_host(bamcl_pf* pf, bamcl_map2d* map, const bamcl_planar_model* model, const double* ranges,
                                const double* angles, int range_count, double range_max, int converged, const double* aos,
                                long n, double* total_out)
{
  if (!pf || !aos || n < 0 || n > pf->max_samples)
    return set_error(BAMCL_E_ARG, "bamcl_pf_update_planar_host: need 0 <= n <= max_samples");
  const bool skip = model && model->model_type == BAMCL_PLANAR_LF_PROB && model->do_beamskip && converged;
  // chunks of the upload pipeline (1..7; BAMCL_HOST_CHUNKS overrides). Measured at 1M x 1080 (profiles/r02_e2e_chunks.txt):
  // update half 1.96 / 1.71 / 1.74 / 1.91 / 1.82 ms with 1 / 2 / 4 / 6 / 7 chunks - every chunk is one more launch of the
  // scoring kernel with its table prologue. Queueing the chunk copies BEFORE the scan is staged was tried and is slower
  // (2.06 ms): the scan's small copies then wait behind the chunks on the one host-to-device copy engine.
  static const int CHUNKS = [] {
    const char* e = getenv("BAMCL_HOST_CHUNKS");
    const int v = e ? atoi(e) : 4;
    return v < 1 ? 1 : (v > 7 ? 7 : v);
  }();
  if (n < 4 * 65536 || skip || (model && model->model_type == BAMCL_PLANAR_BEAM) || (model && model->max_beams < 2))
  {
    // small sets, and models that look at the whole set before weighting it: upload, then update
    BAMCL_TRY(bamcl_pf_upload_samples(pf, aos, n));
    return bamcl_pf_update_planar(pf, map, model, ranges, angles, range_count, range_max, converged, total_out);
  }
  BAMCL_TRY(stage_scan(pf, map, model, ranges, angles, range_count, range_max));
  BAMCL_TRY(pf_ensure_streams(pf));
  pf->set_converged = converged;
  bamcl_ctx* ctx = pf->ctx;
  SampleSet& s = pf->sets[pf->current];
  s.sample_count = (int)n;
  BAMCL_TRY(pf->d_stage.reserve(sizeof(double) * 4 * (size_t)n));
  // the copy stream may only overwrite the staging buffer after the main stream's earlier readers are done
  BAMCL_CUDA_OK(cudaEventRecord(pf->ev_copy[CHUNKS], ctx->stream));
  BAMCL_CUDA_OK(cudaStreamWaitEvent(pf->copy_stream, pf->ev_copy[CHUNKS], 0));
  int rc = BAMCL_OK;
  struct WindowReset  // whatever path leaves this function, the filter is back to its whole-set window
  {
    bamcl_pf* p;
    ~WindowReset()
    {
      p->win_off = 0;
      p->win_cnt = -1;
    }
  } window_reset{pf};
  for (int c = 0; c < CHUNKS && rc == BAMCL_OK; c++)
  {
    const long off = n * c / CHUNKS, cnt = n * (c + 1) / CHUNKS - off;
    double* d = pf->d_stage.as<double>() + 4 * off;
    BAMCL_CUDA_OK(cudaMemcpyAsync(d, aos + 4 * off, sizeof(double) * 4 * (size_t)cnt, cudaMemcpyHostToDevice, pf->copy_stream));
    BAMCL_CUDA_OK(cudaEventRecord(pf->ev_copy[c], pf->copy_stream));
    BAMCL_CUDA_OK(cudaStreamWaitEvent(ctx->stream, pf->ev_copy[c], 0));
    BAMCL_LAUNCH(ctx, k_aos_to_soa, (unsigned)((cnt + 255) / 256), 256, 0, reinterpret_cast<const double4*>(d),
                 s.x.as<double>() + off, s.y.as<double>() + off, s.th.as<double>() + off, s.w.as<double>() + off, (int)cnt);
    pf->win_off = off;  // score this chunk while the next one is on the bus
    pf->win_cnt = (int)cnt;
    rc = launch_score(pf, map);
  }
  pf->win_off = 0;
  pf->win_cnt = -1;
  if (rc != BAMCL_OK)
    return rc;
  BAMCL_TRY(pf_reduce_total(pf, pf->scalars.as<double>()));
  BAMCL_LAUNCH(ctx, k_normalise, (unsigned)((n + 255) / 256), 256, 0, s.w.as<double>(), (int)n, pf->scalars.as<double>(),
               s.sample_count);
  if (total_out)
    return read_scalar(pf, 0, total_out);
  return BAMCL_OK;
}

int bamcl_pf_update_planar(bamcl_pf* pf, bamcl_map2d* map, const bamcl_planar_model* model, const double* ranges,
                           const double* angles, int range_count, double range_max, int converged,
                           double* total_out)
{
  BAMCL_TRY(stage_scan(pf, map, model, ranges, angles, range_count, range_max));
  return bamcl_pf_update_planar_staged(pf, map, converged, total_out);
}

}  // extern "C"
