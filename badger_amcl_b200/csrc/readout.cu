// Pose read-out: Node2D::getMaxWeightPose (src/amcl/node/node_2d.cpp:588-617) over the device-resident cluster
// statistics, and the covariance packing of Node::updatePose (src/amcl/node/node.cpp:381-395).
//
// The reference scans the clusters in order and keeps the first one whose weight is strictly greater than the running
// maximum (starting at 0). After a resample all weights are 1/n, so clusters with equal sample counts have bit-equal
// weights there; the device sums cluster weights in a different order (kdtree.cu), so "equal" is taken as within 1e-12
// relative of the maximum and the lowest such cluster index wins -- the reference's choice among equal counts.
#include "pf.cuh"

using namespace bamcl;

namespace
{
constexpr double TIE_REL = 1e-12;

// stats: 10 doubles per cluster (count, weight, mean[3], cov...); out4: weight, mean[3]; out_idx: cluster or -1
__global__ void __launch_bounds__(1024) k_max_weight_cluster(const double* __restrict__ stats, int C, double* __restrict__ out4,
                                                              int* __restrict__ out_idx)
{
  __shared__ double s_w[32];
  __shared__ int s_i[32];
  __shared__ double s_max;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double best = 0.0;
  for (int c = threadIdx.x; c < C; c += blockDim.x)
    best = fmax(best, stats[(size_t)c * 10 + 1]);
  for (int o = 16; o; o >>= 1)
    best = fmax(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0)
    s_w[warp] = best;
  __syncthreads();
  if (threadIdx.x == 0)
  {
    double m = 0.0;
    for (int k = 0; k < (int)(blockDim.x >> 5); k++)
      m = fmax(m, s_w[k]);
    s_max = m;
  }
  __syncthreads();
  const double mx = s_max;
  int first = INT_MAX;
  if (mx > 0.0)
    for (int c = threadIdx.x; c < C; c += blockDim.x)
    {
      const double w = stats[(size_t)c * 10 + 1];
      if (w > 0.0 && w >= mx * (1.0 - TIE_REL))
        first = min(first, c);
    }
  for (int o = 16; o; o >>= 1)
    first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
  if (lane == 0)
    s_i[warp] = first;
  __syncthreads();
  if (threadIdx.x == 0)
  {
    int f = INT_MAX;
    for (int k = 0; k < (int)(blockDim.x >> 5); k++)
      f = min(f, s_i[k]);
    if (f == INT_MAX)
    {
      *out_idx = -1;
      return;
    }
    *out_idx = f;
    const double* r = stats + (size_t)f * 10;
    out4[0] = r[1];
    out4[1] = r[2];
    out4[2] = r[3];
    out4[3] = r[4];
  }
}
}  // namespace

extern "C" int bamcl_pf_max_weight_pose(bamcl_pf* pf, double* max_weight, double* pose3, int* cluster_index_or_null)
{
  if (!pf || !max_weight || !pose3)
    return set_error(BAMCL_E_ARG, "bamcl_pf_max_weight_pose: null argument");
  bamcl_ctx* ctx = pf->ctx;
  SampleSet& set = pf->sets[pf->current];
  if (set.cluster_count <= 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_max_weight_pose: the current set has no cluster statistics (resample or "
                                    "init first)");
  double* d4 = pf->scalars.as<double>() + 40;
  int* d_idx = reinterpret_cast<int*>(d4 + 4);
  BAMCL_LAUNCH(ctx, k_max_weight_cluster, 1, 1024, 0, set.cluster_stats.as<double>(), set.cluster_count, d4, d_idx);
  double h[5];
  BAMCL_CUDA_OK(cudaMemcpyAsync(h, d4, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  int idx;
  memcpy(&idx, &h[4], sizeof(int));
  if (idx < 0)
    // the reference indexes hyps[-1] here (node_2d.cpp:616, undefined behaviour); the drop-in refuses instead
    return set_error(BAMCL_E_STATE, "bamcl_pf_max_weight_pose: no cluster has positive weight");
  *max_weight = h[0];
  pose3[0] = h[1];
  pose3[1] = h[2];
  pose3[2] = h[3];
  if (cluster_index_or_null)
    *cluster_index_or_null = idx;
  return BAMCL_OK;
}

extern "C" int bamcl_pose_covariance(const bamcl_resample_result* result, double* cov36)
{
  if (!result || !cov36)
    return set_error(BAMCL_E_ARG, "bamcl_pose_covariance: null argument");
  for (int i = 0; i < 36; i++)
    cov36[i] = 0.0;
  // node.cpp:383-395: the filter's overall covariance; result->cov = (0,0),(0,1),(1,0),(1,1),(2,2)
  cov36[6 * 0 + 0] = result->cov[0];
  cov36[6 * 0 + 1] = result->cov[1];
  cov36[6 * 1 + 0] = result->cov[2];
  cov36[6 * 1 + 1] = result->cov[3];
  cov36[6 * 5 + 5] = result->cov[4];
  return BAMCL_OK;
}
