// bamcl_weights_normalise_noalloc: the weight total + normalisation of ParticleFilter::updateSensor for callers that
// must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). It has no
// counterpart in the reference.
//
// The arithmetic is the one of pf.cu (k_sum_partials -> k_sum_final -> k_normalise: a fixed-shape two-level tree, blocks of
// 256 threads x 8 items, then one block of 1024 threads striding over the partials), so the total and the weights are
// bit-identical to what bamcl_pf_score_planar / bamcl_pf_normalise produce for the same input. The only difference is
// where the block partials live: instead of a cudaMalloc'ed buffer of the filter they go to the file-scope array
// `g_noalloc_partials` below, which every GPU that loads this library holds until the process exits (64 KB).
//
// Host code never hands the array's NAME to a kernel (on the host that name is a stand-in, not a device address): the
// device address is looked up once per device with cudaGetSymbolAddress, with that device current, and cached.
#include <mutex>

#include "common.cuh"

using namespace bamcl;

namespace
{
constexpr int NA_BLOCK = 256;  // = RED_BLOCK of pf.cu
constexpr int NA_ITEMS = 8;    // = RED_ITEMS of pf.cu
constexpr long NA_MAX_SAMPLES = BAMCL_NOALLOC_MAX_SAMPLES;
constexpr int NA_MAX_PARTIALS = (int)((NA_MAX_SAMPLES + NA_BLOCK * NA_ITEMS - 1) / (NA_BLOCK * NA_ITEMS));
constexpr int NA_MAX_DEVICES = 64;

// the scratch: one partial per block of the first level. Nothing in it has to start at zero: a call writes every partial
// it later reads (blocks 0..m-1) before reading it, in stream order, and never reads beyond m.
static __device__ __align__(256) double g_noalloc_partials[NA_MAX_PARTIALS];

__global__ void __launch_bounds__(NA_BLOCK) k_na_sum_partials(const double* __restrict__ w, int n, double* __restrict__ partials)
{
  __shared__ double s[NA_BLOCK / 32];
  const int base = blockIdx.x * NA_BLOCK * NA_ITEMS;
  double v = 0.0;
#pragma unroll
  for (int t = 0; t < NA_ITEMS; t++)
  {
    const int i = base + t * NA_BLOCK + threadIdx.x;
    if (i < n)
      v += w[i];
  }
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0)
    s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32)
  {
    double t = threadIdx.x < NA_BLOCK / 32 ? s[threadIdx.x] : 0.0;
    t = warp_sum(t);
    if (threadIdx.x == 0)
      partials[blockIdx.x] = t;
  }
}
__global__ void __launch_bounds__(1024) k_na_sum_final(const double* __restrict__ partials, int m, double* __restrict__ total)
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
// ParticleFilter::updateSensor, particle_filter.cpp:237-266: w / total, or 1 / n where the total is not positive
__global__ void __launch_bounds__(256) k_na_normalise(const double* __restrict__ w, int n, const double* __restrict__ total_p,
                                                      double* __restrict__ out)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  const double total = *total_p;
  out[i] = total > 0.0 ? w[i] / total : 1.0 / n;
}

std::mutex g_lock;
double* g_partials_addr[NA_MAX_DEVICES];  // device address of g_noalloc_partials on each device, looked up on first use

// set-up, once per device, with that device current
int scratch_address(int device, double** out)
{
  if (device < 0 || device >= NA_MAX_DEVICES)
    return set_error(BAMCL_E_ARG, "bamcl_weights_normalise_noalloc: device %d outside 0..%d", device, NA_MAX_DEVICES - 1);
  std::lock_guard<std::mutex> guard(g_lock);
  if (!g_partials_addr[device])
  {
    int current = -1;
    BAMCL_CUDA_OK(cudaGetDevice(&current));
    if (current != device)
      BAMCL_CUDA_OK(cudaSetDevice(device));
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_partials);
    if (current != device && current >= 0)
      cudaSetDevice(current);
    if (e != cudaSuccess || !p)
      return set_error(BAMCL_E_CUDA, "bamcl_weights_normalise_noalloc: cudaGetSymbolAddress failed on device %d: %s", device,
                       cudaGetErrorString(e));
    g_partials_addr[device] = static_cast<double*>(p);
  }
  *out = g_partials_addr[device];
  return BAMCL_OK;
}
}  // namespace

extern "C" int bamcl_weights_normalise_noalloc(bamcl_ctx* ctx, const double* weights_dev, long n, double* weights_out_dev,
                                               double* total_out_dev)
{
  if (!ctx || !total_out_dev || n < 0 || (n > 0 && (!weights_dev || !weights_out_dev)))
    return set_error(BAMCL_E_ARG, "bamcl_weights_normalise_noalloc: null argument or negative count");
  if (n > NA_MAX_SAMPLES)
    return set_error(BAMCL_E_ARG, "bamcl_weights_normalise_noalloc: %ld weights exceed the static scratch's limit of %ld "
                                  "(BAMCL_NOALLOC_MAX_SAMPLES); use bamcl_pf_normalise / the fused update entry points",
                     n, NA_MAX_SAMPLES);
  double* partials = nullptr;
  BAMCL_TRY(scratch_address(ctx->device, &partials));
  const int count = (int)n;
  const int blocks = (count + NA_BLOCK * NA_ITEMS - 1) / (NA_BLOCK * NA_ITEMS);
  if (blocks > 0)
    BAMCL_LAUNCH(ctx, k_na_sum_partials, blocks, NA_BLOCK, 0, weights_dev, count, partials);
  BAMCL_LAUNCH(ctx, k_na_sum_final, 1, 1024, 0, partials, blocks, total_out_dev);
  if (count > 0)
    BAMCL_LAUNCH(ctx, k_na_normalise, (count + 255) / 256, 256, 0, weights_dev, count, total_out_dev, weights_out_dev);
  return BAMCL_OK;
}
