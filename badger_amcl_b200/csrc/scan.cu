// Device-wide exclusive scan of int32 (three launches: tile sums, one-block scan of the sums, tile scan).
#include "scan.cuh"

using namespace bamcl;

namespace
{
constexpr int SCAN_BLOCK = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_BLOCK * SCAN_ITEMS;

__device__ __forceinline__ int block_exclusive_scan(int v, int* s_warp, int* block_total)
{
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1)
  {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o)
      inc += t;
  }
  if (lane == 31)
    s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0)
  {
    int w = lane < (int)(blockDim.x >> 5) ? s_warp[lane] : 0;
    int winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
      const int t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o)
        winc += t;
    }
    s_warp[lane] = winc - w;  // exclusive warp offsets
    if (lane == 31)
      s_warp[32] = winc;
  }
  __syncthreads();
  const int off = s_warp[warp];
  *block_total = s_warp[32];
  __syncthreads();
  return off + inc - v;
}

__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_tile_sums(const int* __restrict__ in, long n, int* __restrict__ sums)
{
  __shared__ int s_warp[33];
  const long base = (long)blockIdx.x * SCAN_TILE;
  int v = 0;
#pragma unroll
  for (int t = 0; t < SCAN_ITEMS; t++)
  {
    const long i = base + t * SCAN_BLOCK + threadIdx.x;
    if (i < n)
      v += in[i];
  }
  v = warp_sum_i(v);
  if ((threadIdx.x & 31) == 0)
    s_warp[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32)
  {
    int t = threadIdx.x < SCAN_BLOCK / 32 ? s_warp[threadIdx.x] : 0;
    t = warp_sum_i(t);
    if (threadIdx.x == 0)
      sums[blockIdx.x] = t;
  }
}

// one block: exclusive scan of m tile sums in place, total written to *total
__global__ void __launch_bounds__(1024) k_scan_sums(int* __restrict__ sums, int m, int* __restrict__ total)
{
  __shared__ int s_warp[33];
  int carry = 0;
  for (int base = 0; base < m; base += 1024)
  {
    const int i = base + threadIdx.x;
    const int v = i < m ? sums[i] : 0;
    int chunk_total;
    const int ex = block_exclusive_scan(v, s_warp, &chunk_total);
    if (i < m)
      sums[i] = carry + ex;
    carry += chunk_total;
  }
  if (threadIdx.x == 0 && total)
    *total = carry;
}

__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_tiles(const int* __restrict__ in, int* __restrict__ out, long n,
                                                           const int* __restrict__ sums)
{
  __shared__ int s_warp[33];
  const long base = (long)blockIdx.x * SCAN_TILE + (long)threadIdx.x * SCAN_ITEMS;
  int v[SCAN_ITEMS];
  int local = 0;
#pragma unroll
  for (int t = 0; t < SCAN_ITEMS; t++)
  {
    v[t] = (base + t < n) ? in[base + t] : 0;
    local += v[t];
  }
  int block_total;
  int run = sums[blockIdx.x] + block_exclusive_scan(local, s_warp, &block_total);
#pragma unroll
  for (int t = 0; t < SCAN_ITEMS; t++)
  {
    if (base + t < n)
      out[base + t] = run;
    run += v[t];
  }
}
}  // namespace

namespace bamcl
{
size_t scan_scratch_ints(long n) { return (size_t)((n + SCAN_TILE - 1) / SCAN_TILE) + 1; }

int exclusive_scan_i32(bamcl_ctx* ctx, const int* in, int* out, long n, int* scratch, int* total_out)
{
  if (n <= 0)
  {
    if (total_out)
      BAMCL_CUDA_OK(cudaMemsetAsync(total_out, 0, sizeof(int), ctx->stream));
    return BAMCL_OK;
  }
  const int tiles = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
  BAMCL_LAUNCH(ctx, k_scan_tile_sums, tiles, SCAN_BLOCK, 0, in, n, scratch);
  BAMCL_LAUNCH(ctx, k_scan_sums, 1, 1024, 0, scratch, tiles, total_out);
  BAMCL_LAUNCH(ctx, k_scan_tiles, tiles, SCAN_BLOCK, 0, in, out, n, scratch);
  return BAMCL_OK;
}
}  // namespace bamcl
