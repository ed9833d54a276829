// Shared device/host helpers for libbamcl_cuda.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <string>
#include <vector>

#include "../../include/bamcl.h"

namespace bamcl
{
extern thread_local std::string g_last_error;
int set_error(int code, const char* fmt, ...);

#define BAMCL_CUDA_OK(expr)                                                                             \
  do                                                                                                    \
  {                                                                                                     \
    cudaError_t _e = (expr);                                                                            \
    if (_e != cudaSuccess)                                                                              \
      return bamcl::set_error(BAMCL_E_CUDA, "%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,        \
                              cudaGetErrorString(_e));                                                  \
  } while (0)

#define BAMCL_TRY(expr)   \
  do                      \
  {                       \
    int _r = (expr);      \
    if (_r != BAMCL_OK)   \
      return _r;          \
  } while (0)

constexpr int kNumSMs = 148;  // B200

// A device buffer that only ever grows.
struct DevBuf
{
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes)
  {
    if (bytes <= cap)
      return BAMCL_OK;
    if (p)
      cudaFree(p);
    p = nullptr;
    cap = 0;
    BAMCL_CUDA_OK(cudaMalloc(&p, bytes));
    cap = bytes;
    return BAMCL_OK;
  }
  void release()
  {
    if (p)
      cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const
  {
    return static_cast<T*>(p);
  }
};

// Pinned host staging buffer that only ever grows.
struct PinBuf
{
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes)
  {
    if (bytes <= cap)
      return BAMCL_OK;
    if (p)
      cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    BAMCL_CUDA_OK(cudaMallocHost(&p, bytes));
    cap = bytes;
    return BAMCL_OK;
  }
  void release()
  {
    if (p)
      cudaFreeHost(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const
  {
    return static_cast<T*>(p);
  }
};
}  // namespace bamcl

struct bamcl_ctx
{
  int device = 0;
  cudaStream_t stream = nullptr;
  int stream_priority = 0, low_priority = 0;  // the main (+ side) streams' priority and the lowest one of the device
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  long launches = 0;
  bamcl::DevBuf flush;  // > L2 scratch for bamcl_ctx_flush_l2
  int coop_max_blocks_per_sm_checked = 0;
};

#define BAMCL_LAUNCH(ctx, kernel, grid, block, smem, ...)                             \
  do                                                                                  \
  {                                                                                   \
    kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);                  \
    (ctx)->launches++;                                                                \
    BAMCL_CUDA_OK(cudaGetLastError());                                                \
  } while (0)

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
namespace bamcl
{
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ long long warp_sum_ll(long long v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// floor(v) as int32 via the round-down magic add (valid for |v| < 2^31): one DADD instead of F2I.F64
__device__ __forceinline__ int floor_to_int(double v) { return __double2loint(__dadd_rd(v, 6755399441055744.0)); }

// cell of a world coordinate exactly as OccupancyMap::convertWorldToMap (occupancy_map.cpp:90-98)
__device__ __forceinline__ int world_to_cell(double x, float origin, double resolution, int size)
{
  return (int)floor((x - (double)origin) / resolution + 0.5) + size / 2;
}
}  // namespace bamcl
