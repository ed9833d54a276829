// Context, error reporting, timing helpers.
#include <algorithm>
#include <cstdarg>
#include <cstdlib>

#include "common.cuh"

namespace bamcl
{
thread_local std::string g_last_error;
int set_error(int code, const char* fmt, ...)
{
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}
}  // namespace bamcl

using namespace bamcl;

namespace
{
__global__ void k_flush(float4* __restrict__ p, size_t n)
{
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    p[i] = make_float4(1.f, 2.f, 3.f, 4.f);
}
// measurement aid: random 2-byte gathers, 16 independent loads in flight per thread (the access pattern of the
// likelihood-field scorer without its arithmetic)
__global__ void __launch_bounds__(256) k_gather_bench(const unsigned short* __restrict__ table, unsigned long long n_elems,
                                                      int iters, unsigned* __restrict__ sink)
{
  unsigned long long s = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ull + 12345ull;
  unsigned acc = 0;
  for (int it = 0; it < iters; it++)
  {
    unsigned v[16];
#pragma unroll
    for (int k = 0; k < 16; k++)
    {
      s = s * 6364136223846793005ull + 1442695040888963407ull;
      const unsigned long long idx = ((s >> 32) * n_elems) >> 32;  // uniform in [0, n_elems), n_elems < 2^32
      v[k] = __ldg(&table[idx]);
    }
#pragma unroll
    for (int k = 0; k < 16; k++)
      acc += v[k];
  }
  if (acc == 0xFFFFFFFFu)
    sink[0] = acc;
}
}  // namespace

extern "C" {

int bamcl_bench_gather(bamcl_ctx* ctx, size_t table_bytes, long n_gathers, float* ms_out)
{
  if (!ctx || !ms_out || table_bytes < 64 || n_gathers < 1)
    return set_error(BAMCL_E_ARG, "bamcl_bench_gather: bad arguments");
  bamcl::DevBuf table, sink;
  BAMCL_TRY(table.reserve(table_bytes));
  BAMCL_TRY(sink.reserve(64));
  BAMCL_CUDA_OK(cudaMemsetAsync(table.p, 1, table_bytes, ctx->stream));
  const int blocks = kNumSMs * 8;
  const int iters = (int)std::max<long>(1, n_gathers / ((long)blocks * 256 * 16));
  int rc = BAMCL_OK;
  float best = 1e30f;
  for (int rep = 0; rep < 4 && rc == BAMCL_OK; rep++)  // first pass warms L2
  {
    cudaEventRecord(ctx->ev0, ctx->stream);
    k_gather_bench<<<blocks, 256, 0, ctx->stream>>>(table.as<unsigned short>(), table_bytes / 2, iters, sink.as<unsigned>());
    ctx->launches++;
    cudaEventRecord(ctx->ev1, ctx->stream);
    if (cudaEventSynchronize(ctx->ev1) != cudaSuccess)
      rc = set_error(BAMCL_E_CUDA, "bamcl_bench_gather: kernel failed");
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    if (rep > 0 && ms < best)
      best = ms;
  }
  table.release();
  sink.release();
  // report per gather actually issued
  *ms_out = best * (float)((double)n_gathers / ((double)blocks * 256 * 16 * iters));
  return rc;
}

const char* bamcl_last_error(void) { return g_last_error.c_str(); }
int bamcl_version(void) { return 100; }

int bamcl_ctx_create(int device, bamcl_ctx** out)
{
  if (!out)
    return set_error(BAMCL_E_ARG, "bamcl_ctx_create: out is null");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return set_error(BAMCL_E_CUDA, "no CUDA device available (%s); this library has no CPU fallback",
                     e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  if (device < 0 || device >= count)
    return set_error(BAMCL_E_ARG, "device %d out of range (0..%d)", device, count - 1);
  BAMCL_CUDA_OK(cudaSetDevice(device));
  bamcl_ctx* ctx = new bamcl_ctx();
  ctx->device = device;
  // The context's stream (and the resampler's bucket side streams) run at the highest priority; the resampler's clustering
  // chain goes to a side stream of the same or of the lowest priority, chosen per call (build_tree_and_stats, kdtree.cu).
  int prio_lo = 0, prio_hi = 0;
  BAMCL_CUDA_OK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  ctx->stream_priority = prio_hi;
  ctx->low_priority = prio_lo;
  BAMCL_CUDA_OK(cudaStreamCreateWithPriority(&ctx->stream, cudaStreamNonBlocking, ctx->stream_priority));
  BAMCL_CUDA_OK(cudaEventCreate(&ctx->ev0));
  BAMCL_CUDA_OK(cudaEventCreate(&ctx->ev1));
  *out = ctx;
  return BAMCL_OK;
}

int bamcl_ctx_destroy(bamcl_ctx* ctx)
{
  if (!ctx)
    return BAMCL_OK;
  cudaStreamSynchronize(ctx->stream);
  ctx->flush.release();
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
  return BAMCL_OK;
}

int bamcl_ctx_sync(bamcl_ctx* ctx)
{
  if (!ctx)
    return set_error(BAMCL_E_ARG, "null ctx");
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  return BAMCL_OK;
}

void* bamcl_ctx_stream(bamcl_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int bamcl_timer_start(bamcl_ctx* ctx)
{
  if (!ctx)
    return set_error(BAMCL_E_ARG, "null ctx");
  BAMCL_CUDA_OK(cudaEventRecord(ctx->ev0, ctx->stream));
  return BAMCL_OK;
}

int bamcl_timer_stop_ms(bamcl_ctx* ctx, float* ms)
{
  if (!ctx || !ms)
    return set_error(BAMCL_E_ARG, "bamcl_timer_stop_ms: bad arguments");
  BAMCL_CUDA_OK(cudaEventRecord(ctx->ev1, ctx->stream));
  BAMCL_CUDA_OK(cudaEventSynchronize(ctx->ev1));
  BAMCL_CUDA_OK(cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
  return BAMCL_OK;
}

int bamcl_host_alloc(size_t bytes, void** out)
{
  if (!out || bytes == 0)
    return set_error(BAMCL_E_ARG, "bamcl_host_alloc: bad arguments");
  BAMCL_CUDA_OK(cudaMallocHost(out, bytes));
  return BAMCL_OK;
}

int bamcl_host_free(void* p)
{
  if (p)
    BAMCL_CUDA_OK(cudaFreeHost(p));
  return BAMCL_OK;
}

int bamcl_host_register(void* p, size_t bytes)
{
  if (!p || bytes == 0)
    return set_error(BAMCL_E_ARG, "bamcl_host_register: bad arguments");
  BAMCL_CUDA_OK(cudaHostRegister(p, bytes, cudaHostRegisterDefault));
  return BAMCL_OK;
}

int bamcl_host_unregister(void* p)
{
  if (p)
    BAMCL_CUDA_OK(cudaHostUnregister(p));
  return BAMCL_OK;
}

long bamcl_ctx_launch_count(bamcl_ctx* ctx) { return ctx ? ctx->launches : 0; }

int bamcl_ctx_flush_l2(bamcl_ctx* ctx)
{
  if (!ctx)
    return set_error(BAMCL_E_ARG, "null ctx");
  const size_t bytes = (size_t)256 << 20;  // 256 MiB > 126 MB of L2
  BAMCL_TRY(ctx->flush.reserve(bytes));
  k_flush<<<kNumSMs * 4, 256, 0, ctx->stream>>>(ctx->flush.as<float4>(), bytes / sizeof(float4));
  BAMCL_CUDA_OK(cudaGetLastError());
  return BAMCL_OK;
}

}  // extern "C"
