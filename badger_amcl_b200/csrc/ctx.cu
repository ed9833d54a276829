// Context, error reporting, timing helpers.
#include <cstdarg>

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
}  // namespace

extern "C" {

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
  BAMCL_CUDA_OK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
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
