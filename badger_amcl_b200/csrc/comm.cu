// Collectives of the sharded filter: NCCL (resolved with dlopen at first use) and the in-process "local" test back
// end. See comm.cuh. Everything is enqueued on the owning context's stream.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <chrono>
#include <cstring>
#include <condition_variable>
#include <mutex>
#include <thread>

#include "comm.cuh"

using namespace bamcl;

namespace
{
// ---- NCCL entry points, bound at run time: torch's bundled libnccl when the process already loaded it, else the
// system one. No link-time dependency, so libbamcl_cuda.so loads on hosts without NCCL.
struct NcclApi
{
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;
};

NcclApi* nccl_api()
{
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (!h)
      h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h)
      h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h)
    {
      const char* e = dlerror();
      api.error = e ? e : "libnccl.so.2 not found";
      return;
    }
    api.handle = h;
    bool ok = true;
    auto bind = [&](const char* name) {
      void* p = dlsym(h, name);
      if (!p)
      {
        ok = false;
        api.error = std::string("missing NCCL symbol ") + name;
      }
      return p;
    };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(bind("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(bind("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(bind("ncclCommDestroy"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(bind("ncclAllGather"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(bind("ncclAllReduce"));
    api.Send = reinterpret_cast<decltype(api.Send)>(bind("ncclSend"));
    api.Recv = reinterpret_cast<decltype(api.Recv)>(bind("ncclRecv"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(bind("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(bind("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(bind("ncclGetErrorString"));
    if (!ok)
      api.handle = nullptr;
  });
  return &api;
}

#define BAMCL_NCCL_OK(api, expr)                                                                              \
  do                                                                                                          \
  {                                                                                                           \
    ncclResult_t _r = (expr);                                                                                 \
    if (_r != ncclSuccess)                                                                                    \
      return bamcl::set_error(BAMCL_E_COMM, "%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,              \
                              (api)->GetErrorString ? (api)->GetErrorString(_r) : "nccl error");              \
  } while (0)

template <typename T>
__global__ void k_reduce_ranks(const T* __restrict__ gathered, T* __restrict__ out, size_t count, int n_ranks, int op)
{
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
  {
    T acc = gathered[i];
    for (int r = 1; r < n_ranks; r++)
    {
      const T v = gathered[(size_t)r * count + i];
      acc = op == CO_SUM ? acc + v : (op == CO_MIN ? (v < acc ? v : acc) : (v > acc ? v : acc));
    }
    out[i] = acc;
  }
}
}  // namespace

// ---- the local back end: rendezvous of G threads
namespace bamcl
{
struct LocalGroup
{
  int n = 0;
  std::mutex m;
  std::condition_variable cv;
  int arrived = 0;
  long generation = 0;
  bool broken = false;
  int refs = 0;
  std::vector<const void*> ptr;
  std::vector<std::vector<P2P>> sends;

  // false = a peer failed or did not arrive in time (the group stays broken: every later collective fails fast)
  bool barrier()
  {
    std::unique_lock<std::mutex> lk(m);
    if (broken)
      return false;
    const long g = generation;
    if (++arrived == n)
    {
      arrived = 0;
      generation++;
      cv.notify_all();
      return true;
    }
    const bool ok = cv.wait_for(lk, std::chrono::seconds(120), [&] { return generation != g || broken; });
    if (!ok || broken)
    {
      broken = true;
      cv.notify_all();
      return false;
    }
    return true;
  }
  void abort()
  {
    std::lock_guard<std::mutex> lk(m);
    broken = true;
    cv.notify_all();
  }
};
}  // namespace bamcl

namespace
{
size_t type_bytes(CommType t) { return t == CT_I32 ? 4 : 8; }

int local_fail(bamcl_comm* c, const char* what)
{
  c->local->abort();
  return set_error(BAMCL_E_COMM, "local comm: %s (a peer rank failed or timed out)", what);
}
}  // namespace

namespace bamcl
{
int comm_all_gather(bamcl_comm* c, const void* send, void* recv, size_t bytes)
{
  if (!c || !send || !recv)
    return set_error(BAMCL_E_ARG, "comm_all_gather: bad arguments");
  cudaStream_t st = c->ctx->stream;
  c->n_collectives++;
  if (c->n_ranks == 1)
  {
    if (send != recv && bytes)
      BAMCL_CUDA_OK(cudaMemcpyAsync(recv, send, bytes, cudaMemcpyDeviceToDevice, st));
    return BAMCL_OK;
  }
  if (c->nccl)
  {
    NcclApi* api = nccl_api();
    BAMCL_NCCL_OK(api, api->AllGather(send, recv, bytes, ncclInt8, (ncclComm_t)c->nccl, st));
    return BAMCL_OK;
  }
  LocalGroup* g = c->local;
  BAMCL_CUDA_OK(cudaStreamSynchronize(st));  // my operand is complete before a peer reads it
  g->ptr[c->rank] = send;
  if (!g->barrier())
    return local_fail(c, "all_gather");
  for (int r = 0; r < c->n_ranks; r++)
  {
    void* dst = static_cast<char*>(recv) + (size_t)r * bytes;
    if (dst != g->ptr[r] && bytes)
      BAMCL_CUDA_OK(cudaMemcpyAsync(dst, g->ptr[r], bytes, cudaMemcpyDeviceToDevice, st));
  }
  BAMCL_CUDA_OK(cudaStreamSynchronize(st));
  if (!g->barrier())  // nobody reuses an operand while a peer still copies from it
    return local_fail(c, "all_gather");
  return BAMCL_OK;
}

int comm_all_reduce(bamcl_comm* c, void* buf, size_t count, CommType type, CommOp op)
{
  if (!c || !buf)
    return set_error(BAMCL_E_ARG, "comm_all_reduce: bad arguments");
  cudaStream_t st = c->ctx->stream;
  c->n_collectives++;
  if (c->n_ranks == 1 || count == 0)
    return BAMCL_OK;
  if (c->nccl)
  {
    NcclApi* api = nccl_api();
    const ncclDataType_t dt = type == CT_I32 ? ncclInt32 : (type == CT_I64 ? ncclInt64 : ncclFloat64);
    const ncclRedOp_t ro = op == CO_SUM ? ncclSum : (op == CO_MIN ? ncclMin : ncclMax);
    BAMCL_NCCL_OK(api, api->AllReduce(buf, buf, count, dt, ro, (ncclComm_t)c->nccl, st));
    return BAMCL_OK;
  }
  LocalGroup* g = c->local;
  const size_t bytes = count * type_bytes(type);
  BAMCL_TRY(c->scratch.reserve(bytes * (size_t)c->n_ranks));
  BAMCL_CUDA_OK(cudaStreamSynchronize(st));
  g->ptr[c->rank] = buf;
  if (!g->barrier())
    return local_fail(c, "all_reduce");
  for (int r = 0; r < c->n_ranks; r++)
    BAMCL_CUDA_OK(cudaMemcpyAsync(c->scratch.as<char>() + (size_t)r * bytes, g->ptr[r], bytes, cudaMemcpyDeviceToDevice, st));
  BAMCL_CUDA_OK(cudaStreamSynchronize(st));
  if (!g->barrier())
    return local_fail(c, "all_reduce");
  const unsigned blocks = (unsigned)std::min<size_t>((count + 255) / 256, (size_t)kNumSMs * 8);
  if (type == CT_I32)
    BAMCL_LAUNCH(c->ctx, k_reduce_ranks<int>, blocks, 256, 0, c->scratch.as<int>(), static_cast<int*>(buf), count, c->n_ranks,
                 (int)op);
  else if (type == CT_I64)
    BAMCL_LAUNCH(c->ctx, k_reduce_ranks<long long>, blocks, 256, 0, c->scratch.as<long long>(), static_cast<long long*>(buf),
                 count, c->n_ranks, (int)op);
  else
    BAMCL_LAUNCH(c->ctx, k_reduce_ranks<double>, blocks, 256, 0, c->scratch.as<double>(), static_cast<double*>(buf), count,
                 c->n_ranks, (int)op);
  return BAMCL_OK;
}

int comm_exchange(bamcl_comm* c, const P2P* sends, int n_sends, const P2P* recvs, int n_recvs)
{
  if (!c || (n_sends && !sends) || (n_recvs && !recvs))
    return set_error(BAMCL_E_ARG, "comm_exchange: bad arguments");
  cudaStream_t st = c->ctx->stream;
  c->n_collectives++;
  if (c->nccl)
  {
    NcclApi* api = nccl_api();
    BAMCL_NCCL_OK(api, api->GroupStart());
    for (int i = 0; i < n_sends; i++)
      if (sends[i].bytes)
        BAMCL_NCCL_OK(api, api->Send(sends[i].ptr, sends[i].bytes, ncclInt8, sends[i].peer, (ncclComm_t)c->nccl, st));
    for (int i = 0; i < n_recvs; i++)
      if (recvs[i].bytes)
        BAMCL_NCCL_OK(api, api->Recv(recvs[i].ptr, recvs[i].bytes, ncclInt8, recvs[i].peer, (ncclComm_t)c->nccl, st));
    BAMCL_NCCL_OK(api, api->GroupEnd());
    return BAMCL_OK;
  }
  if (c->n_ranks == 1)
  {
    // self sends only
    int k = 0;
    for (int i = 0; i < n_recvs; i++)
    {
      while (k < n_sends && sends[k].peer != 0)
        k++;
      if (k >= n_sends || sends[k].bytes != recvs[i].bytes)
        return set_error(BAMCL_E_ARG, "comm_exchange: unmatched self transfer");
      if (recvs[i].bytes)
        BAMCL_CUDA_OK(cudaMemcpyAsync(recvs[i].ptr, sends[k].ptr, recvs[i].bytes, cudaMemcpyDeviceToDevice, st));
      k++;
    }
    return BAMCL_OK;
  }
  LocalGroup* g = c->local;
  BAMCL_CUDA_OK(cudaStreamSynchronize(st));
  g->sends[c->rank].assign(sends, sends + n_sends);
  if (!g->barrier())
    return local_fail(c, "exchange");
  std::vector<int> next(c->n_ranks, 0);  // per peer: how many of its sends to me are consumed
  for (int i = 0; i < n_recvs; i++)
  {
    const int p = recvs[i].peer;
    if (p < 0 || p >= c->n_ranks)
      return local_fail(c, "exchange: peer out of range");
    const std::vector<P2P>& ps = g->sends[p];
    int k = next[p];
    while (k < (int)ps.size() && ps[k].peer != c->rank)
      k++;
    if (k >= (int)ps.size() || ps[k].bytes != recvs[i].bytes)
      return local_fail(c, "exchange: a receive has no matching send");
    next[p] = k + 1;
    if (recvs[i].bytes)
      BAMCL_CUDA_OK(cudaMemcpyAsync(recvs[i].ptr, ps[k].ptr, recvs[i].bytes, cudaMemcpyDeviceToDevice, st));
  }
  BAMCL_CUDA_OK(cudaStreamSynchronize(st));
  if (!g->barrier())
    return local_fail(c, "exchange");
  return BAMCL_OK;
}
}  // namespace bamcl

extern "C" {

int bamcl_comm_unique_id(void* id_out)
{
  if (!id_out)
    return set_error(BAMCL_E_ARG, "bamcl_comm_unique_id: null output");
  NcclApi* api = nccl_api();
  if (!api->handle)
    return set_error(BAMCL_E_COMM, "NCCL is not available: %s", api->error.c_str());
  static_assert(sizeof(ncclUniqueId) == BAMCL_UNIQUE_ID_BYTES, "ncclUniqueId size");
  ncclUniqueId id;
  BAMCL_NCCL_OK(api, api->GetUniqueId(&id));
  memcpy(id_out, &id, sizeof(id));
  return BAMCL_OK;
}

int bamcl_comm_create(bamcl_ctx* ctx, const void* id, int rank, int n_ranks, bamcl_comm** out)
{
  if (!ctx || !out || n_ranks < 1 || rank < 0 || rank >= n_ranks || (n_ranks > 1 && !id))
    return set_error(BAMCL_E_ARG, "bamcl_comm_create: bad arguments");
  bamcl_comm* c = new bamcl_comm();
  c->ctx = ctx;
  c->rank = rank;
  c->n_ranks = n_ranks;
  if (n_ranks > 1)
  {
    NcclApi* api = nccl_api();
    if (!api->handle)
    {
      delete c;
      return set_error(BAMCL_E_COMM, "NCCL is not available: %s", api->error.c_str());
    }
    BAMCL_CUDA_OK(cudaSetDevice(ctx->device));
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    ncclComm_t comm = nullptr;
    ncclResult_t r = api->CommInitRank(&comm, n_ranks, uid, rank);
    if (r != ncclSuccess)
    {
      delete c;
      return set_error(BAMCL_E_COMM, "ncclCommInitRank failed: %s", api->GetErrorString(r));
    }
    c->nccl = comm;
  }
  *out = c;
  return BAMCL_OK;
}

int bamcl_comm_create_file(bamcl_ctx* ctx, const char* path, int rank, int n_ranks, int timeout_s, bamcl_comm** out)
{
  if (!ctx || !out || !path || n_ranks < 1 || rank < 0 || rank >= n_ranks)
    return set_error(BAMCL_E_ARG, "bamcl_comm_create_file: bad arguments");
  unsigned char id[BAMCL_UNIQUE_ID_BYTES];
  memset(id, 0, sizeof(id));
  if (n_ranks > 1)
  {
    if (rank == 0)
    {
      BAMCL_TRY(bamcl_comm_unique_id(id));
      const std::string tmp = std::string(path) + ".tmp";
      FILE* f = fopen(tmp.c_str(), "wb");
      if (!f || fwrite(id, 1, sizeof(id), f) != sizeof(id))
      {
        if (f)
          fclose(f);
        return set_error(BAMCL_E_COMM, "bamcl_comm_create_file: cannot write %s", tmp.c_str());
      }
      fclose(f);
      if (rename(tmp.c_str(), path) != 0)  // atomic: readers never see a partial id
        return set_error(BAMCL_E_COMM, "bamcl_comm_create_file: cannot publish %s", path);
    }
    else
    {
      const auto t0 = std::chrono::steady_clock::now();
      for (;;)
      {
        FILE* f = fopen(path, "rb");
        if (f)
        {
          const size_t got = fread(id, 1, sizeof(id), f);
          fclose(f);
          if (got == sizeof(id))
            break;
        }
        if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s)
          return set_error(BAMCL_E_COMM, "bamcl_comm_create_file: no id at %s after %d s", path, timeout_s);
        std::this_thread::sleep_for(std::chrono::milliseconds(20));
      }
    }
  }
  const int rc = bamcl_comm_create(ctx, id, rank, n_ranks, out);
  return rc;
}

int bamcl_comm_all_reduce_i32(bamcl_comm* c, int* host_inout, int n, int op)
{
  if (!c || !host_inout || n < 0 || op < 0 || op > 2)
    return set_error(BAMCL_E_ARG, "bamcl_comm_all_reduce_i32: bad arguments");
  if (n == 0 || c->n_ranks == 1)
    return BAMCL_OK;
  cudaStream_t st = c->ctx->stream;
  bamcl::DevBuf tmp;
  BAMCL_TRY(tmp.reserve(sizeof(int) * (size_t)n));
  int rc = BAMCL_OK;
  if (cudaMemcpyAsync(tmp.p, host_inout, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, st) != cudaSuccess)
    rc = set_error(BAMCL_E_CUDA, "bamcl_comm_all_reduce_i32: upload failed");
  if (rc == BAMCL_OK)
    rc = comm_all_reduce(c, tmp.p, (size_t)n, CT_I32, (CommOp)op);
  if (rc == BAMCL_OK && (cudaMemcpyAsync(host_inout, tmp.p, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
                         cudaStreamSynchronize(st) != cudaSuccess))
    rc = set_error(BAMCL_E_CUDA, "bamcl_comm_all_reduce_i32: download failed");
  cudaStreamSynchronize(st);
  tmp.release();
  return rc;
}

int bamcl_comm_local_group(int n_ranks, void** group_out)
{
  if (n_ranks < 1 || !group_out)
    return set_error(BAMCL_E_ARG, "bamcl_comm_local_group: bad arguments");
  LocalGroup* g = new LocalGroup();
  g->n = n_ranks;
  g->refs = 1;  // the creator's reference, dropped by bamcl_comm_local_group_free
  g->ptr.assign(n_ranks, nullptr);
  g->sends.resize(n_ranks);
  *group_out = g;
  return BAMCL_OK;
}

int bamcl_comm_local_group_free(void* group)
{
  LocalGroup* g = static_cast<LocalGroup*>(group);
  if (!g)
    return BAMCL_OK;
  bool last;
  {
    std::lock_guard<std::mutex> lk(g->m);
    last = --g->refs == 0;
  }
  if (last)
    delete g;
  return BAMCL_OK;
}

int bamcl_comm_create_local(bamcl_ctx* ctx, void* group, int rank, bamcl_comm** out)
{
  LocalGroup* g = static_cast<LocalGroup*>(group);
  if (!ctx || !g || !out || rank < 0 || rank >= g->n)
    return set_error(BAMCL_E_ARG, "bamcl_comm_create_local: bad arguments");
  bamcl_comm* c = new bamcl_comm();
  c->ctx = ctx;
  c->rank = rank;
  c->n_ranks = g->n;
  c->local = g;
  {
    std::lock_guard<std::mutex> lk(g->m);
    g->refs++;
  }
  *out = c;
  return BAMCL_OK;
}

int bamcl_comm_destroy(bamcl_comm* c)
{
  if (!c)
    return BAMCL_OK;
  cudaStreamSynchronize(c->ctx->stream);
  if (c->nccl)
  {
    NcclApi* api = nccl_api();
    if (api->handle)
      api->CommDestroy((ncclComm_t)c->nccl);
  }
  if (c->local)
    bamcl_comm_local_group_free(c->local);
  c->scratch.release();
  delete c;
  return BAMCL_OK;
}

int bamcl_comm_all_reduce_sum_i32(bamcl_comm* c, int* host_inout, int n) { return bamcl_comm_all_reduce_i32(c, host_inout, n, 0); }

int bamcl_comm_rank(bamcl_comm* c, int* rank, int* n_ranks)
{
  if (!c)
    return set_error(BAMCL_E_ARG, "bamcl_comm_rank: null comm");
  if (rank)
    *rank = c->rank;
  if (n_ranks)
    *n_ranks = c->n_ranks;
  return BAMCL_OK;
}

}  // extern "C"
