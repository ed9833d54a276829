// Rank-to-rank exchange for the sharded filter (one process per GPU): the few collectives the resampler
// needs, issued on the context's stream. Two back ends behind one interface:
//   * NCCL over NVLink / NVSwitch (libnccl.so.2 resolved at run time, so a single-GPU user never needs it);
//   * "local": G ranks = G threads of ONE process on ONE GPU, collectives as stream-ordered device copies between the
//     ranks' buffers. It exists so the multi-rank path can be parity-tested on a single-GPU box; it is not a fallback
//     (bamcl_comm_create never selects it by itself).
#pragma once
#include "common.cuh"

namespace bamcl
{
enum CommType
{
  CT_I32 = 0,
  CT_I64 = 1,
  CT_F64 = 2
};
enum CommOp
{
  CO_SUM = 0,
  CO_MIN = 1,
  CO_MAX = 2
};
struct P2P
{
  int peer;
  void* ptr;
  size_t bytes;
};
struct LocalGroup;
}  // namespace bamcl

struct bamcl_comm
{
  bamcl_ctx* ctx = nullptr;
  int rank = 0, n_ranks = 1;
  void* nccl = nullptr;               // ncclComm_t
  bamcl::LocalGroup* local = nullptr;  // test back end
  bamcl::DevBuf scratch;               // local back end: gathered operands of an all-reduce
  long n_collectives = 0;
};

namespace bamcl
{
// every rank contributes `bytes` at `send`; recv gets n_ranks * bytes in rank order (send may alias its own slot)
int comm_all_gather(bamcl_comm* c, const void* send, void* recv, size_t bytes);
// in place over `count` elements
int comm_all_reduce(bamcl_comm* c, void* buf, size_t count, CommType type, CommOp op);
// grouped point-to-point: the k-th send to a peer matches that peer's k-th receive from this rank
int comm_exchange(bamcl_comm* c, const P2P* sends, int n_sends, const P2P* recvs, int n_recvs);
}  // namespace bamcl
