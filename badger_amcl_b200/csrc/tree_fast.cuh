// Fast exact leaf count of the reference's insertion-order pose-bin tree (pf_kdtree.cpp:97-150), included by
// kdtree.cu (shares TreeOut / pack_key / unpack_key / NONE).
//
// The level-synchronous grid kernel (k_tree_build) needs one grid-wide barrier pair per tree level (~50-60
// levels for 1M bins). This path removes the grid barriers:
//   1. one CTA builds the tree of the first m <= M_TOP keys in shared memory. Because the reference's tree only
//      ever grows downwards, that is exactly the top of the final tree: every pivot it fixes and every child
//      it links is final;
//   2. every later key walks that top tree read-only (no synchronisation) until it falls off at an empty child
//      slot (u, side), or reaches a node u whose pivot is still unset: its bucket;
//   3. buckets are grouped (count, scan, scatter);
//   4. one CTA per bucket sorts its keys by creation index and grows the bucket's subtree in shared memory
//      with the same level-synchronous rule, using __syncthreads only. Subtrees of different buckets never
//      interact: everything routed into a slot stays below it.
// leaf count = childless nodes over all bucket subtrees + top nodes with an unset pivot and an empty bucket.
// A bucket larger than FAST_CAP_L (adversarial insertion orders, e.g. spatially sorted keys) makes the fast path
// give up (TreeOut::fast_ok stays 0) and k_tree_build does the work instead.
#pragma once

// keys in the top tree: the largest bucket holds about ln(M_TOP) / M_TOP of the keys, so sets beyond ~1.5M samples
// (sharded filter: every rank builds the global tree) take the large top tree to stay inside FAST_CAP_L.
constexpr int M_TOP_SMALL = 2048, M_TOP_LARGE = 8192;
constexpr int FAST_CAP_S = 512, FAST_CAP_M = 2048, FAST_CAP_L = 6144;

__device__ __forceinline__ int pivot_dim(unsigned long long ku, unsigned long long kf)
{
  int u[3], f[3];
  unpack_key(ku, &u[0], &u[1], &u[2]);
  unpack_key(kf, &f[0], &f[1], &f[2]);
  int pivot = 0, max_split = 0;  // pf_kdtree.cpp:135-144: largest |delta|, lowest dimension on ties
#pragma unroll
  for (int d = 0; d < 3; d++)
  {
    const int split = abs(f[d] - u[d]);
    if (split > max_split)
    {
      max_split = split;
      pivot = d;
    }
  }
  return pivot;
}
__device__ __forceinline__ int key_dim(unsigned long long k, int d)
{
  return d == 0 ? (int)(k & 0x1FFFFF) : d == 1 ? (int)((k >> 21) & 0x1FFFFF) : (int)((k >> 42) & 0x3FFFFF);
}

// Level-synchronous growth of the tree over key[0..n) (creation order, element 0 = root) in shared memory.
// cur/child/first_at are int[n], int[2n], int[n]. Returns the number of levels. All threads of the CTA call it.
__device__ int smem_tree(const unsigned long long* key, int n, int* cur, int* child, int* first_at)
{
  for (int v = threadIdx.x; v < n; v += blockDim.x)
  {
    child[2 * v] = child[2 * v + 1] = NONE;
    first_at[v] = (v == 0 && n > 1) ? 1 : NONE;
    cur[v] = v == 0 ? -1 : 0;
  }
  __syncthreads();
  int level = 0;
  if (n < 2)
    return 0;
  for (;;)
  {
    for (int v = threadIdx.x; v < n; v += blockDim.x)
    {
      const int u = cur[v];
      if (u < 0)
        continue;
      const unsigned long long ku = key[u];
      const int p = pivot_dim(ku, key[first_at[u]]);
      const int sd = key_dim(key[v], p) > key_dim(ku, p) ? 1 : 0;  // biased fields compare like the signed keys
      cur[v] = u | (sd << 30);
      atomicMin(&child[2 * u + sd], v);
    }
    __syncthreads();
    int any = 0;
    for (int v = threadIdx.x; v < n; v += blockDim.x)
    {
      const int cu = cur[v];
      if (cu < 0)
        continue;
      const int u = cu & 0x3FFFFFFF, sd = cu >> 30;
      const int c = child[2 * u + sd];
      if (c == v)
        cur[v] = -1;
      else
      {
        cur[v] = c;
        atomicMin(&first_at[c], v);
        any = 1;
      }
    }
    level++;
    if (!__syncthreads_or(any))
      break;
  }
  return level;
}

struct FastArgs
{
  TreeOut* out;
  const unsigned long long* nkey;  // [K] creation order
  int* tpivot;                     // [M_TOP] pivot dimension of a top node, -1 while unset
  int* tchild;                     // [2 * M_TOP] children inside the top tree (NONE = empty slot)
  int* bcount;                     // [N_BUCKETS + 1]
  int* boff;                       // [N_BUCKETS + 1]
  int* bslot;                      // [K] bucket of every key >= m
  int* bpos;                       // [K] arrival position inside the bucket
  int* bids;                       // [K] keys grouped by bucket
  int part_rank, part_n;           // distributed filter: this rank grows the subtrees of buckets b with (b >> 1) % part_n == part_rank
};

// ---- 1. top tree (one CTA)
template <int M_TOP>
__global__ void __launch_bounds__(1024) k_fast_top(const FastArgs a)
{
  constexpr int N_BUCKETS = 2 * M_TOP;
  extern __shared__ __align__(16) unsigned char s_raw[];
  unsigned long long* key = reinterpret_cast<unsigned long long*>(s_raw);
  int* cur = reinterpret_cast<int*>(key + M_TOP);
  int* child = cur + M_TOP;
  int* first_at = child + 2 * M_TOP;
  const int K = a.out->K;
  const int m = min(K, M_TOP);
  for (int v = threadIdx.x; v < m; v += blockDim.x)
    key[v] = a.nkey[v];
  for (int b = threadIdx.x; b <= N_BUCKETS; b += blockDim.x)
    a.bcount[b] = 0;
  __syncthreads();
  const int levels = smem_tree(key, m, cur, child, first_at);
  for (int v = threadIdx.x; v < M_TOP; v += blockDim.x)
  {
    int p = -1, c0 = NONE, c1 = NONE;
    if (v < m)
    {
      c0 = child[2 * v];
      c1 = child[2 * v + 1];
      if (first_at[v] != NONE)
        p = pivot_dim(key[v], key[first_at[v]]);
    }
    a.tpivot[v] = p;
    a.tchild[2 * v] = c0;
    a.tchild[2 * v + 1] = c1;
  }
  if (threadIdx.x == 0)
    a.out->levels = levels;
}

// ---- 2. every later key finds its bucket
template <int M_TOP>
__global__ void __launch_bounds__(256) k_fast_descend(const FastArgs a)
{
  extern __shared__ __align__(16) unsigned char s_raw[];  // 17 * M_TOP bytes
  unsigned long long* s_key = reinterpret_cast<unsigned long long*>(s_raw);
  int* s_child = reinterpret_cast<int*>(s_key + M_TOP);
  signed char* s_pivot = reinterpret_cast<signed char*>(s_child + 2 * M_TOP);
  const int K = a.out->K;
  const int m = min(K, M_TOP);
  if (K <= M_TOP)
    return;
  for (int v = threadIdx.x; v < m; v += blockDim.x)
  {
    s_key[v] = a.nkey[v];
    s_pivot[v] = (signed char)a.tpivot[v];
    s_child[2 * v] = a.tchild[2 * v];
    s_child[2 * v + 1] = a.tchild[2 * v + 1];
  }
  __syncthreads();
  for (int v = M_TOP + blockIdx.x * blockDim.x + threadIdx.x; v < K; v += gridDim.x * blockDim.x)
  {
    const unsigned long long kv = a.nkey[v];
    int u = 0, b;
    for (;;)
    {
      const int p = s_pivot[u];
      if (p < 0)
      {
        b = 2 * u;  // a top node that is still childless: the whole bucket hangs under it
        break;
      }
      const int sd = key_dim(kv, p) > key_dim(s_key[u], p) ? 1 : 0;
      const int c = s_child[2 * u + sd];
      if (c == NONE)
      {
        b = 2 * u + sd;
        break;
      }
      u = c;
    }
    a.bslot[v] = b;
    a.bpos[v] = atomicAdd(&a.bcount[b], 1);
  }
}

// ---- 3. bucket offsets (one CTA) and the grouping scatter
template <int M_TOP>
__global__ void __launch_bounds__(1024) k_fast_offsets(const FastArgs a)
{
  constexpr int N_BUCKETS = 2 * M_TOP;
  __shared__ int s_warp[32];
  __shared__ int s_over;
  constexpr int PER = N_BUCKETS / 1024;
  if (threadIdx.x == 0)
    s_over = 0;
  __syncthreads();
  int c[PER], sum = 0;
#pragma unroll
  for (int q = 0; q < PER; q++)
  {
    c[q] = a.bcount[threadIdx.x * PER + q];
    sum += c[q];
    if (c[q] + 1 > FAST_CAP_L)
      s_over = 1;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = sum;
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
    int w = s_warp[lane], winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
      const int t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o)
        winc += t;
    }
    s_warp[lane] = winc - w;
  }
  __syncthreads();
  int run = s_warp[warp] + inc - sum;
#pragma unroll
  for (int q = 0; q < PER; q++)
  {
    a.boff[threadIdx.x * PER + q] = run;
    run += c[q];
  }
  if (threadIdx.x == 1023)
  {
    a.boff[N_BUCKETS] = run;
    a.out->fast_ok = s_over ? 0 : 1;
  }
}

template <int M_TOP>
__global__ void __launch_bounds__(256) k_fast_scatter(const FastArgs a)
{
  const int K = a.out->K;
  if (!a.out->fast_ok)
    return;
  for (int v = M_TOP + blockIdx.x * blockDim.x + threadIdx.x; v < K; v += gridDim.x * blockDim.x)
    a.bids[a.boff[a.bslot[v]] + a.bpos[v]] = v;
}

// ---- 4. one CTA per bucket: sort by creation index, grow the subtree, count its childless nodes
template <int CAP, int CAP_BELOW, int THREADS, int M_TOP>
__global__ void __launch_bounds__(THREADS) k_fast_bucket(const FastArgs a)
{
  extern __shared__ __align__(16) unsigned char s_raw[];
  unsigned long long* key = reinterpret_cast<unsigned long long*>(s_raw);
  int* ids = reinterpret_cast<int*>(key + CAP);
  int* cur = ids + CAP;
  int* child = cur + CAP;
  int* first_at = child + 2 * CAP;
  __shared__ int s_leaves;
  if (!a.out->fast_ok)
    return;
  const int K = a.out->K;
  const int m = min(K, M_TOP);
  const int b = blockIdx.x, u = b >> 1, sd = b & 1;
  if (u >= m || u % a.part_n != a.part_rank)
    return;
  const bool unset = a.tpivot[u] < 0;
  if (unset ? sd == 1 : a.tchild[b] != NONE)
    return;  // no bucket hangs here
  const int off = a.boff[b], cnt = a.boff[b + 1] - off;
  const int n = cnt + (unset ? 1 : 0);
  if (n <= CAP_BELOW || n > CAP)
    return;  // another size class handles it
  if (n == 1)
  {
    // a childless top node with an empty bucket (n == 1 && unset), or a single key in a slot: one leaf
    if (threadIdx.x == 0)
      atomicAdd(&a.out->leaf, 1);
    return;
  }
  int np2 = 1;
  while (np2 < n)
    np2 <<= 1;
  for (int i = threadIdx.x; i < np2; i += THREADS)
  {
    int id = INT_MAX;
    if (unset)
      id = i == 0 ? u : (i <= cnt ? a.bids[off + i - 1] : INT_MAX);
    else if (i < cnt)
      id = a.bids[off + i];
    ids[i] = id;
  }
  if (threadIdx.x == 0)
    s_leaves = 0;
  __syncthreads();
  // bitonic sort of the creation indices
  for (int k = 2; k <= np2; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1)
    {
      for (int i = threadIdx.x; i < np2; i += THREADS)
      {
        const int l = i ^ j;
        if (l > i)
        {
          const int x = ids[i], y = ids[l];
          const bool up = (i & k) == 0;
          if ((x > y) == up)
          {
            ids[i] = y;
            ids[l] = x;
          }
        }
      }
      __syncthreads();
    }
  for (int i = threadIdx.x; i < n; i += THREADS)
    key[i] = a.nkey[ids[i]];
  __syncthreads();
  const int levels = smem_tree(key, n, cur, child, first_at);
  int leaves = 0;
  for (int v = threadIdx.x; v < n; v += THREADS)
    leaves += (child[2 * v] == NONE && child[2 * v + 1] == NONE) ? 1 : 0;
  leaves = warp_sum_i(leaves);
  if ((threadIdx.x & 31) == 0 && leaves)
    atomicAdd(&s_leaves, leaves);
  __syncthreads();
  if (threadIdx.x == 0)
  {
    atomicAdd(&a.out->leaf, s_leaves);
    atomicMax(&a.out->bucket_levels, levels);
  }
}
