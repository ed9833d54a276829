// Fast exact leaf count of the reference's insertion-order pose-bin tree (pf_kdtree.cpp:97-150), included by
// kdtree.cu (shares TreeOut / pack_key / unpack_key / NONE).
//
// The level-synchronous grid kernel (k_tree_build) needs one grid-wide barrier pair per tree level (~50-60
// levels for 1M bins). This path has no grid barriers. Because the reference's tree only ever grows downwards, the
// tree of the first m keys is exactly the top of the final tree: every pivot it fixes and every child it links is
// final. So the tree is grown from the top in three steps, each an order of magnitude wider than the one before:
//   A. one CTA builds the tree of the first M1 keys in shared memory and writes it to the global node arrays;
//   B. keys [M1, M2) walk that top tree read-only to the empty child slot (or childless node) they fall off at --
//      their bucket; one CTA per bucket grows the bucket's subtree in shared memory and WRITES it to the node arrays,
//      which now hold the tree of the first M2 keys (M2 ~ K / 32, so the last step's buckets stay small);
//   C. all remaining keys walk the M2-key tree to their bucket; one warp (<= 64 keys) or CTA per bucket grows the
//      subtree in shared memory and only COUNTS its childless nodes. Subtrees of different buckets never interact:
//      everything routed into a slot stays below it. On a distributed filter the ranks split the buckets of step C.
// leaf count = childless nodes over all step-C subtrees + top nodes with an unset pivot and an empty bucket.
// A bucket larger than FAST_CAP_L (adversarial insertion orders, e.g. spatially sorted keys) makes the fast path
// give up (TreeOut::fast_ok = 0) and k_tree_build does the work instead.
#pragma once

constexpr int M_TOP1 = 2048;         // keys of step A
constexpr int M_TOP2_MAX = 1 << 18;  // most keys the node arrays of steps A + B hold
constexpr int FAST_CAP_W = 64, FAST_CAP_S = 512, FAST_CAP_M = 2048, FAST_CAP_L = 6144;
constexpr int FAST_BYTES_PER_KEY = 28;  // shared memory per key in the bucket kernels

#ifndef TOP2_DIV
#define TOP2_DIV 32
#endif
// keys in the written-out top tree for a tree of K keys (same arithmetic on host and device, monotone in K)
__host__ __device__ inline int top2_size(int K)
{
  if (K <= 16 * M_TOP1)
    return M_TOP1;  // small trees: the step-A tree is the whole top
  int m = M_TOP1;
  while (m < K / TOP2_DIV && m < M_TOP2_MAX)
    m <<= 1;
  return m;
}

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
// cur/child/first_at are int[n], int[2n], int[n]. Returns the number of levels. WARP = the caller is one warp (lanes
// stride the keys, warp barriers); else all threads of the CTA call it.
template <bool WARP>
__device__ int smem_tree(const unsigned long long* key, int n, int* cur, int* child, int* first_at)
{
  const int t0 = WARP ? (threadIdx.x & 31) : threadIdx.x, step = WARP ? 32 : blockDim.x;
  for (int v = t0; v < n; v += step)
  {
    child[2 * v] = child[2 * v + 1] = NONE;
    first_at[v] = (v == 0 && n > 1) ? 1 : NONE;
    cur[v] = v == 0 ? -1 : 0;
  }
  if (WARP)
    __syncwarp();
  else
    __syncthreads();
  int level = 0;
  if (n < 2)
    return 0;
  for (;;)
  {
    for (int v = t0; v < n; v += step)
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
    if (WARP)
      __syncwarp();
    else
      __syncthreads();
    int any = 0;
    for (int v = t0; v < n; v += step)
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
    if (WARP)
    {
      __syncwarp();
      if (!__any_sync(0xffffffffu, any))
        break;
    }
    else if (!__syncthreads_or(any))
      break;
  }
  return level;
}

// A written-out node: one 16-byte load per level of the walk. c0 carries the pivot code in its top two bits
// (0 = unset: the node is still childless; 1..3 = dimension 0..2); child ids are 30 bits, TOP_NONE = empty slot.
struct __align__(16) TopNode
{
  unsigned long long key;
  unsigned c0, c1;
};
constexpr unsigned TOP_NONE = 0x3FFFFFFFu;
__device__ __forceinline__ int top_pivot(unsigned c0) { return (int)(c0 >> 30) - 1; }  // -1 = unset
__device__ __forceinline__ unsigned top_child(const TopNode& nd, int sd) { return (sd ? nd.c1 : nd.c0) & TOP_NONE; }
__device__ __forceinline__ TopNode top_make(unsigned long long key, int pivot, int child0, int child1)
{
  TopNode nd;
  nd.key = key;
  nd.c0 = (child0 == NONE ? TOP_NONE : (unsigned)child0) | ((unsigned)(pivot + 1) << 30);
  nd.c1 = child1 == NONE ? TOP_NONE : (unsigned)child1;
  return nd;
}

struct FastArgs
{
  TreeOut* out;
  const unsigned long long* nkey;  // [K] creation order
  TopNode* top;                    // [M2] the written-out tree
  int* bcount;                     // [2 * M2 + 1] keys per bucket; scanned in place into offsets
  int* lists;                      // bucket ids by size class: S at 0, M at list_m, L at list_l; counts in cls[0..2]
  int* cls;
  int list_m, list_l;
  int* bslot;                      // [K] bucket of every key routed in this step
  int* bpos;                       // [K] arrival position inside the bucket
  int* bids;                       // [K] keys grouped by bucket
  int part_rank, part_n;           // distributed filter: this rank counts the buckets under nodes u, u % part_n == part_rank
};

// keys [lo, hi) of the step and the size m of the tree they walk: step B = [M1, M2) over M1, step C = [M2, K) over M2
__device__ __forceinline__ void step_range(int K, bool step_c, int* lo, int* hi, int* m)
{
  const int m2 = min(K, top2_size(K));
  if (step_c)
  {
    *lo = m2;
    *hi = K;
    *m = m2;
  }
  else
  {
    *lo = min(K, M_TOP1);
    *hi = m2;
    *m = min(K, M_TOP1);
  }
}

// ---- A. top tree of the first M_TOP1 keys (one CTA), written to the node arrays
__global__ void __launch_bounds__(1024) k_fast_top(const FastArgs a)
{
  extern __shared__ __align__(16) unsigned char s_raw[];
  unsigned long long* key = reinterpret_cast<unsigned long long*>(s_raw);
  int* cur = reinterpret_cast<int*>(key + M_TOP1);
  int* child = cur + M_TOP1;
  int* first_at = child + 2 * M_TOP1;
  const int K = a.out->K;
  const int m = min(K, M_TOP1);
  for (int v = threadIdx.x; v < m; v += blockDim.x)
    key[v] = a.nkey[v];
  __syncthreads();
  const int levels = smem_tree<false>(key, m, cur, child, first_at);
  for (int v = threadIdx.x; v < m; v += blockDim.x)
    a.top[v] = top_make(key[v], first_at[v] != NONE ? pivot_dim(key[v], key[first_at[v]]) : -1, child[2 * v], child[2 * v + 1]);
  if (threadIdx.x == 0)
  {
    a.out->levels = levels;
    a.out->fast_ok = 1;
  }
}

// ---- every key of the step finds its bucket: a read-only walk down the written-out tree (L1 / L2 resident)
template <bool STEP_C>
__global__ void __launch_bounds__(256) k_fast_descend(const FastArgs a)
{
  int lo, hi, m;
  step_range(a.out->K, STEP_C, &lo, &hi, &m);
  for (int v = lo + blockIdx.x * blockDim.x + threadIdx.x; v < hi; v += gridDim.x * blockDim.x)
  {
    const unsigned long long kv = a.nkey[v];
    int u = 0, b;
    for (;;)
    {
      const uint4 raw = __ldg(reinterpret_cast<const uint4*>(&a.top[u]));
      const int p = top_pivot(raw.z);
      if (p < 0)
      {
        b = 2 * u;  // a node that is still childless: the whole bucket hangs under it
        break;
      }
      const unsigned long long ku = (unsigned long long)raw.x | ((unsigned long long)raw.y << 32);
      const int sd = key_dim(kv, p) > key_dim(ku, p) ? 1 : 0;
      const unsigned c = (sd ? raw.w : raw.z) & TOP_NONE;
      if (c == TOP_NONE)
      {
        b = 2 * u + sd;
        break;
      }
      u = (int)c;
    }
    a.bslot[v] = b;
    a.bpos[v] = atomicAdd(&a.bcount[b], 1);
  }
}

template <bool STEP_C>
__global__ void __launch_bounds__(256) k_fast_scatter(const FastArgs a)
{
  int lo, hi, m;
  step_range(a.out->K, STEP_C, &lo, &hi, &m);
  for (int v = lo + blockIdx.x * blockDim.x + threadIdx.x; v < hi; v += gridDim.x * blockDim.x)
    a.bids[a.bcount[a.bslot[v]] + a.bpos[v]] = v;  // bcount holds the exclusive offsets by now
}

// which keys make up bucket b of a tree of m written-out nodes: returns n (0: no bucket hangs here) and (off, cnt,
// unset). Only called between the kernels that write the tree.
__device__ __forceinline__ int bucket_shape(const FastArgs& a, int b, int m, int* off, int* cnt, bool* unset)
{
  const int u = b >> 1, sd = b & 1;
  if (u >= m)
    return 0;
  const TopNode nd = a.top[u];
  *unset = top_pivot(nd.c0) < 0;
  if (*unset ? sd == 1 : top_child(nd, sd) != TOP_NONE)
    return 0;
  *off = a.bcount[b];
  *cnt = a.bcount[b + 1] - *off;
  return *cnt + (*unset ? 1 : 0);
}

// ---- buckets by size class (a CTA of the right size per bucket, and no CTA at all for the many empty slots)
template <bool STEP_C>
__global__ void __launch_bounds__(256) k_fast_classify(const FastArgs a)
{
  int lo, hi, m;
  step_range(a.out->K, STEP_C, &lo, &hi, &m);
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= 2 * m)
    return;
  if (STEP_C && (b >> 1) % a.part_n != a.part_rank)
    return;
  int off = 0, cnt = 0;
  bool unset = false;
  const int n = bucket_shape(a, b, m, &off, &cnt, &unset);
  if (n > FAST_CAP_L)
    a.out->fast_ok = 0;  // outgrew shared memory: the grid-synchronous build takes over
  else if (n > FAST_CAP_M)
    a.lists[a.list_l + atomicAdd(&a.cls[2], 1)] = b;
  else if (n > FAST_CAP_S)
    a.lists[a.list_m + atomicAdd(&a.cls[1], 1)] = b;
  else if (STEP_C ? n > FAST_CAP_W : cnt > 0)
    a.lists[atomicAdd(&a.cls[0], 1)] = b;
}

// ---- one CTA per listed bucket: sort by creation index, grow the subtree; step B writes it out, step C counts its leaves
template <int CAP, int CLS, int THREADS, bool STEP_C>
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
  int lo, hi, m;
  step_range(a.out->K, STEP_C, &lo, &hi, &m);
  if ((int)blockIdx.x >= a.cls[CLS])
    return;
  const int b = a.lists[(CLS == 0 ? 0 : CLS == 1 ? a.list_m : a.list_l) + blockIdx.x], u = b >> 1;
  int off = 0, cnt = 0;
  bool unset = false;
  const int n = bucket_shape(a, b, m, &off, &cnt, &unset);
  if (n > CAP || n < 1)
    return;  // (cannot happen: the class lists are built from the same shapes)
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
  const int levels = smem_tree<false>(key, n, cur, child, first_at);
  if (!STEP_C)
  {
    // the subtree joins the written-out tree: node ids[j] gets its pivot and children, the slot gets its root
    for (int j = threadIdx.x; j < n; j += THREADS)
      a.top[ids[j]] = top_make(key[j], first_at[j] != NONE ? pivot_dim(key[j], key[first_at[j]]) : -1,
                               child[2 * j] == NONE ? NONE : ids[child[2 * j]],
                               child[2 * j + 1] == NONE ? NONE : ids[child[2 * j + 1]]);
    if (!unset && threadIdx.x == 0)
    {
      // the parent's pivot is final; the two slots of a node are different words, each written by its own bucket's CTA
      if (b & 1)
        a.top[u].c1 = (unsigned)ids[0];
      else
        a.top[u].c0 = (a.top[u].c0 & ~TOP_NONE) | (unsigned)ids[0];
    }
    if (threadIdx.x == 0)
      atomicMax(&a.out->levels, levels);
    return;
  }
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

// ---- step C, buckets of at most FAST_CAP_W keys: one WARP per bucket, eight buckets per CTA
constexpr int FAST_W_WARPS = 8;
__global__ void __launch_bounds__(FAST_W_WARPS * 32) k_fast_bucket_warp(const FastArgs a)
{
  __shared__ unsigned long long s_key[FAST_W_WARPS][FAST_CAP_W];
  __shared__ int s_cur[FAST_W_WARPS][FAST_CAP_W], s_first[FAST_W_WARPS][FAST_CAP_W];
  __shared__ int s_child[FAST_W_WARPS][2 * FAST_CAP_W];
  if (!a.out->fast_ok)
    return;
  int lo, hi, m;
  step_range(a.out->K, true, &lo, &hi, &m);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * FAST_W_WARPS + warp, u = b >> 1;
  if (u % a.part_n != a.part_rank)
    return;
  int off = 0, cnt = 0;
  bool unset = false;
  const int n = bucket_shape(a, b, m, &off, &cnt, &unset);
  if (n == 0 || n > FAST_CAP_W)
    return;
  if (n == 1)
  {
    if (lane == 0)
      atomicAdd(&a.out->leaf, 1);
    return;
  }
  unsigned long long* key = s_key[warp];
  // rank sort of the (distinct) creation indices, two per lane
  int mine[2];
#pragma unroll
  for (int q = 0; q < 2; q++)
  {
    const int i = lane + 32 * q;
    int id = INT_MAX;
    if (i < n)
      id = unset ? (i == 0 ? u : a.bids[off + i - 1]) : a.bids[off + i];
    mine[q] = id;
    s_cur[warp][i] = id;
  }
  __syncwarp();
#pragma unroll
  for (int q = 0; q < 2; q++)
  {
    if (lane + 32 * q < n)
    {
      int r = 0;
      for (int j = 0; j < n; j++)
        r += s_cur[warp][j] < mine[q] ? 1 : 0;
      key[r] = a.nkey[mine[q]];
    }
  }
  __syncwarp();
  const int levels = smem_tree<true>(key, n, s_cur[warp], s_child[warp], s_first[warp]);
  int leaves = 0;
  for (int v = lane; v < n; v += 32)
    leaves += (s_child[warp][2 * v] == NONE && s_child[warp][2 * v + 1] == NONE) ? 1 : 0;
  leaves = warp_sum_i(leaves);
  if (lane == 0)
  {
    atomicAdd(&a.out->leaf, leaves);
    atomicMax(&a.out->bucket_levels, levels);
  }
}
