// K-d3/K-d4: the pose histogram "kd-tree" of a sample set, its KLD leaf count, cluster labels and the cluster /
// set statistics. Reference: src/amcl/pf/pf_kdtree.cpp:49-194, src/amcl/pf/particle_filter.cpp:170-220, 505-636.
//
// What the reference's insertion-order BST observably produces:
//   * one node per distinct integer key floor(pose / (0.5, 0.5, 10 deg)), in order of first occurrence;
//   * leaf_count = number of nodes that no later key was ever routed through. A node fixes its pivot dimension
//     from the first foreign key that reaches it (largest |delta|, lowest dimension on ties) and routes by
//     key[pivot] > node.key[pivot];
//   * cluster labels = connected components under the 26-neighbourhood, numbered by their earliest node.
// Device construction: a hash table gives every distinct key its first-occurrence sample index; a scan turns
// that into creation-order node ids; the tree is then grown for ALL keys at once, one level per step (every
// key still travelling sits at the node it has reached; the smallest id among the keys at a node is that node's
// first foreign key, the smallest id per side becomes the child) inside one cooperative kernel; components
// come from a lock-free union-find that always hooks the larger root under the smaller, so a root is its
// component's earliest node and ranking the roots reproduces the reference's numbering.
#include <cooperative_groups.h>

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "pf.cuh"
#include "scan.cuh"

namespace cg = cooperative_groups;
using namespace bamcl;

namespace
{
constexpr unsigned long long EMPTY_KEY = ~0ull;
constexpr int NONE = INT_MAX;
constexpr int ROW = 10;  // per cluster: count, weight, m0, m1, m2, m3, c00, c01, c10, c11

struct TreeOut
{
  int K;     // nodes
  int C;     // clusters
  int leaf;  // childless nodes
  int conv;  // samples within dist_threshold of the mean
  int err;   // bit0 key out of the packed range, bit1 cluster rows exhausted
  int level_flags[3];
  int levels;
  int fast_ok;        // the bucketed shared-memory build (tree_fast.cuh) produced the leaf count
  int bucket_levels;  // deepest bucket subtree of that build
  int pad;
  double set_sums[12];  // weight, m0..m3, c00..c11, count, sum x, sum y
};

__host__ __device__ __forceinline__ unsigned long long pack_key(int kx, int ky, int kt)
{
  return (unsigned long long)(unsigned)(kx + (1 << 20)) | ((unsigned long long)(unsigned)(ky + (1 << 20)) << 21) |
         ((unsigned long long)(unsigned)(kt + (1 << 21)) << 42);
}
__device__ __forceinline__ void unpack_key(unsigned long long k, int* kx, int* ky, int* kt)
{
  *kx = (int)(k & 0x1FFFFF) - (1 << 20);
  *ky = (int)((k >> 21) & 0x1FFFFF) - (1 << 20);
  *kt = (int)((k >> 42) & 0x3FFFFF) - (1 << 21);
}
__device__ __forceinline__ unsigned long long mix64(unsigned long long z)
{
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}

// ---- 1. distinct keys and their first occurrence
__global__ void __launch_bounds__(256) k_hash_insert(const double* __restrict__ x, const double* __restrict__ y,
                                                     const double* __restrict__ th, int n, double cx, double cy, double ct,
                                                     unsigned long long* __restrict__ hkeys, int* __restrict__ hfirst,
                                                     unsigned hmask, int* __restrict__ slot_out, TreeOut* out)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  // pf_kdtree.cpp:51-54
  const double fx = floor(x[i] / cx), fy = floor(y[i] / cy), ft = floor(th[i] / ct);
  if (!(fabs(fx) < 1048575.0) || !(fabs(fy) < 1048575.0) || !(fabs(ft) < 2097151.0))
  {
    atomicOr(&out->err, 1);
    slot_out[i] = 0;
    return;
  }
  const unsigned long long key = pack_key((int)fx, (int)fy, (int)ft);
  unsigned s = (unsigned)mix64(key) & hmask;
  for (;;)
  {
    const unsigned long long cur = hkeys[s];
    if (cur == key)
      break;
    if (cur == EMPTY_KEY)
    {
      const unsigned long long old = atomicCAS(&hkeys[s], EMPTY_KEY, key);
      if (old == EMPTY_KEY || old == key)
        break;
    }
    s = (s + 1) & hmask;
  }
  atomicMin(&hfirst[s], i);
  slot_out[i] = (int)s;
}

__global__ void __launch_bounds__(256) k_flag_first(const int* __restrict__ slot, const int* __restrict__ hfirst,
                                                    int* __restrict__ flag, int n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    flag[i] = hfirst[slot[i]] == i ? 1 : 0;
}

__global__ void __launch_bounds__(256) k_make_nodes(const int* __restrict__ slot, const int* __restrict__ hfirst,
                                                    const int* __restrict__ rank, const unsigned long long* __restrict__ hkeys,
                                                    int* __restrict__ hnode, unsigned long long* __restrict__ nkey,
                                                    int* __restrict__ nfirst, int* __restrict__ parent, int n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  const int s = slot[i];
  if (hfirst[s] != i)
    return;
  const int id = rank[i];
  nkey[id] = hkeys[s];
  nfirst[id] = i;
  hnode[s] = id;
  parent[id] = id;  // union-find forest of the clustering step: every node its own root
}

// ---- 1b. the same, on a dense grid of pose bins. When the occupied bins fit a box of at most DENSE_CAP cells (a 200 m
// map at 0.5 m x 0.5 m x 10 deg is 5.8M cells) a bin is addressed directly: no probing, no CAS, the 26 neighbour probes
// of the clustering step are plain loads, and on a distributed filter ONE all-reduce (min) of the grid of first
// occurrences gives every rank the global creation order.
constexpr long DENSE_CAP = 1l << 26;
constexpr int BOX_MARGIN = 2;
struct Box
{
  int mn[3], dim[3];
};
__device__ __forceinline__ bool pose_key(double x, double y, double th, double cx, double cy, double ct, int* kx, int* ky, int* kt)
{
  // pf_kdtree.cpp:51-54
  const double fx = floor(x / cx), fy = floor(y / cy), ft = floor(th / ct);
  if (!(fabs(fx) < 1048575.0) || !(fabs(fy) < 1048575.0) || !(fabs(ft) < 2097151.0))
    return false;
  *kx = (int)fx;
  *ky = (int)fy;
  *kt = (int)ft;
  return true;
}
// box6 = {min kx, min ky, min kt, -max kx, -max ky, -max kt}: one MIN reduction serves both ends
__global__ void __launch_bounds__(256) k_key_bbox(const double* __restrict__ x, const double* __restrict__ y,
                                                  const double* __restrict__ th, int n, double cx, double cy, double ct,
                                                  int* __restrict__ box6)
{
  int v[6] = {INT_MAX, INT_MAX, INT_MAX, INT_MAX, INT_MAX, INT_MAX};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
  {
    int k[3];
    if (!pose_key(x[i], y[i], th[i], cx, cy, ct, &k[0], &k[1], &k[2]))
      continue;
#pragma unroll
    for (int d = 0; d < 3; d++)
    {
      v[d] = min(v[d], k[d]);
      v[3 + d] = min(v[3 + d], -k[d]);
    }
  }
#pragma unroll
  for (int d = 0; d < 6; d++)
  {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      v[d] = min(v[d], __shfl_xor_sync(0xffffffffu, v[d], o));
    if ((threadIdx.x & 31) == 0 && v[d] != INT_MAX)
      atomicMin(&box6[d], v[d]);
  }
}
__global__ void __launch_bounds__(256) k_grid_first(const double* __restrict__ x, const double* __restrict__ y,
                                                    const double* __restrict__ th, int n, long off, double cx, double cy, double ct,
                                                    const Box box, int* __restrict__ g_first, int* __restrict__ cell_out,
                                                    TreeOut* out)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  int k[3];
  if (!pose_key(x[i], y[i], th[i], cx, cy, ct, &k[0], &k[1], &k[2]))
  {
    atomicOr(&out->err, 1);
    cell_out[i] = 0;
    return;
  }
  const int a = k[0] - box.mn[0], b = k[1] - box.mn[1], c = k[2] - box.mn[2];
  if ((unsigned)a >= (unsigned)box.dim[0] || (unsigned)b >= (unsigned)box.dim[1] || (unsigned)c >= (unsigned)box.dim[2])
  {
    atomicOr(&out->err, 4);  // outside the sticky box: the host re-fits it and repeats
    cell_out[i] = 0;
    return;
  }
  const int cell = a + box.dim[0] * (b + box.dim[1] * c);
  atomicMin(&g_first[cell], (int)(off + i));
  cell_out[i] = cell;
}
// bit f of `bits` = "global sample f is the first occurrence of its bin"
__global__ void __launch_bounds__(256) k_grid_mark(const int* __restrict__ g_first, long cells, int n_global,
                                                   unsigned* __restrict__ bits)
{
  const long c = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cells)
    return;
  const int f = g_first[c];
  if (f < n_global)
    atomicOr(&bits[f >> 5], 1u << (f & 31));
}
__global__ void __launch_bounds__(256) k_popc(const unsigned* __restrict__ bits, int words, int* __restrict__ cnt)
{
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w < words)
    cnt[w] = __popc(bits[w]);
}
// creation-order node id of every occupied bin = number of first occurrences before its own
__global__ void __launch_bounds__(256) k_grid_nodes(const int* __restrict__ g_first, long cells, int n_global,
                                                    const unsigned* __restrict__ bits, const int* __restrict__ pre, const Box box,
                                                    int* __restrict__ g_node, unsigned long long* __restrict__ nkey,
                                                    int* __restrict__ nfirst, int* __restrict__ parent)
{
  const long c = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cells)
    return;
  const int f = g_first[c];
  if (f >= n_global)
  {
    g_node[c] = -1;
    return;
  }
  const int id = pre[f >> 5] + __popc(bits[f >> 5] & ((1u << (f & 31)) - 1u));
  g_node[c] = id;
  const int a = (int)(c % box.dim[0]), b = (int)((c / box.dim[0]) % box.dim[1]), t = (int)(c / ((long)box.dim[0] * box.dim[1]));
  nkey[id] = pack_key(a + box.mn[0], b + box.mn[1], t + box.mn[2]);
  nfirst[id] = f;
  parent[id] = id;
}

// ---- 2. the tree, one level per step
struct TreeArgs
{
  TreeOut* out;
  const unsigned long long* nkey;
  int* cur;       // node a key currently sits at, -1 once it has become a tree node itself
  int* child;     // [2 * K]
  int* first_at;  // smallest id among the keys that reached a node = its first foreign key
  signed char* side;
};

#include "tree_fast.cuh"

// General level-synchronous build over the whole grid: only runs when the fast path gave up.
__global__ void __launch_bounds__(256) k_tree_build(const TreeArgs a)
{
  cg::grid_group grid = cg::this_grid();
  if (a.out->fast_ok)
    return;
  const int K = a.out->K;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  if (tid == 0)
    a.out->leaf = a.out->bucket_levels = 0;  // whatever an abandoned fast build had counted
  for (int v = tid; v < K; v += nth)
  {
    a.child[2 * v] = a.child[2 * v + 1] = NONE;
    a.first_at[v] = (v == 0 && K > 1) ? 1 : NONE;
    a.cur[v] = v == 0 ? -1 : 0;
  }
  grid.sync();
  int level = 0;
  for (;;)
  {
    // every travelling key picks its side at the node it sits at (pf_kdtree.cpp:131-150)
    for (int v = tid; v < K; v += nth)
    {
      const int u = a.cur[v];
      if (u < 0)
        continue;
      const int f = a.first_at[u];
      int ku[3], kf[3], kv[3];
      unpack_key(a.nkey[u], &ku[0], &ku[1], &ku[2]);
      unpack_key(a.nkey[f], &kf[0], &kf[1], &kf[2]);
      unpack_key(a.nkey[v], &kv[0], &kv[1], &kv[2]);
      int pivot = 0, max_split = 0;
#pragma unroll
      for (int d = 0; d < 3; d++)
      {
        const int split = abs(kf[d] - ku[d]);
        if (split > max_split)
        {
          max_split = split;
          pivot = d;
        }
      }
      const int sd = kv[pivot] > ku[pivot] ? 1 : 0;
      a.side[v] = (signed char)sd;
      atomicMin(&a.child[2 * u + sd], v);
    }
    grid.sync();
    bool any = false;
    for (int v = tid; v < K; v += nth)
    {
      const int u = a.cur[v];
      if (u < 0)
        continue;
      const int c = a.child[2 * u + a.side[v]];
      if (c == v)
        a.cur[v] = -1;
      else
      {
        a.cur[v] = c;
        atomicMin(&a.first_at[c], v);
        any = true;
      }
    }
    if (any)
      a.out->level_flags[level % 3] = 1;
    grid.sync();
    const int more = *((volatile int*)&a.out->level_flags[level % 3]);
    if (tid == 0)
      a.out->level_flags[(level + 1) % 3] = 0;
    level++;
    if (!more)
      break;
  }
  int cnt = 0;
  for (int v = tid; v < K; v += nth)
    cnt += (a.child[2 * v] == NONE && a.child[2 * v + 1] == NONE) ? 1 : 0;
  cnt = warp_sum_i(cnt);
  if ((threadIdx.x & 31) == 0 && cnt)
    atomicAdd(&a.out->leaf, cnt);
  if (tid == 0)
    a.out->levels = level;
}

// ---- 3. connected components (pf_kdtree.cpp:58-76, 169-194)
// CA = read the forest through L1 (ld.global.ca). A stale entry is an older parent pointer, which is still an ancestor
// (roots only ever get hooked under smaller ids, compression only skips ancestors), and every hook is decided by an
// atomicCAS on the true value, so staleness costs at most a retry; it spreads the reads of a large component's root
// over the SMs' L1s instead of one L2 slice.
template <bool CA>
__device__ __forceinline__ int uf_load(const int* p)
{
  return CA ? __ldca(p) : *((const volatile int*)p);
}
template <bool CA>
__device__ __forceinline__ int uf_find(int* parent, int v)
{
  int cur = uf_load<CA>(&parent[v]);
  if (cur != v)
  {
    int prev = v, next;
    while (cur > (next = uf_load<CA>(&parent[cur])))
    {
      parent[prev] = next;
      prev = cur;
      cur = next;
    }
  }
  return cur;
}
// find(v) when parent[v] has been loaded already (the caller issued that load early, next to independent ones)
template <bool CA>
__device__ __forceinline__ int uf_find_from(int* parent, int v, int cur)
{
  if (cur != v)
  {
    int prev = v, next;
    while (cur > (next = uf_load<CA>(&parent[cur])))
    {
      parent[prev] = next;
      prev = cur;
      cur = next;
    }
  }
  return cur;
}
template <bool CA>
__device__ __forceinline__ void uf_hook(int* parent, int ra, int rb)
{
  bool repeat;
  do
  {
    repeat = false;
    if (ra != rb)
    {
      int ret;
      if (ra < rb)
      {
        if ((ret = atomicCAS(&parent[rb], rb, ra)) != rb)
        {
          rb = ret;
          repeat = true;
        }
      }
      else
      {
        if ((ret = atomicCAS(&parent[ra], ra, rb)) != ra)
        {
          ra = ret;
          repeat = true;
        }
      }
    }
  } while (repeat);
}
template <bool CA>
__device__ __forceinline__ void uf_union(int* parent, int a, int b)
{
  // Hook the larger root under the smaller, so a root is always its component's earliest node. A failed CAS returns
  // the parent the loser now has: climbing from there one CAS at a time measured FASTER than re-running find through
  // L2 (the dense retry traffic on the forest is what costs), so that is what the loop does.
  int ra = uf_find<CA>(parent, a), rb = uf_find<CA>(parent, b);
  bool repeat;
  do
  {
    repeat = false;
    if (ra != rb)
    {
      int ret;
      if (ra < rb)
      {
        if ((ret = atomicCAS(&parent[rb], rb, ra)) != rb)
        {
          rb = ret;
          repeat = true;
        }
      }
      else
      {
        if ((ret = atomicCAS(&parent[ra], ra, rb)) != ra)
        {
          ra = ret;
          repeat = true;
        }
      }
    }
  } while (repeat);
}
template <bool CA>
__global__ void __launch_bounds__(256) k_cc_union(const TreeOut* __restrict__ out, const unsigned long long* __restrict__ nkey,
                                                  const unsigned long long* __restrict__ hkeys, const int* __restrict__ hnode,
                                                  unsigned hmask, int* parent, int n_max)
{
  const int K = out->K;
  // one thread per (node, neighbour offset in the positive half space): 13 offsets
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int v = (int)(t / 13);
  const int o = (int)(t % 13) + 14;  // offsets 14..26 of the 27 = lexicographically after (0,0,0)
  if (v >= K || v >= n_max)
    return;
  int kx, ky, kt;
  unpack_key(nkey[v], &kx, &ky, &kt);
  const int dx = o / 9 - 1, dy = (o % 9) / 3 - 1, dt = (o % 9) % 3 - 1;
  const int nx = kx + dx, ny = ky + dy, nt = kt + dt;
  if (abs(nx) >= (1 << 20) - 1 || abs(ny) >= (1 << 20) - 1 || abs(nt) >= (1 << 21) - 1)
    return;
  const unsigned long long key = pack_key(nx, ny, nt);
  unsigned s = (unsigned)mix64(key) & hmask;
  for (;;)
  {
    const unsigned long long cur = hkeys[s];
    if (cur == EMPTY_KEY)
      return;
    if (cur == key)
      break;
    s = (s + 1) & hmask;
  }
  uf_union<CA>(parent, v, hnode[s]);
}

// mode 0: every edge (single GPU). Distributed filter: the grid is cut into slabs of rows (second dimension), one per
// rank -- mode 1: only the edges with BOTH ends in slab `me` (each rank unites its own slab); the ranks' forests are then
// merged with an all-reduce (min) of the parent array (a rank only ever lowers the parents of its own slab's nodes, every
// other entry still points at itself); mode 2: the edges that cross a slab boundary, united by every rank on the merged
// forest. A root is always its component's smallest node id, so the result does not depend on the split.
template <bool CA>
__global__ void __launch_bounds__(256) k_cc_union_dense(const TreeOut* __restrict__ out, const unsigned long long* __restrict__ nkey,
                                                        const Box box, const int* __restrict__ g_node, int* parent, int n_max,
                                                        int mode, int me, int rows_per_slab)
{
  const int K = out->K;
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int v = (int)(t / 13);
  const int o = (int)(t % 13) + 14;  // offsets 14..26 of the 27 = lexicographically after (0,0,0)
  if (v >= K || v >= n_max)
    return;
  const int pv = uf_load<CA>(&parent[v]);  // independent of the key: both loads travel together
  int kx, ky, kt;
  unpack_key(nkey[v], &kx, &ky, &kt);
  const int b0 = ky - box.mn[1];
  if (mode == 1 && b0 / rows_per_slab != me)
    return;
  const int a = kx - box.mn[0] + (o / 9 - 1), b = b0 + ((o % 9) / 3 - 1), c = kt - box.mn[2] + ((o % 9) % 3 - 1);
  if ((unsigned)a >= (unsigned)box.dim[0] || (unsigned)b >= (unsigned)box.dim[1] || (unsigned)c >= (unsigned)box.dim[2])
    return;
  if (mode != 0)
  {
    const bool same = b / rows_per_slab == b0 / rows_per_slab;
    if ((mode == 1) != same)
      return;
  }
  const int u = __ldg(&g_node[a + box.dim[0] * (b + box.dim[1] * c)]);
  if (u < 0)
    return;
  // both first hops in flight together (pv was requested at the top, next to the key), then the two walks
  const int pu = uf_load<CA>(&parent[u]);
  const int ra = uf_find_from<CA>(parent, v, pv), rb = uf_find_from<CA>(parent, u, pu);
  uf_hook<CA>(parent, ra, rb);
}

// Head start for the unions (the initialisation of ECL-CC): every node points at its smallest occupied neighbour with a
// smaller id. Pointers only ever go to smaller ids, so this is a valid forest whose roots are the local minima, and the
// union pass starts from a few large trees instead of K singletons. mode / me / rows_per_slab as in k_cc_union_dense.
__global__ void __launch_bounds__(256) k_cc_init(const TreeOut* __restrict__ out, const unsigned long long* __restrict__ nkey,
                                                 const Box box, const int* __restrict__ g_node, int* __restrict__ parent, int n_max,
                                                 int mode, int me, int rows_per_slab)
{
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= out->K || v >= n_max)
    return;
  int kx, ky, kt;
  unpack_key(nkey[v], &kx, &ky, &kt);
  const int a0 = kx - box.mn[0], b0 = ky - box.mn[1], c0 = kt - box.mn[2];
  if (mode == 1 && b0 / rows_per_slab != me)
    return;
  int best = v;
#pragma unroll
  for (int o = 0; o < 27; o++)
  {
    if (o == 13)
      continue;
    const int a = a0 + (o / 9 - 1), b = b0 + ((o % 9) / 3 - 1), c = c0 + ((o % 9) % 3 - 1);
    if ((unsigned)a >= (unsigned)box.dim[0] || (unsigned)b >= (unsigned)box.dim[1] || (unsigned)c >= (unsigned)box.dim[2])
      continue;
    if (mode == 1 && b / rows_per_slab != me)
      continue;
    const int u = __ldg(&g_node[a + box.dim[0] * (b + box.dim[1] * c)]);
    if (u >= 0 && u < best)
      best = u;
  }
  parent[v] = best;
}

// The same unions in bin order: one thread per grid cell, the 13 forward neighbours are plain (coalesced, L1-shared)
// loads of the neighbouring rows. Used when the grid is reasonably full; the node-order kernel above serves sparse sets.
template <bool CA>
__global__ void __launch_bounds__(256) k_cc_union_cells(const Box box, const int* __restrict__ g_node, long cells, int* parent)
{
  const long c = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cells)
    return;
  const int v = __ldg(&g_node[c]);
  if (v < 0)
    return;
  const int d0 = box.dim[0], d1 = box.dim[1], d2 = box.dim[2];
  const int a = (int)(c % d0), b = (int)((c / d0) % d1), t = (int)(c / ((long)d0 * d1));
#pragma unroll
  for (int o = 14; o < 27; o++)
  {
    const int da = o / 9 - 1, db = (o % 9) / 3 - 1, dt = (o % 9) % 3 - 1;
    if ((unsigned)(a + da) >= (unsigned)d0 || (unsigned)(b + db) >= (unsigned)d1 || (unsigned)(t + dt) >= (unsigned)d2)
      continue;
    const int u = __ldg(&g_node[c + da + (long)d0 * (db + (long)d1 * dt)]);
    if (u >= 0)
      uf_union<CA>(parent, v, u);
  }
}

// Read-only find: a compressing find here would race with other threads' finds (a stale intermediate ancestor
// written over an already final root), so the final roots go to a separate array.
__global__ void __launch_bounds__(256) k_cc_roots(const TreeOut* __restrict__ out, const int* __restrict__ parent,
                                                  int* __restrict__ root, int* __restrict__ is_root, int n)
{
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n)
    return;
  const int K = out->K;
  if (v >= K)
  {
    is_root[v] = 0;
    return;
  }
  int r = v, up;
  while ((up = parent[r]) != r)
    r = up;
  root[v] = r;
  is_root[v] = r == v ? 1 : 0;
}

// ---- 4. labels + statistics (particle_filter.cpp:505-636)
struct StatsArgs
{
  const double *x, *y, *th, *w;
  int n;
  const int* slot;
  const int* hnode;
  const int* parent;
  const int* cid;
  int* label;
  double* rows;  // [cap][ROW]
  int cap;
  TreeOut* out;
  double* set_part;  // [blocks][12] per-block set sums (reduced in a fixed order by k_set_sums_final)
};
constexpr int ST_BLOCK = 256;
constexpr int ST_CHUNKS = 8;  // chunks of 32 samples per warp

__device__ __forceinline__ void flush_rows(double* rows, int cap, int label, const double (&v)[ROW], TreeOut* out)
{
  if (label < 0)
    return;
  if (label >= cap)
  {
    atomicOr(&out->err, 2);
    return;
  }
#pragma unroll
  for (int q = 0; q < ROW; q++)
    atomicAdd(&rows[(size_t)label * ROW + q], v[q]);
}

__global__ void __launch_bounds__(ST_BLOCK) k_labels_stats(const StatsArgs a)
{
  __shared__ double s_set[ST_BLOCK / 32][12];
  __shared__ double s_open[ST_BLOCK / 32][ROW];
  __shared__ int s_open_label[ST_BLOCK / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int warp_base = (blockIdx.x * (ST_BLOCK / 32) + warp) * (32 * ST_CHUNKS);
  double setv[12];
#pragma unroll
  for (int q = 0; q < 12; q++)
    setv[q] = 0.0;
  double open[ROW];
#pragma unroll
  for (int q = 0; q < ROW; q++)
    open[q] = 0.0;
  int open_label = -1;  // warp-uniform
  for (int ch = 0; ch < ST_CHUNKS; ch++)
  {
    const int i = warp_base + ch * 32 + lane;
    const bool live = i < a.n;
    int label = -1;
    double v[ROW];
#pragma unroll
    for (int q = 0; q < ROW; q++)
      v[q] = 0.0;
    if (live)
    {
      // (node < 0: the sample's bin lies outside the sticky dense box; the build is flagged and repeated with a re-fitted box)
      const int node = a.hnode[a.slot[i]];
      label = node >= 0 ? a.cid[a.parent[node]] : -1;
      a.label[i] = label;
      const double px = a.x[i], py = a.y[i], pt = a.th[i], pw = a.w[i];
      double sn, cs;
      sincos(pt, &sn, &cs);
      v[0] = 1.0;
      v[1] = pw;
      v[2] = pw * px;
      v[3] = pw * py;
      v[4] = pw * cs;
      v[5] = pw * sn;
      v[6] = pw * px * px;
      v[7] = pw * px * py;
      v[8] = pw * py * px;
      v[9] = pw * py * py;
      setv[0] += pw;
      setv[1] += v[2];
      setv[2] += v[3];
      setv[3] += v[4];
      setv[4] += v[5];
      setv[5] += v[6];
      setv[6] += v[7];
      setv[7] += v[8];
      setv[8] += v[9];
      setv[9] += 1.0;
      setv[10] += px;
      setv[11] += py;
    }
    const unsigned live_mask = __ballot_sync(0xffffffffu, live);
    if (live_mask == 0)
      break;
    const int first_label = __shfl_sync(0xffffffffu, label, __ffs(live_mask) - 1);
    const bool uniform = __all_sync(0xffffffffu, !live || label == first_label);
    if (uniform && (open_label == first_label || open_label < 0))
    {
      open_label = first_label;
#pragma unroll
      for (int q = 0; q < ROW; q++)
        open[q] += v[q];
      continue;
    }
    // mixed chunk: one warp reduction per distinct label. The first label seen becomes the warp's open row, so a
    // dominant cluster is accumulated in registers instead of hammering one row with atomics.
    if (open_label < 0)
      open_label = first_label;
    unsigned todo = live_mask;
    while (todo)
    {
      const int lead = __ffs(todo) - 1;
      const int lab = __shfl_sync(0xffffffffu, label, lead);
      const bool mine = live && label == lab;
      double r[ROW];
#pragma unroll
      for (int q = 0; q < ROW; q++)
        r[q] = warp_sum(mine ? v[q] : 0.0);
      if (lab == open_label)
      {
        if (lane == 0)
        {
#pragma unroll
          for (int q = 0; q < ROW; q++)
            open[q] += r[q];
        }
      }
      else if (lane == 0)
        flush_rows(a.rows, a.cap, lab, r, a.out);
      todo &= ~__ballot_sync(0xffffffffu, mine);
    }
  }
  // warp totals
#pragma unroll
  for (int q = 0; q < ROW; q++)
    open[q] = warp_sum(open[q]);
#pragma unroll
  for (int q = 0; q < 12; q++)
    setv[q] = warp_sum(setv[q]);
  if (lane == 0)
  {
#pragma unroll
    for (int q = 0; q < ROW; q++)
      s_open[warp][q] = open[q];
    s_open_label[warp] = open_label;
#pragma unroll
    for (int q = 0; q < 12; q++)
      s_set[warp][q] = setv[q];
  }
  __syncthreads();
  if (warp == 0)
  {
    // merge the warps' open rows that share a label (normally all of them)
    if (lane < 12)
    {
      double t = 0.0;
      for (int k = 0; k < ST_BLOCK / 32; k++)
        t += s_set[k][lane];
      a.set_part[(size_t)blockIdx.x * 12 + lane] = t;
    }
    if (lane < ROW)
    {
      for (int k = 0; k < ST_BLOCK / 32; k++)
      {
        const int lab = s_open_label[k];
        if (lab < 0)
          continue;
        bool seen = false;
        for (int k2 = 0; k2 < k; k2++)
          seen |= s_open_label[k2] == lab;
        if (seen)
          continue;
        double t = 0.0;
        for (int k2 = k; k2 < ST_BLOCK / 32; k2++)
          if (s_open_label[k2] == lab)
            t += s_open[k2][lane];
        if (lab >= a.cap)
          atomicOr(&a.out->err, 2);
        else
          atomicAdd(&a.rows[(size_t)lab * ROW + lane], t);
      }
    }
  }
}

// the 12 set sums in a fixed order (run-to-run identical): warp q owns column q
__global__ void __launch_bounds__(384) k_set_sums_final(const double* __restrict__ part, int blocks, double* __restrict__ sums12)
{
  const int q = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double t = 0.0;
  for (int b = lane; b < blocks; b += 32)
    t += part[(size_t)b * 12 + q];
  t = warp_sum(t);
  if (lane == 0)
    sums12[q] = t;
}

// updateConverged (particle_filter.cpp:170-220); n_set = samples of the whole set (x, y, n: this rank's window)
__global__ void __launch_bounds__(256) k_converged_count(const double* __restrict__ x, const double* __restrict__ y, int n,
                                                         double dist_threshold, TreeOut* out, int n_set)
{
  const double mean_x = out->set_sums[10] / n_set, mean_y = out->set_sums[11] / n_set;
  int cnt = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    cnt += (fabs(x[i] - mean_x) <= dist_threshold && fabs(y[i] - mean_y) <= dist_threshold) ? 1 : 0;
  cnt = warp_sum_i(cnt);
  if ((threadIdx.x & 31) == 0 && cnt)
    atomicAdd(&out->conv, cnt);
}

// distributed filter: the integer results that are sums over ranks (leaf count of this rank's bucket subtrees, its
// convergence count, error bits, the resampler's "target beyond the CDF" flag) -> one all-reduce -> back
__global__ void k_ints_pack(const TreeOut* __restrict__ out, const int* __restrict__ pf_flags, int* __restrict__ v)
{
  v[0] = out->leaf;
  v[1] = out->conv;
  v[2] = out->err & 1;
  v[3] = (out->err >> 1) & 1;
  v[4] = (out->err >> 2) & 1;
  v[5] = *pf_flags & 1;
  v[6] = (*pf_flags >> 1) & 1;
  v[7] = 0;
}
__global__ void k_ints_unpack(const int* __restrict__ v, TreeOut* out, int* pf_flags)
{
  out->leaf = v[0];
  out->conv = v[1];
  out->err = (v[2] ? 1 : 0) | (v[3] ? 2 : 0) | (v[4] ? 4 : 0);
  *pf_flags = (v[5] ? 1 : 0) | (v[6] ? 2 : 0);
}

// normalizeCluster (particle_filter.cpp:548-562) -> count, weight, mean[3], cov(0,0),(0,1),(1,0),(1,1),(2,2)
__global__ void __launch_bounds__(256) k_cluster_finalize(const TreeOut* __restrict__ out, const double* __restrict__ rows,
                                                          int cap, double* __restrict__ stats)
{
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int C = min(out->C, cap);
  if (c >= C)
    return;
  const double* r = rows + (size_t)c * ROW;
  double* o = stats + (size_t)c * 10;
  const double weight = r[1];
  const double mean0 = r[2] / weight, mean1 = r[3] / weight;
  o[0] = r[0];
  o[1] = weight;
  o[2] = mean0;
  o[3] = mean1;
  o[4] = atan2(r[5], r[4]);
  o[5] = r[6] / weight - mean0 * mean0;
  o[6] = r[7] / weight - mean0 * mean1;
  o[7] = r[8] / weight - mean1 * mean0;
  o[8] = r[9] / weight - mean1 * mean1;
  o[9] = -2 * log(sqrt(r[4] * r[4] + r[5] * r[5]));
}
}  // namespace

namespace
{
// the bucketed shared-memory leaf count (tree_fast.cuh) of a tree of at most k_max keys (the device knows K)
int launch_fast_tree(bamcl_pf* pf, TreeOut* out, int k_max)
{
  bamcl_ctx* ctx = pf->ctx;
  const int m2 = top2_size(k_max);  // >= the device's top2_size(K)
  const size_t n_b = 2 * (size_t)m2 + 2;
  const size_t scan_ints = scan_scratch_ints((long)n_b);
  // class lists: a bucket of more than c keys exists at most (keys + nodes) / c times
  const long mass = (long)k_max + m2;
  const int cap_s = (int)std::min<long>(2l * m2, std::max<long>(2l * M_TOP1, mass / FAST_CAP_W + 1));
  const int cap_m = (int)std::min<long>(2l * m2, mass / FAST_CAP_S + 1), cap_l = (int)std::min<long>(2l * m2, mass / FAST_CAP_M + 1);
  BAMCL_TRY(pf->t_fast.reserve(sizeof(TopNode) * (size_t)m2 +
                               sizeof(int) * (n_b + scan_ints + (size_t)cap_s + cap_m + cap_l + 32)));
  FastArgs fa;
  fa.out = out;
  fa.nkey = pf->t_nkey.as<unsigned long long>();
  fa.top = pf->t_fast.as<TopNode>();
  fa.bcount = reinterpret_cast<int*>(fa.top + m2);
  int* scan_scratch = fa.bcount + n_b;
  fa.cls = scan_scratch + scan_ints;
  fa.lists = fa.cls + 8;
  fa.list_m = cap_s;
  fa.list_l = cap_s + cap_m;
  fa.bslot = pf->t_cur.as<int>();
  fa.bpos = pf->t_first_at.as<int>();
  fa.bids = pf->t_child.as<int>();
  fa.part_rank = pf_rank(pf);
  fa.part_n = pf_ranks(pf);
  const int per = FAST_BYTES_PER_KEY;
  static bool attr_done = false;
  if (!attr_done)
  {
    BAMCL_CUDA_OK(cudaFuncSetAttribute(k_fast_top, cudaFuncAttributeMaxDynamicSharedMemorySize, 24 * M_TOP1));
    BAMCL_CUDA_OK(cudaFuncSetAttribute((k_fast_bucket<FAST_CAP_M, 1, 512, false>), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       per * FAST_CAP_M));
    BAMCL_CUDA_OK(cudaFuncSetAttribute((k_fast_bucket<FAST_CAP_L, 2, 1024, false>), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       per * FAST_CAP_L));
    BAMCL_CUDA_OK(cudaFuncSetAttribute((k_fast_bucket<FAST_CAP_M, 1, 512, true>), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       per * FAST_CAP_M));
    BAMCL_CUDA_OK(cudaFuncSetAttribute((k_fast_bucket<FAST_CAP_L, 2, 1024, true>), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       per * FAST_CAP_L));
    attr_done = true;
  }
  auto walk_blocks = [](long keys) { return (unsigned)std::max<long>(1, std::min<long>((long)kNumSMs * 8, (keys + 255) / 256)); };
  BAMCL_CUDA_OK(cudaMemsetAsync(fa.bcount, 0, sizeof(int) * n_b, ctx->stream));
  BAMCL_CUDA_OK(cudaMemsetAsync(fa.cls, 0, sizeof(int) * 8, ctx->stream));
  BAMCL_LAUNCH(ctx, k_fast_top, 1, 1024, 24 * M_TOP1, fa);
  if (m2 > M_TOP1)
  {
    // ---- step B: keys [M1, M2) extend the written-out tree
    const int nb1 = 2 * M_TOP1;
    const long keys_b = m2 - M_TOP1;
    BAMCL_LAUNCH(ctx, k_fast_descend<false>, walk_blocks(keys_b), 256, 0, fa);
    BAMCL_TRY(exclusive_scan_i32(ctx, fa.bcount, fa.bcount, nb1 + 1, scan_scratch, nullptr));
    BAMCL_LAUNCH(ctx, k_fast_scatter<false>, walk_blocks(keys_b), 256, 0, fa);
    BAMCL_LAUNCH(ctx, k_fast_classify<false>, (nb1 + 255) / 256, 256, 0, fa);
    BAMCL_LAUNCH(ctx, (k_fast_bucket<FAST_CAP_S, 0, 256, false>), std::min<long>(nb1, cap_s), 256, per * FAST_CAP_S, fa);
    BAMCL_LAUNCH(ctx, (k_fast_bucket<FAST_CAP_M, 1, 512, false>), std::min<long>(cap_m, keys_b / FAST_CAP_S + 1), 512,
                 per * FAST_CAP_M, fa);
    BAMCL_LAUNCH(ctx, (k_fast_bucket<FAST_CAP_L, 2, 1024, false>), std::min<long>(cap_l, keys_b / FAST_CAP_M + 1), 1024,
                 per * FAST_CAP_L, fa);
    BAMCL_CUDA_OK(cudaMemsetAsync(fa.bcount, 0, sizeof(int) * n_b, ctx->stream));
    BAMCL_CUDA_OK(cudaMemsetAsync(fa.cls, 0, sizeof(int) * 8, ctx->stream));
  }
  // ---- step C: every remaining key, leaves counted per bucket
  const int nb2 = 2 * m2;
  if (k_max > M_TOP1)
  {
    BAMCL_LAUNCH(ctx, k_fast_descend<true>, walk_blocks(k_max), 256, 0, fa);
    BAMCL_TRY(exclusive_scan_i32(ctx, fa.bcount, fa.bcount, nb2 + 1, scan_scratch, nullptr));
    BAMCL_LAUNCH(ctx, k_fast_scatter<true>, walk_blocks(k_max), 256, 0, fa);
  }
  if (k_max < M_TOP1 + FAST_CAP_W)
  {
    BAMCL_LAUNCH(ctx, k_fast_bucket_warp, (nb2 + FAST_W_WARPS - 1) / FAST_W_WARPS, FAST_W_WARPS * 32, 0, fa);
    return BAMCL_OK;
  }
  BAMCL_LAUNCH(ctx, k_fast_classify<true>, (nb2 + 255) / 256, 256, 0, fa);
  // the size classes touch disjoint buckets: large and small CTA buckets go to side streams, the rest stays here
  cudaStream_t here = ctx->stream;
  BAMCL_CUDA_OK(cudaEventRecord(pf->ev_bucket_fork, here));
  {
    struct StreamSwap
    {
      bamcl_ctx* c;
      cudaStream_t keep;
      ~StreamSwap() { c->stream = keep; }
    } swap{ctx, here};
    for (int i = 0; i < 2; i++)
    {
      BAMCL_CUDA_OK(cudaStreamWaitEvent(pf->bucket_stream[i], pf->ev_bucket_fork, 0));
      ctx->stream = pf->bucket_stream[i];
      if (i == 0)
        BAMCL_LAUNCH(ctx, (k_fast_bucket<FAST_CAP_L, 2, 1024, true>), cap_l, 1024, per * FAST_CAP_L, fa);
      else
        BAMCL_LAUNCH(ctx, (k_fast_bucket<FAST_CAP_S, 0, 256, true>), cap_s, 256, per * FAST_CAP_S, fa);
      BAMCL_CUDA_OK(cudaEventRecord(pf->ev_bucket_join[i], pf->bucket_stream[i]));
    }
  }
  BAMCL_LAUNCH(ctx, k_fast_bucket_warp, (nb2 + FAST_W_WARPS - 1) / FAST_W_WARPS, FAST_W_WARPS * 32, 0, fa);
  BAMCL_LAUNCH(ctx, (k_fast_bucket<FAST_CAP_M, 1, 512, true>), cap_m, 512, per * FAST_CAP_M, fa);
  for (int i = 0; i < 2; i++)
    BAMCL_CUDA_OK(cudaStreamWaitEvent(here, pf->ev_bucket_join[i], 0));
  return BAMCL_OK;
}
}  // namespace

namespace bamcl
{
static int coop_blocks_for(bamcl_pf* pf)
{
  if (pf->coop_blocks == 0)
  {
    int per_sm = 0, sms = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_tree_build, 256, 0) != cudaSuccess || per_sm < 1)
      per_sm = 1;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, pf->ctx->device) != cudaSuccess || sms < 1)
      sms = kNumSMs;
    pf->coop_blocks = per_sm * sms;
  }
  return pf->coop_blocks;
}

// pose-bin box of the (global) set: bounding box of its keys plus a margin, identical on every rank
static int fit_box(bamcl_pf* pf, SampleSet& set, long off, int n, double cx, double cy, double ct)
{
  bamcl_ctx* ctx = pf->ctx;
  BAMCL_TRY(pf->t_ints.reserve(sizeof(int) * 32));
  int* box6 = pf->t_ints.as<int>() + 16;
  BAMCL_CUDA_OK(cudaMemsetAsync(box6, 0x7F, sizeof(int) * 6, ctx->stream));
  if (n > 0)
    BAMCL_LAUNCH(ctx, k_key_bbox, std::min((n + 255) / 256, kNumSMs * 8), 256, 0, set.x.as<double>() + off, set.y.as<double>() + off,
                 set.th.as<double>() + off, n, cx, cy, ct, box6);
  if (pf_ranks(pf) > 1)
    BAMCL_TRY(comm_all_reduce(pf->comm, box6, 6, CT_I32, CO_MIN));
  int h[6];
  BAMCL_CUDA_OK(cudaMemcpyAsync(h, box6, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  pf->box_valid = true;
  pf->box_dense = false;
  pf->box_cells = 0;
  const bool force_hash = getenv("BAMCL_TREE_HASH") != nullptr;  // tests: exercise the hashed path
  if (force_hash)  // (a distributed filter then takes build_replicated, as it would in a world too large for the grid)
    return BAMCL_OK;
  long cells = 1;
  for (int d = 0; d < 3; d++)
  {
    if (h[d] == 0x7F7F7F7F || h[3 + d] == 0x7F7F7F7F)
      return BAMCL_OK;  // no representable key: the hashed path reports it
    pf->box_min[d] = h[d] - BOX_MARGIN;
    const long dim = (long)(-h[3 + d]) + BOX_MARGIN - pf->box_min[d] + 1;
    if (dim <= 0 || dim > DENSE_CAP)
      return BAMCL_OK;
    pf->box_dim[d] = (int)dim;
    cells *= dim;
    if (cells > DENSE_CAP)
      return BAMCL_OK;
  }
  pf->box_dense = true;
  pf->box_cells = cells;
  return BAMCL_OK;
}

// Distributed filter whose bins do not fit the dense grid (a very large world): every rank receives the whole set and
// builds the tree for itself, as a single-GPU filter would. Rare; cost is proportional to the global set.
int build_tree_and_stats(bamcl_pf* pf, int set_index, double dist_threshold, double convergence_threshold,
                         bamcl_resample_result* res);
static int build_replicated(bamcl_pf* pf, int set_index, double dist_threshold, double convergence_threshold,
                            bamcl_resample_result* res)
{
  SampleSet& set = pf->sets[set_index];
  bamcl_comm* comm = pf->comm;
  const int G = comm->n_ranks, me = comm->rank;
  const long n = set.sample_count;
  std::vector<P2P> sends, recvs;
  double* arr[4] = {set.x.as<double>(), set.y.as<double>(), set.th.as<double>(), set.w.as<double>()};
  const long o0 = shard_begin(n, G, me), c0 = shard_begin(n, G, me + 1) - o0;
  for (int q = 0; q < G; q++)
  {
    if (q == me)
      continue;
    const long o = shard_begin(n, G, q), c = shard_begin(n, G, q + 1) - o;
    for (int k = 0; k < 4; k++)
    {
      sends.push_back(P2P{q, arr[k] + o0, sizeof(double) * (size_t)c0});
      recvs.push_back(P2P{q, arr[k] + o, sizeof(double) * (size_t)c});
    }
  }
  BAMCL_TRY(comm_exchange(comm, sends.data(), (int)sends.size(), recvs.data(), (int)recvs.size()));
  const long keep_off = pf->win_off;
  const int keep_cnt = pf->win_cnt;
  pf->comm = nullptr;
  pf->win_off = 0;
  pf->win_cnt = -1;
  pf->box_valid = false;
  const int rc = build_tree_and_stats(pf, set_index, dist_threshold, convergence_threshold, res);
  pf->comm = comm;
  pf->win_off = keep_off;
  pf->win_cnt = keep_cnt;
  pf->box_valid = false;
  return rc;
}

int build_tree_and_stats(bamcl_pf* pf, int set_index, double dist_threshold, double convergence_threshold,
                         bamcl_resample_result* res)
{
  bamcl_ctx* ctx = pf->ctx;
  SampleSet& set = pf->sets[set_index];
  const int G = pf_ranks(pf), me = pf_rank(pf);
  const int n_set = set.sample_count;  // the whole (global) set
  // the samples this rank holds: its window of a distributed set, else everything
  const long off = G > 1 ? shard_begin(n_set, G, me) : 0;
  const int n = G > 1 ? (int)(shard_begin(n_set, G, me + 1) - off) : n_set;
  set.leaf_count = set.node_count = set.cluster_count = 0;
  if (n_set <= 0)
    return BAMCL_OK;
  const size_t cap_n = (size_t)pf->max_samples;
  BAMCL_TRY(pf->t_slot.reserve(sizeof(int) * cap_n));
  BAMCL_TRY(pf->t_flag.reserve(sizeof(int) * cap_n));
  BAMCL_TRY(pf->t_nkey.reserve(sizeof(unsigned long long) * cap_n));
  BAMCL_TRY(pf->t_cur.reserve(sizeof(int) * cap_n));
  BAMCL_TRY(pf->t_child.reserve(sizeof(int) * 2 * cap_n));
  BAMCL_TRY(pf->t_first_at.reserve(sizeof(int) * cap_n));
  BAMCL_TRY(pf->t_side.reserve(cap_n));
  BAMCL_TRY(pf->t_parent.reserve(sizeof(int) * cap_n));
  BAMCL_TRY(pf->t_cid.reserve(sizeof(int) * cap_n));
  BAMCL_TRY(pf->t_nfirst.reserve(sizeof(int) * cap_n));
  BAMCL_TRY(pf->t_scan.reserve(sizeof(int) * scan_scratch_ints((long)cap_n)));
  BAMCL_TRY(pf->t_out.reserve(sizeof(TreeOut)));
  BAMCL_TRY(pf->t_hout.reserve(sizeof(TreeOut)));
  BAMCL_TRY(pf->t_root.reserve(sizeof(int) * cap_n));
  BAMCL_TRY(pf->t_ints.reserve(sizeof(int) * 32));
  const int per_block = ST_BLOCK * ST_CHUNKS;
  const int stat_blocks = (n + per_block - 1) / per_block;
  BAMCL_TRY(pf->t_setpart.reserve(sizeof(double) * 12 * (size_t)std::max(stat_blocks, 1)));
  BAMCL_TRY(pf_ensure_streams(pf));
  if (pf->t_cluster_cap == 0)
    pf->t_cluster_cap = 4096;
  TreeOut* out = pf->t_out.as<TreeOut>();
  TreeOut* h_out = pf->t_hout.as<TreeOut>();
  const int blocks_n = (n + 255) / 256;
  // cell sizes exactly as pf_kdtree.cpp:35-37 writes them
  const double cx = 0.50, cy = 0.50, ct = (10 * M_PI / 180);
  static const bool force_general = getenv("BAMCL_TREE_GENERAL") != nullptr;  // tests: exercise k_tree_build
  const bool general_only = force_general || pf->force_general_tree;

  for (int attempt = 0; attempt < 6; attempt++)
  {
    if (!pf->box_valid)
      BAMCL_TRY(fit_box(pf, set, off, n, cx, cy, ct));
    const bool dense = pf->box_dense;
    if (!dense && G > 1)
      return build_replicated(pf, set_index, dist_threshold, convergence_threshold, res);
    pf->tree_dense = dense ? 1 : 0;
    // node-indexed arrays hold at most k_max entries
    const int k_max = dense ? (int)std::min<long>(n_set, pf->box_cells) : n_set;
    const int blocks_k = (k_max + 255) / 256;
    BAMCL_TRY(pf->t_sums.reserve(sizeof(double) * (ROW * pf->t_cluster_cap + 16)));
    double* set_tail = pf->t_sums.as<double>() + ROW * pf->t_cluster_cap;  // the 12 set sums ride behind the cluster rows
    BAMCL_TRY(set.cluster_stats.reserve(sizeof(double) * 10 * pf->t_cluster_cap));
    BAMCL_CUDA_OK(cudaMemsetAsync(out, 0, sizeof(TreeOut), ctx->stream));
    BAMCL_CUDA_OK(cudaMemsetAsync(pf->t_sums.p, 0, sizeof(double) * (ROW * pf->t_cluster_cap + 16), ctx->stream));
    Box box;
    unsigned hmask = 0;
    if (dense)
    {
      for (int d = 0; d < 3; d++)
      {
        box.mn[d] = pf->box_min[d];
        box.dim[d] = pf->box_dim[d];
      }
      const long cells = pf->box_cells;
      const int words = (n_set + 31) / 32;
      BAMCL_TRY(pf->g_first.reserve(sizeof(int) * (size_t)cells));
      BAMCL_TRY(pf->g_node.reserve(sizeof(int) * (size_t)cells));
      BAMCL_TRY(pf->t_bits.reserve(sizeof(unsigned) * (size_t)(words + 1)));
      BAMCL_TRY(pf->t_bpre.reserve(sizeof(int) * (size_t)(words + 1)));
      BAMCL_CUDA_OK(cudaMemsetAsync(pf->g_first.p, 0x7F, sizeof(int) * (size_t)cells, ctx->stream));
      BAMCL_CUDA_OK(cudaMemsetAsync(pf->t_bits.p, 0, sizeof(unsigned) * (size_t)words, ctx->stream));
      if (n > 0)
        BAMCL_LAUNCH(ctx, k_grid_first, blocks_n, 256, 0, set.x.as<double>() + off, set.y.as<double>() + off,
                     set.th.as<double>() + off, n, off, cx, cy, ct, box, pf->g_first.as<int>(), pf->t_slot.as<int>(), out);
      pf_mark(pf, "grid_first");
      if (G > 1)  // global first occurrence of every bin
        BAMCL_TRY(comm_all_reduce(pf->comm, pf->g_first.p, (size_t)cells, CT_I32, CO_MIN));
      pf_mark(pf, "allreduce_grid");
      const unsigned cell_blocks = (unsigned)((cells + 255) / 256);
      BAMCL_LAUNCH(ctx, k_grid_mark, cell_blocks, 256, 0, pf->g_first.as<int>(), cells, n_set, pf->t_bits.as<unsigned>());
      BAMCL_LAUNCH(ctx, k_popc, (words + 255) / 256, 256, 0, pf->t_bits.as<unsigned>(), words, pf->t_bpre.as<int>());
      BAMCL_TRY(exclusive_scan_i32(ctx, pf->t_bpre.as<int>(), pf->t_bpre.as<int>(), words, pf->t_scan.as<int>(), &out->K));
      if (G > 1)  // the parent array is all-reduced over [0, k_max): entries past K must be neutral for min
        BAMCL_CUDA_OK(cudaMemsetAsync(pf->t_parent.p, 0x7F, sizeof(int) * (size_t)k_max, ctx->stream));
      BAMCL_LAUNCH(ctx, k_grid_nodes, cell_blocks, 256, 0, pf->g_first.as<int>(), cells, n_set, pf->t_bits.as<unsigned>(),
                   pf->t_bpre.as<int>(), box, pf->g_node.as<int>(), pf->t_nkey.as<unsigned long long>(), pf->t_nfirst.as<int>(),
                   pf->t_parent.as<int>());
    }
    else
    {
      // hash capacity: power of two >= 2n
      size_t H = 1024;
      while (H < (size_t)n * 2)
        H <<= 1;
      hmask = (unsigned)(H - 1);
      BAMCL_TRY(pf->t_hkeys.reserve(sizeof(unsigned long long) * H));
      BAMCL_TRY(pf->t_hfirst.reserve(sizeof(int) * H));
      BAMCL_TRY(pf->t_hnode.reserve(sizeof(int) * H));
      BAMCL_CUDA_OK(cudaMemsetAsync(pf->t_hkeys.p, 0xFF, sizeof(unsigned long long) * H, ctx->stream));
      BAMCL_CUDA_OK(cudaMemsetAsync(pf->t_hfirst.p, 0x7F, sizeof(int) * H, ctx->stream));
      BAMCL_LAUNCH(ctx, k_hash_insert, blocks_n, 256, 0, set.x.as<double>(), set.y.as<double>(), set.th.as<double>(), n, cx, cy,
                   ct, pf->t_hkeys.as<unsigned long long>(), pf->t_hfirst.as<int>(), hmask, pf->t_slot.as<int>(), out);
      BAMCL_LAUNCH(ctx, k_flag_first, blocks_n, 256, 0, pf->t_slot.as<int>(), pf->t_hfirst.as<int>(), pf->t_flag.as<int>(), n);
      BAMCL_TRY(exclusive_scan_i32(ctx, pf->t_flag.as<int>(), pf->t_flag.as<int>(), n, pf->t_scan.as<int>(), &out->K));
      BAMCL_LAUNCH(ctx, k_make_nodes, blocks_n, 256, 0, pf->t_slot.as<int>(), pf->t_hfirst.as<int>(), pf->t_flag.as<int>(),
                   pf->t_hkeys.as<unsigned long long>(), pf->t_hnode.as<int>(), pf->t_nkey.as<unsigned long long>(),
                   pf->t_nfirst.as<int>(), pf->t_parent.as<int>(), n);
    }
    pf_mark(pf, "nodes");
    // ---- fork: the cluster chain (needs only the nodes and the bin index) goes to the auxiliary stream and overlaps the
    // leaf-count chain on the main stream; both join before anything is exchanged or copied back
    cudaStream_t main_stream = ctx->stream;
    BAMCL_CUDA_OK(cudaEventRecord(pf->ev_fork, main_stream));
    // Which side stream: with up to ~2M bins the tree chain is many short kernels, and behind the clustering pass's huge
    // grid each of them would wait for a free slot -- the clustering chain takes the LOW-priority stream, the tree chain
    // is served first and the two really overlap. With many more bins the tree's own kernels fill the machine and would
    // starve the (longer) clustering chain: equal priorities there. Resample ms, low / equal priority (scripts/
    // ab_priority.sh): 1M samples 0.78 / 0.83, 2M 1.12 / 1.17, 3M 1.75 / 1.79, 8M 2.84 / 2.72.
    // BAMCL_PRIO_MODE = low | flat overrides.
    static const char* prio_mode = getenv("BAMCL_PRIO_MODE");
    const bool aux_low = prio_mode ? prio_mode[0] == 'l' : k_max <= 4000000;
    cudaStream_t aux = aux_low ? pf->aux_stream_lo : pf->aux_stream;
    BAMCL_CUDA_OK(cudaStreamWaitEvent(aux, pf->ev_fork, 0));
    {
      struct StreamSwap  // BAMCL_LAUNCH and the scan use ctx->stream; restored on every exit path
      {
        bamcl_ctx* c;
        cudaStream_t keep;
        ~StreamSwap() { c->stream = keep; }
      } swap{ctx, main_stream};
      ctx->stream = aux;
      const long threads = (long)k_max * 13;
      static const bool cc_volatile = getenv("BAMCL_CC_VOLATILE") != nullptr;  // A/B switch: forest reads bypass L1
      // (Measured and dropped: collecting the neighbour pairs first and uniting them in a second, dense launch. The
      // unions then all contend at once: 0.73 ms for ~1M pairs against 0.38 ms for this kernel with the unions spread
      // over the probes. Also measured and dropped: one thread per node with its 13 probes issued together and its own
      // root carried through the unions (the ECL-CC shape) -- resample 1.14 ms against 0.89 ms at 1M samples / 243k
      // bins: 13x fewer threads leave the forest walks latency-bound.)
      const unsigned cc_blocks = (unsigned)((threads + 255) / 256);
      // (Measured and dropped as the default: one thread per grid cell with the 13 neighbours as coalesced loads,
      // k_cc_union_cells. Neighbouring cells belong to the same component and all race for the same root at once:
      // 0.77 ms against 0.26 ms for node order at 1M samples / 780k bins. Kept behind a switch for sparse experiments.)
      static const bool cc_cells = getenv("BAMCL_CC_CELLS") != nullptr;
      if (dense && cc_cells)
      {
        const unsigned cell_blocks = (unsigned)((pf->box_cells + 255) / 256);
        if (cc_volatile)
          BAMCL_LAUNCH(ctx, k_cc_union_cells<false>, cell_blocks, 256, 0, box, pf->g_node.as<int>(), pf->box_cells,
                       pf->t_parent.as<int>());
        else
          BAMCL_LAUNCH(ctx, k_cc_union_cells<true>, cell_blocks, 256, 0, box, pf->g_node.as<int>(), pf->box_cells,
                       pf->t_parent.as<int>());
      }
      else if (dense)
      {
        // (Measured and dropped again on the dense grid: one thread per node with its 13 neighbour loads in flight
        // together -- resample 0.99 ms against 0.85 ms at 1M samples, 3.48 against 2.76 ms at 8M. The unions are a latency
        // chain per warp; more, shorter-lived threads hide it better.)
        auto unite = [&](int mode, int rows_per_slab) -> int {
          if (cc_volatile)
            BAMCL_LAUNCH(ctx, k_cc_union_dense<false>, cc_blocks, 256, 0, out, pf->t_nkey.as<unsigned long long>(), box,
                         pf->g_node.as<int>(), pf->t_parent.as<int>(), k_max, mode, me, rows_per_slab);
          else
            BAMCL_LAUNCH(ctx, k_cc_union_dense<true>, cc_blocks, 256, 0, out, pf->t_nkey.as<unsigned long long>(), box,
                         pf->g_node.as<int>(), pf->t_parent.as<int>(), k_max, mode, me, rows_per_slab);
          return BAMCL_OK;
        };
        // (Measured and dropped as the default: the ECL-CC head start, k_cc_init. It takes 0.10 ms at 780k bins and
        // shortens the union pass from 0.29 to 0.21 ms: resample 0.91 ms against 0.84 without it; 3.04 against 2.73 ms at
        // 8M samples / 3.9M bins. Behind a switch for other occupancies.)
        static const bool cc_init = getenv("BAMCL_CC_INIT") != nullptr;
        auto head_start = [&](int mode, int rows_per_slab) -> int {
          if (cc_init)
            BAMCL_LAUNCH(ctx, k_cc_init, blocks_k, 256, 0, out, pf->t_nkey.as<unsigned long long>(), box, pf->g_node.as<int>(),
                         pf->t_parent.as<int>(), k_max, mode, me, rows_per_slab);
          return BAMCL_OK;
        };
        static const bool cc_replicated = getenv("BAMCL_CC_REPLICATED") != nullptr;  // A/B switch
        // (A/B on one 8 x B200 box, 1M particles per GPU, step ms distributed / replicated: G = 4: 3.33 / 3.47, G = 8: 3.84 /
        // 3.74, 64M over 8: 13.9 / 13.5 -- the all-reduce of the parent array, the boundary pass and the rank skew it
        // waits for outgrow the 1/G saving beyond four ranks)
        static const bool cc_distributed = getenv("BAMCL_CC_DISTRIBUTED") != nullptr;
        if (G > 1 && !cc_replicated && (G <= 4 || cc_distributed) && box.dim[1] >= 2 * G)
        {
          // every rank unites one slab of the grid; forests merged by min; slab-crossing edges united by everyone
          const int rows_per_slab = (box.dim[1] + G - 1) / G;
          BAMCL_TRY(head_start(1, rows_per_slab));
          BAMCL_TRY(unite(1, rows_per_slab));
          BAMCL_TRY(comm_all_reduce(pf->comm, pf->t_parent.p, (size_t)k_max, CT_I32, CO_MIN));
          BAMCL_TRY(unite(2, rows_per_slab));
        }
        else
        {
          // (Measured and dropped: unions inside 16 x 16 x 8 tiles in shared memory first (one CTA per tile, lock-free
          // union-find over local cell indices), then only the tile-crossing edges on the global forest -- 0.19 + 0.10 ms
          // against 0.29 ms for the single pass, resample 0.95 against 0.83 ms at 1M samples, 3.60 against 2.73 ms at 8M:
          // the ~500 unions of a tile contend for a handful of local roots and serialise.)
          BAMCL_TRY(head_start(0, 1));
          BAMCL_TRY(unite(0, 1));
        }
      }
      else if (cc_volatile)
        BAMCL_LAUNCH(ctx, k_cc_union<false>, cc_blocks, 256, 0, out, pf->t_nkey.as<unsigned long long>(),
                     pf->t_hkeys.as<unsigned long long>(), pf->t_hnode.as<int>(), hmask, pf->t_parent.as<int>(), k_max);
      else
        BAMCL_LAUNCH(ctx, k_cc_union<true>, cc_blocks, 256, 0, out, pf->t_nkey.as<unsigned long long>(),
                     pf->t_hkeys.as<unsigned long long>(), pf->t_hnode.as<int>(), hmask, pf->t_parent.as<int>(), k_max);
      BAMCL_LAUNCH(ctx, k_cc_roots, blocks_k, 256, 0, out, pf->t_parent.as<int>(), pf->t_root.as<int>(), pf->t_cid.as<int>(), k_max);
      BAMCL_TRY(exclusive_scan_i32(ctx, pf->t_cid.as<int>(), pf->t_cid.as<int>(), k_max, pf->t_scan.as<int>(), &out->C));
      StatsArgs sa;
      sa.x = set.x.as<double>() + off;
      sa.y = set.y.as<double>() + off;
      sa.th = set.th.as<double>() + off;
      sa.w = set.w.as<double>() + off;
      sa.n = n;
      sa.slot = pf->t_slot.as<int>();
      sa.hnode = dense ? pf->g_node.as<int>() : pf->t_hnode.as<int>();
      sa.parent = pf->t_root.as<int>();  // final roots (k_cc_roots)
      sa.cid = pf->t_cid.as<int>();
      sa.label = set.label.as<int>() + off;
      sa.rows = pf->t_sums.as<double>();
      sa.cap = (int)pf->t_cluster_cap;
      sa.out = out;
      sa.set_part = pf->t_setpart.as<double>();
      if (stat_blocks > 0)
        BAMCL_LAUNCH(ctx, k_labels_stats, stat_blocks, ST_BLOCK, 0, sa);
      BAMCL_LAUNCH(ctx, k_set_sums_final, 1, 384, 0, pf->t_setpart.as<double>(), stat_blocks, set_tail);
      BAMCL_CUDA_OK(cudaEventRecord(pf->ev_join, aux));
    }
    // ---- main stream: the leaf count
    auto launch_general = [&]() -> int {
      TreeArgs ta;
      ta.out = out;
      ta.nkey = pf->t_nkey.as<unsigned long long>();
      ta.cur = pf->t_cur.as<int>();
      ta.child = pf->t_child.as<int>();
      ta.first_at = pf->t_first_at.as<int>();
      ta.side = pf->t_side.as<signed char>();
      int blocks = std::min(coop_blocks_for(pf), std::max(1, (k_max + 511) / 512));
      void* args[] = {&ta};
      BAMCL_CUDA_OK(cudaLaunchCooperativeKernel((void*)k_tree_build, dim3(blocks), dim3(256), args, 0, ctx->stream));
      ctx->launches++;
      return BAMCL_OK;
    };
    if (!general_only)
    {
      BAMCL_TRY(launch_fast_tree(pf, out, k_max));
    }
    else
      BAMCL_TRY(launch_general());
    BAMCL_CUDA_OK(cudaStreamWaitEvent(main_stream, pf->ev_join, 0));
    pf_mark(pf, "tree||clusters");
    // ---- joined: cluster rows + set sums over all ranks, then the convergence count against the global mean
    if (G > 1)
      BAMCL_TRY(comm_all_reduce(pf->comm, pf->t_sums.p, ROW * pf->t_cluster_cap + 12, CT_F64, CO_SUM));
    BAMCL_CUDA_OK(cudaMemcpyAsync(out->set_sums, set_tail, sizeof(double) * 12, cudaMemcpyDeviceToDevice, ctx->stream));
    if (n > 0)
      BAMCL_LAUNCH(ctx, k_converged_count, std::min(blocks_n, kNumSMs * 8), 256, 0, set.x.as<double>() + off,
                   set.y.as<double>() + off, n, dist_threshold, out, n_set);
    if (G > 1)
    {
      int* ints = pf->t_ints.as<int>();
      BAMCL_LAUNCH(ctx, k_ints_pack, 1, 1, 0, out, pf_flags(pf), ints);
      BAMCL_TRY(comm_all_reduce(pf->comm, ints, 8, CT_I32, CO_SUM));
      BAMCL_LAUNCH(ctx, k_ints_unpack, 1, 1, 0, ints, out, pf_flags(pf));
    }
    BAMCL_LAUNCH(ctx, k_cluster_finalize, (unsigned)((pf->t_cluster_cap + 255) / 256), 256, 0, out, pf->t_sums.as<double>(),
                 (int)pf->t_cluster_cap, set.cluster_stats.as<double>());
    BAMCL_CUDA_OK(cudaMemcpyAsync(h_out, out, sizeof(TreeOut), cudaMemcpyDeviceToHost, ctx->stream));
    BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    pf_mark(pf, "sums+counts");
    if (h_out->err & 1)
      return set_error(BAMCL_E_UNSUPPORTED, "a pose is outside the kd-tree key range (|x|,|y| < 5e5 m, |theta| < 3.6e5 rad)");
    if (h_out->err & 4)
    {
      pf->box_valid = false;  // a bin outside the sticky box (the cloud moved): re-fit and repeat
      continue;
    }
    if (!general_only && !h_out->fast_ok)
    {
      // a bucket outgrew shared memory (adversarial insertion order): the grid-synchronous build takes over. Every rank
      // of a distributed filter sees the same flag and builds the whole tree (the partial sums above were all zero).
      BAMCL_TRY(launch_general());
      BAMCL_CUDA_OK(cudaMemcpyAsync(h_out, out, sizeof(TreeOut), cudaMemcpyDeviceToHost, ctx->stream));
      BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    }
    if ((h_out->err & 2) || (size_t)h_out->C > pf->t_cluster_cap)
    {
      // more clusters than rows: grow and redo (rare: only widely scattered sets have thousands of components)
      size_t want = pf->t_cluster_cap;
      while (want < (size_t)h_out->C)
        want <<= 1;
      pf->t_cluster_cap = want;
      continue;
    }
    // a box far larger than the occupied bins (the cloud converged) costs a pass over its cells every time: re-fit next time
    if (dense && pf->box_cells > (1l << 20) && (long)h_out->K * 64 < pf->box_cells)
      pf->box_valid = false;
    break;
  }
  if (h_out->err & 4)
    return set_error(BAMCL_E_STATE, "pose-bin box could not be fitted");
  set.leaf_count = h_out->leaf;
  set.node_count = h_out->K;
  set.cluster_count = h_out->C;
  pf->tree_levels = h_out->levels + h_out->bucket_levels;
  pf->tree_fast = h_out->fast_ok;
  // computeSetStats (particle_filter.cpp:612-628)
  const double* s = h_out->set_sums;
  const double weight = s[0];
  res->mean[0] = s[1] / weight;
  res->mean[1] = s[2] / weight;
  res->mean[2] = std::atan2(s[4], s[3]);
  res->cov[0] = s[5] / weight - res->mean[0] * res->mean[0];
  res->cov[1] = s[6] / weight - res->mean[0] * res->mean[1];
  res->cov[2] = s[7] / weight - res->mean[1] * res->mean[0];
  res->cov[3] = s[8] / weight - res->mean[1] * res->mean[1];
  res->cov[4] = -2 * std::log(std::sqrt(s[3] * s[3] + s[4] * s[4]));
  res->leaf_count = h_out->leaf;
  res->node_count = h_out->K;
  res->cluster_count = h_out->C;
  // updateConverged :206-219 (percentage in float, as the reference computes it)
  const double percent = static_cast<float>(h_out->conv) / static_cast<float>(n_set) * 100;
  res->converged = percent >= convergence_threshold ? 1 : 0;
  return BAMCL_OK;
}
}  // namespace bamcl

extern "C" {

int bamcl_pf_sample_clusters(bamcl_pf* pf, int* labels, long n)
{
  long off = 0, cnt = 0;
  if (pf)
    bamcl::pf_owned(pf, &off, &cnt);
  if (!pf || !labels || n != cnt)
    return set_error(BAMCL_E_ARG, "bamcl_pf_sample_clusters: n must equal sample_count (a distributed filter: its window)");
  if (n == 0)
    return BAMCL_OK;
  BAMCL_CUDA_OK(cudaMemcpyAsync(labels, pf->sets[pf->current].label.as<int>() + off, sizeof(int) * (size_t)n,
                                cudaMemcpyDeviceToHost, pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  return BAMCL_OK;
}

int bamcl_pf_cluster_stats(bamcl_pf* pf, double* out, int max_clusters)
{
  if (!pf || !out || max_clusters < 0)
    return set_error(BAMCL_E_ARG, "bamcl_pf_cluster_stats: bad arguments");
  bamcl::SampleSet& set = pf->sets[pf->current];
  const int c = std::min(max_clusters, set.cluster_count);
  if (c == 0)
    return BAMCL_OK;
  BAMCL_CUDA_OK(cudaMemcpyAsync(out, set.cluster_stats.p, sizeof(double) * 10 * (size_t)c, cudaMemcpyDeviceToHost,
                                pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  return BAMCL_OK;
}

}  // extern "C"
