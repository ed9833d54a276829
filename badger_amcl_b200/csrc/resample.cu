// K-d1/K-d2: sequential-rounding CDF, systematic target sequence, selection + pose gather, and the
// updateResample / initWithPoseFn entry points. Reference: src/amcl/pf/particle_filter.cpp:269-354, 423-502.
//
// Bit-exactness. The reference builds c[i+1] = fl(c[i] + w[i]) and target += delta one rounding
// at a time. A tree-shaped scan rounds differently and moves ~1e-10-wide CDF boundaries, enough to flip
// resample indices. While the running sum s stays in one binade [2^e, 2^(e+1)) it is a multiple of the ulp
// u = 2^(e-52), so fl(s + w) = s + R_u(w) with R_u = round-to-nearest-multiple-of-u (ties go to the even
// multiple, i.e. they depend on the parity of s/u). Inside a binade the sequential sum is therefore an INTEGER
// prefix sum of per-element functions parity -> increment, and composing those functions is associative:
//   * k_cdf_tile_sums / k_cdf_classify : approximate tile start values; a tile whose whole range provably stays
//     inside one binade is "fast";
//   * k_cdf_tiles : single-pass chained scan (decoupled look-back). Fast tiles publish their aggregate function
//     before knowing their exact start value; the few tiles that contain a binade crossing wait for the exact
//     start value, scan up to the crossing, take the crossing step as a real FP add, and continue.
// The systematic targets get the same treatment on the host in closed form (<= ~100 binade segments).
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "pf.cuh"
#include "scan.cuh"

using namespace bamcl;

namespace bamcl
{
int build_tree_and_stats(bamcl_pf* pf, int set_index, double dist_threshold, double convergence_threshold,
                         bamcl_resample_result* result);  // kdtree.cu
}

namespace
{
#ifndef CDF_BLOCK_THREADS
#define CDF_BLOCK_THREADS 256
#endif
#ifndef CDF_ITEMS_PER_THREAD
#define CDF_ITEMS_PER_THREAD 8
#endif
constexpr int CDF_BLOCK = CDF_BLOCK_THREADS;
constexpr int CDF_ITEMS = CDF_ITEMS_PER_THREAD;
constexpr int CDF_TILE = CDF_BLOCK * CDF_ITEMS;
constexpr long long SAT = 1ll << 56;
constexpr long long TWO53 = 1ll << 53;
enum
{
  ST_EMPTY = 0,
  ST_AGG = 1,
  ST_INCL = 2
};

struct Fn
{
  long long d0, d1;  // increment (in ulps of the binade) if the incoming sum/ulp is even / odd
};
__device__ __forceinline__ long long sat(long long v) { return v > SAT ? SAT : v; }
__device__ __forceinline__ long long fn_at(const Fn& f, long long m) { return (m & 1) ? f.d1 : f.d0; }
// a first, then b
__device__ __forceinline__ Fn compose(const Fn& a, const Fn& b)
{
  Fn c;
  c.d0 = sat(a.d0 + ((a.d0 & 1) ? b.d1 : b.d0));
  c.d1 = sat(a.d1 + (((1 + a.d1) & 1) ? b.d1 : b.d0));
  return c;
}
// R_u(w) for u = 2^(e-52) as a function of the parity of the running sum
__device__ __forceinline__ Fn elem_fn(double w, int e)
{
  const unsigned long long b = (unsigned long long)__double_as_longlong(w);
  Fn f;
  f.d0 = f.d1 = 0;
  if ((b << 1) == 0)
    return f;
  const int E = (int)((b >> 52) & 0x7ff);
  unsigned long long m = b & ((1ull << 52) - 1);
  int q;
  if (E == 0)
    q = -1074;
  else
  {
    m |= 1ull << 52;
    q = E - 1075;
  }
  const int sh = (e - 52) - q;
  if (sh <= 0)
  {
    f.d0 = f.d1 = (sh < -3) ? SAT : (long long)(m << (-sh));
    return f;
  }
  if (sh > 54)
    return f;
  unsigned long long base = m >> sh;
  const unsigned long long rem = m & ((1ull << sh) - 1), half = 1ull << (sh - 1);
  if (rem > half)
    base++;
  else if (rem == half)
  {
    f.d0 = (long long)(base + (base & 1));
    f.d1 = (long long)(base + ((1 + base) & 1));
    return f;
  }
  f.d0 = f.d1 = (long long)base;
  return f;
}
__device__ __forceinline__ int exponent_of(double s) { return (int)((__double_as_longlong(s) >> 52) & 0x7ff) - 1023; }
__device__ __forceinline__ long long mantissa_of(double s)
{
  return (long long)((__double_as_longlong(s) & ((1ll << 52) - 1)) | (1ll << 52));
}
__device__ __forceinline__ double from_units(long long m, int e) { return scalbn((double)m, e - 52); }

__device__ __forceinline__ Fn shfl_up_fn(const Fn& f, int d)
{
  Fn r;
  r.d0 = __shfl_up_sync(0xffffffffu, f.d0, d);
  r.d1 = __shfl_up_sync(0xffffffffu, f.d1, d);
  return r;
}

// exclusive block scan of per-thread functions; also returns the block aggregate
__device__ __forceinline__ Fn block_scan_fn(const Fn& mine, Fn* s_warp /*[9]*/, Fn* aggregate)
{
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  Fn inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1)
  {
    const Fn t = shfl_up_fn(inc, o);
    if (lane >= o)
      inc = compose(t, inc);
  }
  if (lane == 31)
    s_warp[warp] = inc;
  __syncthreads();
  Fn ex = shfl_up_fn(inc, 1);
  if (lane == 0)
    ex.d0 = ex.d1 = 0;
  Fn warp_off;
  warp_off.d0 = warp_off.d1 = 0;
  for (int k = 0; k < warp; k++)
    warp_off = compose(warp_off, s_warp[k]);
  if (aggregate)
  {
    Fn all = warp_off;
    for (int k = warp; k < CDF_BLOCK / 32; k++)
      all = compose(all, s_warp[k]);
    *aggregate = all;
  }
  __syncthreads();
  return compose(warp_off, ex);
}

__global__ void __launch_bounds__(CDF_BLOCK) k_cdf_tile_sums(const double* __restrict__ w, int n, double* __restrict__ sums)
{
  __shared__ double s[CDF_BLOCK / 32];
  const int base = blockIdx.x * CDF_TILE;
  double v = 0.0;
#pragma unroll
  for (int t = 0; t < CDF_ITEMS; t++)
  {
    const int i = base + t * CDF_BLOCK + threadIdx.x;
    if (i < n)
      v += w[i];
  }
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0)
    s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32)
  {
    double t = threadIdx.x < CDF_BLOCK / 32 ? s[threadIdx.x] : 0.0;
    t = warp_sum(t);
    if (threadIdx.x == 0)
      sums[blockIdx.x] = t;
  }
}

// ---------------------------------------------------------------------------------------------
// The CDF in three stream-ordered steps (no kernel ever waits for another block):
//   k_cdf_tile_fns  every fast tile's aggregate function; slow tiles are listed (and, on a sharded set, their raw
//                   weights copied into the summary that is all-gathered);
//   k_cdf_resolve   ONE block walks the ranks' tile summaries in order: runs of fast tiles by a block-wide scan of the
//                   aggregate functions, slow tiles by the sequential-rounding loop below. Out come the exact start
//                   value of every tile of this rank and the exact CDF value at every rank boundary ("bounds": the
//                   exclusive scan over ranks that resampleSystematic needs, particle_filter.cpp:280-285);
//   k_cdf_apply     every tile re-derives its elements from its exact start value and writes c[].
// ---------------------------------------------------------------------------------------------
struct CdfHdr  // what a rank publishes before tiles can be classified
{
  double local_sum;  // tree-shaped (approximate) sum of the window's weights
  long long n;       // samples in the window
  int n_tiles;
  int pad[11];
};
static_assert(sizeof(CdfHdr) == 64, "CdfHdr is exchanged as 64 bytes");

// per-rank tile summary in ONE buffer (one all-gather ships it):
//   d0[T] d1[T] s0[T] s1[T] wk[T] | slow_w[w_cap * CDF_TILE] | tile_e[T] split_e[T] | slow_tile[cap] | meta[4]
// meta = {n_slow, overflow, 0, 0}. A tile is fast (tile_e = its binade: (d0, d1) is its aggregate function), split
// (tile_e = INT_MIN, split_e = the binade it starts in: exactly one binade crossing at a known element -- (d0, d1) = the
// elements before it, wk = the crossing element's weight, (s0, s1) = the elements after it, in the next binade), or slow
// (both INT_MIN: its raw weights are walked by the sequential-rounding loop).
struct CdfPack
{
  long long *d0, *d1, *s0, *s1;
  double *wk, *slow_w;
  int *tile_e, *split_e, *slow_tile, *meta;
};
__host__ __device__ inline size_t cdf_pack_bytes(int T, int cap, int w_cap)
{
  const size_t b = (size_t)T * 40 + (size_t)w_cap * CDF_TILE * 8 + (size_t)T * 8 + (size_t)cap * 4 + 16;
  return (b + 15) & ~(size_t)15;
}
__host__ __device__ inline CdfPack cdf_pack_at(void* base, int T, int cap, int w_cap)
{
  CdfPack p;
  p.d0 = static_cast<long long*>(base);
  p.d1 = p.d0 + T;
  p.s0 = p.d1 + T;
  p.s1 = p.s0 + T;
  p.wk = reinterpret_cast<double*>(p.s1 + T);
  p.slow_w = p.wk + T;
  p.tile_e = reinterpret_cast<int*>(p.slow_w + (size_t)w_cap * CDF_TILE);
  p.split_e = p.tile_e + T;
  p.slow_tile = p.split_e + T;
  p.meta = p.slow_tile + cap;
  return p;
}

__global__ void __launch_bounds__(1024) k_cdf_hdr(const double* __restrict__ sums, int n_tiles, long long n, CdfHdr* hdr)
{
  __shared__ double s_part[32];
  const int per = (n_tiles + 1023) / 1024;
  const int lo = min(n_tiles, (int)threadIdx.x * per), hi = min(n_tiles, lo + per);
  double acc = 0.0;
  for (int i = lo; i < hi; i++)
    acc += sums[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0)
    s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32)
  {
    double t = warp_sum(s_part[threadIdx.x]);
    if (threadIdx.x == 0)
    {
      hdr->local_sum = t;
      hdr->n = n;
      hdr->n_tiles = n_tiles;
    }
  }
}

// one block: approximate start value of every tile of THIS rank (the ranks before it enter through their published
// sums) and the fast/slow decision; slow tiles are listed in tile order
__global__ void __launch_bounds__(1024) k_cdf_classify(const double* __restrict__ sums, int n_tiles, const CdfHdr* __restrict__ hdrs,
                                                       int rank, int n_ranks, double* __restrict__ starts, CdfPack pack, int cap,
                                                       int* __restrict__ tile_slot)
{
  __shared__ double s_part[1024];
  __shared__ int s_cnt[1024];
  __shared__ double s_total;
  double c0 = 0.0, total_all = 0.0, units = 64.0 + n_ranks;
  for (int q = 0; q < n_ranks; q++)
  {
    if (q == rank)
      c0 = total_all;
    total_all += hdrs[q].local_sum;
    units += (double)hdrs[q].n + (double)hdrs[q].n_tiles;
  }
  // chunked serial-by-thread scan: thread t owns a contiguous run of tiles
  const int per = (n_tiles + 1023) / 1024;
  const int lo = min(n_tiles, (int)threadIdx.x * per), hi = min(n_tiles, lo + per);
  double acc = 0.0;
  for (int i = lo; i < hi; i++)
    acc += sums[i];
  s_part[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x == 0)
  {
    double run = c0;  // approximate value of the CDF before this rank's first element
    for (int i = 0; i < 1024; i++)
    {
      const double t = s_part[i];
      s_part[i] = run;
      run += t;
    }
    s_total = run;
  }
  __syncthreads();
  double run = s_part[threadIdx.x];
  for (int i = lo; i < hi; i++)
  {
    starts[i] = run;
    run += sums[i];
  }
  if (threadIdx.x == 0)
    starts[n_tiles] = s_total;
  __syncthreads();
  // every approximate value is within err of the sequentially rounded one (n roundings of at most u * total each, and
  // as many again for the tree-shaped sums)
  const double err = 2.0 * units * 1.1102230246251565e-16 * total_all;
  if (threadIdx.x == 0)
    starts[n_tiles + 1] = err;  // k_cdf_tile_fns needs the same bound
  int slow = 0;
  for (int i = lo; i < hi; i++)
  {
    const double a = starts[i] - err, b = starts[i + 1] + err;
    int e = INT_MIN;  // slow
    if (a > 2.2250738585072014e-308 && isfinite(b) && exponent_of(a) == exponent_of(b))
      e = exponent_of(a);
    pack.tile_e[i] = e;
    slow += e == INT_MIN ? 1 : 0;
  }
  // ordered compaction of the slow tiles
  s_cnt[threadIdx.x] = slow;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1)
  {
    const int v = threadIdx.x >= o ? s_cnt[threadIdx.x - o] : 0;
    __syncthreads();
    s_cnt[threadIdx.x] += v;
    __syncthreads();
  }
  int at = s_cnt[threadIdx.x] - slow;
  for (int i = lo; i < hi; i++)
  {
    int slot = -1;
    if (pack.tile_e[i] == INT_MIN)
    {
      slot = at++;
      if (slot < cap)
        pack.slow_tile[slot] = i;
    }
    tile_slot[i] = slot;
  }
  if (threadIdx.x == 1023)
  {
    const int n_slow = s_cnt[1023];
    pack.meta[0] = min(n_slow, cap);
    pack.meta[1] = n_slow > cap ? 1 : 0;
    pack.meta[2] = pack.meta[3] = 0;
  }
}

struct SlowShared
{
  Fn warp[CDF_BLOCK / 32 + 1];
  int first;
  double val;
};

// The reference's loop c[i+1] = c[i] + w[i] over one tile whose running sum may change binade: scan to the first
// crossing inside the current binade, take the crossing step as a real FP add, continue. Thread t holds elements
// [t * CDF_ITEMS, (t + 1) * CDF_ITEMS) of the tile in wv. c_tile (may be null) receives c_tile[i + 1] for element i.
// Returns the value after the last element. All CDF_BLOCK threads call it.
__device__ double slow_tile_scan(double s, const double (&wv)[CDF_ITEMS], int tile_n, double* c_tile, SlowShared& sh, int* flags)
{
  int start = 0;  // first element of the tile not yet finalised
  const int local0 = threadIdx.x * CDF_ITEMS;
  while (start < tile_n)
  {
    __syncthreads();
    if (threadIdx.x == 0)
      sh.first = INT_MAX;
    __syncthreads();
    if (s == 0.0)
    {
      // fl(0 + w) = w: skip zeros up to the first non-zero weight
      int first = INT_MAX;
#pragma unroll
      for (int k = CDF_ITEMS - 1; k >= 0; k--)
        if (local0 + k >= start && local0 + k < tile_n && wv[k] != 0.0)
          first = local0 + k;
      if (first != INT_MAX)
        atomicMin(&sh.first, first);
      __syncthreads();
      const int fz = min(sh.first, tile_n);
#pragma unroll
      for (int k = 0; k < CDF_ITEMS; k++)
      {
        const int li = local0 + k;
        if (c_tile && li >= start && li < fz)
          c_tile[li + 1] = 0.0;
        if (li == fz && li < tile_n)
        {
          if (c_tile)
            c_tile[li + 1] = wv[k];
          sh.val = wv[k];
        }
      }
      __syncthreads();
      if (fz < tile_n)
        s = sh.val;
      start = fz + 1;
      continue;
    }
    if (!(s >= 2.2250738585072014e-308) || !isfinite(s))
    {
      if (threadIdx.x == 0)
        atomicOr(flags, 2);
      break;
    }
    const int e = exponent_of(s);
    Fn f[CDF_ITEMS];
    Fn mine;
    mine.d0 = mine.d1 = 0;
#pragma unroll
    for (int k = 0; k < CDF_ITEMS; k++)
    {
      const int li = local0 + k;
      if (li >= start && li < tile_n)
        f[k] = elem_fn(wv[k], e);
      else
        f[k].d0 = f[k].d1 = 0;
      mine = compose(mine, f[k]);
    }
    const Fn ex = block_scan_fn(mine, sh.warp, nullptr);
    const long long m_in = mantissa_of(s);
    long long m = m_in + fn_at(ex, m_in);
    long long mk[CDF_ITEMS];
    int cross = INT_MAX;
    long long prev = m;
#pragma unroll
    for (int k = 0; k < CDF_ITEMS; k++)
    {
      prev = m;
      m += fn_at(f[k], m);
      m = sat(m);
      mk[k] = m;
      const int li = local0 + k;
      if (li >= start && li < tile_n && m >= TWO53 && cross == INT_MAX)
      {
        cross = li;
        mk[k] = prev;  // value before the crossing add
      }
    }
    if (cross != INT_MAX)
      atomicMin(&sh.first, cross);
    __syncthreads();
    const int cr = sh.first;
    const int stop = min(cr, tile_n);
#pragma unroll
    for (int k = 0; k < CDF_ITEMS; k++)
    {
      const int li = local0 + k;
      if (c_tile && li >= start && li < stop)
        c_tile[li + 1] = from_units(mk[k], e);
      if (li == cr)
      {
        // the crossing add itself, as the reference does it
        const double before = from_units(mk[k], e);
        const double after = before + wv[k];
        if (c_tile)
          c_tile[li + 1] = after;
        sh.val = after;
      }
      else if (cr == INT_MAX && li == tile_n - 1)
        sh.val = from_units(mk[k], e);
    }
    __syncthreads();
    s = sh.val;
    start = (cr == INT_MAX) ? tile_n : cr + 1;
  }
  __syncthreads();
  return s;
}

__global__ void __launch_bounds__(CDF_BLOCK) k_cdf_tile_fns(const double* __restrict__ w, int n, CdfPack pack, int w_cap,
                                                            const int* __restrict__ tile_slot, const double* __restrict__ starts,
                                                            int n_tiles)
{
  __shared__ Fn s_warp[CDF_BLOCK / 32 + 1];
  __shared__ double s_wsum[CDF_BLOCK / 32];
  __shared__ int s_k;
  __shared__ double s_pk;
  const int t = blockIdx.x;
  const int idx0 = t * CDF_TILE + threadIdx.x * CDF_ITEMS;
  double wv[CDF_ITEMS];
#pragma unroll
  for (int k = 0; k < CDF_ITEMS; k++)
    wv[k] = (idx0 + k < n) ? w[idx0 + k] : 0.0;
  const int e = pack.tile_e[t];
  if (e != INT_MIN)
  {
    Fn mine;
    mine.d0 = mine.d1 = 0;
#pragma unroll
    for (int k = 0; k < CDF_ITEMS; k++)
      mine = compose(mine, elem_fn(wv[k], e));
    Fn agg;
    block_scan_fn(mine, s_warp, &agg);
    if (threadIdx.x == 0)
    {
      pack.d0[t] = agg.d0;
      pack.d1[t] = agg.d1;
      pack.split_e[t] = INT_MIN;
    }
    return;
  }
  // ---- a tile that may cross a binade. If it crosses exactly ONE boundary, at an element the approximate prefix sums
  // place beyond doubt, it is summarised as (function before, crossing weight, function after) and costs the resolve
  // step O(1); otherwise its raw weights go to the sequential-rounding loop.
  const double err = starts[n_tiles + 1];
  const double a0 = starts[t] - err, b0 = starts[t + 1] + err;
  bool candidate = a0 > 2.2250738585072014e-308 && isfinite(b0) && exponent_of(b0) == exponent_of(a0) + 1;
  const int e_lo = candidate ? exponent_of(a0) : 0;
  const double boundary = candidate ? scalbn(1.0, e_lo + 1) : 0.0;
  // approximate inclusive prefix of every element: thread-local running sums + a block scan of the thread totals
  double run[CDF_ITEMS], tot = 0.0;
#pragma unroll
  for (int k = 0; k < CDF_ITEMS; k++)
  {
    tot += wv[k];
    run[k] = tot;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double inc = tot;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1)
  {
    const double v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o)
      inc += v;
  }
  if (lane == 31)
    s_wsum[warp] = inc;
  if (threadIdx.x == 0)
    s_k = INT_MAX;
  __syncthreads();
  double base = starts[t] + (inc - tot);
  for (int q = 0; q < warp; q++)
    base += s_wsum[q];
  const int tile_n = min(CDF_TILE, n - t * CDF_TILE);
  int first = INT_MAX;
#pragma unroll
  for (int k = CDF_ITEMS - 1; k >= 0; k--)
    if (threadIdx.x * CDF_ITEMS + k < tile_n && base + run[k] + err >= boundary)
      first = threadIdx.x * CDF_ITEMS + k;
  if (candidate && first != INT_MAX)
    atomicMin(&s_k, first);
  __syncthreads();
  const int kc = s_k;
  if (kc != INT_MAX && kc / CDF_ITEMS == (int)threadIdx.x)
    s_pk = base + run[kc % CDF_ITEMS];
  __syncthreads();
  // beyond doubt: every earlier element stays below the boundary even with +err (by construction of kc), and element kc
  // is above it even with -err
  const bool split = candidate && kc != INT_MAX && s_pk - err >= boundary;
  if (split)
  {
    Fn before, after;
    before.d0 = before.d1 = after.d0 = after.d1 = 0;
#pragma unroll
    for (int k = 0; k < CDF_ITEMS; k++)
    {
      const int li = threadIdx.x * CDF_ITEMS + k;
      if (li < kc)
        before = compose(before, elem_fn(wv[k], e_lo));
      else if (li > kc && li < tile_n)
        after = compose(after, elem_fn(wv[k], e_lo + 1));
    }
    Fn agg_b, agg_a;
    block_scan_fn(before, s_warp, &agg_b);
    block_scan_fn(after, s_warp, &agg_a);
    if (kc / CDF_ITEMS == (int)threadIdx.x)
      pack.wk[t] = wv[kc % CDF_ITEMS];
    if (threadIdx.x == 0)
    {
      pack.d0[t] = agg_b.d0;
      pack.d1[t] = agg_b.d1;
      pack.s0[t] = agg_a.d0;
      pack.s1[t] = agg_a.d1;
      pack.split_e[t] = e_lo;
    }
    return;
  }
  if (threadIdx.x == 0)
  {
    pack.d0[t] = pack.d1[t] = 0;
    pack.split_e[t] = INT_MIN;
  }
  const int slot = tile_slot[t];
  if (slot >= 0 && slot < w_cap)
  {
    double* dst = pack.slow_w + (size_t)slot * CDF_TILE + threadIdx.x * CDF_ITEMS;
#pragma unroll
    for (int k = 0; k < CDF_ITEMS; k++)
      dst[k] = wv[k];
  }
}

struct ResolveArgs
{
  int n_ranks, me, T, cap, w_cap;
  const CdfHdr* hdrs;  // [n_ranks]
  char* packs;         // [n_ranks] summaries, pack_bytes apart
  size_t pack_bytes;
  const double* w;  // this rank's window (its slow tiles are read in place)
  double* bounds;   // [n_ranks + 1] exact CDF value at every rank boundary
  double* starts;   // [tiles of this rank] exact start value of every tile
  int* flags;       // bit1 unsupported running sum, bit2 a rank listed more slow tiles than fit its summary
};

__global__ void __launch_bounds__(CDF_BLOCK) k_cdf_resolve(const ResolveArgs a)
{
  __shared__ SlowShared sh;
  for (int q = 0; q < a.n_ranks; q++)
    if (cdf_pack_at(a.packs + (size_t)q * a.pack_bytes, a.T, a.cap, a.w_cap).meta[1])
    {
      if (threadIdx.x == 0)
        atomicOr(a.flags, 4);
      return;
    }
  double s = 0.0;  // the exact running value, identical in every thread
  for (int q = 0; q < a.n_ranks; q++)
  {
    if (threadIdx.x == 0)
      a.bounds[q] = s;
    const CdfPack p = cdf_pack_at(a.packs + (size_t)q * a.pack_bytes, a.T, a.cap, a.w_cap);
    const int tiles = a.hdrs[q].n_tiles;
    const long long n_q = a.hdrs[q].n;
    const int n_slow = p.meta[0];
    int t = 0, k = 0;
    while (t < tiles)
    {
      const int next_slow = k < n_slow ? p.slow_tile[k] : tiles;
      while (t < next_slow)
      {
        // a run of fast tiles: all in the binade of s
        const int chunk = min(CDF_BLOCK, next_slow - t);
        Fn mine;
        mine.d0 = mine.d1 = 0;
        int e_t = INT_MIN;
        if ((int)threadIdx.x < chunk)
        {
          mine.d0 = p.d0[t + threadIdx.x];
          mine.d1 = p.d1[t + threadIdx.x];
          e_t = p.tile_e[t + threadIdx.x];
        }
        Fn agg;
        const Fn ex = block_scan_fn(mine, sh.warp, &agg);
        const bool normal = s >= 2.2250738585072014e-308 && isfinite(s);
        const int e = normal ? exponent_of(s) : 0;
        const long long m_in = normal ? mantissa_of(s) : 0;
        if ((int)threadIdx.x < chunk && (!normal || e_t != e))
          atomicOr(a.flags, 2);
        if (q == a.me && (int)threadIdx.x < chunk)
          a.starts[t + threadIdx.x] = from_units(m_in + fn_at(ex, m_in), e);
        const long long m_out = m_in + fn_at(agg, m_in);
        if (m_out >= TWO53 && threadIdx.x == 0)
          atomicOr(a.flags, 2);
        s = normal ? from_units(m_out, e) : s;
        t += chunk;
      }
      if (t < tiles && p.split_e[t] != INT_MIN)
      {
        // a split tile: function before the crossing, the crossing add as a real FP add, function after it
        if (q == a.me && threadIdx.x == 0)
          a.starts[t] = s;
        const int e0 = p.split_e[t];
        bool ok = s >= 2.2250738585072014e-308 && isfinite(s) && exponent_of(s) == e0;
        double v = s;
        if (ok)
        {
          Fn fb, fa;
          fb.d0 = p.d0[t];
          fb.d1 = p.d1[t];
          fa.d0 = p.s0[t];
          fa.d1 = p.s1[t];
          const long long m0 = mantissa_of(s);
          const long long m1 = m0 + fn_at(fb, m0);
          ok = m1 < TWO53;
          const double before = from_units(m1, e0);
          const double after = before + p.wk[t];  // the crossing step, as the reference takes it
          ok = ok && isfinite(after) && exponent_of(after) == e0 + 1;
          const long long m2 = mantissa_of(after);
          const long long m3 = m2 + fn_at(fa, m2);
          ok = ok && m3 < TWO53;
          v = from_units(m3, e0 + 1);
        }
        if (!ok && threadIdx.x == 0)
          atomicOr(a.flags, 2);  // the summary's assumptions do not hold for the exact value: report, never guess
        s = v;
        t++;
        k++;
      }
      else if (t < tiles)
      {
        // the slow tile t: sequential-rounding loop over its raw weights
        if (q == a.me && threadIdx.x == 0)
          a.starts[t] = s;
        const long long base = (long long)t * CDF_TILE;
        const int tile_n = (int)min((long long)CDF_TILE, n_q - base);
        const double* src = q == a.me ? a.w + base : p.slow_w + (size_t)k * CDF_TILE;
        double wv[CDF_ITEMS];
#pragma unroll
        for (int j = 0; j < CDF_ITEMS; j++)
        {
          const int li = threadIdx.x * CDF_ITEMS + j;
          wv[j] = li < tile_n ? src[li] : 0.0;
        }
        s = slow_tile_scan(s, wv, tile_n, nullptr, sh, a.flags);
        t++;
        k++;
      }
    }
  }
  if (threadIdx.x == 0)
    a.bounds[a.n_ranks] = s;
}

__global__ void __launch_bounds__(CDF_BLOCK) k_cdf_apply(const double* __restrict__ w, int n, double* __restrict__ c,
                                                         const int* __restrict__ tile_e, const double* __restrict__ starts,
                                                         int* flags)
{
  __shared__ SlowShared sh;
  const int t = blockIdx.x;
  const int idx0 = t * CDF_TILE + threadIdx.x * CDF_ITEMS;
  double wv[CDF_ITEMS];
#pragma unroll
  for (int k = 0; k < CDF_ITEMS; k++)
    wv[k] = (idx0 + k < n) ? w[idx0 + k] : 0.0;
  const double s_in = starts[t];
  if (t == 0 && threadIdx.x == 0)
    c[0] = s_in;
  const int e_fast = tile_e[t];
  if (e_fast == INT_MIN)
  {
    slow_tile_scan(s_in, wv, min(CDF_TILE, n - t * CDF_TILE), c + (size_t)t * CDF_TILE, sh, flags);
    return;
  }
  // ---- fast tile: the whole tile stays in binade e_fast
  Fn f[CDF_ITEMS];
  Fn mine;
  mine.d0 = mine.d1 = 0;
#pragma unroll
  for (int k = 0; k < CDF_ITEMS; k++)
  {
    f[k] = elem_fn(wv[k], e_fast);
    mine = compose(mine, f[k]);
  }
  const Fn ex = block_scan_fn(mine, sh.warp, nullptr);
  if (exponent_of(s_in) != e_fast && threadIdx.x == 0)
    atomicOr(flags, 2);
  const long long m_in = mantissa_of(s_in);
  long long m = m_in + fn_at(ex, m_in);
#pragma unroll
  for (int k = 0; k < CDF_ITEMS; k++)
  {
    m += fn_at(f[k], m);
    if (idx0 + k < n)
      c[idx0 + k + 1] = from_units(m, e_fast);
  }
  if (threadIdx.x == CDF_BLOCK - 1 && m >= TWO53)
    atomicOr(flags, 2);
}

// ---- systematic selection (particle_filter.cpp:309-351)
struct TargetSeg
{
  long long k0;  // first output of the segment
  long long m0;  // target = (m0 + (k - k0) * d) * 2^eu
  long long d;
  int eu;
  int pad;
};
constexpr int MAX_SEGS = 512;

struct SelectArgs
{
  const double* c;  // P + 1
  int P;
  const double *ax, *ay, *ath;
  double *bx, *by, *bth, *bw;
  int* src;
  int n, n_rand;
  const double* random_poses;
  const TargetSeg* segs;
  int n_segs;
  double weight;  // 1/total with total = n (particle_filter.cpp:457-462)
  int* flags;
};

__global__ void __launch_bounds__(256) k_select(const SelectArgs a)
{
  __shared__ TargetSeg s_segs[MAX_SEGS];
  for (int i = threadIdx.x; i < a.n_segs; i += blockDim.x)
    s_segs[i] = a.segs[i];
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n)
    return;
  a.bw[i] = a.weight;
  if (i < a.n_rand)
  {
    a.bx[i] = a.random_poses[3 * i];
    a.by[i] = a.random_poses[3 * i + 1];
    a.bth[i] = a.random_poses[3 * i + 2];
    a.src[i] = -1;
    return;
  }
  const long long k = i - a.n_rand;
  int lo = 0, hi = a.n_segs - 1;  // last segment with k0 <= k
  while (lo < hi)
  {
    const int mid = (lo + hi + 1) >> 1;
    if (s_segs[mid].k0 <= k)
      lo = mid;
    else
      hi = mid - 1;
  }
  const TargetSeg sg = s_segs[lo];
  const double target = scalbn((double)(sg.m0 + (k - sg.k0) * sg.d), sg.eu);
  // j with c[j] <= target < c[j+1]  ==  (number of entries of c[1..P] that are <= target)
  int L = 0, H = a.P;  // search in c[1..P]
  while (L < H)
  {
    const int mid = (L + H) >> 1;
    if (__ldg(&a.c[mid + 1]) <= target)
      L = mid + 1;
    else
      H = mid;
  }
  int j = L;
  if (j >= a.P)
  {
    // target >= c[P]: the reference never terminates here; take the last particle that has weight
    atomicOr(a.flags, 1);
    const double top = a.c[a.P];
    int l2 = 0, h2 = a.P;
    while (l2 < h2)
    {
      const int mid = (l2 + h2) >> 1;
      if (a.c[mid + 1] < top)
        l2 = mid + 1;
      else
        h2 = mid;
    }
    j = min(l2, a.P - 1);
  }
  a.src[i] = j;
  a.bx[i] = a.ax[j];
  a.by[i] = a.ay[j];
  a.bth[i] = a.ath[j];
}
}  // namespace

// =============================================================================================
// host: systematic target sequence in closed form per binade
// =============================================================================================
namespace
{
inline int h_exponent(double s)
{
  uint64_t b;
  memcpy(&b, &s, 8);
  return (int)((b >> 52) & 0x7ff) - 1023;
}
inline long long h_mantissa(double s)
{
  uint64_t b;
  memcpy(&b, &s, 8);
  return (long long)((b & ((1ull << 52) - 1)) | (1ull << 52));
}
// delta rounded to units of 2^(e-52): base, and whether it is an exact tie
inline void h_round_units(double w, int e, long long* base_out, bool* tie, bool* huge)
{
  uint64_t b;
  memcpy(&b, &w, 8);
  *tie = false;
  *huge = false;
  *base_out = 0;
  if ((b << 1) == 0)
    return;
  const int E = (int)((b >> 52) & 0x7ff);
  uint64_t m = b & ((1ull << 52) - 1);
  int q;
  if (E == 0)
    q = -1074;
  else
  {
    m |= 1ull << 52;
    q = E - 1075;
  }
  const int sh = (e - 52) - q;
  if (sh <= 0)
  {
    if (sh < -3)
      *huge = true;
    else
      *base_out = (long long)(m << (-sh));
    return;
  }
  if (sh > 54)
    return;
  uint64_t base = m >> sh;
  const uint64_t rem = m & ((1ull << sh) - 1), half = 1ull << (sh - 1);
  if (rem > half)
    base++;
  else if (rem == half)
    *tie = true;
  *base_out = (long long)base;
}
inline double wrap_target(double t) { return t > 1.0 ? t - 1.0 : t; }

// t_0 = start, t_{k+1} = wrap(fl(t_k + delta)) for k < n, as segments of exact arithmetic progressions
int build_target_segments(double start, double delta, long long n, std::vector<TargetSeg>& segs)
{
  segs.clear();
  double t = start;
  long long k = 0;
  while (k < n)
  {
    if ((int)segs.size() >= MAX_SEGS - 2)
      return -1;
    TargetSeg sg;
    sg.pad = 0;
    sg.k0 = k;
    if (t == 0.0 || !(t >= 2.2250738585072014e-308))
    {
      // zero (or subnormal) target: a segment of one, then a real FP step
      uint64_t b;
      memcpy(&b, &t, 8);
      sg.m0 = (long long)b;  // subnormal payload = value / 2^-1074
      sg.d = 0;
      sg.eu = -1074;
      segs.push_back(sg);
      t = wrap_target(t + delta);
      k++;
      continue;
    }
    const int e = h_exponent(t);
    long long m = h_mantissa(t);
    long long base;
    bool tie, huge;
    h_round_units(delta, e, &base, &tie, &huge);
    sg.eu = e - 52;
    sg.m0 = m;
    // t == 1.0 can be reached without wrapping (the test is target > 1.0); every later step wraps
    if (huge || base >= TWO53 || t >= 1.0)
    {
      sg.d = 0;
      segs.push_back(sg);
      t = wrap_target(t + delta);
      k++;
      continue;
    }
    long long d = base;
    if (tie)
    {
      // tie: the increment is base + ((sum + base) & 1) so that the result is even
      if ((m & 1) != 0)
      {
        // odd sum: one parity-dependent step, after which the sum is even and the increment constant
        const long long first = base + ((1 + base) & 1);
        sg.d = 0;
        segs.push_back(sg);
        k++;
        if (m + first >= TWO53)
        {
          t = wrap_target(t + delta);
          continue;
        }
        m += first;
        t = std::ldexp((double)m, e - 52);
        if (k >= n)
          break;
        sg.k0 = k;
        sg.m0 = m;
      }
      d = base + (base & 1);
    }
    // values m, m+d, ... stay below 2^53 (inside the binade) for `count` elements
    long long count;
    if (d == 0)
      count = n - k;
    else
      count = (TWO53 - 1 - m) / d + 1;
    if (count > n - k)
      count = n - k;
    sg.d = d;
    segs.push_back(sg);
    const long long m_last = m + (count - 1) * d;
    k += count;
    if (k >= n)
      break;
    const double t_last = std::ldexp((double)m_last, e - 52);
    t = wrap_target(t_last + delta);  // the crossing step, as the reference takes it
  }
  return 0;
}
}  // namespace

namespace bamcl
{
// c[0..n] of this rank's window of the current set -> pf->cdf (device), exact sequential rounding with the GLOBAL
// values (c[0] = exact end of the previous rank's window); pf->cdf_bounds = the exact value at every rank boundary.
// Collective on a distributed filter (two all-gathers); nothing synchronises.
int pf_build_cdf(bamcl_pf* pf)
{
  SampleSet& set = pf->sets[pf->current];
  bamcl_comm* comm = pf->comm;
  const int G = pf_ranks(pf), me = pf_rank(pf);
  const int n = pf_win_cnt(pf);
  const double* w = set.w.as<double>() + pf_win_off(pf);
  bamcl_ctx* ctx = pf->ctx;
  BAMCL_TRY(pf->cdf.reserve(sizeof(double) * ((size_t)n + 1)));
  const int tiles = (n + CDF_TILE - 1) / CDF_TILE;
  // summary geometry: identical on every rank (windows differ by at most one sample)
  const long max_win = G > 1 ? ((long)set.sample_count + G - 1) / G : (long)n;
  const int T = std::max(1, (int)((max_win + CDF_TILE - 1) / CDF_TILE));
  const int cap = G > 1 ? std::min(pf->cdf_slow_cap, T) : T;
  const int w_cap = G > 1 ? cap : 0;
  const size_t pack_bytes = cdf_pack_bytes(T, cap, w_cap);
  // scratch: hdrs[G] | packs[G] | sums[T] starts[T+2] exact[T] bounds[G+1] | tile_slot[T]
  const size_t off_packs = (size_t)G * sizeof(CdfHdr);
  const size_t off_d = off_packs + (size_t)G * pack_bytes;
  const size_t n_d = (size_t)T * 3 + 2 + (size_t)G + 1;
  const size_t off_i = off_d + n_d * sizeof(double);
  BAMCL_TRY(pf->scratch.reserve(off_i + sizeof(int) * (size_t)T + 64));
  char* base = pf->scratch.as<char>();
  CdfHdr* hdrs = reinterpret_cast<CdfHdr*>(base);
  char* packs = base + off_packs;
  double* sums = reinterpret_cast<double*>(base + off_d);
  double* starts = sums + T;
  double* exact = starts + T + 2;
  double* bounds = exact + T;
  int* tile_slot = reinterpret_cast<int*>(base + off_i);
  int* flags = pf_flags(pf);
  const CdfPack mine = cdf_pack_at(packs + (size_t)me * pack_bytes, T, cap, w_cap);
  pf->cdf_bounds = bounds;
  BAMCL_CUDA_OK(cudaMemsetAsync(flags, 0, sizeof(int), ctx->stream));
  if (tiles > 0)
    BAMCL_LAUNCH(ctx, k_cdf_tile_sums, tiles, CDF_BLOCK, 0, w, n, sums);
  BAMCL_LAUNCH(ctx, k_cdf_hdr, 1, 1024, 0, sums, tiles, (long long)n, hdrs + me);
  if (G > 1)
    BAMCL_TRY(comm_all_gather(comm, hdrs + me, hdrs, sizeof(CdfHdr)));
  BAMCL_LAUNCH(ctx, k_cdf_classify, 1, 1024, 0, sums, tiles, hdrs, me, G, starts, mine, cap, tile_slot);
  if (tiles > 0)
    BAMCL_LAUNCH(ctx, k_cdf_tile_fns, tiles, CDF_BLOCK, 0, w, n, mine, w_cap, tile_slot, starts, tiles);
  if (G > 1)
    BAMCL_TRY(comm_all_gather(comm, packs + (size_t)me * pack_bytes, packs, pack_bytes));
  ResolveArgs ra;
  ra.n_ranks = G;
  ra.me = me;
  ra.T = T;
  ra.cap = cap;
  ra.w_cap = w_cap;
  ra.hdrs = hdrs;
  ra.packs = packs;
  ra.pack_bytes = pack_bytes;
  ra.w = w;
  ra.bounds = bounds;
  ra.starts = exact;
  ra.flags = flags;
  BAMCL_LAUNCH(ctx, k_cdf_resolve, 1, CDF_BLOCK, 0, ra);
  if (tiles > 0)
    BAMCL_LAUNCH(ctx, k_cdf_apply, tiles, CDF_BLOCK, 0, w, n, pf->cdf.as<double>(), mine.tile_e, exact, flags);
  else
    BAMCL_CUDA_OK(cudaMemcpyAsync(pf->cdf.p, bounds + me, sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
  return BAMCL_OK;
}
}  // namespace bamcl

static int resample_limit(int k, int min_samples, int max_samples, double pop_err, double pop_z)
{
  // particle_filter.cpp:475-502
  if (k <= 1)
    return max_samples;
  const double kd = static_cast<double>(k);
  const double a = 1;
  const double b = 2 / (9 * (kd - 1));
  const double c = std::sqrt(2 / (9 * (kd - 1))) * pop_z;
  const double x = a - b + c;
  const int n = static_cast<int>(std::ceil((k - 1) / (2 * pop_err) * x * x * x));
  if (n < min_samples)
    return min_samples;
  if (n > max_samples)
    return max_samples;
  return n;
}

static inline uint64_t drand48_next(uint64_t x) { return (0x5DEECE66DULL * x + 0xBULL) & ((1ULL << 48) - 1); }

namespace bamcl
{
int resample_multinomial(bamcl_pf* pf, const bamcl_resample_params* prm, bamcl_resample_result* res, const double* host_draws,
                         const double* host_poses, long n_host_poses);  // multinomial.cu
int device_random_poses(bamcl_pf* pf, unsigned long long state0, int n, double* out3_dev,
                        unsigned long long* state_out);  // init.cu
}

extern "C" {

int bamcl_resample_limit(int k, int min_samples, int max_samples, double pop_err, double pop_z)
{
  return resample_limit(k, min_samples, max_samples, pop_err, pop_z);
}

int bamcl_systematic_targets(double start, double delta, long n, double* out)
{
  if (!out || n < 0)
    return set_error(BAMCL_E_ARG, "bamcl_systematic_targets: bad arguments");
  std::vector<TargetSeg> segs;
  if (build_target_segments(start, delta, n, segs) != 0)
    return set_error(BAMCL_E_UNSUPPORTED, "target sequence needs more than %d binade segments", MAX_SEGS);
  for (size_t s = 0; s < segs.size(); s++)
  {
    const long long k1 = (s + 1 < segs.size()) ? segs[s + 1].k0 : n;
    for (long long k = segs[s].k0; k < k1; k++)
      out[k] = std::ldexp((double)(segs[s].m0 + (k - segs[s].k0) * segs[s].d), segs[s].eu);
  }
  return BAMCL_OK;
}

// pf_build_cdf + the host's view of it: bounds_out[0 .. n_ranks] (may be null). A rank whose window crosses more binades
// than its summary holds makes EVERY rank see the overflow flag; all of them grow their summaries and repeat.
static int build_cdf_checked(bamcl_pf* pf, double* bounds_out)
{
  const int G = pf_ranks(pf);
  std::vector<double> tmp((size_t)G + 1, 0.0);
  for (int attempt = 0;; attempt++)
  {
    BAMCL_TRY(pf_build_cdf(pf));
    int flags = 0;
    BAMCL_CUDA_OK(cudaMemcpyAsync(tmp.data(), pf->cdf_bounds, sizeof(double) * (size_t)(G + 1), cudaMemcpyDeviceToHost,
                                  pf->ctx->stream));
    BAMCL_CUDA_OK(cudaMemcpyAsync(&flags, pf_flags(pf), sizeof(int), cudaMemcpyDeviceToHost, pf->ctx->stream));
    BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
    if (flags & 4)
    {
      if (attempt >= 8)
        return set_error(BAMCL_E_UNSUPPORTED, "CDF: the ranks' tile summaries keep overflowing");
      pf->cdf_slow_cap *= 8;
      continue;
    }
    if (flags & 2)
      return set_error(BAMCL_E_UNSUPPORTED, "CDF: a running sum left the supported range (subnormal or non-finite weights)");
    break;
  }
  if (bounds_out)
    memcpy(bounds_out, tmp.data(), sizeof(double) * (size_t)(G + 1));
  return BAMCL_OK;
}

int bamcl_pf_cdf(bamcl_pf* pf, double* c_out, long n_plus_1)
{
  if (!pf || !c_out)
    return set_error(BAMCL_E_ARG, "bamcl_pf_cdf: bad arguments");
  long off = 0, n = 0;
  pf_owned(pf, &off, &n);
  if (n_plus_1 != n + 1)
    return set_error(BAMCL_E_ARG, "bamcl_pf_cdf: need sample_count + 1 = %ld entries (a distributed filter: its window)", n + 1);
  pf_apply_shard_window(pf);
  BAMCL_TRY(build_cdf_checked(pf, nullptr));
  BAMCL_CUDA_OK(cudaMemcpyAsync(c_out, pf->cdf.p, sizeof(double) * (size_t)n_plus_1, cudaMemcpyDeviceToHost, pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  return BAMCL_OK;
}

static int resample_impl(bamcl_pf* pf, const bamcl_resample_params* prm, bamcl_resample_result* res, double* aos_out,
                         long aos_cap);

int bamcl_pf_resample(bamcl_pf* pf, const bamcl_resample_params* prm, bamcl_resample_result* res)
{
  return resample_impl(pf, prm, res, nullptr, 0);
}

int bamcl_pf_resample_host(bamcl_pf* pf, const bamcl_resample_params* prm, bamcl_resample_result* res, double* aos_out,
                           long aos_capacity)
{
  if (!aos_out || aos_capacity < 0)
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample_host: bad output buffer");
  return resample_impl(pf, prm, res, aos_out, aos_capacity);
}

int bamcl_pf_resample_multinomial_drawn(bamcl_pf* pf, const bamcl_resample_params* prm, const double* draws,
                                        bamcl_resample_result* res)
{
  if (!pf || !prm || !res || !draws)
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample_multinomial_drawn: bad arguments");
  if (prm->resample_systematic)
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample_multinomial_drawn: the systematic model takes its random poses through "
                                  "bamcl_resample_params::random_poses");
  if (prm->max_samples > pf->max_samples || prm->max_samples <= 0)
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample_multinomial_drawn: max_samples exceeds the filter's capacity");
  if (pf_ranks(pf) > 1 || pf->win_cnt >= 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_resample_multinomial_drawn: not available on a distributed / windowed filter");
  if (prm->n_random_poses > 0 && !prm->random_poses)
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample_multinomial_drawn: random_poses missing");
  for (int k = 0; k < prm->max_samples; k++)  // every -(1 + i) must name a handed-over pose
    if (draws[k] < 0.0 && (long)(-draws[k]) - 1 >= prm->n_random_poses)
      return set_error(BAMCL_E_ARG, "bamcl_pf_resample_multinomial_drawn: draws[%d] names pose %ld of %ld", k,
                       (long)(-draws[k]) - 1, prm->n_random_poses);
  memset(res, 0, sizeof(*res));
  BAMCL_TRY(resample_multinomial(pf, prm, res, draws, prm->random_poses, prm->n_random_poses));
  res->drand48_state = prm->drand48_state;  // nothing was drawn here
  return BAMCL_OK;
}

static int resample_impl(bamcl_pf* pf, const bamcl_resample_params* prm, bamcl_resample_result* res, double* aos_out,
                         long aos_cap)
{
  if (!pf || !prm || !res)
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample: bad arguments");
  if (prm->max_samples > pf->max_samples || prm->max_samples <= 0)
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample: max_samples exceeds the filter's capacity");
  if (!(prm->w_diff >= 0.0))  // NaN when no sensor update preceded the resample
    return set_error(BAMCL_E_STATE, "bamcl_pf_resample: w_diff is NaN/negative (updateSensor must succeed first)");
  memset(res, 0, sizeof(*res));
  if (pf_ranks(pf) > 1)
    return set_error(BAMCL_E_STATE, "bamcl_pf_resample: this filter is one shard of a distributed set; use bamcl_pf_shard_resample");
  if (pf->win_cnt >= 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_resample: a scoring window is set (bamcl_pf_set_window); clear it first");
  if (!prm->resample_systematic)
  {
    BAMCL_TRY(resample_multinomial(pf, prm, res, nullptr, nullptr, 0));
    if (aos_out)
    {
      if (res->sample_count > aos_cap)
        return set_error(BAMCL_E_ARG, "bamcl_pf_resample_host: %d samples do not fit the output buffer", res->sample_count);
      return bamcl_pf_download_samples(pf, aos_out, res->sample_count);
    }
    return BAMCL_OK;
  }

  bamcl_ctx* ctx = pf->ctx;
  SampleSet& set_a = pf->sets[pf->current];
  SampleSet& set_b = pf->sets[1 - pf->current];
  const int P = set_a.sample_count;
  if (P <= 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_resample: the current set is empty");
  BAMCL_TRY(pf_build_cdf(pf));

  // particle_filter.cpp:291-306
  int new_count = resample_limit(set_a.leaf_count, prm->min_samples, prm->max_samples, prm->pop_err, prm->pop_z);
  const double w_diff = prm->w_diff;
  if (w_diff > 0.0)
  {
    new_count *= (1.0 + w_diff);
    if (new_count > prm->max_samples)
      new_count = prm->max_samples;
  }
  const int n_rand = w_diff * new_count;
  const int n_sys = new_count - n_rand;
  // random poses: handed over by the caller (what its random_pose_fn_ returned), or -- when none are and the node's pose
  // function is registered (bamcl_pf_set_random_pose_source) -- generated on the device from the stream, after the
  // start draw (:309 then :316-324)
  PoseSource rp_probe;
  const bool device_poses = n_rand > 0 && !prm->random_poses && pf_pose_source(pf, &rp_probe);
  if (n_rand > 0 && !device_poses && (!prm->random_poses || prm->n_random_poses < n_rand))
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample: %d random poses needed, %ld supplied", n_rand, prm->n_random_poses);
  // :309-310
  const uint64_t state = drand48_next(prm->drand48_state);
  const double start = std::ldexp((double)state, -48);
  const double delta = 1.0 / n_sys;
  std::vector<TargetSeg> segs;
  if (n_sys > 0 && build_target_segments(start, delta, n_sys, segs) != 0)
    return set_error(BAMCL_E_UNSUPPORTED, "target sequence needs more than %d binade segments", MAX_SEGS);

  BAMCL_TRY(pf->src.reserve(sizeof(int) * (size_t)pf->max_samples));
  BAMCL_TRY(pf->h_stage.reserve(sizeof(TargetSeg) * MAX_SEGS));
  BAMCL_TRY(pf->seg_buf.reserve(sizeof(TargetSeg) * MAX_SEGS));
  if (!segs.empty())
  {
    memcpy(pf->h_stage.p, segs.data(), sizeof(TargetSeg) * segs.size());
    BAMCL_CUDA_OK(cudaMemcpyAsync(pf->seg_buf.p, pf->h_stage.p, sizeof(TargetSeg) * segs.size(), cudaMemcpyHostToDevice,
                                  ctx->stream));
  }
  unsigned long long state_after = state;
  if (n_rand > 0)
  {
    BAMCL_TRY(pf->random_poses.reserve(sizeof(double) * 3 * (size_t)n_rand));
    if (device_poses)
      BAMCL_TRY(device_random_poses(pf, state, n_rand, pf->random_poses.as<double>(), &state_after));
    else
      BAMCL_CUDA_OK(cudaMemcpyAsync(pf->random_poses.p, prm->random_poses, sizeof(double) * 3 * (size_t)n_rand,
                                    cudaMemcpyHostToDevice, ctx->stream));
  }
  SelectArgs a;
  a.c = pf->cdf.as<double>();
  a.P = P;
  a.ax = set_a.x.as<double>();
  a.ay = set_a.y.as<double>();
  a.ath = set_a.th.as<double>();
  a.bx = set_b.x.as<double>();
  a.by = set_b.y.as<double>();
  a.bth = set_b.th.as<double>();
  a.bw = set_b.w.as<double>();
  a.src = pf->src.as<int>();
  a.n = new_count;
  a.n_rand = n_rand;
  a.random_poses = pf->random_poses.as<double>();
  a.segs = pf->seg_buf.as<TargetSeg>();
  a.n_segs = (int)segs.size();
  // :343-346 total += 1.0 per sample, then :457-462 weight /= total
  a.weight = 1.0 / (double)new_count;
  a.flags = pf_flags(pf);
  if (aos_out && new_count > aos_cap)
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample_host: %d samples do not fit the output buffer", new_count);
  if (new_count > 0)
    BAMCL_LAUNCH(ctx, k_select, (new_count + 255) / 256, 256, 0, a);
  set_b.sample_count = new_count;
  pf->current = 1 - pf->current;
  // the new set is final once k_select has run: its copy to the host overlaps the pose-bin tree and the clustering
  if (aos_out && new_count > 0)
    BAMCL_TRY(pf_download_async(pf, aos_out, new_count));
  // the pinned segment buffer is reused by later calls
  BAMCL_TRY(build_tree_and_stats(pf, pf->current, prm->dist_threshold, prm->convergence_threshold, res));
  if (aos_out && new_count > 0)
    BAMCL_CUDA_OK(cudaStreamSynchronize(pf->copy_stream));
  int flags = 0;
  BAMCL_CUDA_OK(cudaMemcpyAsync(&flags, pf_flags(pf), sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  if (flags & 2)
    return set_error(BAMCL_E_UNSUPPORTED, "resample: a CDF running sum left the supported range");
  res->sample_count = new_count;
  res->random_poses_used = n_rand;
  res->status_flags = flags & 1;
  res->drand48_state = state_after;
  return BAMCL_OK;
}

int bamcl_pf_init_stats(bamcl_pf* pf, bamcl_resample_result* res)
{
  if (!pf || !res)
    return set_error(BAMCL_E_ARG, "bamcl_pf_init_stats: bad arguments");
  memset(res, 0, sizeof(*res));
  BAMCL_TRY(build_tree_and_stats(pf, pf->current, 0.5, 1e300, res));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  res->sample_count = pf->sets[pf->current].sample_count;
  res->converged = 0;  // initConverged, particle_filter.cpp:163-167
  return BAMCL_OK;
}

int bamcl_pf_resample_sources(bamcl_pf* pf, int* src, long n)
{
  long off = 0, cnt = 0;
  if (pf)
    pf_owned(pf, &off, &cnt);
  if (!pf || !src || n != cnt || pf->src.cap < sizeof(int) * (size_t)(off + n))
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample_sources: n must equal sample_count after a resample (a distributed "
                                  "filter: its window)");
  if (n == 0)
    return BAMCL_OK;
  BAMCL_CUDA_OK(cudaMemcpyAsync(src, pf->src.as<int>() + off, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  return BAMCL_OK;
}

}  // extern "C"

// =============================================================================================
// Distributed filter: updateResample over ONE global set whose samples are owned rank-major by G ranks.
// Global order is rank-major, so every order-dependent reference semantic (sequential CDF, target sequence, first
// occurrence of pose bins) is the single-set one.
// =============================================================================================
namespace
{
struct ShardSelArgs
{
  const double* c;  // cnt + 1, global values
  int P;            // window size
  long off;         // window offset = global index of local particle 0
  const double *ax, *ay, *ath;  // window base
  double *ox, *oy, *oth;        // staging rows (SoA)
  int* osrc;                    // global source index per row
  long k_begin, k_end, row0;    // outputs [k_begin, k_end) of the systematic sequence; row = row0 + k
  const TargetSeg* segs;
  int n_segs;
  int* flags;
};

__global__ void __launch_bounds__(256) k_select_shard(const ShardSelArgs a)
{
  __shared__ TargetSeg s_segs[MAX_SEGS];
  for (int i = threadIdx.x; i < a.n_segs; i += blockDim.x)
    s_segs[i] = a.segs[i];
  __syncthreads();
  const long long k = a.k_begin + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= a.k_end)
    return;
  int lo = 0, hi = a.n_segs - 1;
  while (lo < hi)
  {
    const int mid = (lo + hi + 1) >> 1;
    if (s_segs[mid].k0 <= k)
      lo = mid;
    else
      hi = mid - 1;
  }
  const TargetSeg sg = s_segs[lo];
  const double target = scalbn((double)(sg.m0 + (k - sg.k0) * sg.d), sg.eu);
  int L = 0, H = a.P;
  while (L < H)
  {
    const int mid = (L + H) >> 1;
    if (__ldg(&a.c[mid + 1]) <= target)
      L = mid + 1;
    else
      H = mid;
  }
  int j = L;
  if (j >= a.P)
  {
    // only the rank that owns the end of the CDF is given such targets (bamcl_systematic_plan): SURVEY Q10
    atomicOr(a.flags, 1);
    const double top = a.c[a.P];
    int l2 = 0, h2 = a.P;
    while (l2 < h2)
    {
      const int mid = (l2 + h2) >> 1;
      if (a.c[mid + 1] < top)
        l2 = mid + 1;
      else
        h2 = mid;
    }
    j = min(l2, a.P - 1);
  }
  const long long row = a.row0 + k;
  a.ox[row] = a.ax[j];
  a.oy[row] = a.ay[j];
  a.oth[row] = a.ath[j];
  a.osrc[row] = (int)(a.off + j);
}

__global__ void k_fill_range(double* __restrict__ p, double v, long n)
{
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    p[i] = v;
}
__global__ void k_fill_range_i(int* __restrict__ p, int v, long n)
{
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    p[i] = v;
}
// rows [r0, r0 + n) of packed (x, y, theta) triples -> SoA at the same global rows
__global__ void k_rows3_to_soa(const double* __restrict__ in3, long r0, long n, double* __restrict__ x, double* __restrict__ y,
                               double* __restrict__ th)
{
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  x[r0 + i] = in3[3 * (r0 + i)];
  y[r0 + i] = in3[3 * (r0 + i) + 1];
  th[r0 + i] = in3[3 * (r0 + i) + 2];
}

// received blocks (x | y | theta | source index of `rows` rows each) -> output slots; one thread per received row
constexpr int UNPACK_MAX = 48;
struct UnpackArgs
{
  int n;
  long prefix[UNPACK_MAX + 1];  // rows before piece i among this launch's pieces
  const char* block[UNPACK_MAX];
  long rows[UNPACK_MAX];  // rows of the whole block the piece sits in
  long q[UNPACK_MAX];     // first row of the piece inside its block
  long o[UNPACK_MAX];     // first output slot of the piece
  double *bx, *by, *bth;
  int* bsrc;
};
__global__ void __launch_bounds__(256) k_unpack_moves(const UnpackArgs a)
{
  const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= a.prefix[a.n])
    return;
  int lo = 0, hi = a.n - 1;  // last piece with prefix <= t
  while (lo < hi)
  {
    const int mid = (lo + hi + 1) >> 1;
    if (a.prefix[mid] <= t)
      lo = mid;
    else
      hi = mid - 1;
  }
  const long r = a.q[lo] + (t - a.prefix[lo]), slot = a.o[lo] + (t - a.prefix[lo]);
  const double* x = reinterpret_cast<const double*>(a.block[lo]);
  const long c = a.rows[lo];
  a.bx[slot] = x[r];
  a.by[slot] = x[c + r];
  a.bth[slot] = x[2 * c + r];
  a.bsrc[slot] = reinterpret_cast<const int*>(x + 3 * c)[r];
}

struct Move
{
  long src, dst, kb, ke, pos;
};

// the transfers of the pose exchange, in the order both sides issue them (see bamcl_shard_exchange_plan)
int exchange_plan(const long* runs, int n_runs, long n_random, long n, int G, std::vector<Move>& out)
{
  out.clear();
  const long total = n_random + n;
  std::vector<long> pos(G, 0);
  for (int i = 0; i < n_runs; i++)
  {
    const long kb = runs[3 * i], ke = runs[3 * i + 1], s = runs[3 * i + 2];
    if (s < 0 || s >= G || kb > ke)
      return -1;
    for (int d = 0; d < G; d++)
    {
      const long ob = shard_begin(total, G, d), oe = shard_begin(total, G, d + 1);
      const long a = std::max(kb, ob - n_random), b = std::min(ke, oe - n_random);
      if (a < b)
        out.push_back(Move{s, (long)d, a, b, pos[s] + (a - kb)});
    }
    pos[s] += ke - kb;
  }
  return 0;
}
}  // namespace

extern "C" {

int bamcl_pf_set_window(bamcl_pf* pf, long offset, long count)
{
  if (!pf)
    return set_error(BAMCL_E_ARG, "bamcl_pf_set_window: null pf");
  if (pf_ranks(pf) > 1)
    return set_error(BAMCL_E_STATE, "bamcl_pf_set_window: a distributed filter manages its window itself");
  const int n = pf->sets[pf->current].sample_count;
  if (count < 0)
  {
    pf->win_off = 0;
    pf->win_cnt = -1;
    return BAMCL_OK;
  }
  if (offset < 0 || offset + count > n)
    return set_error(BAMCL_E_ARG, "bamcl_pf_set_window: [%ld, %ld) is outside the set of %d samples", offset, offset + count, n);
  pf->win_off = offset;
  pf->win_cnt = (int)count;
  return BAMCL_OK;
}

int bamcl_systematic_plan(double start, double delta, long n, const double* bounds, int n_ranks, long* runs, int max_runs,
                          int* n_runs)
{
  if (!bounds || !runs || !n_runs || n_ranks < 1 || n < 0 || max_runs < 1)
    return set_error(BAMCL_E_ARG, "bamcl_systematic_plan: bad arguments");
  for (int r = 0; r < n_ranks; r++)
    if (!(bounds[r] <= bounds[r + 1]))
      return set_error(BAMCL_E_ARG, "bamcl_systematic_plan: bounds must be non-decreasing");
  std::vector<TargetSeg> segs;
  if (n > 0 && build_target_segments(start, delta, n, segs) != 0)
    return set_error(BAMCL_E_UNSUPPORTED, "target sequence needs more than %d binade segments", MAX_SEGS);
  int last_mass = -1;  // last rank that holds weight: owner of targets at/after the end of the CDF
  for (int r = 0; r < n_ranks; r++)
    if (bounds[r] < bounds[r + 1])
      last_mass = r;
  if (last_mass < 0 && n > 0)
    return set_error(BAMCL_E_STATE, "bamcl_systematic_plan: the CDF is empty (all weights zero)");
  int cnt = 0;
  for (size_t s = 0; s < segs.size(); s++)
  {
    const long long k0 = segs[s].k0, k1 = (s + 1 < segs.size()) ? segs[s + 1].k0 : (long long)n;
    auto value = [&](long long k) { return std::ldexp((double)(segs[s].m0 + (k - k0) * segs[s].d), segs[s].eu); };
    long long prev_cut = k0;
    for (int i = 1; i <= n_ranks + 1; i++)
    {
      // keys in [prev_cut, cut) have exactly i-1 of bounds[1..G] <= target
      long long cut = k1;
      if (i <= n_ranks)
      {
        long long lo = prev_cut, hi = k1;  // first k with value(k) >= bounds[i]
        while (lo < hi)
        {
          const long long mid = lo + (hi - lo) / 2;
          if (value(mid) >= bounds[i])
            hi = mid;
          else
            lo = mid + 1;
        }
        cut = lo;
      }
      if (cut > prev_cut)
      {
        const int owner = (i - 1 < n_ranks) ? i - 1 : last_mass;
        if (cnt > 0 && runs[3 * (cnt - 1) + 2] == owner && runs[3 * (cnt - 1) + 1] == prev_cut)
          runs[3 * (cnt - 1) + 1] = cut;
        else
        {
          if (cnt >= max_runs)
            return set_error(BAMCL_E_ARG, "bamcl_systematic_plan: more than %d runs", max_runs);
          runs[3 * cnt] = prev_cut;
          runs[3 * cnt + 1] = cut;
          runs[3 * cnt + 2] = owner;
          cnt++;
        }
      }
      prev_cut = cut;
    }
  }
  *n_runs = cnt;
  return BAMCL_OK;
}

int bamcl_shard_exchange_plan(const long* runs, int n_runs, long n_random, long n, int n_ranks, long* moves, int max_moves,
                              int* n_moves)
{
  if (!runs || !moves || !n_moves || n_runs < 0 || n_ranks < 1 || n_random < 0 || n < 0)
    return set_error(BAMCL_E_ARG, "bamcl_shard_exchange_plan: bad arguments");
  std::vector<Move> mv;
  if (exchange_plan(runs, n_runs, n_random, n, n_ranks, mv) != 0)
    return set_error(BAMCL_E_ARG, "bamcl_shard_exchange_plan: malformed runs");
  if ((int)mv.size() > max_moves)
    return set_error(BAMCL_E_ARG, "bamcl_shard_exchange_plan: %zu transfers do not fit %d", mv.size(), max_moves);
  for (size_t i = 0; i < mv.size(); i++)
  {
    moves[5 * i] = mv[i].src;
    moves[5 * i + 1] = mv[i].dst;
    moves[5 * i + 2] = mv[i].kb;
    moves[5 * i + 3] = mv[i].ke;
    moves[5 * i + 4] = mv[i].pos;
  }
  *n_moves = (int)mv.size();
  return BAMCL_OK;
}

int bamcl_pf_shard_attach(bamcl_pf* pf, bamcl_comm* comm)
{
  if (!pf)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_attach: null pf");
  if (comm && comm->ctx != pf->ctx)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_attach: the communicator belongs to another context");
  pf->comm = comm;
  pf->win_off = 0;
  pf->win_cnt = -1;
  pf->box_valid = false;
  if (const char* cap = getenv("BAMCL_CDF_SLOW_CAP"))  // tests: start with tiny summaries to exercise the regrow path
    pf->cdf_slow_cap = std::max(1, atoi(cap));
  pf_apply_shard_window(pf);
  return BAMCL_OK;
}

int bamcl_pf_shard_window(bamcl_pf* pf, long* window_offset, long* window_count)
{
  if (!pf)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_window: null pf");
  if (window_offset)
    *window_offset = pf_win_off(pf);
  if (window_count)
    *window_count = pf_win_cnt(pf);
  return BAMCL_OK;
}

int bamcl_pf_shard_set_count(bamcl_pf* pf, long n_global, long* window_offset, long* window_count)
{
  if (!pf || n_global < 0 || n_global > pf->max_samples)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_set_count: need 0 <= n_global <= max_samples");
  pf->sets[pf->current].sample_count = (int)n_global;
  pf->win_off = 0;
  pf->win_cnt = -1;
  pf_apply_shard_window(pf);
  return bamcl_pf_shard_window(pf, window_offset, window_count);
}

int bamcl_pf_shard_init_stats(bamcl_pf* pf, bamcl_resample_result* res)
{
  if (!pf || !res)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_init_stats: bad arguments");
  memset(res, 0, sizeof(*res));
  pf_apply_shard_window(pf);
  BAMCL_TRY(build_tree_and_stats(pf, pf->current, 0.5, 1e300, res));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  res->sample_count = pf->sets[pf->current].sample_count;
  res->converged = 0;  // initConverged, particle_filter.cpp:163-167
  return BAMCL_OK;
}

int bamcl_pf_shard_bounds(bamcl_pf* pf, double* bounds, int n_plus_1)
{
  if (!pf || !bounds || n_plus_1 != pf_ranks(pf) + 1 || !pf->cdf_bounds)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_bounds: need n_ranks + 1 entries after a resample");
  BAMCL_CUDA_OK(cudaMemcpyAsync(bounds, pf->cdf_bounds, sizeof(double) * (size_t)n_plus_1, cudaMemcpyDeviceToHost, pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  return BAMCL_OK;
}

int bamcl_pf_shard_resample(bamcl_pf* pf, const bamcl_resample_params* prm, bamcl_resample_result* res)
{
  if (!pf || !prm || !res)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_resample: bad arguments");
  if (prm->max_samples > pf->max_samples || prm->max_samples <= 0)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_resample: max_samples exceeds the filter's capacity");
  if (!(prm->w_diff >= 0.0))
    return set_error(BAMCL_E_STATE, "bamcl_pf_shard_resample: w_diff is NaN/negative (updateSensor must succeed first)");
  if (!prm->resample_systematic)
    return set_error(BAMCL_E_UNSUPPORTED, "bamcl_pf_shard_resample: the distributed filter resamples systematically "
                                          "(multinomial stops on a running leaf count, pf_kdtree state per output)");
  memset(res, 0, sizeof(*res));
  bamcl_ctx* ctx = pf->ctx;
  bamcl_comm* comm = pf->comm;
  const int G = pf_ranks(pf), me = pf_rank(pf);
  SampleSet& set_a = pf->sets[pf->current];
  SampleSet& set_b = pf->sets[1 - pf->current];
  if (set_a.sample_count <= 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_shard_resample: the current set is empty");
  pf_apply_shard_window(pf);

  // particle_filter.cpp:291-306 (host arithmetic, identical on every rank)
  int new_count = resample_limit(set_a.leaf_count, prm->min_samples, prm->max_samples, prm->pop_err, prm->pop_z);
  const double w_diff = prm->w_diff;
  if (w_diff > 0.0)
  {
    new_count *= (1.0 + w_diff);
    if (new_count > prm->max_samples)
      new_count = prm->max_samples;
  }
  const int n_rand = w_diff * new_count;
  const int n_sys = new_count - n_rand;
  PoseSource rp_probe;
  const bool device_poses = n_rand > 0 && !prm->random_poses && pf_pose_source(pf, &rp_probe);
  if (n_rand > 0 && !device_poses && (!prm->random_poses || prm->n_random_poses < n_rand))
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_resample: %d random poses needed, %ld supplied", n_rand, prm->n_random_poses);
  // :309-310
  const uint64_t state = drand48_next(prm->drand48_state);
  const double start = std::ldexp((double)state, -48);
  const double delta = 1.0 / n_sys;
  std::vector<TargetSeg> segs;
  if (n_sys > 0 && build_target_segments(start, delta, n_sys, segs) != 0)
    return set_error(BAMCL_E_UNSUPPORTED, "target sequence needs more than %d binade segments", MAX_SEGS);

  // ---- exact CDF of every window and the value at every rank boundary (:280-285)
  std::vector<double> bounds(G + 1, 0.0);
  pf_mark(pf, nullptr);
  BAMCL_TRY(build_cdf_checked(pf, bounds.data()));
  pf_mark(pf, "cdf+bounds");

  // ---- who sources which output slot, and who owns it (host arithmetic, identical on every rank)
  std::vector<long> runs((size_t)3 * (4 * G + 2 * MAX_SEGS + 16));
  int n_runs = 0;
  if (n_sys > 0)
    BAMCL_TRY(bamcl_systematic_plan(start, delta, n_sys, bounds.data(), G, runs.data(), (int)(runs.size() / 3), &n_runs));
  std::vector<Move> moves;
  if (exchange_plan(runs.data(), n_runs, n_rand, n_sys, G, moves) != 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_shard_resample: malformed plan");
  long my_rows = 0;
  for (int i = 0; i < n_runs; i++)
    if (runs[3 * i + 2] == me)
      my_rows += runs[3 * i + 1] - runs[3 * i];
  const int P = pf_win_cnt(pf);
  if (my_rows > 0 && P <= 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_shard_resample: this rank's window is empty but it was given outputs");

  // ---- select + exchange (:326-351 writes set_b in slot order). Every transfer of the plan is one contiguous piece of
  // output slots from one source rank to one owner. Pieces this rank keeps are selected straight into the new set; pieces
  // for another rank are selected into ONE block per destination (x | y | theta | source index, SoA), so the exchange is a
  // single send and a single receive per peer (large messages: NVLink rate instead of per-message latency), and one
  // kernel scatters the received blocks into the new set.
  BAMCL_TRY(pf->src.reserve(sizeof(int) * (size_t)pf->max_samples));
  BAMCL_TRY(pf->h_stage.reserve(sizeof(TargetSeg) * MAX_SEGS));
  BAMCL_TRY(pf->seg_buf.reserve(sizeof(TargetSeg) * MAX_SEGS));
  if (!segs.empty())
  {
    memcpy(pf->h_stage.p, segs.data(), sizeof(TargetSeg) * segs.size());
    BAMCL_CUDA_OK(cudaMemcpyAsync(pf->seg_buf.p, pf->h_stage.p, sizeof(TargetSeg) * segs.size(), cudaMemcpyHostToDevice,
                                  ctx->stream));
  }
  double* bx = set_b.x.as<double>();
  double* by = set_b.y.as<double>();
  double* bth = set_b.th.as<double>();
  int* bsrc = pf->src.as<int>();
  std::vector<long> c_send(G, 0), c_recv(G, 0);
  for (const Move& m : moves)
  {
    if (m.src == me && m.dst != me)
      c_send[m.dst] += m.ke - m.kb;
    if (m.dst == me && m.src != me)
      c_recv[m.src] += m.ke - m.kb;
  }
  auto block_bytes = [](long rows) { return ((size_t)rows * 28 + 15) & ~(size_t)15; };
  std::vector<size_t> s_off(G + 1, 0), r_off(G + 1, 0);
  for (int q = 0; q < G; q++)
  {
    s_off[q + 1] = s_off[q] + block_bytes(c_send[q]);
    r_off[q + 1] = r_off[q] + block_bytes(c_recv[q]);
  }
  BAMCL_TRY(pf->stage_x.reserve(std::max<size_t>(s_off[G], 16)));
  BAMCL_TRY(pf->stage_y.reserve(std::max<size_t>(r_off[G], 16)));
  char* send_base = pf->stage_x.as<char>();
  char* recv_base = pf->stage_y.as<char>();
  {
    std::vector<long> filled(G, 0);
    for (const Move& m : moves)
    {
      if (m.src != me || m.ke == m.kb)
        continue;
      ShardSelArgs a;
      a.c = pf->cdf.as<double>();
      a.P = P;
      a.off = pf_win_off(pf);
      a.ax = set_a.x.as<double>() + a.off;
      a.ay = set_a.y.as<double>() + a.off;
      a.ath = set_a.th.as<double>() + a.off;
      if (m.dst == me)
      {
        a.ox = bx;
        a.oy = by;
        a.oth = bth;
        a.osrc = bsrc;
        a.row0 = n_rand;  // row = output slot
      }
      else
      {
        const long c = c_send[m.dst];
        a.ox = reinterpret_cast<double*>(send_base + s_off[m.dst]);
        a.oy = a.ox + c;
        a.oth = a.ox + 2 * c;
        a.osrc = reinterpret_cast<int*>(a.ox + 3 * c);
        a.row0 = filled[m.dst] - m.kb;
        filled[m.dst] += m.ke - m.kb;
      }
      a.k_begin = m.kb;
      a.k_end = m.ke;
      a.segs = pf->seg_buf.as<TargetSeg>();
      a.n_segs = (int)segs.size();
      a.flags = pf_flags(pf);
      BAMCL_LAUNCH(ctx, k_select_shard, (unsigned)((m.ke - m.kb + 255) / 256), 256, 0, a);
    }
  }
  pf_mark(pf, "plan+select");
  if (G > 1)
  {
    std::vector<P2P> sends, recvs;
    for (int q = 0; q < G; q++)
    {
      if (c_send[q] > 0)
        sends.push_back(P2P{q, send_base + s_off[q], (size_t)c_send[q] * 28});
      if (c_recv[q] > 0)
        recvs.push_back(P2P{q, recv_base + r_off[q], (size_t)c_recv[q] * 28});
    }
    BAMCL_TRY(comm_exchange(comm, sends.data(), (int)sends.size(), recvs.data(), (int)recvs.size()));
    // received blocks -> output slots
    UnpackArgs ua;
    ua.bx = bx;
    ua.by = by;
    ua.bth = bth;
    ua.bsrc = bsrc;
    ua.n = 0;
    ua.prefix[0] = 0;
    std::vector<long> taken(G, 0);
    auto flush = [&]() -> int {
      if (ua.n > 0 && ua.prefix[ua.n] > 0)
        BAMCL_LAUNCH(ctx, k_unpack_moves, (unsigned)((ua.prefix[ua.n] + 255) / 256), 256, 0, ua);
      ua.n = 0;
      ua.prefix[0] = 0;
      return BAMCL_OK;
    };
    for (const Move& m : moves)
    {
      if (m.dst != me || m.src == me || m.ke == m.kb)
        continue;
      if (ua.n == UNPACK_MAX)
        BAMCL_TRY(flush());
      ua.block[ua.n] = recv_base + r_off[m.src];
      ua.rows[ua.n] = c_recv[m.src];
      ua.q[ua.n] = taken[m.src];
      ua.o[ua.n] = n_rand + m.kb;
      ua.prefix[ua.n + 1] = ua.prefix[ua.n] + (m.ke - m.kb);
      taken[m.src] += m.ke - m.kb;
      ua.n++;
    }
    BAMCL_TRY(flush());
  }

  pf_mark(pf, "exchange");
  // ---- this rank's output window: weights 1/n (:343-346, 457-462) and the random poses among its slots (:316-324)
  const long ob = shard_begin(new_count, G, me), oe = shard_begin(new_count, G, me + 1);
  if (oe > ob)
    BAMCL_LAUNCH(ctx, k_fill_range, (unsigned)((oe - ob + 255) / 256), 256, 0, set_b.w.as<double>() + ob, 1.0 / (double)new_count,
                 oe - ob);
  unsigned long long state_after = state;
  if (n_rand > 0)
  {
    // every rank holds all n_rand poses (host-supplied, or generated from the same stream state) and keeps its own rows
    BAMCL_TRY(pf->random_poses.reserve(sizeof(double) * 3 * (size_t)n_rand));
    if (device_poses)
      BAMCL_TRY(device_random_poses(pf, state, n_rand, pf->random_poses.as<double>(), &state_after));
    else
      BAMCL_CUDA_OK(cudaMemcpyAsync(pf->random_poses.p, prm->random_poses, sizeof(double) * 3 * (size_t)n_rand,
                                    cudaMemcpyHostToDevice, ctx->stream));
    const long r0 = ob, r1 = std::min<long>(oe, n_rand);
    if (r1 > r0)
    {
      BAMCL_LAUNCH(ctx, k_rows3_to_soa, (unsigned)((r1 - r0 + 255) / 256), 256, 0, pf->random_poses.as<double>(), r0, r1 - r0, bx,
                   by, bth);
      BAMCL_LAUNCH(ctx, k_fill_range_i, (unsigned)((r1 - r0 + 255) / 256), 256, 0, bsrc + r0, -1, r1 - r0);
    }
  }
  set_b.sample_count = new_count;
  pf->current = 1 - pf->current;
  pf->win_off = 0;
  pf->win_cnt = -1;
  pf_apply_shard_window(pf);
  BAMCL_TRY(build_tree_and_stats(pf, pf->current, prm->dist_threshold, prm->convergence_threshold, res));
  int flags = 0;
  BAMCL_CUDA_OK(cudaMemcpyAsync(&flags, pf_flags(pf), sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  // (bit0 -- a target at/after the end of the CDF, SURVEY Q10 -- is raised on the rank that selected it and was summed
  // over the ranks with the tree's integer results)
  res->sample_count = new_count;
  res->random_poses_used = n_rand;
  res->status_flags = flags & 1;
  res->drand48_state = state_after;
  pf_mark_report(pf);
  return BAMCL_OK;
}

int bamcl_pf_leaf_count(bamcl_pf* pf, int* leaf_count)
{
  if (!pf || !leaf_count)
    return set_error(BAMCL_E_ARG, "bamcl_pf_leaf_count: bad arguments");
  *leaf_count = pf->sets[pf->current].leaf_count;
  return BAMCL_OK;
}

}  // extern "C"
