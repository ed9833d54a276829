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
#include <climits>
#include <cmath>
#include <cstring>

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

__device__ __forceinline__ Fn shfl_fn(const Fn& f, int src)
{
  Fn r;
  r.d0 = __shfl_sync(0xffffffffu, f.d0, src);
  r.d1 = __shfl_sync(0xffffffffu, f.d1, src);
  return r;
}
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

// one block: approximate start value of every tile and the fast/slow decision
__global__ void __launch_bounds__(1024) k_cdf_classify(const double* __restrict__ sums, int n_tiles, int n,
                                                       double* __restrict__ starts, int* __restrict__ tile_e, double c0)
{
  __shared__ double s_part[1024];
  __shared__ double s_total;
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
    double run = c0;  // value of the CDF before the first element (0 unless this is a later shard of a sharded set)
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
  const double err = 2.0 * ((double)n + (double)n_tiles + 64.0) * 1.1102230246251565e-16 * s_total;
  for (int i = lo; i < hi; i++)
  {
    const double a = starts[i] - err, b = starts[i + 1] + err;
    int e = INT_MIN;  // slow
    if (a > 2.2250738585072014e-308 && isfinite(b) && exponent_of(a) == exponent_of(b))
      e = exponent_of(a);
    tile_e[i] = e;
  }
}

struct CdfArgs
{
  const double* w;
  int n;
  double* c;  // n + 1
  const int* tile_e;
  int* status;
  long long* d0;
  long long* d1;
  double* incl;
  int* counter;
  int* flags;  // bit1: unsupported value (subnormal running sum / inconsistent binade)
  double c0;   // c[0]: 0, or the exact end value of the previous shard (sharded filter)
};

// exact inclusive value of tile t-1, by warp 0
__device__ double cdf_lookback(const CdfArgs& a, int t)
{
  const int lane = threadIdx.x & 31;
  Fn G;
  G.d0 = G.d1 = 0;
  int j = t - 1;
  for (;;)
  {
    const int jj = j - lane;
    int st = ST_INCL;
    if (jj >= 0)
    {
      do
      {
        st = *((volatile int*)&a.status[jj]);
      } while (st == ST_EMPTY);
    }
    __threadfence();
    Fn F;
    F.d0 = F.d1 = 0;
    double base = a.c0;  // "tile -1" ends at c[0]
    if (jj >= 0)
    {
      if (st == ST_AGG)
      {
        F.d0 = *((volatile long long*)&a.d0[jj]);
        F.d1 = *((volatile long long*)&a.d1[jj]);
      }
      else
        base = *((volatile double*)&a.incl[jj]);
    }
    const unsigned incl_mask = __ballot_sync(0xffffffffu, st == ST_INCL);
    const int L = incl_mask ? (__ffs(incl_mask) - 1) : 32;
    Fn acc;
    acc.d0 = acc.d1 = 0;
    for (int k = L - 1; k >= 0; k--)
      acc = compose(acc, shfl_fn(F, k));
    G = compose(acc, G);
    if (incl_mask)
    {
      const double b = __shfl_sync(0xffffffffu, base, L);
      if (G.d0 == 0 && G.d1 == 0)
        return b;
      if (!(b >= 2.2250738585072014e-308))
      {
        if (lane == 0)
          atomicOr(a.flags, 2);
        return b;
      }
      const long long m = mantissa_of(b);
      const long long m2 = m + fn_at(G, m);
      if (m2 >= TWO53 && lane == 0)
        atomicOr(a.flags, 2);
      return from_units(m2, exponent_of(b));
    }
    j -= 32;
  }
}

__global__ void __launch_bounds__(CDF_BLOCK) k_cdf_tiles(const CdfArgs a)
{
  __shared__ Fn s_warp[CDF_BLOCK / 32 + 1];
  __shared__ int s_tile, s_first;
  __shared__ double s_val;
  if (threadIdx.x == 0)
    s_tile = atomicAdd(a.counter, 1);
  __syncthreads();
  const int t = s_tile;
  const int idx0 = t * CDF_TILE + threadIdx.x * CDF_ITEMS;
  double wv[CDF_ITEMS];
#pragma unroll
  for (int k = 0; k < CDF_ITEMS; k++)
    wv[k] = (idx0 + k < a.n) ? a.w[idx0 + k] : 0.0;
  if (t == 0 && threadIdx.x == 0)
    a.c[0] = a.c0;
  const int e_fast = a.tile_e[t];
  const bool last_thread = threadIdx.x == CDF_BLOCK - 1;

  if (e_fast != INT_MIN)
  {
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
    Fn agg;
    const Fn ex = block_scan_fn(mine, s_warp, &agg);
    if (threadIdx.x == 0)
    {
      a.d0[t] = agg.d0;
      a.d1[t] = agg.d1;
      __threadfence();
      *((volatile int*)&a.status[t]) = ST_AGG;
    }
    if (threadIdx.x < 32)
    {
      const double v = cdf_lookback(a, t);
      if (threadIdx.x == 0)
        s_val = v;
    }
    __syncthreads();
    const double s_in = s_val;
    if (exponent_of(s_in) != e_fast && threadIdx.x == 0)
      atomicOr(a.flags, 2);
    const long long m_in = mantissa_of(s_in);
    long long m = m_in + fn_at(ex, m_in);
#pragma unroll
    for (int k = 0; k < CDF_ITEMS; k++)
    {
      m += fn_at(f[k], m);
      if (idx0 + k < a.n)
        a.c[idx0 + k + 1] = from_units(m, e_fast);
    }
    if (last_thread)
    {
      if (m >= TWO53)
        atomicOr(a.flags, 2);
      a.incl[t] = from_units(m, e_fast);
      __threadfence();
      *((volatile int*)&a.status[t]) = ST_INCL;
    }
    return;
  }

  // ---- slow tile: needs the exact start value first
  if (threadIdx.x < 32)
  {
    const double v = (t == 0) ? a.c0 : cdf_lookback(a, t);
    if (threadIdx.x == 0)
      s_val = v;
  }
  __syncthreads();
  double s = s_val;
  int start = 0;  // first element of the tile not yet finalised
  const int tile_n = min(CDF_TILE, a.n - t * CDF_TILE);
  const int local0 = threadIdx.x * CDF_ITEMS;
  while (start < tile_n)
  {
    __syncthreads();
    if (threadIdx.x == 0)
      s_first = INT_MAX;
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
        atomicMin(&s_first, first);
      __syncthreads();
      const int fz = min(s_first, tile_n);
#pragma unroll
      for (int k = 0; k < CDF_ITEMS; k++)
      {
        const int li = local0 + k;
        if (li >= start && li < fz)
          a.c[idx0 + k + 1] = 0.0;
        if (li == fz && li < tile_n)
        {
          a.c[idx0 + k + 1] = wv[k];
          s_val = wv[k];
        }
      }
      __syncthreads();
      if (fz < tile_n)
        s = s_val;
      start = fz + 1;
      continue;
    }
    if (!(s >= 2.2250738585072014e-308) || !isfinite(s))
    {
      if (threadIdx.x == 0)
        atomicOr(a.flags, 2);
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
    const Fn ex = block_scan_fn(mine, s_warp, nullptr);
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
      atomicMin(&s_first, cross);
    __syncthreads();
    const int cr = s_first;
    const int stop = min(cr, tile_n);
#pragma unroll
    for (int k = 0; k < CDF_ITEMS; k++)
    {
      const int li = local0 + k;
      if (li >= start && li < stop)
        a.c[idx0 + k + 1] = from_units(mk[k], e);
      if (li == cr)
      {
        // the crossing add itself, as the reference does it
        const double before = from_units(mk[k], e);
        const double after = before + wv[k];
        a.c[idx0 + k + 1] = after;
        s_val = after;
      }
      else if (cr == INT_MAX && li == tile_n - 1)
        s_val = from_units(mk[k], e);
    }
    __syncthreads();
    s = s_val;
    start = (cr == INT_MAX) ? tile_n : cr + 1;
  }
  if (threadIdx.x == 0)
  {
    a.incl[t] = s;
    __threadfence();
    *((volatile int*)&a.status[t]) = ST_INCL;
  }
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
// c[0..n] of the current set's window -> pf->cdf (device), exact sequential rounding, c[0] = c0
int pf_build_cdf(bamcl_pf* pf, double c0)
{
  SampleSet& set = pf->sets[pf->current];
  const int n = pf_win_cnt(pf);
  const double* w = set.w.as<double>() + pf_win_off(pf);
  bamcl_ctx* ctx = pf->ctx;
  BAMCL_TRY(pf->cdf.reserve(sizeof(double) * ((size_t)n + 1)));
  if (n == 0)
  {
    BAMCL_CUDA_OK(cudaMemcpyAsync(pf->cdf.p, &c0, sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    return BAMCL_OK;
  }
  const int tiles = (n + CDF_TILE - 1) / CDF_TILE;
  // scratch layout: sums[tiles] starts[tiles+1] incl[tiles] d0[tiles] d1[tiles] | tile_e[tiles] status[tiles] counter flags
  const size_t n8 = (size_t)tiles * 5 + 2;
  BAMCL_TRY(pf->scratch.reserve(sizeof(double) * n8 + sizeof(int) * ((size_t)tiles * 2 + 4)));
  double* sums = pf->scratch.as<double>();
  double* starts = sums + tiles;
  double* incl = starts + tiles + 1;
  long long* d0 = reinterpret_cast<long long*>(incl + tiles);
  long long* d1 = d0 + tiles;
  int* tile_e = reinterpret_cast<int*>(sums + n8);
  int* status = tile_e + tiles;
  int* counter = status + tiles;
  int* flags = pf_flags(pf);
  BAMCL_CUDA_OK(cudaMemsetAsync(status, 0, sizeof(int) * ((size_t)tiles + 2), ctx->stream));
  BAMCL_CUDA_OK(cudaMemsetAsync(flags, 0, sizeof(int), ctx->stream));
  BAMCL_LAUNCH(ctx, k_cdf_tile_sums, tiles, CDF_BLOCK, 0, w, n, sums);
  BAMCL_LAUNCH(ctx, k_cdf_classify, 1, 1024, 0, sums, tiles, n, starts, tile_e, c0);
  CdfArgs a;
  a.w = w;
  a.c0 = c0;
  a.n = n;
  a.c = pf->cdf.as<double>();
  a.tile_e = tile_e;
  a.status = status;
  a.d0 = d0;
  a.d1 = d1;
  a.incl = incl;
  a.counter = counter;
  a.flags = flags;
  BAMCL_LAUNCH(ctx, k_cdf_tiles, tiles, CDF_BLOCK, 0, a);
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
int resample_multinomial(bamcl_pf* pf, const bamcl_resample_params* prm, bamcl_resample_result* res);  // multinomial.cu
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

int bamcl_pf_cdf(bamcl_pf* pf, double* c_out, long n_plus_1)
{
  if (!pf || !c_out)
    return set_error(BAMCL_E_ARG, "bamcl_pf_cdf: bad arguments");
  const int n = pf->sets[pf->current].sample_count;
  if (n_plus_1 != (long)n + 1)
    return set_error(BAMCL_E_ARG, "bamcl_pf_cdf: need sample_count + 1 = %d entries", n + 1);
  BAMCL_TRY(pf_build_cdf(pf));
  int flags = 0;
  BAMCL_CUDA_OK(cudaMemcpyAsync(c_out, pf->cdf.p, sizeof(double) * (size_t)n_plus_1, cudaMemcpyDeviceToHost, pf->ctx->stream));
  if (n > 0)
    BAMCL_CUDA_OK(cudaMemcpyAsync(&flags, pf_flags(pf), sizeof(int), cudaMemcpyDeviceToHost, pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  if (flags & 2)
    return set_error(BAMCL_E_UNSUPPORTED, "CDF: a running sum left the supported range (subnormal or non-finite weights)");
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
  if (pf->win_cnt >= 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_resample: a shard window is set; use the bamcl_pf_shard_* phases");
  if (!prm->resample_systematic)
  {
    BAMCL_TRY(resample_multinomial(pf, prm, res));
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
  const bool device_poses = n_rand > 0 && !prm->random_poses && pf->rp_map && pf->rp_map->free_count > 0;
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
  if (new_count > 0)
    BAMCL_LAUNCH(ctx, k_select, (new_count + 255) / 256, 256, 0, a);
  set_b.sample_count = new_count;
  pf->current = 1 - pf->current;
  if (aos_out)
  {
    // the new set is final once k_select has run: its copy to the host overlaps the pose-bin tree and the clustering
    if (new_count > aos_cap)
      return set_error(BAMCL_E_ARG, "bamcl_pf_resample_host: %d samples do not fit the output buffer", new_count);
    if (new_count > 0)
      BAMCL_TRY(pf_download_async(pf, aos_out, new_count));
  }
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
  if (!pf || !src || n != pf->sets[pf->current].sample_count || pf->src.cap < sizeof(int) * (size_t)n)
    return set_error(BAMCL_E_ARG, "bamcl_pf_resample_sources: n must equal sample_count after a resample");
  BAMCL_CUDA_OK(cudaMemcpyAsync(src, pf->src.p, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  return BAMCL_OK;
}

}  // extern "C"

// =============================================================================================
// Sharded filter: the phases of resampleSystematic split at the points where ranks exchange data.
// Global order is rank-major, so every order-dependent reference semantic (sequential CDF, target sequence,
// first-occurrence order of pose bins) is the single-set one. Rank r owns the window [off, off + cnt) of the
// current set; c_local[0] = exact end value of rank r-1's CDF, so the concatenation of the local CDFs IS the
// reference's c[0..P].
// =============================================================================================
namespace
{
struct ShardSelArgs
{
  const double* c;  // cnt + 1, global values
  int P;            // window size
  long off;         // window offset = global index of local particle 0
  const double *ax, *ay, *ath;  // window base
  double* out3;     // global AoS buffer, row = output slot, 3 doubles (x, y, theta)
  int* src_out;     // optional: global source index per output slot
  long k_begin, k_end, row0;  // outputs [k_begin, k_end) of the systematic sequence; row = row0 + k
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
  a.out3[3 * row] = a.ax[j];
  a.out3[3 * row + 1] = a.ay[j];
  a.out3[3 * row + 2] = a.ath[j];
  if (a.src_out)
    a.src_out[row] = (int)(a.off + j);
}

__global__ void k_aos3_to_soa(const double* __restrict__ in3, double* __restrict__ x, double* __restrict__ y,
                              double* __restrict__ th, double* __restrict__ w, double weight, int n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  x[i] = in3[3 * (size_t)i];
  y[i] = in3[3 * (size_t)i + 1];
  th[i] = in3[3 * (size_t)i + 2];
  w[i] = weight;
}
}  // namespace

extern "C" {

int bamcl_pf_set_window(bamcl_pf* pf, long offset, long count)
{
  if (!pf)
    return set_error(BAMCL_E_ARG, "bamcl_pf_set_window: null pf");
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

int bamcl_pf_shard_cdf(bamcl_pf* pf, double c0, double* end_out)
{
  if (!pf || !end_out)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_cdf: bad arguments");
  if (!(c0 >= 0.0))
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_cdf: c0 must be >= 0");
  BAMCL_TRY(pf_build_cdf(pf, c0));
  const int n = pf_win_cnt(pf);
  int flags = 0;
  BAMCL_CUDA_OK(cudaMemcpyAsync(end_out, pf->cdf.as<double>() + n, sizeof(double), cudaMemcpyDeviceToHost, pf->ctx->stream));
  if (n > 0)
    BAMCL_CUDA_OK(cudaMemcpyAsync(&flags, pf_flags(pf), sizeof(int), cudaMemcpyDeviceToHost, pf->ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(pf->ctx->stream));
  if (flags & 2)
    return set_error(BAMCL_E_UNSUPPORTED, "CDF: a running sum left the supported range (subnormal or non-finite weights)");
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

int bamcl_pf_shard_select(bamcl_pf* pf, double start, double delta, long n_sys, long row0, long k_begin, long k_end,
                          double* out_aos3_dev, int* src_out_dev_or_null)
{
  if (!pf || !out_aos3_dev || k_begin < 0 || k_end > n_sys || k_begin > k_end)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_select: bad arguments");
  if (k_begin == k_end)
    return BAMCL_OK;
  const int P = pf_win_cnt(pf);
  if (P <= 0)
    return set_error(BAMCL_E_STATE, "bamcl_pf_shard_select: this rank's window is empty but it was given outputs");
  bamcl_ctx* ctx = pf->ctx;
  SampleSet& set_a = pf->sets[pf->current];
  std::vector<TargetSeg> segs;
  if (build_target_segments(start, delta, n_sys, segs) != 0)
    return set_error(BAMCL_E_UNSUPPORTED, "target sequence needs more than %d binade segments", MAX_SEGS);
  BAMCL_TRY(pf->h_stage.reserve(sizeof(TargetSeg) * MAX_SEGS));
  BAMCL_TRY(pf->seg_buf.reserve(sizeof(TargetSeg) * MAX_SEGS));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));  // the pinned segment buffer may still feed an earlier copy
  memcpy(pf->h_stage.p, segs.data(), sizeof(TargetSeg) * segs.size());
  BAMCL_CUDA_OK(cudaMemcpyAsync(pf->seg_buf.p, pf->h_stage.p, sizeof(TargetSeg) * segs.size(), cudaMemcpyHostToDevice,
                                ctx->stream));
  ShardSelArgs a;
  a.c = pf->cdf.as<double>();
  a.P = P;
  a.off = pf_win_off(pf);
  a.ax = set_a.x.as<double>() + a.off;
  a.ay = set_a.y.as<double>() + a.off;
  a.ath = set_a.th.as<double>() + a.off;
  a.out3 = out_aos3_dev;
  a.src_out = src_out_dev_or_null;
  a.k_begin = k_begin;
  a.k_end = k_end;
  a.row0 = row0;
  a.segs = pf->seg_buf.as<TargetSeg>();
  a.n_segs = (int)segs.size();
  a.flags = pf_flags(pf);
  const long long cnt = k_end - k_begin;
  BAMCL_LAUNCH(ctx, k_select_shard, (unsigned)((cnt + 255) / 256), 256, 0, a);
  return BAMCL_OK;
}

int bamcl_pf_shard_install(bamcl_pf* pf, const double* aos3_dev, long n, const bamcl_resample_params* prm,
                           bamcl_resample_result* res)
{
  if (!pf || !aos3_dev || !prm || !res || n <= 0 || n > pf->max_samples)
    return set_error(BAMCL_E_ARG, "bamcl_pf_shard_install: bad arguments");
  memset(res, 0, sizeof(*res));
  bamcl_ctx* ctx = pf->ctx;
  SampleSet& set_b = pf->sets[1 - pf->current];
  BAMCL_LAUNCH(ctx, k_aos3_to_soa, (unsigned)((n + 255) / 256), 256, 0, aos3_dev, set_b.x.as<double>(), set_b.y.as<double>(),
               set_b.th.as<double>(), set_b.w.as<double>(), 1.0 / (double)n, (int)n);
  set_b.sample_count = (int)n;
  pf->current = 1 - pf->current;
  pf->win_off = 0;
  pf->win_cnt = -1;
  BAMCL_TRY(build_tree_and_stats(pf, pf->current, prm->dist_threshold, prm->convergence_threshold, res));
  int flags = 0;
  BAMCL_CUDA_OK(cudaMemcpyAsync(&flags, pf_flags(pf), sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  res->sample_count = (int)n;
  res->status_flags = flags & 1;
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
