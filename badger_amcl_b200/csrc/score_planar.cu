// K-a: likelihood-field scoring of every particle against a planar scan (+ coordAdd + recalcWeight), and
// the scan staging that feeds it. Reference: src/amcl/sensors/planar_scanner.cpp:141-164, 236-323,
// 325-533, 552-640, 642-682, 693-701.
//
// Design (B200): the scan is staged once per call as the list of beams that survive the reference's
// step / max-range / NaN filter, in cell units: (r cos b, r sin b) / resolution. A beam's end point for a
// particle is then a 2x2 rotation in FP64 (4 DFMA; FP32 would flip ~2e-4 of the cell look-ups, SURVEY
// H5) + a round-down magic add for the floor. The per-cell value comes from two small look-ups:
//   k   = tiled u16 field (squared cell distance to the nearest obstacle), 32 MB for 4000^2 -> L2 resident
//   val = table[k] in shared memory, built on the host with glibc's exp/log in the reference's own
//         expression order (so every per-beam term is bit-identical to the reference's pz^3 / pz / log pz)
// A warp owns a batch of PB particles and walks the beams in passes of 32*BPL: lane l holds BPL beams in
// registers (adjacent lanes = adjacent beams = neighbouring cells -> shared sectors), particle data is
// broadcast from shared memory, partial sums stay in registers until the batch ends.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include <vector>

#include "pf.cuh"

using namespace bamcl;

namespace
{
// tuning knobs (overridable at build time for sweeps: BAMCL_EXTRA_NVCC="-DSCORE_BPL=8 ...")
#ifndef SCORE_PB
#define SCORE_PB 16
#endif
#ifndef SCORE_BPL
#define SCORE_BPL 2
#endif
#ifndef SCORE_GB
#define SCORE_GB 8
#endif
#ifndef SCORE_WARPS
#define SCORE_WARPS 8
#endif
#ifdef SCORE_MINB
#define SCORE_BOUNDS __launch_bounds__(SCORE_WARPS * 32, SCORE_MINB)
#else
#define SCORE_BOUNDS __launch_bounds__(SCORE_WARPS * 32)
#endif
constexpr int PB = SCORE_PB;        // particles per warp batch
constexpr int BPL = SCORE_BPL;      // beams per lane per pass
constexpr int GB = SCORE_GB;        // particles per gather batch (GB * BPL loads in flight per lane)
constexpr int WARPS = SCORE_WARPS;  // warps per block
constexpr int PASS_BEAMS = 32 * BPL;

struct ScoreArgs
{
  const double* x;
  const double* y;
  const double* th;
  double* w;
  int n;
  const double2* beams;
  int n_beams;
  const double* table;
  int table_n;
  const int* table_q;  // round(table * 2^q_shift) for the fixed-point path
  double q_scale_inv;  // 2^-q_shift
  const uint16_t* kmap;
  int tiles_x, size_x, size_y;
  const int8_t* cells;
  const float* dist;
  double resolution;
  float origin_x, origin_y;
  double scan_x, scan_y, cos_sa, sin_sa;  // scanner mount (planar_scanner_pose_)
  int model;
  double gompertz_a, gompertz_b, gompertz_c, input_shift, input_scale, output_shift;
  double off_map_factor, non_free_space_factor, non_free_space_radius;
};

// TABLE: 2 = shared memory, 16 interleaved copies (entry r, copy c at double index r*16 + c; lane l reads copy
// l & 15, so both half-warps of a 64-bit load are bank-conflict free whatever the ranks are); 1 = one copy in
// shared memory; 0 = global memory.
//         3 = shared memory, 32 interleaved copies of the table in 32-bit fixed point (entry r, copy l at int
//             index r*32 + l: one wavefront per warp load); terms are summed as integers, so a particle's sum
//             does not depend on the order or grouping of its beams.
template <bool TAIL, int LAYOUT, int TABLE, typename Acc>
__device__ __forceinline__ void score_pass(const ScoreArgs& a, const double* __restrict__ tab, const double4* stash,
                                           int beam0, int lane, Acc (&acc)[PB])
{
  double bx[BPL], by[BPL];
  bool live[BPL];
#pragma unroll
  for (int q = 0; q < BPL; q++)
  {
    const int b = beam0 + q * 32 + lane;
    live[q] = !TAIL || b < a.n_beams;
    const double2 v = live[q] ? __ldg(&a.beams[b]) : make_double2(0.0, 0.0);
    bx[q] = v.x;
    by[q] = v.y;
  }
  const unsigned k_off = (unsigned)(a.table_n - 1);  // rank of "off the map"
  const unsigned copy = TABLE == 2 ? (unsigned)(lane & 15) : 0u;
  // GB particles x BPL beams = 16 gathers are issued back to back before the first one is consumed: the field
  // lives in L2 (random poses leave nothing for L1), so a warp must keep many loads in flight to hide its latency.
#pragma unroll
  for (int g0 = 0; g0 < PB; g0 += GB)
  {
    unsigned kk[GB][BPL];
#pragma unroll
    for (int gg = 0; gg < GB; gg++)
    {
      const double4 p = stash[g0 + gg];  // cos, sin, X, Y (cell units, +0.5 and +size/2 folded in)
#pragma unroll
      for (int q = 0; q < BPL; q++)
      {
        const double gx = fma(p.x, bx[q], fma(-p.y, by[q], p.z));
        const double gy = fma(p.y, bx[q], fma(p.x, by[q], p.w));
        const int ix = floor_to_int(gx);
        const int iy = floor_to_int(gy);
        const bool inb = ((unsigned)ix < (unsigned)a.size_x) && ((unsigned)iy < (unsigned)a.size_y);
        unsigned k = k_off;
        if (inb)
          k = __ldg(&a.kmap[tiled_offset_t<LAYOUT>(ix, iy, a.tiles_x)]);
        kk[gg][q] = k;
      }
    }
    if (TABLE == 3)
    {
      const int* tabq = reinterpret_cast<const int*>(tab);
#pragma unroll
      for (int gg = 0; gg < GB; gg++)
      {
        int part = 0;  // BPL terms fit 32 bits by construction of q_shift
#pragma unroll
        for (int q = 0; q < BPL; q++)
        {
          const int v = tabq[(kk[gg][q] << 5) | (unsigned)lane];
          part += (!TAIL || live[q]) ? v : 0;
        }
        acc[g0 + gg] += part;
      }
    }
    else
    {
#pragma unroll
      for (int gg = 0; gg < GB; gg++)
#pragma unroll
        for (int q = 0; q < BPL; q++)
        {
          const unsigned k = kk[gg][q];
          const double v = TABLE == 2 ? tab[(k << 4) | copy] : tab[k];
          if (!TAIL || live[q])
            acc[g0 + gg] += v;
        }
    }
  }
}

template <int TABLE, int LAYOUT>
__global__ void SCORE_BOUNDS k_score_lf(const ScoreArgs a)
{
  extern __shared__ __align__(16) unsigned char s_raw[];
  double4* s_stash = reinterpret_cast<double4*>(s_raw);                       // [WARPS][PB]
  double* s_tab = reinterpret_cast<double*>(s_raw + sizeof(double4) * WARPS * PB);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (TABLE == 3)
  {
    int* s_q = reinterpret_cast<int*>(s_tab);
    for (int t = threadIdx.x; t < a.table_n * 32; t += blockDim.x)
      s_q[t] = a.table_q[t >> 5];
    __syncthreads();
  }
  else if (TABLE == 2)
  {
    for (int t = threadIdx.x; t < a.table_n * 16; t += blockDim.x)
      s_tab[t] = a.table[t >> 4];
    __syncthreads();
  }
  else if (TABLE == 1)
  {
    for (int t = threadIdx.x; t < a.table_n; t += blockDim.x)
      s_tab[t] = a.table[t];
    __syncthreads();
  }
  const double* tab = TABLE != 0 ? s_tab : a.table;
  double4* stash = s_stash + warp * PB;
  const int n_batches = (a.n + PB - 1) / PB;
  const int warps_total = gridDim.x * WARPS;
  const int n_full = a.n_beams / PASS_BEAMS;

  for (int batch = blockIdx.x * WARPS + warp; batch < n_batches; batch += warps_total)
  {
    // ---- prologue: lane l < PB prepares particle batch*PB + l
    const int idx = batch * PB + lane;
    const bool mine = lane < PB && idx < a.n;
    double w = 0.0, factor = 1.0;
    if (lane < PB)
    {
      const int src = min(idx, a.n - 1);
      const double px = a.x[src], py = a.y[src], pth = a.th[src];
      w = a.w[src];
      double s, c;
      sincos(pth, &s, &c);
      // coordAdd, planar_scanner.cpp:693-701 (the angle is only used through its sine and cosine)
      const double sx = px + a.scan_x * c - a.scan_y * s;
      const double sy = py + a.scan_x * s + a.scan_y * c;
      const double c2 = c * a.cos_sa - s * a.sin_sa;
      const double s2 = s * a.cos_sa + c * a.sin_sa;
      double4 st;
      st.x = c2;
      st.y = s2;
      st.z = (sx - (double)a.origin_x) / a.resolution + 0.5 + (double)(a.size_x / 2);
      st.w = (sy - (double)a.origin_y) / a.resolution + 0.5 + (double)(a.size_y / 2);
      // a non-finite (or absurdly distant) pose is off the map for every beam, as (int)floor(NaN) = INT_MIN makes it in
      // the reference (occupancy_map.cpp:96-97); the magic-number floor below needs |v| < 2^31
      if (!(fabs(st.z) < 1e9) || !(fabs(st.w) < 1e9) || !(fabs(c2) <= 1.0) || !(fabs(s2) <= 1.0))
      {
        st.x = st.y = 0.0;
        st.z = st.w = -1e9;
      }
      stash[lane] = st;
      // recalcWeight, planar_scanner.cpp:642-682 (uses the robot pose, not the scanner pose)
      const int ci = world_to_cell(px, a.origin_x, a.resolution, a.size_x);
      const int cj = world_to_cell(py, a.origin_y, a.resolution, a.size_y);
      if (!(ci >= 0 && ci < a.size_x && cj >= 0 && cj < a.size_y))
        factor = a.off_map_factor;
      else if (a.cells[(size_t)cj * a.size_x + ci] != -1)
        factor = a.non_free_space_factor;
      else
      {
        const double distance = (double)a.dist[(size_t)cj * a.size_x + ci];
        if (distance < a.non_free_space_radius)
        {
          const double delta_d = distance / a.non_free_space_radius;
          factor = a.non_free_space_factor + delta_d * (1.0 - a.non_free_space_factor);
        }
      }
    }
    __syncwarp();

    // ---- beams
    typedef typename std::conditional<TABLE == 3, long long, double>::type Acc;
    Acc acc[PB];
#pragma unroll
    for (int g = 0; g < PB; g++)
      acc[g] = 0;
    for (int pass = 0; pass < n_full; pass++)
      score_pass<false, LAYOUT, TABLE, Acc>(a, tab, stash, pass * PASS_BEAMS, lane, acc);
    if (n_full * PASS_BEAMS < a.n_beams)
      score_pass<true, LAYOUT, TABLE, Acc>(a, tab, stash, n_full * PASS_BEAMS, lane, acc);

    // ---- reduce the PB partial sums across lanes; lane g keeps particle g's sum
    double sum = 0.0;
#pragma unroll
    for (int g = 0; g < PB; g++)
    {
      double v;
      if (TABLE == 3)
        v = (double)warp_sum_ll((long long)acc[g]) * a.q_scale_inv;  // exact: |total| < 2^53, power-of-two scale
      else
        v = warp_sum((double)acc[g]);
      if (lane == g)
        sum = v;
    }
    __syncwarp();

    // ---- epilogue: model-specific combination + map factor
    if (mine)
    {
      double p;
      if (a.model == BAMCL_PLANAR_LF)
        p = 1.0 + sum;  // planar_scanner.cpp:259,315
      else if (a.model == BAMCL_PLANAR_LF_GOMPERTZ)
      {
        if (a.n_beams > 0)  // :624-633
        {
          p = sum / a.n_beams;
          p = p * a.input_scale + a.input_shift;  // applyGompertz :540-550
          p = a.gompertz_a * exp(-1.0 * a.gompertz_b * exp(-1.0 * a.gompertz_c * p));
          p += a.output_shift;
        }
        else
          p = 1.0;
      }
      else
        p = exp(sum);  // :477
      w *= p;
      w *= factor;  // 1.0 where the reference leaves the weight alone (same value)
      a.w[idx] = w;
    }
  }
}

// ---- beam skipping, pass 1 (planar_scanner.cpp:398-451): per staged beam, the number of particles whose end point
// lies on the map within beam_skip_distance of an obstacle. Lane l loads particle l of a group of 32; the group's
// particles are then broadcast one by one and the lanes stride over the beams. Counts are kept per CTA in shared
// memory and flushed once.
__global__ void __launch_bounds__(256) k_skip_count(const ScoreArgs a, const unsigned char* __restrict__ agree,
                                                    int* __restrict__ counts, int layout)
{
  extern __shared__ int s_cnt[];
  for (int b = threadIdx.x; b < a.n_beams; b += blockDim.x)
    s_cnt[b] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  const unsigned k_off = (unsigned)(a.table_n - 1);
  for (int base = warp * 32; base < a.n; base += n_warps * 32)
  {
    const int idx = base + lane;
    double4 st = make_double4(0.0, 0.0, 0.0, 0.0);
    if (idx < a.n)
    {
      const double px = a.x[idx], py = a.y[idx], pth = a.th[idx];
      double sn, cs;
      sincos(pth, &sn, &cs);
      const double sx = px + a.scan_x * cs - a.scan_y * sn;
      const double sy = py + a.scan_x * sn + a.scan_y * cs;
      st.x = cs * a.cos_sa - sn * a.sin_sa;
      st.y = sn * a.cos_sa + cs * a.sin_sa;
      st.z = (sx - (double)a.origin_x) / a.resolution + 0.5 + (double)(a.size_x / 2);
      st.w = (sy - (double)a.origin_y) / a.resolution + 0.5 + (double)(a.size_y / 2);
    }
    const int live = min(32, a.n - base);
    for (int g = 0; g < live; g++)
    {
      double4 p;
      p.x = __shfl_sync(0xffffffffu, st.x, g);
      p.y = __shfl_sync(0xffffffffu, st.y, g);
      p.z = __shfl_sync(0xffffffffu, st.z, g);
      p.w = __shfl_sync(0xffffffffu, st.w, g);
      for (int b = lane; b < a.n_beams; b += 32)
      {
        const double2 v = __ldg(&a.beams[b]);
        const double gx = fma(p.x, v.x, fma(-p.y, v.y, p.z));
        const double gy = fma(p.y, v.x, fma(p.x, v.y, p.w));
        const int ix = floor_to_int(gx), iy = floor_to_int(gy);
        if (((unsigned)ix < (unsigned)a.size_x) && ((unsigned)iy < (unsigned)a.size_y))
        {
          const unsigned k = __ldg(&a.kmap[tiled_offset(ix, iy, a.tiles_x, layout)]);
          if (k != k_off && agree[k])
            atomicAdd(&s_cnt[b], 1);
        }
      }
    }
  }
  __syncthreads();
  for (int b = threadIdx.x; b < a.n_beams; b += blockDim.x)
    if (s_cnt[b])
      atomicAdd(&counts[b], s_cnt[b]);
}
}  // namespace

// =============================================================================================
// host side: scan staging
// =============================================================================================
namespace bamcl
{
// value of the model's per-beam term for a cached distance z, in the reference's expression order
static double lf_term(const bamcl_planar_model& m, double z, double range_max)
{
  const double z_hit_denom = 2 * m.sigma_hit * m.sigma_hit;
  double pz = 0.0;
  if (m.model_type == BAMCL_PLANAR_LF)
  {
    const double z_rand_mult = 1.0 / range_max;
    pz += m.z_hit * std::exp(-(z * z) / z_hit_denom);
    pz += m.z_rand * z_rand_mult;
    return pz * pz * pz;
  }
  if (m.model_type == BAMCL_PLANAR_LF_GOMPERTZ)
  {
    pz += m.z_hit * std::exp(-(z * z) / z_hit_denom);
    pz += m.z_rand;
    return pz;
  }
  const double z_rand_mult = 1.0 / range_max;
  pz += m.z_hit * std::exp(-(z * z) / z_hit_denom);
  pz += m.z_rand * z_rand_mult;
  return std::log(pz);
}

int stage_scan(bamcl_pf* pf, bamcl_map2d* map, const bamcl_planar_model* model, const double* ranges,
               const double* angles, int range_count, double range_max)
{
  if (!pf || !map || !model || range_count < 0 || (range_count > 0 && (!ranges || !angles)))
    return set_error(BAMCL_E_ARG, "stage_scan: bad arguments");
  StagedScan& sc = pf->scan;
  sc.valid = false;
  if (!map->has_dist)  // every planar model reads the cache through recalcWeight
    return set_error(BAMCL_E_STATE, "the map has no distance cache (updateDistancesLUT / upload_distances first)");
  sc.model = *model;
  sc.range_max = range_max;
  sc.map_version = map->version;
  sc.range_count = range_count;
  if (model->max_beams < 2)  // planar_scanner.cpp:128,144: nothing is applied
  {
    sc.n_beams = 0;
    sc.valid = true;
    return BAMCL_OK;
  }
  bamcl_ctx* ctx = pf->ctx;
  const int type = model->model_type;
  int step;
  if (type == BAMCL_PLANAR_LF_PROB)
    step = (int)std::ceil(range_count / static_cast<double>(model->max_beams));  // :339
  else
    step = (range_count - 1) / (model->max_beams - 1);  // :193,:265,:578
  if (type == BAMCL_PLANAR_BEAM)
  {
    if (step < 1)  // the reference never terminates here
      return set_error(BAMCL_E_ARG, "beam model needs range_count-1 >= max_beams-1 (reference loops forever)");
  }
  else if (step < 1)
    step = 1;
  sc.step = step;

  if (type == BAMCL_PLANAR_BEAM)
  {
    BAMCL_TRY(sc.ranges.reserve(sizeof(double) * (size_t)(range_count + 1)));
    BAMCL_TRY(sc.angles.reserve(sizeof(double) * (size_t)(range_count + 1)));
    BAMCL_CUDA_OK(cudaMemcpyAsync(sc.ranges.p, ranges, sizeof(double) * range_count, cudaMemcpyHostToDevice, ctx->stream));
    BAMCL_CUDA_OK(cudaMemcpyAsync(sc.angles.p, angles, sizeof(double) * range_count, cudaMemcpyHostToDevice, ctx->stream));
    BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    sc.valid = true;
    return BAMCL_OK;
  }

  // beams that the reference integrates (:271-282), in cell units
  const size_t max_beams = (size_t)(range_count / step + 2);
  BAMCL_TRY(pf->h_stage.reserve(sizeof(double) * 2 * max_beams + sizeof(double) * 100000));
  double* hb = pf->h_stage.as<double>();
  int nb = 0, slot = 0;
  sc.beam_slot.clear();
  for (int i = 0; i < range_count; i += step, slot++)
  {
    const double obs_range = ranges[i], obs_bearing = angles[i];
    if (obs_range >= range_max)
      continue;
    if (obs_range != obs_range)
      continue;
    hb[2 * nb] = obs_range * std::cos(obs_bearing) / map->resolution;
    hb[2 * nb + 1] = obs_range * std::sin(obs_bearing) / map->resolution;
    sc.beam_slot.push_back(slot);
    nb++;
  }
  sc.n_beams = nb;
  sc.n_slots = slot;
  sc.h_beams.assign(hb, hb + 2 * nb);
  // value table over the rank of k = squared cell distance (map2d.cuh); [n_rep] = in-map cell beyond the
  // radius; [n_rep + 1] = off the map
  const int n_rep = (int)map->rep_k.size();
  const int tn = n_rep + 2;
  double* ht = hb + 2 * max_beams;
  for (int r = 0; r < n_rep; r++)
    ht[r] = lf_term(*model, (double)(float)(std::sqrt((double)map->rep_k[r]) * map->resolution), range_max);
  ht[n_rep] = lf_term(*model, (double)(float)map->max_distance, range_max);
  if (type == BAMCL_PLANAR_LF_PROB)
  {
    // :350,:443: off-map beams use exp(-max^2/denom) with the un-narrowed max distance
    const double z_hit_denom = 2 * model->sigma_hit * model->sigma_hit;
    const double max_dist_prob = std::exp(-(map->max_distance * map->max_distance) / z_hit_denom);
    double pz = 0.0;
    pz += model->z_hit * max_dist_prob;
    pz += model->z_rand * (1.0 / range_max);
    ht[n_rep + 1] = std::log(pz);
  }
  else
    ht[n_rep + 1] = lf_term(*model, map->max_distance, range_max);  // :298,:613: z = getMaxDistanceToObject()
  // fixed-point copy: the largest shift that keeps BPL terms inside 32 bits; unusable if a term is not finite
  {
    double maxabs = 0.0;
    bool finite = true;
    for (int r = 0; r < tn; r++)
    {
      finite = finite && std::isfinite(ht[r]);
      maxabs = std::max(maxabs, std::fabs(ht[r]));
    }
    int* hq = reinterpret_cast<int*>(ht + tn);
    sc.q_shift = -1;
    if (finite && maxabs > 0.0 && maxabs < 1e6)
    {
      int e;
      std::frexp(maxabs * BPL, &e);  // maxabs * BPL < 2^e
      sc.q_shift = 30 - e;
      for (int r = 0; r < tn; r++)
        hq[r] = (int)std::llrint(std::ldexp(ht[r], sc.q_shift));
      BAMCL_TRY(sc.table_q.reserve(sizeof(int) * (size_t)tn));
      BAMCL_CUDA_OK(cudaMemcpyAsync(sc.table_q.p, hq, sizeof(int) * tn, cudaMemcpyHostToDevice, ctx->stream));
    }
  }
  sc.table_n = tn;
  BAMCL_TRY(sc.beams.reserve(sizeof(double) * 2 * (size_t)(nb + 1)));
  BAMCL_TRY(sc.table.reserve(sizeof(double) * (size_t)tn));
  if (nb > 0)
    BAMCL_CUDA_OK(cudaMemcpyAsync(sc.beams.p, hb, sizeof(double) * 2 * nb, cudaMemcpyHostToDevice, ctx->stream));
  BAMCL_CUDA_OK(cudaMemcpyAsync(sc.table.p, ht, sizeof(double) * tn, cudaMemcpyHostToDevice, ctx->stream));
  if (type == BAMCL_PLANAR_LF_PROB && model->do_beamskip)
  {
    // :448-451: an in-map beam agrees with the map when its cached distance is below beam_skip_distance
    unsigned char* ha = reinterpret_cast<unsigned char*>(ht + tn) + sizeof(int) * (size_t)tn;
    for (int r = 0; r < n_rep; r++)
      ha[r] = (double)(float)(std::sqrt((double)map->rep_k[r]) * map->resolution) < model->beam_skip_distance ? 1 : 0;
    ha[n_rep] = (double)(float)map->max_distance < model->beam_skip_distance ? 1 : 0;
    ha[n_rep + 1] = 0;  // off the map: never counted
    BAMCL_TRY(sc.agree.reserve((size_t)tn));
    BAMCL_CUDA_OK(cudaMemcpyAsync(sc.agree.p, ha, (size_t)tn, cudaMemcpyHostToDevice, ctx->stream));
  }
  // the pinned staging buffer is reused by the next call: wait for the copies
  BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  sc.valid = true;
  return BAMCL_OK;
}

int launch_beam_model(bamcl_pf* pf, bamcl_map2d* map);  // score_beam.cu

// enqueue the scoring of the current set with the staged scan (no synchronisation)
int launch_score(bamcl_pf* pf, bamcl_map2d* map)
{
  StagedScan& sc = pf->scan;
  if (!sc.valid)
    return set_error(BAMCL_E_STATE, "no scan staged");
  if (sc.map_version != map->version)
    return set_error(BAMCL_E_STATE, "the map's distance cache changed after the scan was staged");
  SampleSet& set = pf->sets[pf->current];
  if (set.sample_count == 0 || sc.model.max_beams < 2)
    return BAMCL_OK;
  if (sc.model.model_type == BAMCL_PLANAR_BEAM)
    return launch_beam_model(pf, map);
  bamcl_ctx* ctx = pf->ctx;
  ScoreArgs a;
  a.x = set.x.as<double>() + pf_win_off(pf);
  a.y = set.y.as<double>() + pf_win_off(pf);
  a.th = set.th.as<double>() + pf_win_off(pf);
  a.w = set.w.as<double>() + pf_win_off(pf);
  a.n = pf_win_cnt(pf);
  a.beams = sc.beams.as<double2>();
  a.n_beams = sc.n_beams;
  a.table = sc.table.as<double>();
  a.table_n = sc.table_n;
  a.table_q = sc.table_q.as<int>();
  a.q_scale_inv = std::ldexp(1.0, -sc.q_shift);
  a.kmap = map->kmap.as<uint16_t>();
  a.tiles_x = map->tiles_x;
  a.size_x = map->size_x;
  a.size_y = map->size_y;
  a.cells = map->cells.as<int8_t>();
  a.dist = map->dist.as<float>();
  a.resolution = map->resolution;
  a.origin_x = map->origin_x;
  a.origin_y = map->origin_y;
  a.scan_x = sc.model.scanner_pose[0];
  a.scan_y = sc.model.scanner_pose[1];
  a.cos_sa = std::cos(sc.model.scanner_pose[2]);
  a.sin_sa = std::sin(sc.model.scanner_pose[2]);
  a.model = sc.model.model_type;
  a.gompertz_a = sc.model.gompertz_a;
  a.gompertz_b = sc.model.gompertz_b;
  a.gompertz_c = sc.model.gompertz_c;
  a.input_shift = sc.model.input_shift;
  a.input_scale = sc.model.input_scale;
  a.output_shift = sc.model.output_shift;
  a.off_map_factor = sc.model.off_map_factor;
  a.non_free_space_factor = sc.model.non_free_space_factor;
  a.non_free_space_radius = sc.model.non_free_space_radius;

  // ---- beam skipping (planar_scanner.cpp:361-395, 480-529): only for a converged set
  if (sc.model.model_type == BAMCL_PLANAR_LF_PROB && sc.model.do_beamskip && pf->set_converged)
  {
    // On a sharded set the agreement counts are over ALL ranks' windows and the ratio is against the whole set
    // (:483-494): a counts-only pass (bamcl_pf_skip_counts) returns this window's counts, the caller sums them over
    // the ranks and hands the sums back (bamcl_pf_set_skip_counts) before the scoring call.
    const bool counts_only = pf->skip_counts_out != nullptr;
    if (pf->win_cnt >= 0 && !counts_only && !pf->skip_global_set)
      return set_error(BAMCL_E_UNSUPPORTED,
                       "beam skipping on a sharded filter needs the agreement counts of all ranks: bamcl_pf_skip_counts, "
                       "sum over ranks, bamcl_pf_set_skip_counts, then score");
    const bool use_global = pf->skip_global_set && !counts_only;
    const int max_beams = sc.model.max_beams, nb = sc.n_beams;
    const int P = use_global ? (int)pf->skip_global_n : a.n;
    if (use_global && (int)pf->skip_global_counts.size() != nb)
    {
      pf->skip_global_set = false;
      return set_error(BAMCL_E_ARG, "bamcl_pf_set_skip_counts: %d counts supplied, the staged scan has %d beams",
                       (int)pf->skip_global_counts.size(), nb);
    }
    // clearTempData bookkeeping (:378-394, :684-690); temp_obs_ itself is never stored: its entries are recomputed
    if (!counts_only && (pf->skip_max_obs < max_beams || pf->skip_max_samples < P))
    {
      pf->skip_max_obs = max_beams;
      pf->skip_max_samples = std::max(pf->skip_max_samples, P);
      pf->skip_written.assign((size_t)max_beams, 0);
    }
    std::vector<int> counts((size_t)std::max(nb, 1), 0);
    if (use_global)
    {
      for (int b = 0; b < nb; b++)
        counts[b] = pf->skip_global_counts[b];
      pf->skip_global_set = false;  // one scoring call per hand-over
    }
    else if (nb > 0 && a.n > 0)
    {
      BAMCL_TRY(sc.skip_count.reserve(sizeof(int) * (size_t)nb));
      BAMCL_CUDA_OK(cudaMemsetAsync(sc.skip_count.p, 0, sizeof(int) * (size_t)nb, ctx->stream));
      const int cblocks = std::max(1, std::min(kNumSMs * 4, (a.n + 255) / 256));
      BAMCL_LAUNCH(ctx, k_skip_count, cblocks, 256, sizeof(int) * (size_t)nb, a, sc.agree.as<unsigned char>(),
                   sc.skip_count.as<int>(), map->layout);
      BAMCL_CUDA_OK(cudaMemcpyAsync(counts.data(), sc.skip_count.p, sizeof(int) * (size_t)nb, cudaMemcpyDeviceToHost,
                                    ctx->stream));
      BAMCL_CUDA_OK(cudaStreamSynchronize(ctx->stream));
    }
    if (counts_only)
    {
      pf->skip_counts_out->assign(counts.begin(), counts.begin() + nb);
      return BAMCL_OK;
    }
    std::vector<int> obs_count((size_t)max_beams, 0);
    std::vector<char> now((size_t)max_beams, 0);
    for (int b = 0; b < nb; b++)
    {
      obs_count[sc.beam_slot[b]] = counts[b];
      now[sc.beam_slot[b]] = 1;
      pf->skip_written[sc.beam_slot[b]] = std::max(pf->skip_written[sc.beam_slot[b]], P);
    }
    int skipped = 0;
    std::vector<char> mask((size_t)max_beams, 0);
    for (int s = 0; s < max_beams; s++)  // :483-494
    {
      if ((obs_count[s] / static_cast<double>(P)) > sc.model.beam_skip_threshold)
        mask[s] = 1;
      else
        skipped++;
    }
    const bool error = skipped >= (max_beams * sc.model.beam_skip_error_threshold);  // :502-510
    bool zero_all = false;
    if (error)
    {
      // every slot 0..max_beams-1 is integrated (:519-525). A slot this scan did not write still holds what an earlier
      // scan left there: 0.0 if nothing ever did (log 0 = -inf: every weight becomes 0), else that scan's value.
      for (int s = 0; s < max_beams; s++)
        if (!now[s])
        {
          if (pf->skip_written[s] == 0)
            zero_all = true;
          else
            return set_error(BAMCL_E_UNSUPPORTED,
                             "beam skipping: the reference would integrate slot %d's stale temp_obs_ entry of an earlier "
                             "scan (SURVEY Q12); the device recomputes entries and does not keep old scans", s);
        }
    }
    if (zero_all)
    {
      BAMCL_CUDA_OK(cudaMemsetAsync(a.w, 0, sizeof(double) * (size_t)a.n, ctx->stream));  // w *= exp(-inf)
      return BAMCL_OK;
    }
    // the beams that are integrated, in slot order
    std::vector<double> sel;
    sel.reserve(2 * (size_t)nb);
    for (int b = 0; b < nb; b++)
      if (error || mask[sc.beam_slot[b]])
      {
        sel.push_back(sc.h_beams[2 * b]);
        sel.push_back(sc.h_beams[2 * b + 1]);
      }
    const int n_sel = (int)(sel.size() / 2);
    BAMCL_TRY(sc.beams_sel.reserve(sizeof(double) * 2 * (size_t)(n_sel + 1)));
    if (n_sel > 0)
      BAMCL_CUDA_OK(cudaMemcpy(sc.beams_sel.p, sel.data(), sizeof(double) * 2 * (size_t)n_sel, cudaMemcpyHostToDevice));
    a.beams = sc.beams_sel.as<double2>();
    a.n_beams = n_sel;
  }

  const int n_batches = (a.n + PB - 1) / PB;
  int blocks = (n_batches + WARPS - 1) / WARPS;
  static const int per_sm = getenv("BAMCL_SCORE_CTAS") ? atoi(getenv("BAMCL_SCORE_CTAS")) : 4;
  const int max_blocks = kNumSMs * per_sm;
  if (blocks > max_blocks)
    blocks = max_blocks;
  const size_t stash_bytes = sizeof(double4) * WARPS * PB;
  const size_t tab_bytes = sizeof(double) * (size_t)a.table_n;
  // bamcl_pf_set_exact_terms(pf, 1) / BAMCL_EXACT_TERMS=1 keeps every per-beam term a bit-exact f64 (modes 2/1/0); the default fixed-point table
  // is within 2^-28 relative of the largest term per beam (DESIGN.md "K-a").
  static const bool env_exact = getenv("BAMCL_EXACT_TERMS") != nullptr && atoi(getenv("BAMCL_EXACT_TERMS")) != 0;
  const bool exact_terms = env_exact || pf->exact_terms;
  const bool can_fixed = sc.q_shift >= 0 && !exact_terms && tab_bytes * 16 <= 64 * 1024;
  const int table_mode = can_fixed ? 3 : tab_bytes * 16 <= 64 * 1024 ? 2 : tab_bytes <= 40 * 1024 ? 1 : 0;
  const size_t smem_bytes = stash_bytes + (table_mode >= 2 ? tab_bytes * 16 : table_mode == 1 ? tab_bytes : 0);
  BAMCL_CUDA_OK(cudaEventRecord(pf->score_ev0, ctx->stream));
#define BAMCL_SCORE_CASE(L)                                                                              \
  if (table_mode == 3)                                                                                   \
  {                                                                                                      \
    BAMCL_CUDA_OK(cudaFuncSetAttribute((k_score_lf<3, L>), cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                       (int)(stash_bytes + 64 * 1024)));                                 \
    BAMCL_LAUNCH(ctx, (k_score_lf<3, L>), blocks, WARPS * 32, smem_bytes, a);                            \
  }                                                                                                      \
  else if (table_mode == 2)                                                                                   \
  {                                                                                                      \
    BAMCL_CUDA_OK(cudaFuncSetAttribute((k_score_lf<2, L>), cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                       (int)(stash_bytes + 64 * 1024)));                                 \
    BAMCL_LAUNCH(ctx, (k_score_lf<2, L>), blocks, WARPS * 32, smem_bytes, a);                            \
  }                                                                                                      \
  else if (table_mode == 1)                                                                              \
    BAMCL_LAUNCH(ctx, (k_score_lf<1, L>), blocks, WARPS * 32, smem_bytes, a);                            \
  else                                                                                                   \
    BAMCL_LAUNCH(ctx, (k_score_lf<0, L>), blocks, WARPS * 32, smem_bytes, a);
  if (map->layout == 0)
  {
    BAMCL_SCORE_CASE(0)
  }
  else if (map->layout == 1)
  {
    BAMCL_SCORE_CASE(1)
  }
  else
  {
    BAMCL_SCORE_CASE(2)
  }
#undef BAMCL_SCORE_CASE
  BAMCL_CUDA_OK(cudaEventRecord(pf->score_ev1, ctx->stream));
  pf->score_timed = true;
  return BAMCL_OK;
}
}  // namespace bamcl

// ---- sharded beam skipping (SURVEY 8e collective (2)): the per-beam agreement counts cross the ranks
extern "C" int bamcl_pf_skip_counts(bamcl_pf* pf, bamcl_map2d* map, int converged, int* counts_out, int capacity, int* n_out)
{
  if (!pf || !map || !n_out || (capacity > 0 && !counts_out))
    return set_error(BAMCL_E_ARG, "bamcl_pf_skip_counts: bad arguments");
  *n_out = 0;
  StagedScan& sc = pf->scan;
  if (!sc.valid)
    return set_error(BAMCL_E_STATE, "no scan staged");
  if (!(sc.model.model_type == BAMCL_PLANAR_LF_PROB && sc.model.do_beamskip && converged))
    return BAMCL_OK;  // no beam skipping in this update: nothing to exchange
  if (sc.n_beams > capacity)
    return set_error(BAMCL_E_ARG, "bamcl_pf_skip_counts: %d counts, capacity %d", sc.n_beams, capacity);
  std::vector<int> counts;
  pf->set_converged = true;
  pf->skip_counts_out = &counts;
  const int rc = launch_score(pf, map);
  pf->skip_counts_out = nullptr;
  if (rc != BAMCL_OK)
    return rc;
  for (size_t b = 0; b < counts.size(); b++)
    counts_out[b] = counts[b];
  *n_out = (int)counts.size();
  return BAMCL_OK;
}

extern "C" int bamcl_pf_set_skip_counts(bamcl_pf* pf, const int* counts, int n, long global_sample_count)
{
  if (!pf || n < 0 || (n > 0 && !counts) || global_sample_count <= 0)
    return set_error(BAMCL_E_ARG, "bamcl_pf_set_skip_counts: bad arguments");
  pf->skip_global_counts.assign(counts, counts + n);
  pf->skip_global_n = global_sample_count;
  pf->skip_global_set = true;
  return BAMCL_OK;
}
