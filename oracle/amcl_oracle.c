/* oracle/amcl_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C CPU restatement of the reference's hot path; see amcl_oracle.h for scope and pinning.
 * Citations are file:line under /root/reference. The statement order inside every floating-point
 * expression follows the reference so that results are bit-identical to the compiled reference
 * (both are built with -ffp-contract=off).
 */
#define _GNU_SOURCE
#include "amcl_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ====================================================================== drand48 */
/* POSIX: X <- (0x5DEECE66D * X + 0xB) mod 2^48; srand48(s): X = (s << 16) | 0x330E; value X / 2^48 */
void orc_srand48(orc_rng* r, long seed) { r->x = ((((uint64_t)seed) & 0xFFFFFFFFull) << 16) | 0x330Eull; }
double orc_drand48(orc_rng* r)
{
  r->x = (0x5DEECE66Dull * r->x + 0xBull) & 0xFFFFFFFFFFFFull;
  return ldexp((double)r->x, -48);
}
/* pdf_gaussian.cpp:77-97, polar Box-Muller */
double orc_gaussian_draw(orc_rng* rng, double sigma)
{
  double x1, x2, w, r;
  do
  {
    do
    {
      r = orc_drand48(rng);
    } while (r == 0.0);
    x1 = 2.0 * r - 1.0;
    do
    {
      r = orc_drand48(rng);
    } while (r == 0.0);
    x2 = 2.0 * r - 1.0;
    w = x1 * x1 + x2 * x2;
  } while (w > 1.0 || w == 0.0);
  return (sigma * x2 * sqrt(-2.0 * log(w) / w));
}

/* ====================================================================== OccupancyMap */
/* occupancy_map.cpp:90-98 (origin_ members are float, promoted to double in the expression) */
void orc_map2d_world_to_map(const orc_map2d* m, double x, double y, int* i, int* j)
{
  *i = (int)(floor((x - m->origin_x) / m->resolution + 0.5) + m->size_x / 2);
  *j = (int)(floor((y - m->origin_y) / m->resolution + 0.5) + m->size_y / 2);
}
/* occupancy_map.cpp:75-88 */
void orc_map2d_map_to_world(const orc_map2d* m, int i, int j, double* x, double* y)
{
  *x = m->origin_x + (i - m->size_x / 2) * m->resolution;
  *y = m->origin_y + (j - m->size_y / 2) * m->resolution;
}
/* occupancy_map.cpp:100-105 */
int orc_map2d_is_valid(const orc_map2d* m, int i, int j) { return (i >= 0) && (i < m->size_x) && (j >= 0) && (j < m->size_y); }
static unsigned cell_index(const orc_map2d* m, int i, int j) { return (unsigned)i + (unsigned)j * (unsigned)m->size_x; } /* :107-110 */
/* occupancy_map.cpp:64-73; returns float (max distance narrowed to float when off the map) */
float orc_map2d_distance(const orc_map2d* m, int i, int j)
{
  if (orc_map2d_is_valid(m, i, j))
    return m->distances[cell_index(m, i, j)];
  return (float)m->max_distance;
}

/* occupancy_map.cpp:257-364 */
double orc_map2d_calc_range(const orc_map2d* m, double ox, double oy, double oa, double max_range, long* steps_out)
{
  int x0, x1, y0, y1, x, y, xstep, ystep, steep, tmp, deltax, deltay, error, deltaerr;
  long steps = 0;
  orc_map2d_world_to_map(m, ox, oy, &x0, &y0);
  orc_map2d_world_to_map(m, ox + max_range * cos(oa), oy + max_range * sin(oa), &x1, &y1);
  if (steps_out)
    *steps_out = 0;
  if (x0 == x1 && y0 == y1)
    return max_range;
  steep = abs(y1 - y0) > abs(x1 - x0);
  if (steep)
  {
    tmp = x0; x0 = y0; y0 = tmp;
    tmp = x1; x1 = y1; y1 = tmp;
  }
  deltax = abs(x1 - x0);
  deltay = abs(y1 - y0);
  error = 0;
  deltaerr = deltay;
  x = x0;
  y = y0;
  xstep = (x0 < x1) ? 1 : -1;
  ystep = (y0 < y1) ? 1 : -1;
#define ORC_BLOCKED(cx, cy) (!orc_map2d_is_valid(m, (cx), (cy)) || m->cells[cell_index(m, (cx), (cy))] != -1)
  steps++;
  if (steep ? ORC_BLOCKED(y, x) : ORC_BLOCKED(x, y))
  {
    if (steps_out)
      *steps_out = steps;
    return sqrt((double)((x - x0) * (x - x0) + (y - y0) * (y - y0))) * m->resolution;
  }
  while (x != (x1 + xstep * 1))
  {
    x += xstep;
    error += deltaerr;
    if (2 * error >= deltax)
    {
      y += ystep;
      error -= deltax;
    }
    steps++;
    if (steep ? ORC_BLOCKED(y, x) : ORC_BLOCKED(x, y))
    {
      if (steps_out)
        *steps_out = steps;
      return sqrt((double)((x - x0) * (x - x0) + (y - y0) * (y - y0))) * m->resolution;
    }
  }
#undef ORC_BLOCKED
  if (steps_out)
    *steps_out = steps;
  return max_range;
}

/* --- brushfire, occupancy_map.cpp:138-242. The queue is std::priority_queue<OccupancyMapCellData>
 * whose operator< (occupancy_map.h:111-114) orders by the cell's *current cached distance*, so the
 * pop order among equal distances is libstdc++'s binary-heap order; push_heap/pop_heap are restated
 * (bits/stl_heap.h __push_heap/__adjust_heap) to keep that order. */
typedef struct
{
  int i, j, src_i, src_j;
} bf_cell;
typedef struct
{
  bf_cell* v;
  long n, cap;
  const float* dist;
  const orc_map2d* m;
} bf_heap;
static float bf_d(const bf_heap* h, const bf_cell* c) { return h->dist[cell_index(h->m, c->i, c->j)]; }
/* comp(a, b) == (a < b) == dist(a) > dist(b) */
static int bf_less(const bf_heap* h, const bf_cell* a, const bf_cell* b) { return bf_d(h, a) > bf_d(h, b); }
static void bf_push_heap(bf_heap* h, long hole, long top, bf_cell value)
{
  long parent = (hole - 1) / 2;
  while (hole > top && bf_less(h, &h->v[parent], &value))
  {
    h->v[hole] = h->v[parent];
    hole = parent;
    parent = (hole - 1) / 2;
  }
  h->v[hole] = value;
}
static void bf_push(bf_heap* h, bf_cell c)
{
  if (h->n == h->cap)
  {
    h->cap = h->cap ? h->cap * 2 : 1024;
    h->v = (bf_cell*)realloc(h->v, sizeof(bf_cell) * (size_t)h->cap);
  }
  h->v[h->n++] = c;
  bf_push_heap(h, h->n - 1, 0, c);
}
static void bf_pop(bf_heap* h)
{
  if (h->n > 1)
  {
    bf_cell value = h->v[h->n - 1];
    long len = h->n - 1, hole = 0, second = 0;
    h->v[h->n - 1] = h->v[0];
    while (second < (len - 1) / 2)
    {
      second = 2 * (second + 1);
      if (bf_less(h, &h->v[second], &h->v[second - 1]))
        second--;
      h->v[hole] = h->v[second];
      hole = second;
    }
    if ((len & 1) == 0 && second == (len - 2) / 2)
    {
      second = 2 * (second + 1);
      h->v[hole] = h->v[second - 1];
      hole = second - 1;
    }
    bf_push_heap(h, hole, 0, value);
  }
  h->n--;
}

void orc_map2d_update_distances(const orc_map2d* m, double max_distance, float* out)
{
  const int sx = m->size_x, sy = m->size_y;
  unsigned char* marked;
  double* sq; /* CachedDistanceOccupancyMap :122-135 */
  int cell_radius, i, j, k;
  bf_heap h;
  if (max_distance == 0.0) /* :141-145 */
    return;
  marked = (unsigned char*)calloc((size_t)sx * sy, 1);
  cell_radius = (int)floor(max_distance / m->resolution);
  sq = (double*)malloc(sizeof(double) * (size_t)(cell_radius + 2) * (cell_radius + 2));
  for (i = 0; i <= cell_radius + 1; i++)
    for (j = 0; j <= cell_radius + 1; j++)
      sq[i * (cell_radius + 2) + j] = sqrt((double)(i * i + j * j));
  memset(&h, 0, sizeof(h));
  h.dist = out;
  h.m = m;
  /* iterateObstacleCells :162-185 (i outer, j inner) */
  for (i = 0; i < sx; i++)
    for (j = 0; j < sy; j++)
    {
      if (m->cells[cell_index(m, i, j)] == 1)
      {
        bf_cell c = {i, j, i, j};
        out[cell_index(m, i, j)] = 0.0f;
        marked[cell_index(m, i, j)] = 1;
        bf_push(&h, c);
      }
      else
        out[cell_index(m, i, j)] = (float)max_distance;
    }
  /* iterateEmptyCells :187-211: neighbours are offered while the current cell is still queued */
  while (h.n > 0)
  {
    bf_cell cur = h.v[0];
    const int ni[4] = {cur.i - 1, cur.i, cur.i + 1, cur.i};
    const int nj[4] = {cur.j, cur.j - 1, cur.j, cur.j + 1};
    const int ok[4] = {cur.i > 0, cur.j > 0, cur.i < sx - 1, cur.j < sy - 1};
    for (k = 0; k < 4; k++)
    {
      unsigned idx;
      if (!ok[k])
        continue;
      idx = cell_index(m, ni[k], nj[k]);
      if (!marked[idx]) /* updateNode :213-222, enqueue :224-242 */
      {
        int di = abs(ni[k] - cur.src_i), dj = abs(nj[k] - cur.src_j);
        double distance = sq[di * (cell_radius + 2) + dj];
        if (distance <= cell_radius)
        {
          bf_cell c = {ni[k], nj[k], cur.src_i, cur.src_j};
          out[idx] = (float)(distance * m->resolution);
          bf_push(&h, c);
          marked[idx] = 1;
        }
      }
    }
    bf_pop(&h);
  }
  free(h.v);
  free(sq);
  free(marked);
}

/* Exact capped EDT with the reference's value formula: (float)(sqrt(di^2+dj^2) * res) if
 * sqrt(di^2+dj^2) <= floor(max/res) else (float)max (occupancy_map.cpp:125,181,229-232). */
void orc_map2d_exact_edt(const orc_map2d* m, double max_distance, float* out)
{
  const int sx = m->size_x, sy = m->size_y;
  const int R = (int)floor(max_distance / m->resolution);
  const int BIG = 1 << 28;
  int* g = (int*)malloc(sizeof(int) * (size_t)sx * sy); /* distance along j to nearest occupied cell in column i */
  int i, j;
  for (i = 0; i < sx; i++)
  {
    int last = -BIG;
    for (j = 0; j < sy; j++)
    {
      if (m->cells[cell_index(m, i, j)] == 1)
        last = j;
      g[cell_index(m, i, j)] = (j - last > R) ? BIG : j - last;
    }
    last = BIG;
    for (j = sy - 1; j >= 0; j--)
    {
      if (m->cells[cell_index(m, i, j)] == 1)
        last = j;
      if (last - j <= R && last - j < g[cell_index(m, i, j)])
        g[cell_index(m, i, j)] = last - j;
    }
  }
  for (j = 0; j < sy; j++)
    for (i = 0; i < sx; i++)
    {
      long best = (long)BIG * 4;
      int d;
      for (d = -R; d <= R; d++)
      {
        int ii = i + d;
        long gv, k2;
        if (ii < 0 || ii >= sx)
          continue;
        gv = g[cell_index(m, ii, j)];
        if (gv >= BIG)
          continue;
        k2 = (long)d * d + gv * gv;
        if (k2 < best)
          best = k2;
      }
      if (best <= (long)R * R && sqrt((double)best) <= R)
        out[cell_index(m, i, j)] = (float)(sqrt((double)best) * m->resolution);
      else
        out[cell_index(m, i, j)] = (float)max_distance;
    }
  free(g);
}

/* ====================================================================== PlanarScanner */
struct orc_planar_state
{
  int max_samples, max_obs;
  double* temp_obs; /* [max_samples][max_obs] */
};
orc_planar_state* orc_planar_state_create(void) { return (orc_planar_state*)calloc(1, sizeof(orc_planar_state)); }
void orc_planar_state_destroy(orc_planar_state* s)
{
  if (s)
  {
    free(s->temp_obs);
    free(s);
  }
}
/* third-party ros/angles (not vendored, unpinned): fmod(a + pi, 2 pi) form, see oracle/shim/angles */
double orc_normalize_angle(double a)
{
  const double result = fmod(a + M_PI, 2.0 * M_PI);
  if (result <= 0.0)
    return result + M_PI;
  return result - M_PI;
}
/* planar_scanner.cpp:693-701 */
void orc_coord_add(const double a[3], const double b[3], double c[3])
{
  c[0] = b[0] + a[0] * cos(b[2]) - a[1] * sin(b[2]);
  c[1] = b[1] + a[0] * sin(b[2]) + a[1] * cos(b[2]);
  c[2] = b[2] + a[2];
  c[2] = orc_normalize_angle(c[2]);
}
/* planar_scanner.cpp:540-550 */
double orc_apply_gompertz(const orc_planar_model* p, double v)
{
  v = v * p->input_scale + p->input_shift;
  v = p->gompertz_a * exp(-1.0 * p->gompertz_b * exp(-1.0 * p->gompertz_c * v));
  v += p->output_shift;
  return v;
}

/* planar_scanner.cpp:168-234 */
static double calc_beam_model(const orc_map2d* m, const orc_planar_model* P, const double* ranges, const double* angles,
                              int range_count, double range_max, double* s, long n)
{
  double total_weight = 0.0;
  long j;
  int i, step = (range_count - 1) / (P->max_beams - 1);
  if (step < 1) /* the reference loops forever here; the oracle refuses instead */
    return 0.0;
  for (j = 0; j < n; j++)
  {
    double pose[3], p = 1.0;
    orc_coord_add(P->scanner_pose, &s[4 * j], pose);
    for (i = 0; i < range_count; i += step)
    {
      double obs_range = ranges[i], obs_bearing = angles[i];
      double map_range = orc_map2d_calc_range(m, pose[0], pose[1], pose[2] + obs_bearing, range_max, NULL);
      double pz = 0.0, z = obs_range - map_range;
      pz += P->z_hit * exp(-(z * z) / (2 * P->sigma_hit * P->sigma_hit));
      if (z < 0)
        pz += P->z_short * P->lambda_short * exp(-P->lambda_short * obs_range);
      if (obs_range == range_max)
        pz += P->z_max * 1.0;
      if (obs_range < range_max)
        pz += P->z_rand * 1.0 / range_max;
      p += pz * pz * pz;
    }
    s[4 * j + 3] *= p;
    total_weight += s[4 * j + 3];
  }
  return total_weight;
}

/* planar_scanner.cpp:236-323 */
static double calc_lf_model(const orc_map2d* m, const orc_planar_model* P, const double* ranges, const double* angles,
                            int range_count, double range_max, double* s, long n)
{
  double total_weight = 0.0;
  long j;
  int i;
  for (j = 0; j < n; j++)
  {
    double pose[3], p = 1.0;
    double z_hit_denom = 2 * P->sigma_hit * P->sigma_hit;
    double z_rand_mult = 1.0 / range_max;
    int step = (range_count - 1) / (P->max_beams - 1);
    orc_coord_add(P->scanner_pose, &s[4 * j], pose);
    if (step < 1)
      step = 1;
    for (i = 0; i < range_count; i += step)
    {
      double obs_range = ranges[i], obs_bearing = angles[i], pz, z, hx, hy;
      int mi, mj;
      if (obs_range >= range_max)
        continue;
      if (obs_range != obs_range)
        continue;
      pz = 0.0;
      hx = pose[0] + obs_range * cos(pose[2] + obs_bearing);
      hy = pose[1] + obs_range * sin(pose[2] + obs_bearing);
      orc_map2d_world_to_map(m, hx, hy, &mi, &mj);
      if (!orc_map2d_is_valid(m, mi, mj))
        z = m->max_distance;
      else
        z = orc_map2d_distance(m, mi, mj);
      pz += P->z_hit * exp(-(z * z) / z_hit_denom);
      pz += P->z_rand * z_rand_mult;
      p += pz * pz * pz;
    }
    s[4 * j + 3] *= p;
    total_weight += s[4 * j + 3];
  }
  return total_weight;
}

/* planar_scanner.cpp:684-690 */
static void clear_temp_data(orc_planar_state* st, int new_max_samples, int new_max_obs)
{
  st->max_obs = new_max_obs;
  st->max_samples = (int)fmax(st->max_samples, new_max_samples);
  free(st->temp_obs);
  st->temp_obs = (double*)calloc((size_t)st->max_samples * (size_t)st->max_obs, sizeof(double));
}

/* planar_scanner.cpp:325-533 */
static double calc_lf_prob_model(const orc_map2d* m, const orc_planar_model* P, orc_planar_state* st,
                                 const double* ranges, const double* angles, int range_count, double range_max,
                                 double* s, long n, int converged)
{
  double total_weight = 0.0;
  int step = (int)ceil((range_count) / (double)(P->max_beams));
  double z_hit_denom = 2 * P->sigma_hit * P->sigma_hit;
  double z_rand_mult = 1.0 / range_max;
  double max_distance_to_object = m->max_distance;
  double max_dist_prob = exp(-(max_distance_to_object * max_distance_to_object) / z_hit_denom);
  int do_beamskip = P->do_beamskip;
  int* obs_count = (int*)calloc((size_t)P->max_beams, sizeof(int));
  unsigned char* obs_mask = (unsigned char*)calloc((size_t)P->max_beams, 1);
  int beam_ind = 0, i;
  long j;
  if (step < 1)
    step = 1;
  if (do_beamskip && !converged)
    do_beamskip = 0;
  if (do_beamskip)
  {
    int clear_temp = 0;
    if (st->max_obs < P->max_beams)
      clear_temp = 1;
    if (st->max_samples < n)
      clear_temp = 1;
    if (clear_temp)
      clear_temp_data(st, (int)n, P->max_beams);
  }
  for (j = 0; j < n; j++)
  {
    double pose[3], log_p = 0;
    orc_coord_add(P->scanner_pose, &s[4 * j], pose);
    beam_ind = 0;
    for (i = 0; i < range_count; i += step, beam_ind++)
    {
      double obs_range = ranges[i], obs_bearing = angles[i], pz, z, hx, hy;
      int mi, mj;
      if (obs_range >= range_max)
        continue;
      if (obs_range != obs_range)
        continue;
      pz = 0.0;
      hx = pose[0] + obs_range * cos(pose[2] + obs_bearing);
      hy = pose[1] + obs_range * sin(pose[2] + obs_bearing);
      orc_map2d_world_to_map(m, hx, hy, &mi, &mj);
      if (!orc_map2d_is_valid(m, mi, mj))
        pz += P->z_hit * max_dist_prob;
      else
      {
        z = orc_map2d_distance(m, mi, mj);
        if (z < P->beam_skip_distance)
          obs_count[beam_ind] += 1;
        pz += P->z_hit * exp(-(z * z) / z_hit_denom);
      }
      pz += P->z_rand * z_rand_mult;
      if (!do_beamskip)
        log_p += log(pz);
      else
        st->temp_obs[(size_t)j * st->max_obs + beam_ind] = pz;
    }
    if (!do_beamskip)
    {
      s[4 * j + 3] *= exp(log_p);
      total_weight += s[4 * j + 3];
    }
  }
  if (do_beamskip)
  {
    int skipped_beam_count = 0, error = 0;
    for (beam_ind = 0; beam_ind < P->max_beams; beam_ind++)
    {
      if ((obs_count[beam_ind] / (double)n) > P->beam_skip_threshold)
        obs_mask[beam_ind] = 1;
      else
      {
        obs_mask[beam_ind] = 0;
        skipped_beam_count++;
      }
    }
    if (skipped_beam_count >= (beam_ind * P->beam_skip_error_threshold))
      error = 1;
    for (j = 0; j < n; j++)
    {
      double log_p = 0;
      for (beam_ind = 0; beam_ind < P->max_beams; beam_ind++)
        if (error || obs_mask[beam_ind])
          log_p += log(st->temp_obs[(size_t)j * st->max_obs + beam_ind]);
      s[4 * j + 3] *= exp(log_p);
      total_weight += s[4 * j + 3];
    }
  }
  free(obs_count);
  free(obs_mask);
  return total_weight;
}

/* planar_scanner.cpp:552-640 */
static double calc_lf_gompertz_model(const orc_map2d* m, const orc_planar_model* P, const double* ranges,
                                     const double* angles, int range_count, double range_max, double* s, long n)
{
  double total_weight = 0.0;
  long j;
  int i;
  for (j = 0; j < n; j++)
  {
    double pose[3], p, sum_pz = 0.0;
    double z_hit_denom = 2 * P->sigma_hit * P->sigma_hit;
    int step = (range_count - 1) / (P->max_beams - 1), valid_beams = 0;
    orc_coord_add(P->scanner_pose, &s[4 * j], pose);
    if (step < 1)
      step = 1;
    for (i = 0; i < range_count; i += step)
    {
      double obs_range = ranges[i], obs_bearing = angles[i], pz, z, hx, hy;
      int mi, mj;
      if (obs_range >= range_max)
        continue;
      if (obs_range != obs_range)
        continue;
      valid_beams++;
      pz = 0.0;
      hx = pose[0] + obs_range * cos(pose[2] + obs_bearing);
      hy = pose[1] + obs_range * sin(pose[2] + obs_bearing);
      orc_map2d_world_to_map(m, hx, hy, &mi, &mj);
      if (!orc_map2d_is_valid(m, mi, mj))
        z = m->max_distance;
      else
        z = orc_map2d_distance(m, mi, mj);
      pz += P->z_hit * exp(-(z * z) / z_hit_denom);
      pz += P->z_rand;
      sum_pz += pz;
    }
    if (valid_beams > 0)
    {
      p = sum_pz / valid_beams;
      p = orc_apply_gompertz(P, p);
    }
    else
      p = 1.0;
    s[4 * j + 3] *= p;
    total_weight += s[4 * j + 3];
  }
  return total_weight;
}

/* planar_scanner.cpp:642-682 */
double orc_planar_recalc_weight(const orc_map2d* m, const orc_planar_model* P, double* s, long n)
{
  double rv = 0.0;
  long j;
  for (j = 0; j < n; j++)
  {
    int mi, mj;
    orc_map2d_world_to_map(m, s[4 * j], s[4 * j + 1], &mi, &mj);
    if (!orc_map2d_is_valid(m, mi, mj))
      s[4 * j + 3] *= P->off_map_factor;
    else if (m->cells[cell_index(m, mi, mj)] != -1)
      s[4 * j + 3] *= P->non_free_space_factor;
    else
    {
      double distance = orc_map2d_distance(m, mi, mj);
      if (distance < P->non_free_space_radius)
      {
        double delta_d = orc_map2d_distance(m, mi, mj) / P->non_free_space_radius;
        double f = P->non_free_space_factor;
        f += delta_d * (1.0 - P->non_free_space_factor);
        s[4 * j + 3] *= f;
      }
    }
    rv += s[4 * j + 3];
  }
  return rv;
}

/* planar_scanner.cpp:141-164 */
double orc_planar_apply(const orc_map2d* m, const orc_planar_model* P, orc_planar_state* st, const double* ranges,
                        const double* angles, int range_count, double range_max, double* samples, long n,
                        int converged)
{
  double rv = 0.0;
  if (P->max_beams < 2)
    return 0.0;
  if (P->model_type == ORC_PLANAR_BEAM)
    rv = calc_beam_model(m, P, ranges, angles, range_count, range_max, samples, n);
  else if (P->model_type == ORC_PLANAR_LF)
    rv = calc_lf_model(m, P, ranges, angles, range_count, range_max, samples, n);
  else if (P->model_type == ORC_PLANAR_LF_PROB)
    rv = calc_lf_prob_model(m, P, st, ranges, angles, range_count, range_max, samples, n, converged);
  else if (P->model_type == ORC_PLANAR_LF_GOMPERTZ)
    rv = calc_lf_gompertz_model(m, P, ranges, angles, range_count, range_max, samples, n);
  if (rv > 0.0)
    rv = orc_planar_recalc_weight(m, P, samples, n);
  return rv;
}

/* ====================================================================== PFKDTree */
typedef struct
{
  int depth, pivot_dim, key[3], cluster;
  double value;
  long children[2]; /* node index, -1 = NULL */
} kd_node;
struct orc_kdtree
{
  kd_node* nodes; /* creation order == the reference's std::deque (pf_kdtree.cpp:119) */
  long n, cap;
  long root;
  int leaf_count;
};
orc_kdtree* orc_kdtree_create(void)
{
  orc_kdtree* t = (orc_kdtree*)calloc(1, sizeof(orc_kdtree));
  t->root = -1;
  return t;
}
void orc_kdtree_destroy(orc_kdtree* t)
{
  if (t)
  {
    free(t->nodes);
    free(t);
  }
}
void orc_kdtree_clear(orc_kdtree* t)
{
  t->root = -1;
  t->leaf_count = 0;
  t->n = 0;
}
/* pf_kdtree.cpp:35-37, 51-54 */
void orc_pose_key(const double pose[3], int key[3])
{
  key[0] = (int)floor(pose[0] / 0.50);
  key[1] = (int)floor(pose[1] / 0.50);
  key[2] = (int)floor(pose[2] / (10 * M_PI / 180));
}
static int key_eq(const int a[3], const int b[3]) { return a[0] == b[0] && a[1] == b[1] && a[2] == b[2]; }
/* makeLeafNode :117-129 */
static long kd_make_leaf(orc_kdtree* t, const int key[3], double value, int depth)
{
  kd_node* nd;
  if (t->n == t->cap)
  {
    t->cap = t->cap ? t->cap * 2 : 256;
    t->nodes = (kd_node*)realloc(t->nodes, sizeof(kd_node) * (size_t)t->cap);
  }
  nd = &t->nodes[t->n];
  nd->depth = depth;
  nd->key[0] = key[0];
  nd->key[1] = key[1];
  nd->key[2] = key[2];
  nd->value = value;
  nd->pivot_dim = -1;
  nd->cluster = -1;
  nd->children[0] = nd->children[1] = -1;
  t->leaf_count += 1;
  return t->n++;
}
/* insertNode :97-115 + traverseNode :131-150, iteratively */
void orc_kdtree_insert(orc_kdtree* t, const double pose[3], double value)
{
  int key[3], depth = 0;
  long cur;
  orc_pose_key(pose, key);
  if (t->root < 0)
  {
    t->root = kd_make_leaf(t, key, value, 0);
    return;
  }
  cur = t->root;
  for (;;)
  {
    kd_node* nd = &t->nodes[cur];
    int child;
    if (key_eq(key, nd->key))
    {
      nd->value += value;
      return;
    }
    if (nd->pivot_dim == -1)
    {
      int i, split, max_split = 0;
      for (i = 0; i < 3; i++)
      {
        split = abs(key[i] - nd->key[i]);
        if (split > max_split)
        {
          max_split = split;
          nd->pivot_dim = i;
        }
      }
      t->leaf_count -= 1;
    }
    child = key[nd->pivot_dim] > nd->key[nd->pivot_dim];
    depth++;
    if (nd->children[child] < 0)
    {
      long made = kd_make_leaf(t, key, value, depth); /* may realloc: re-fetch the parent */
      t->nodes[cur].children[child] = made;
      return;
    }
    cur = nd->children[child];
  }
}
/* findNode :152-167 (a childless node ends the search) */
static long kd_find(const orc_kdtree* t, const int key[3])
{
  long cur = t->root;
  while (cur >= 0)
  {
    const kd_node* nd = &t->nodes[cur];
    if (key_eq(key, nd->key))
      return cur;
    if (nd->pivot_dim < 0)
      return -1;
    cur = nd->children[key[nd->pivot_dim] > nd->key[nd->pivot_dim]];
  }
  return -1;
}
/* cluster :58-76 + clusterNode :169-194 (explicit stack instead of recursion; same labels) */
void orc_kdtree_cluster(orc_kdtree* t)
{
  long i, sp, *stack = (long*)malloc(sizeof(long) * (size_t)(t->n + 1));
  int cluster_count = 0;
  for (i = 0; i < t->n; i++)
    t->nodes[i].cluster = -1;
  for (i = 0; i < t->n; i++)
  {
    if (t->nodes[i].cluster != -1)
      continue;
    t->nodes[i].cluster = cluster_count++;
    sp = 0;
    stack[sp++] = i;
    while (sp > 0)
    {
      const long cur = stack[--sp];
      int k;
      for (k = 0; k < 27; k++)
      {
        int nk[3];
        long nb;
        nk[0] = t->nodes[cur].key[0] + (k / 9) - 1;
        nk[1] = t->nodes[cur].key[1] + ((k % 9) / 3) - 1;
        nk[2] = t->nodes[cur].key[2] + ((k % 9) % 3) - 1;
        if (key_eq(t->nodes[cur].key, nk))
          continue;
        nb = kd_find(t, nk);
        if (nb < 0 || t->nodes[nb].cluster >= 0)
          continue;
        t->nodes[nb].cluster = t->nodes[cur].cluster;
        stack[sp++] = nb;
      }
    }
  }
  free(stack);
}
int orc_kdtree_get_cluster(const orc_kdtree* t, const double pose[3])
{
  int key[3];
  long nd;
  orc_pose_key(pose, key);
  nd = kd_find(t, key);
  return nd < 0 ? -1 : t->nodes[nd].cluster;
}
int orc_kdtree_leaf_count(const orc_kdtree* t) { return t->leaf_count; }
long orc_kdtree_node_count(const orc_kdtree* t) { return t->n; }
void orc_kdtree_get_nodes(const orc_kdtree* t, int* keys, int* clusters, int* pivots, double* values)
{
  long k;
  for (k = 0; k < t->n; k++)
  {
    keys[3 * k] = t->nodes[k].key[0];
    keys[3 * k + 1] = t->nodes[k].key[1];
    keys[3 * k + 2] = t->nodes[k].key[2];
    clusters[k] = t->nodes[k].cluster;
    pivots[k] = t->nodes[k].pivot_dim;
    values[k] = t->nodes[k].value;
  }
}

/* ====================================================================== ParticleFilter */
typedef struct
{
  int count;
  double weight, mean[3], cov[9], m[4], c[2][2];
} pf_cluster;
typedef struct
{
  int sample_count;
  double* samples; /* AoS x, y, theta, w */
  orc_kdtree* kdtree;
  int cluster_count, cluster_max_count;
  pf_cluster* clusters;
  long clusters_cap;
  double mean[3], cov[9];
  int converged;
} pf_set;
struct orc_pf
{
  int resample_systematic;
  int min_samples, max_samples;
  double w_slow, w_fast, dist_threshold, conv_threshold, pop_err, pop_z, alpha_slow, alpha_fast;
  int current_set;
  pf_set sets[2];
  int converged;
  double* random_poses;
  long random_n, random_next;
  /* alternative random_pose_fn_: Node::uniformPoseGenerator over a free-space list, drawing from the resampler's rng */
  const orc_map2d* uniform_map;
  const int* uniform_free_ij;
  long uniform_free_n;
  double uniform_threshold, uniform_multiplier;
  int* last_src;
};

orc_pf* orc_pf_create(int min_samples, int max_samples, double alpha_slow, double alpha_fast, double conv_thresh)
{
  orc_pf* pf = (orc_pf*)calloc(1, sizeof(orc_pf));
  int j, i;
  pf->resample_systematic = 0; /* PF_RESAMPLE_MULTINOMIAL :46 */
  pf->min_samples = min_samples;
  pf->max_samples = max_samples;
  pf->conv_threshold = conv_thresh;
  pf->pop_err = 0.01;
  pf->pop_z = 3;
  pf->dist_threshold = 0.5;
  for (j = 0; j < 2; j++)
  {
    pf_set* set = &pf->sets[j];
    set->sample_count = max_samples;
    set->samples = (double*)calloc((size_t)max_samples * 4, sizeof(double));
    for (i = 0; i < max_samples; i++)
      set->samples[4 * i + 3] = 1.0 / max_samples;
    set->kdtree = orc_kdtree_create();
    set->cluster_count = 0;
    set->cluster_max_count = max_samples;
  }
  pf->alpha_slow = alpha_slow;
  pf->alpha_fast = alpha_fast;
  pf->last_src = (int*)malloc(sizeof(int) * (size_t)max_samples);
  return pf;
}
void orc_pf_destroy(orc_pf* pf)
{
  int j;
  if (!pf)
    return;
  for (j = 0; j < 2; j++)
  {
    free(pf->sets[j].samples);
    free(pf->sets[j].clusters);
    orc_kdtree_destroy(pf->sets[j].kdtree);
  }
  free(pf->random_poses);
  free(pf->last_src);
  free(pf);
}
void orc_pf_set_resample_model(orc_pf* pf, int systematic) { pf->resample_systematic = systematic; }
void orc_pf_set_population_params(orc_pf* pf, double pop_err, double pop_z)
{
  pf->pop_err = pop_err;
  pf->pop_z = pop_z;
}
void orc_pf_set_decay_rates(orc_pf* pf, double a_slow, double a_fast)
{
  pf->alpha_slow = a_slow;
  pf->alpha_fast = a_fast;
}
void orc_pf_set_random_poses(orc_pf* pf, const double* poses, long n)
{
  free(pf->random_poses);
  pf->random_poses = (double*)malloc(sizeof(double) * 3 * (size_t)(n > 0 ? n : 1));
  memcpy(pf->random_poses, poses, sizeof(double) * 3 * (size_t)n);
  pf->random_n = n;
  pf->random_next = 0;
}
long orc_pf_random_poses_used(const orc_pf* pf) { return pf->random_next; }
void orc_uniform_poses(const orc_map2d* m, const int* free_ij, long free_n, orc_rng* rng, double starting_weight_threshold,
                       double deweight_multiplier, long n, double* poses3);
void orc_pf_set_uniform_pose_source(orc_pf* pf, const orc_map2d* m, const int* free_ij, long free_n, double threshold,
                                    double multiplier)
{
  pf->uniform_map = m;
  pf->uniform_free_ij = free_ij;
  pf->uniform_free_n = free_n;
  pf->uniform_threshold = threshold;
  pf->uniform_multiplier = multiplier;
  pf->random_next = 0;
}
static void random_pose(orc_pf* pf, orc_rng* rng, double out[3])
{
  out[0] = out[1] = out[2] = 0.0;
  if (pf->uniform_map)
  {
    orc_uniform_poses(pf->uniform_map, pf->uniform_free_ij, pf->uniform_free_n, rng, pf->uniform_threshold,
                      pf->uniform_multiplier, 1, out);
    pf->random_next++;
    return;
  }
  if (pf->random_next < pf->random_n)
  {
    memcpy(out, &pf->random_poses[3 * pf->random_next], sizeof(double) * 3);
    pf->random_next++;
  }
}

/* computeClusterStatsForSet :505-539 with helpers :541-636 */
static void compute_cluster_stats(pf_set* set)
{
  double m[4] = {0.0, 0.0, 0.0, 0.0}, c[4] = {0.0, 0.0, 0.0, 0.0}, weight = 0.0;
  long need;
  int i, j, k;
  orc_kdtree_cluster(set->kdtree);
  /* the reference re-initialises cluster_max_count (= max_samples) clusters; only as many clusters as
   * there are tree nodes can ever be touched, so that many are kept here. */
  need = orc_kdtree_node_count(set->kdtree) + 1;
  if (need > set->clusters_cap)
  {
    set->clusters = (pf_cluster*)realloc(set->clusters, sizeof(pf_cluster) * (size_t)need);
    set->clusters_cap = need;
  }
  memset(set->clusters, 0, sizeof(pf_cluster) * (size_t)need);
  set->cluster_count = 0;
  for (i = 0; i < set->sample_count; i++)
  {
    const double* s = &set->samples[4 * i];
    int cidx = orc_kdtree_get_cluster(set->kdtree, s); /* getClusterIndexOfSampleInSet :570-580 */
    pf_cluster* cl;
    if (cidx >= set->cluster_max_count)
      continue;
    if (cidx < 0)
      continue;
    if (cidx + 1 > set->cluster_count)
      set->cluster_count = cidx + 1;
    cl = &set->clusters[cidx];
    cl->count += 1; /* addSampleStatsToCluster :582-601 */
    cl->weight += s[3];
    cl->m[0] += s[3] * s[0];
    cl->m[1] += s[3] * s[1];
    cl->m[2] += s[3] * cos(s[2]);
    cl->m[3] += s[3] * sin(s[2]);
    for (j = 0; j < 2; j++)
      for (k = 0; k < 2; k++)
        cl->c[j][k] += s[3] * s[j] * s[k];
    weight += s[3]; /* addSampleStatsToSet :603-618 */
    m[0] += s[3] * s[0];
    m[1] += s[3] * s[1];
    m[2] += s[3] * cos(s[2]);
    m[3] += s[3] * sin(s[2]);
    for (j = 0; j < 2; j++)
      for (k = 0; k < 2; k++)
        c[2 * j + k] += s[3] * s[j] * s[k];
  }
  for (i = 0; i < set->cluster_count; i++) /* normalizeCluster :555-568 */
  {
    pf_cluster* cl = &set->clusters[i];
    cl->mean[0] = cl->m[0] / cl->weight;
    cl->mean[1] = cl->m[1] / cl->weight;
    cl->mean[2] = atan2(cl->m[3], cl->m[2]);
    for (j = 0; j < 2; j++)
      for (k = 0; k < 2; k++)
        cl->cov[3 * j + k] = (cl->c[j][k] / cl->weight - cl->mean[j] * cl->mean[k]);
    cl->cov[8] = -2 * log(sqrt(cl->m[2] * cl->m[2] + cl->m[3] * cl->m[3]));
  }
  set->mean[0] = m[0] / weight; /* computeSetStats :620-636 */
  set->mean[1] = m[1] / weight;
  set->mean[2] = atan2(m[3], m[2]);
  for (j = 0; j < 2; j++)
    for (k = 0; k < 2; k++)
      set->cov[3 * j + k] = c[2 * j + k] / weight - set->mean[j] * set->mean[k];
  set->cov[8] = -2 * log(sqrt(m[2] * m[2] + m[3] * m[3]));
}

/* initWithPoseFn :136-162 */
void orc_pf_init_with_poses(orc_pf* pf, const double* poses, long n)
{
  pf_set* set = &pf->sets[pf->current_set];
  int i;
  orc_kdtree_clear(set->kdtree);
  set->sample_count = pf->max_samples;
  for (i = 0; i < set->sample_count; i++)
  {
    double* s = &set->samples[4 * i];
    s[3] = 1.0 / pf->max_samples;
    memcpy(s, &poses[3 * (i % n)], sizeof(double) * 3);
    orc_kdtree_insert(set->kdtree, s, s[3]);
  }
  pf->w_slow = pf->w_fast = 0.0;
  compute_cluster_stats(set);
  set->converged = 0; /* initConverged :164-168 */
  pf->converged = 0;
}
int orc_pf_sample_count(const orc_pf* pf) { return pf->sets[pf->current_set].sample_count; }
void orc_pf_get_samples(const orc_pf* pf, double* aos)
{
  const pf_set* set = &pf->sets[pf->current_set];
  memcpy(aos, set->samples, sizeof(double) * 4 * (size_t)set->sample_count);
}
void orc_pf_set_samples(orc_pf* pf, const double* aos, long n)
{
  pf_set* set = &pf->sets[pf->current_set];
  if (n > set->sample_count)
    n = set->sample_count;
  memcpy(set->samples, aos, sizeof(double) * 4 * (size_t)n);
}
double* orc_pf_samples_ptr(orc_pf* pf) { return pf->sets[pf->current_set].samples; }

/* ====================================================================== global localisation helpers */
/* node_2d.cpp:320-338 */
long orc_map2d_free_space(const orc_map2d* m, double non_free_space_radius, int* ij_out)
{
  long n = 0;
  int i, j;
  for (i = 0; i < m->size_x; i++)
    for (j = 0; j < m->size_y; j++)
      if (m->cells[cell_index(m, i, j)] == -1 && orc_map2d_distance(m, i, j) > non_free_space_radius)
      {
        if (ij_out)
        {
          ij_out[2 * n] = i;
          ij_out[2 * n + 1] = j;
        }
        n++;
      }
  return n;
}
/* node.cpp:823-840 */
static void random_free_space_pose(const orc_map2d* m, const int* free_ij, long free_n, orc_rng* rng, double* p)
{
  const unsigned int k = orc_drand48(rng) * (size_t)free_n;
  orc_map2d_map_to_world(m, free_ij[2 * k], free_ij[2 * k + 1], &p[0], &p[1]);
  p[2] = orc_drand48(rng) * 2 * M_PI - M_PI;
}
/* node.cpp:847-868 with scorePose(p) == 1.0 */
void orc_uniform_poses(const orc_map2d* m, const int* free_ij, long free_n, orc_rng* rng, double starting_weight_threshold,
                       double deweight_multiplier, long n, double* poses3)
{
  long s;
  for (s = 0; s < n; s++)
  {
    double good = starting_weight_threshold;
    double* p = &poses3[3 * s];
    random_free_space_pose(m, free_ij, free_n, rng, p);
    if (good > 0.0 && deweight_multiplier < 1.0 && deweight_multiplier >= 0.0)
      while (1.0 < good)
      {
        random_free_space_pose(m, free_ij, free_n, rng, p);
        good *= deweight_multiplier;
      }
  }
}

/* ====================================================================== Odom::updateAction */
/* odom.cpp:318-321: angleDiff(a, b) = angles::shortest_angular_distance(b, a) = normalize_angle(a - b) */
static double odom_angle_diff(double a, double b) { return orc_normalize_angle(a - b); }
/* odom.cpp:145-148: distance of a rotation from the nearer of 0 and pi */
static double odom_rot_noise(double rot)
{
  const double to_zero = fabs(odom_angle_diff(rot, 0.0)), to_pi = fabs(odom_angle_diff(rot, M_PI));
  return to_pi < to_zero ? to_pi : to_zero;
}

void orc_pf_update_action(orc_pf* pf, orc_rng* rng, int model_type, const double* alpha, const double* pose,
                          const double* delta, const double* absolute_motion)
{
  pf_set* set = &pf->sets[pf->current_set];
  const double a1 = alpha[0], a2 = alpha[1], a3 = alpha[2], a4 = alpha[3], a5 = alpha[4];
  const double old_theta = pose[2] - delta[2]; /* :83-85, only the heading of the old pose is used */
  const double trans = sqrt(delta[0] * delta[0] + delta[1] * delta[1]);
  const double rot = delta[2];
  int i;
  if (model_type == 1 || model_type == 3)
  {
    /* omni :88-123, omni corrected :171-206 (standard deviations instead of variances) */
    double sd_trans = a3 * (trans * trans) + a1 * (rot * rot);
    double sd_rot = a4 * (rot * rot) + a2 * (trans * trans);
    double sd_strafe = a1 * (rot * rot) + a5 * (trans * trans);
    if (model_type == 3)
    {
      sd_trans = sqrt(sd_trans);
      sd_rot = sqrt(sd_rot);
      sd_strafe = sqrt(sd_strafe);
    }
    for (i = 0; i < set->sample_count; i++)
    {
      double* s = &set->samples[4 * i];
      const double bearing = odom_angle_diff(atan2(delta[1], delta[0]), old_theta) + s[2];
      const double cb = cos(bearing), sb = sin(bearing);
      const double trans_hat = trans + orc_gaussian_draw(rng, sd_trans);
      const double rot_hat = rot + orc_gaussian_draw(rng, sd_rot);
      const double strafe_hat = 0 + orc_gaussian_draw(rng, sd_strafe);
      s[0] += (trans_hat * cb + strafe_hat * sb);
      s[1] += (trans_hat * sb - strafe_hat * cb);
      s[2] += rot_hat;
    }
  }
  else if (model_type == 0 || model_type == 2)
  {
    /* diff :125-169, diff corrected :208-255 */
    const double rot1 = trans < 0.01 ? 0.0 : odom_angle_diff(atan2(delta[1], delta[0]), old_theta);
    const double rot2 = odom_angle_diff(delta[2], rot1);
    const double n1 = odom_rot_noise(rot1), n2 = odom_rot_noise(rot2);
    double sd_rot1 = a1 * n1 * n1 + a2 * trans * trans;
    double sd_trans = a3 * trans * trans + a4 * n1 * n1 + a4 * n2 * n2;
    double sd_rot2 = a1 * n2 * n2 + a2 * trans * trans;
    if (model_type == 2)
    {
      sd_rot1 = sqrt(sd_rot1);
      sd_trans = sqrt(sd_trans);
      sd_rot2 = sqrt(sd_rot2);
    }
    for (i = 0; i < set->sample_count; i++)
    {
      double* s = &set->samples[4 * i];
      const double rot1_hat = odom_angle_diff(rot1, orc_gaussian_draw(rng, sd_rot1));
      const double trans_hat = trans - orc_gaussian_draw(rng, sd_trans);
      const double rot2_hat = odom_angle_diff(rot2, orc_gaussian_draw(rng, sd_rot2));
      const double heading = s[2] + rot1_hat;
      s[0] += trans_hat * cos(heading);
      s[1] += trans_hat * sin(heading);
      s[2] += rot1_hat + rot2_hat;
    }
  }
  else
  {
    /* gaussian :257-311: noise from the absolute motion since the last update, applied along the mid heading */
    const double at2 = absolute_motion[0] * absolute_motion[0], as2 = absolute_motion[1] * absolute_motion[1],
                 ar2 = absolute_motion[2] * absolute_motion[2];
    const double sd_rot = sqrt(a1 * ar2 + a2 * at2);
    const double sd_trans = sqrt(a3 * at2 + a4 * ar2);
    const double sd_strafe = sqrt(a4 * ar2 + a5 * as2);
    for (i = 0; i < set->sample_count; i++)
    {
      double* s = &set->samples[4 * i];
      const double heading = s[2] + delta[2] / 2;
      const double ch = cos(heading), sh = sin(heading);
      const double bearing = odom_angle_diff(atan2(delta[1], delta[0]), old_theta) + s[2];
      const double cb = cos(bearing), sb = sin(bearing);
      const double trans_hat = orc_gaussian_draw(rng, sd_trans);
      const double strafe_hat = orc_gaussian_draw(rng, sd_strafe);
      const double rot_hat = orc_gaussian_draw(rng, sd_rot);
      s[0] += (trans * cb);
      s[1] += (trans * sb);
      s[2] += rot;
      s[0] += (trans_hat * ch + strafe_hat * sh);
      s[1] += (trans_hat * sh - strafe_hat * ch);
      s[2] += rot_hat;
    }
  }
}

/* ParticleFilter::updateSensor :237-266 */
void orc_pf_normalise(orc_pf* pf, double total)
{
  pf_set* set = &pf->sets[pf->current_set];
  int i;
  if (total > 0.0)
  {
    double w_avg = 0.0;
    for (i = 0; i < set->sample_count; i++)
    {
      w_avg += set->samples[4 * i + 3];
      set->samples[4 * i + 3] /= total;
    }
    w_avg /= set->sample_count;
    if (pf->w_slow == 0.0)
      pf->w_slow = w_avg;
    else
      pf->w_slow += pf->alpha_slow * (w_avg - pf->w_slow);
    if (pf->w_fast == 0.0)
      pf->w_fast = w_avg;
    else
      pf->w_fast += pf->alpha_fast * (w_avg - pf->w_fast);
  }
  else
  {
    for (i = 0; i < set->sample_count; i++)
      set->samples[4 * i + 3] = 1.0 / set->sample_count;
  }
}
/* planar_scanner.cpp:125-137 */
int orc_pf_update_sensor_planar(orc_pf* pf, const orc_map2d* m, const orc_planar_model* p, orc_planar_state* st,
                                const double* ranges, const double* angles, int range_count, double range_max)
{
  pf_set* set = &pf->sets[pf->current_set];
  double total;
  if (p->max_beams < 2)
    return 0;
  total = orc_planar_apply(m, p, st, ranges, angles, range_count, range_max, set->samples, set->sample_count,
                           set->converged);
  orc_pf_normalise(pf, total);
  return 1;
}

/* resampleLimit :475-502 */
int orc_pf_resample_limit(const orc_pf* pf, int k)
{
  double a, b, c, x, kd;
  int n;
  if (k <= 1)
    return pf->max_samples;
  kd = (double)k;
  a = 1;
  b = 2 / (9 * (kd - 1));
  c = sqrt(2 / (9 * (kd - 1))) * pf->pop_z;
  x = a - b + c;
  n = (int)ceil((k - 1) / (2 * pf->pop_err) * x * x * x);
  if (n < pf->min_samples)
    return pf->min_samples;
  if (n > pf->max_samples)
    return pf->max_samples;
  return n;
}

/* resampleSystematic :269-354 */
static double resample_systematic(orc_pf* pf, double w_diff, orc_rng* rng)
{
  pf_set* set_a = &pf->sets[pf->current_set];
  pf_set* set_b = &pf->sets[(pf->current_set + 1) % 2];
  const int na = set_a->sample_count;
  double* c = (double*)malloc(sizeof(double) * (size_t)(na + 2));
  double total = 0, start, delta, target;
  int i, c_i, new_count, num_random, num_sys;
  c[0] = 0.0;
  for (i = 0; i < na; i++)
    c[i + 1] = c[i] + set_a->samples[4 * i + 3];
  c[na + 1] = c[na]; /* the reference may read one past the table; keep it defined */
  new_count = orc_pf_resample_limit(pf, orc_kdtree_leaf_count(set_a->kdtree));
  if (w_diff > 0.0)
  {
    new_count = (int)(new_count * (1.0 + w_diff));
    if (new_count > pf->max_samples)
      new_count = pf->max_samples;
  }
  set_b->sample_count = new_count;
  num_random = (int)(w_diff * set_b->sample_count);
  num_sys = set_b->sample_count - num_random;
  start = orc_drand48(rng);
  delta = 1.0 / num_sys;
  for (c_i = 0; c_i < na; c_i++)
    if ((c[c_i] <= start) && (start < c[c_i + 1]))
      break;
  for (i = 0; i < num_random; ++i)
  {
    double* sb = &set_b->samples[4 * i];
    random_pose(pf, rng, sb);
    sb[3] = 1.0;
    total += sb[3];
    orc_kdtree_insert(set_b->kdtree, sb, sb[3]);
    pf->last_src[i] = -1;
  }
  target = start;
  for (; i < set_b->sample_count; ++i)
  {
    double* sb = &set_b->samples[4 * i];
    int guard = 0;
    while (!((c[c_i] <= target) && (target < c[c_i + 1])))
    {
      c_i++;
      if (c_i >= na)
      {
        c_i = 0;
        if (++guard > 1) /* the reference hangs here when target >= c[P] (Q10): clamp to the last
                            particle with non-zero weight so that tests terminate */
        {
          for (c_i = na - 1; c_i > 0 && !(c[c_i] < c[c_i + 1]); c_i--)
            ;
          break;
        }
      }
    }
    target += delta;
    if (target > 1.0)
      target -= 1.0;
    memcpy(sb, &set_a->samples[4 * c_i], sizeof(double) * 3);
    sb[3] = 1.0;
    total += sb[3];
    orc_kdtree_insert(set_b->kdtree, sb, sb[3]);
    pf->last_src[i] = c_i;
  }
  free(c);
  return total;
}

/* resampleMultinomial :356-420 */
static double resample_multinomial(orc_pf* pf, double w_diff, orc_rng* rng)
{
  pf_set* set_a = &pf->sets[pf->current_set];
  pf_set* set_b = &pf->sets[(pf->current_set + 1) % 2];
  const int na = set_a->sample_count;
  double* c = (double*)malloc(sizeof(double) * (size_t)(na + 1));
  double total = 0;
  int i;
  c[0] = 0.0;
  for (i = 0; i < na; i++)
    c[i + 1] = c[i] + set_a->samples[4 * i + 3];
  set_b->sample_count = 0;
  while (set_b->sample_count < pf->max_samples)
  {
    const int out = set_b->sample_count++;
    double* sb = &set_b->samples[4 * out];
    if (orc_drand48(rng) < w_diff)
    {
      random_pose(pf, rng, sb);
      pf->last_src[out] = -1;
    }
    else
    {
      double r = orc_drand48(rng);
      for (i = 0; i < na; i++)
        if ((c[i] <= r) && (r < c[i + 1]))
          break;
      if (i >= na) /* ROS_ASSERT(i < sample_count) in the reference */
        i = na - 1;
      memcpy(sb, &set_a->samples[4 * i], sizeof(double) * 3);
      pf->last_src[out] = i;
    }
    sb[3] = 1.0;
    total += sb[3];
    orc_kdtree_insert(set_b->kdtree, sb, sb[3]);
    if (set_b->sample_count > orc_pf_resample_limit(pf, orc_kdtree_leaf_count(set_b->kdtree)))
      break;
  }
  free(c);
  return total;
}

/* updateConverged :170-220 */
static void update_converged(orc_pf* pf)
{
  pf_set* set = &pf->sets[pf->current_set];
  double mean_x = 0, mean_y = 0, percent_converged;
  int particles_converged = 0, k;
  for (k = 0; k < set->sample_count; k++)
  {
    mean_x += set->samples[4 * k];
    mean_y += set->samples[4 * k + 1];
  }
  mean_x /= set->sample_count;
  mean_y /= set->sample_count;
  for (k = 0; k < set->sample_count; k++)
    if (fabs(set->samples[4 * k] - mean_x) <= pf->dist_threshold &&
        fabs(set->samples[4 * k + 1] - mean_y) <= pf->dist_threshold)
      particles_converged++;
  percent_converged = (float)particles_converged / (float)k * 100; /* float arithmetic, :206 */
  if (percent_converged >= pf->conv_threshold)
  {
    set->converged = 1;
    pf->converged = 1;
  }
  else
  {
    set->converged = 0;
    pf->converged = 0;
  }
}

/* updateResample :423-471 */
void orc_pf_update_resample(orc_pf* pf, orc_rng* rng)
{
  pf_set* set_b = &pf->sets[(pf->current_set + 1) % 2];
  double w_diff, total;
  int i;
  orc_kdtree_clear(set_b->kdtree);
  w_diff = 1.0 - pf->w_fast / pf->w_slow;
  if (w_diff < 0.0)
    w_diff = 0.0;
  if (pf->resample_systematic)
    total = resample_systematic(pf, w_diff, rng);
  else
    total = resample_multinomial(pf, w_diff, rng);
  if (w_diff > 0.0)
    pf->w_slow = pf->w_fast = 0.0;
  for (i = 0; i < set_b->sample_count; i++)
    set_b->samples[4 * i + 3] /= total;
  compute_cluster_stats(set_b);
  pf->current_set = (pf->current_set + 1) % 2;
  update_converged(pf);
}
void orc_pf_get_resample_sources(const orc_pf* pf, int* src)
{
  memcpy(src, pf->last_src, sizeof(int) * (size_t)pf->sets[pf->current_set].sample_count);
}
void orc_pf_get_averages(const orc_pf* pf, double* w_slow, double* w_fast)
{
  *w_slow = pf->w_slow;
  *w_fast = pf->w_fast;
}
void orc_pf_set_averages(orc_pf* pf, double w_slow, double w_fast)
{
  pf->w_slow = w_slow;
  pf->w_fast = w_fast;
}
int orc_pf_leaf_count(const orc_pf* pf) { return orc_kdtree_leaf_count(pf->sets[pf->current_set].kdtree); }
long orc_pf_node_count(const orc_pf* pf) { return orc_kdtree_node_count(pf->sets[pf->current_set].kdtree); }
int orc_pf_is_converged(const orc_pf* pf) { return pf->converged; }
int orc_pf_cluster_count(const orc_pf* pf) { return pf->sets[pf->current_set].cluster_count; }
void orc_pf_get_sample_clusters(const orc_pf* pf, int* labels)
{
  const pf_set* set = &pf->sets[pf->current_set];
  int k;
  for (k = 0; k < set->sample_count; k++)
    labels[k] = orc_kdtree_get_cluster(set->kdtree, &set->samples[4 * k]);
}
void orc_pf_get_cluster_stats(const orc_pf* pf, double* out)
{
  const pf_set* set = &pf->sets[pf->current_set];
  int k;
  for (k = 0; k < set->cluster_count; k++)
  {
    const pf_cluster* c = &set->clusters[k];
    double* o = out + 10 * k;
    o[0] = c->count;
    o[1] = c->weight;
    o[2] = c->mean[0];
    o[3] = c->mean[1];
    o[4] = c->mean[2];
    o[5] = c->cov[0];
    o[6] = c->cov[1];
    o[7] = c->cov[3];
    o[8] = c->cov[4];
    o[9] = c->cov[8];
  }
}
void orc_pf_get_set_stats(const orc_pf* pf, double* out)
{
  const pf_set* set = &pf->sets[pf->current_set];
  out[0] = set->mean[0];
  out[1] = set->mean[1];
  out[2] = set->mean[2];
  out[3] = set->cov[0];
  out[4] = set->cov[1];
  out[5] = set->cov[3];
  out[6] = set->cov[4];
  out[7] = set->cov[8];
}
void orc_pf_cdf(const orc_pf* pf, double* c)
{
  const pf_set* set = &pf->sets[pf->current_set];
  int i;
  c[0] = 0.0;
  for (i = 0; i < set->sample_count; i++)
    c[i + 1] = c[i] + set->samples[4 * i + 3];
}

/* ====================================================================== OctoMap */
struct orc_map3d
{
  double resolution, max_distance, max_distance_ratio;
  int* src; /* occupied voxel coordinates (i, j, k) after convertWorldToMap of the leaf centres */
  long n_src;
  int cmin[3], cmax[3];
  int width;
  uint32_t num_poses;
  int num_z;
  uint32_t* pose_indices;
  uint8_t* ratios;
  size_t ratios_n, ratios_cap;
  int lut_created;
};
/* octomap.cpp:98-109 */
void orc_map3d_world_to_map(const double xyz[3], double res, int ijk[3])
{
  ijk[0] = (int)floor(xyz[0] / res + 0.5);
  ijk[1] = (int)floor(xyz[1] / res + 0.5);
  ijk[2] = (int)floor(xyz[2] / res + 0.5);
}
static void m3_recount(orc_map3d* m)
{
  m->width = m->cmax[0] - m->cmin[0] + 1;
  m->num_poses = (uint32_t)(m->width * (m->cmax[1] - m->cmin[1] + 1));
  m->num_z = m->cmax[2] - m->cmin[2] + 1;
}
/* initFromOctree :53-74; getMetricMin/Max as octomap's calcMinMax over the leaves (third-party) */
orc_map3d* orc_map3d_create(double resolution, const int* keys, long n_keys, double max_distance)
{
  orc_map3d* m = (orc_map3d*)calloc(1, sizeof(orc_map3d));
  double mn[3], mx[3];
  long k;
  int a;
  m->resolution = resolution;
  m->max_distance = max_distance;
  m->max_distance_ratio = max_distance / 255;
  m->src = (int*)malloc(sizeof(int) * 3 * (size_t)(n_keys > 0 ? n_keys : 1));
  m->n_src = n_keys;
  for (a = 0; a < 3; a++)
  {
    mn[a] = 1.7976931348623157e308;
    mx[a] = -1.7976931348623157e308;
  }
  for (k = 0; k < n_keys; k++)
  {
    double c[3];
    for (a = 0; a < 3; a++)
    {
      double lo;
      c[a] = (keys[3 * k + a] + 0.5) * resolution;
      lo = c[a] - resolution / 2.0;
      if (lo < mn[a])
        mn[a] = lo;
      if (lo + resolution > mx[a])
        mx[a] = lo + resolution;
    }
    orc_map3d_world_to_map(c, resolution, &m->src[3 * k]); /* iterateObstacleCells :220-227 */
  }
  orc_map3d_world_to_map(mn, resolution, m->cmin);
  orc_map3d_world_to_map(mx, resolution, m->cmax);
  m3_recount(m);
  return m;
}
void orc_map3d_destroy(orc_map3d* m)
{
  if (m)
  {
    free(m->src);
    free(m->pose_indices);
    free(m->ratios);
    free(m);
  }
}
/* :112-123 */
int orc_map3d_is_pose_valid(const orc_map3d* m, int i, int j)
{
  return (i <= m->cmax[0] && i >= m->cmin[0] && j <= m->cmax[1] && j >= m->cmin[1]);
}
static int m3_voxel_valid(const orc_map3d* m, int i, int j, int k)
{
  return orc_map3d_is_pose_valid(m, i, j) && k <= m->cmax[2] && k >= m->cmin[2];
}
/* getDistanceToObject :336-350 */
double orc_map3d_distance(const orc_map3d* m, int i, int j, int k)
{
  uint32_t col;
  if (m->lut_created && !m3_voxel_valid(m, i, j, k))
    return m->max_distance;
  col = m->pose_indices[(j - m->cmin[1]) * m->width + (i - m->cmin[0])];
  return m->ratios[col + (uint32_t)(k - m->cmin[2])] * m->max_distance_ratio;
}
/* setDistanceToObject :315-333 */
static void m3_set(orc_map3d* m, int i, int j, int k, double d)
{
  uint32_t pose_index = (uint32_t)((j - m->cmin[1]) * m->width + (i - m->cmin[0]));
  uint32_t start = m->pose_indices[pose_index];
  if (start == 0)
  {
    start = (uint32_t)m->ratios_n;
    m->pose_indices[pose_index] = start;
    if (m->ratios_n + (size_t)m->num_z > m->ratios_cap)
    {
      m->ratios_cap = (m->ratios_cap ? m->ratios_cap * 2 : 4096) + (size_t)m->num_z;
      m->ratios = (uint8_t*)realloc(m->ratios, m->ratios_cap);
    }
    memset(m->ratios + m->ratios_n, 255, (size_t)m->num_z);
    m->ratios_n += (size_t)m->num_z;
  }
  d = d < m->max_distance ? d : m->max_distance;
  d = d / m->max_distance * 255;
  m->ratios[start + (uint32_t)(k - m->cmin[2])] = (uint8_t)(int)floor(d);
}
static int idx3_desc(const void* a, const void* b)
{
  const int* x = (const int*)a;
  const int* y = (const int*)b;
  int t;
  for (t = 0; t < 3; t++)
    if (x[t] != y[t])
      return x[t] < y[t] ? 1 : -1;
  return 0;
}
typedef struct
{
  int i, j, k, si, sj, sk;
} m3_cell;
/* updateDistancesLUT :174-207, iterateObstacleCells :209-251, iterateEmptyCells :253-287, enqueue :290-312 */
void orc_map3d_update_distances(orc_map3d* m)
{
  static const int SHIFTS[6][3] = {{-1, 0, 0}, {0, -1, 0}, {0, 0, -1}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  int cell_radius, a, b, c, R2;
  double* cd;
  int* ord;
  long n_ord = 0, k, head = 0, tail = 0, qcap;
  m3_cell* q;
  if (m->max_distance == 0.0)
    return;
  free(m->pose_indices);
  m->pose_indices = (uint32_t*)calloc(m->num_poses ? m->num_poses : 1, sizeof(uint32_t));
  m->ratios_n = 0;
  if ((size_t)m->num_z > m->ratios_cap)
  {
    m->ratios_cap = (size_t)m->num_z + 4096;
    m->ratios = (uint8_t*)realloc(m->ratios, m->ratios_cap);
  }
  memset(m->ratios, 255, (size_t)m->num_z); /* shared "nothing near" column at offset 0, :188 */
  m->ratios_n = (size_t)m->num_z;
  m->lut_created = 0;
  cell_radius = (int)floor(m->max_distance / m->resolution); /* CachedDistanceOctoMap :152-170 */
  R2 = cell_radius + 2;
  cd = (double*)malloc(sizeof(double) * (size_t)R2 * R2 * R2);
  for (a = 0; a < R2; a++)
    for (b = 0; b < R2; b++)
      for (c = 0; c < R2; c++)
        cd[(a * R2 + b) * R2 + c] = sqrt((double)(a * a + b * b + c * c)) * m->resolution;
  ord = (int*)malloc(sizeof(int) * 3 * (size_t)(m->n_src > 0 ? m->n_src : 1));
  for (k = 0; k < m->n_src; k++)
  {
    const int* s = &m->src[3 * k];
    if (!m3_voxel_valid(m, s[0], s[1], s[2]))
      continue;
    m3_set(m, s[0], s[1], s[2], 0.0);
    memcpy(&ord[3 * n_ord++], s, sizeof(int) * 3);
  }
  /* std::priority_queue<Index3> pops the largest (i, j, k) first (octomap.h:49-55) */
  qsort(ord, (size_t)n_ord, sizeof(int) * 3, idx3_desc);
  qcap = n_ord * 2 + 1024;
  q = (m3_cell*)malloc(sizeof(m3_cell) * (size_t)qcap);
  for (k = 0; k < n_ord; k++)
  {
    m3_cell cell = {ord[3 * k], ord[3 * k + 1], ord[3 * k + 2], ord[3 * k], ord[3 * k + 1], ord[3 * k + 2]};
    q[tail++] = cell;
  }
  while (head < tail)
  {
    const m3_cell cur = q[head];
    const int ok[6] = {cur.i > m->cmin[0], cur.j > m->cmin[1], cur.k > m->cmin[2],
                       cur.i < m->cmax[0], cur.j < m->cmax[1], cur.k < m->cmax[2]};
    int s;
    for (s = 0; s < 6; s++)
    {
      int i, j, kk, di, dj, dk;
      double new_distance, old_distance;
      if (!ok[s])
        continue;
      i = cur.i + SHIFTS[s][0];
      j = cur.j + SHIFTS[s][1];
      kk = cur.k + SHIFTS[s][2];
      di = abs(i - cur.si);
      dj = abs(j - cur.sj);
      dk = abs(kk - cur.sk);
      /* the reference indexes cached_distances_lut_[di][dj][dk] without a bound (sizes radius+2);
       * relaxation stops one cell beyond the radius, so the indices stay inside the table */
      new_distance = cd[(di * R2 + dj) * R2 + dk];
      old_distance = orc_map3d_distance(m, i, j, kk);
      if (old_distance - new_distance > m->max_distance_ratio)
      {
        m3_cell cell = {i, j, kk, cur.si, cur.sj, cur.sk};
        m3_set(m, i, j, kk, new_distance);
        if (tail == qcap)
        {
          if (head > 0)
          {
            memmove(q, q + head, sizeof(m3_cell) * (size_t)(tail - head));
            tail -= head;
            head = 0;
          }
          if (tail * 2 > qcap)
          {
            qcap *= 2;
            q = (m3_cell*)realloc(q, sizeof(m3_cell) * (size_t)qcap);
          }
        }
        q[tail++] = cell;
      }
    }
    head++;
  }
  free(q);
  free(ord);
  free(cd);
  m->lut_created = 1;
}
/* setMapBounds :128-150 */
void orc_map3d_set_bounds(orc_map3d* m, const double mn[2], const double mx[2])
{
  int a;
  for (a = 0; a < 2; a++)
  {
    int cmin = (int)floor((mn[a] - m->max_distance) / m->resolution + 0.5);
    int cmax = (int)floor((mx[a] + m->max_distance) / m->resolution + 0.5);
    if (cmin > m->cmin[a])
      m->cmin[a] = cmin;
    if (cmax < m->cmax[a])
      m->cmax[a] = cmax;
  }
  m3_recount(m);
  orc_map3d_update_distances(m);
}
void orc_map3d_get_bounds(const orc_map3d* m, int mn[3], int mx[3])
{
  memcpy(mn, m->cmin, sizeof(int) * 3);
  memcpy(mx, m->cmax, sizeof(int) * 3);
}
void orc_map3d_get_dense(const orc_map3d* m, uint8_t* out)
{
  const long w = m->width, h = m->cmax[1] - m->cmin[1] + 1, d = m->num_z;
  long i, j, k;
  for (k = 0; k < d; k++)
    for (j = 0; j < h; j++)
      for (i = 0; i < w; i++)
        out[(k * h + j) * w + i] = m->ratios[m->pose_indices[j * w + i] + k];
}
void orc_map3d_set_dense(orc_map3d* m, const uint8_t* in)
{
  const long w = m->width, h = m->cmax[1] - m->cmin[1] + 1, d = m->num_z;
  long i, j, k;
  free(m->pose_indices);
  m->pose_indices = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(w * h));
  m->ratios_cap = (size_t)(w * h * d + d);
  m->ratios = (uint8_t*)realloc(m->ratios, m->ratios_cap);
  memset(m->ratios, 255, (size_t)d);
  m->ratios_n = m->ratios_cap;
  for (j = 0; j < h; j++)
    for (i = 0; i < w; i++)
    {
      m->pose_indices[j * w + i] = (uint32_t)(d + (j * w + i) * d);
      for (k = 0; k < d; k++)
        m->ratios[d + (j * w + i) * d + k] = in[(k * h + j) * w + i];
    }
  m->lut_created = 1;
}
/* exact quantised EDT: floor(min(sqrt(k2)*res, max)/max*255) over the sources within the radius box */
void orc_map3d_exact_edt(const orc_map3d* m, uint8_t* out)
{
  const long w = m->width, h = m->cmax[1] - m->cmin[1] + 1, d = m->num_z;
  /* any source further than max_distance quantises to 255, so a (R+1) box is enough */
  const int R = (int)floor(m->max_distance / m->resolution) + 1;
  long* best = (long*)malloc(sizeof(long) * (size_t)(w * h * d));
  long n = w * h * d, t, s;
  for (t = 0; t < n; t++)
    best[t] = -1;
  for (s = 0; s < m->n_src; s++)
  {
    const int* p = &m->src[3 * s];
    int a, b, c;
    if (!m3_voxel_valid(m, p[0], p[1], p[2]))
      continue;
    for (c = -R; c <= R; c++)
      for (b = -R; b <= R; b++)
        for (a = -R; a <= R; a++)
        {
          const int i = p[0] + a, j = p[1] + b, k = p[2] + c;
          long idx, k2 = (long)a * a + (long)b * b + (long)c * c;
          if (!m3_voxel_valid(m, i, j, k))
            continue;
          idx = ((long)(k - m->cmin[2]) * h + (j - m->cmin[1])) * w + (i - m->cmin[0]);
          if (best[idx] < 0 || k2 < best[idx])
            best[idx] = k2;
        }
  }
  for (t = 0; t < n; t++)
  {
    double dd;
    if (best[t] < 0)
    {
      out[t] = 255;
      continue;
    }
    dd = sqrt((double)best[t]) * m->resolution;
    dd = dd < m->max_distance ? dd : m->max_distance;
    dd = dd / m->max_distance * 255;
    out[t] = (uint8_t)(int)floor(dd);
  }
  free(best);
}

/* ====================================================================== PointCloudScanner */
/* getMapCloud, point_cloud_scanner.cpp:231-248. Third-party arithmetic (tf2, tf2_sensor_msgs, Eigen;
 * not under /root/reference, versions unpinned) restated from their published sources:
 *   tf2::Quaternion::setRPY(0, 0, yaw)           -> q = (0, 0, sin(yaw/2), cos(yaw/2))
 *   tf2::Transform(q, origin)                    -> basis = Matrix3x3(q) (setRotation), origin
 *   t = footprint_to_map * scanner_to_footprint  -> basis product, origin = R1 * o2 + o1 (f64)
 *   tf2::toMsg                                   -> basis.getRotation(q) (Shepperd branches), origin
 *   tf2::doTransform(PointCloud2)                -> Eigen::Transform<float,3,Affine> T =
 *        Translation3f(float(t)) * Quaternionf(float(q)); per point  p' = T * p  in float:
 *        p'[r] = (R[r][0]*x + R[r][1]*y) + R[r][2]*z + t[r], R = Quaternionf::toRotationMatrix()
 * PARITY UNPINNED: no reference test covers this boundary. */
static void quat_to_basis_d(const double q[4], double R[9]) /* tf2::Matrix3x3::setRotation */
{
  double d = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  double s = 2.0 / d;
  double xs = q[0] * s, ys = q[1] * s, zs = q[2] * s;
  double wx = q[3] * xs, wy = q[3] * ys, wz = q[3] * zs;
  double xx = q[0] * xs, xy = q[0] * ys, xz = q[0] * zs;
  double yy = q[1] * ys, yz = q[1] * zs, zz = q[2] * zs;
  R[0] = 1.0 - (yy + zz); R[1] = xy - wz;         R[2] = xz + wy;
  R[3] = xy + wz;         R[4] = 1.0 - (xx + zz); R[5] = yz - wx;
  R[6] = xz - wy;         R[7] = yz + wx;         R[8] = 1.0 - (xx + yy);
}
static void basis_to_quat_d(const double R[9], double q[4]) /* tf2::Matrix3x3::getRotation */
{
  double trace = R[0] + R[4] + R[8];
  if (trace > 0.0)
  {
    double s = sqrt(trace + 1.0);
    q[3] = s * 0.5;
    s = 0.5 / s;
    q[0] = (R[7] - R[5]) * s;
    q[1] = (R[2] - R[6]) * s;
    q[2] = (R[3] - R[1]) * s;
  }
  else
  {
    int i = R[0] < R[4] ? (R[4] < R[8] ? 2 : 1) : (R[0] < R[8] ? 2 : 0);
    int j = (i + 1) % 3, k = (i + 2) % 3;
    double s = sqrt(R[i * 3 + i] - R[j * 3 + j] - R[k * 3 + k] + 1.0);
    q[i] = s * 0.5;
    s = 0.5 / s;
    q[3] = (R[k * 3 + j] - R[j * 3 + k]) * s;
    q[j] = (R[j * 3 + i] + R[i * 3 + j]) * s;
    q[k] = (R[k * 3 + i] + R[i * 3 + k]) * s;
  }
}
void orc_cloud_transform(const orc_cloud_model* P, const double pose[3], const float* in, long n, float* out)
{
  double q1[4], R1[9], R2[9], R[9], o[3], q[4];
  float qf[4], Rf[9], tf[3];
  double half = pose[2] * 0.5; /* setRPY with roll = pitch = 0 */
  long k;
  int r, c;
  q1[0] = 0.0 * cos(half) - 1.0 * 0.0 * sin(half); /* sin(0)=0, cos(0)=1 products of setRPY */
  q1[1] = 0.0;
  q1[2] = sin(half);
  q1[3] = cos(half);
  quat_to_basis_d(q1, R1);
  quat_to_basis_d(P->mount_rotation, R2);
  for (r = 0; r < 3; r++)
    for (c = 0; c < 3; c++)
      R[r * 3 + c] = R1[r * 3] * R2[c] + R1[r * 3 + 1] * R2[3 + c] + R1[r * 3 + 2] * R2[6 + c];
  {
    const double o1[3] = {pose[0], pose[1], 0.0};
    for (r = 0; r < 3; r++)
      o[r] = R1[r * 3] * P->mount_translation[0] + R1[r * 3 + 1] * P->mount_translation[1] +
             R1[r * 3 + 2] * P->mount_translation[2] + o1[r];
  }
  basis_to_quat_d(R, q);
  for (r = 0; r < 4; r++)
    qf[r] = (float)q[r];
  for (r = 0; r < 3; r++)
    tf[r] = (float)o[r];
  { /* Eigen::Quaternionf::toRotationMatrix */
    const float tx = 2.0f * qf[0], ty = 2.0f * qf[1], tz = 2.0f * qf[2];
    const float twx = tx * qf[3], twy = ty * qf[3], twz = tz * qf[3];
    const float txx = tx * qf[0], txy = ty * qf[0], txz = tz * qf[0];
    const float tyy = ty * qf[1], tyz = tz * qf[1], tzz = tz * qf[2];
    Rf[0] = 1.0f - (tyy + tzz); Rf[1] = txy - twz;          Rf[2] = txz + twy;
    Rf[3] = txy + twz;          Rf[4] = 1.0f - (txx + tzz); Rf[5] = tyz - twx;
    Rf[6] = txz - twy;          Rf[7] = tyz + twx;          Rf[8] = 1.0f - (txx + tyy);
  }
  for (k = 0; k < n; k++)
  {
    const float x = in[3 * k], y = in[3 * k + 1], z = in[3 * k + 2];
    for (r = 0; r < 3; r++)
      out[3 * k + r] = ((Rf[r * 3] * x + Rf[r * 3 + 1] * y) + Rf[r * 3 + 2] * z) + tf[r];
  }
}
/* applyGompertz :250-260 */
static double cloud_gompertz(const orc_cloud_model* p, double v)
{
  v = v * p->input_scale + p->input_shift;
  v = p->gompertz_a * exp(-1.0 * p->gompertz_b * exp(-1.0 * p->gompertz_c * v));
  v += p->output_shift;
  return v;
}
/* applyModelToSampleSet :111-129 -> calcPointCloudModel :132-167 / Gompertz :169-203 -> recalcWeight :205-229 */
double orc_cloud_apply(const orc_map3d* m, const orc_cloud_model* P, const float* points, long n_points,
                       double* s, long n)
{
  double total_weight = 0.0, z_hit_denom = 2 * P->sigma_hit * P->sigma_hit;
  double z_rand_mult = 1.0 / m->max_distance;
  float* mc;
  long j, t;
  if (P->max_beams < 2)
    return 0.0;
  mc = (float*)malloc(sizeof(float) * 3 * (size_t)(n_points > 0 ? n_points : 1));
  for (j = 0; j < n; j++)
  {
    double p = 1.0, sum_pz = 0;
    int count = 0;
    orc_cloud_transform(P, &s[4 * j], points, n_points, mc);
    for (t = 0; t < n_points; t++)
    {
      double w[3] = {mc[3 * t], mc[3 * t + 1], mc[3 * t + 2]}, z, pz;
      int v[3];
      orc_map3d_world_to_map(w, m->resolution, v);
      z = orc_map3d_distance(m, v[0], v[1], v[2]);
      pz = P->z_hit * exp(-(z * z) / z_hit_denom);
      if (P->model_type == 0)
      {
        pz += P->z_rand * z_rand_mult;
        p += pz * pz * pz;
      }
      else
      {
        pz += P->z_rand;
        sum_pz += pz;
        count++;
      }
    }
    if (P->model_type != 0)
    {
      p = sum_pz / count; /* no zero guard in the reference (:197) */
      p = cloud_gompertz(P, p);
    }
    s[4 * j + 3] *= p;
    total_weight += s[4 * j + 3];
  }
  free(mc);
  if (total_weight > 0.0)
  {
    double rv = 0.0;
    for (j = 0; j < n; j++)
    {
      const double w[3] = {s[4 * j], s[4 * j + 1], 0.0};
      int v[3];
      orc_map3d_world_to_map(w, m->resolution, v);
      if (!orc_map3d_is_pose_valid(m, v[0], v[1]))
        s[4 * j + 3] *= P->off_map_factor;
      rv += s[4 * j + 3];
    }
    total_weight = rv;
  }
  return total_weight;
}
