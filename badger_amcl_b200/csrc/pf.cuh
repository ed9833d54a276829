// ParticleFilter device state: two SoA sample sets + per-set kd-tree/cluster state + staged scan.
#pragma once
#include <vector>

#include "comm.cuh"
#include "common.cuh"
#include "map2d.cuh"

namespace bamcl
{
// One sample set (PFSampleSet, particle_filter.h:68-87) in SoA form.
struct SampleSet
{
  DevBuf x, y, th, w;  // f64[max_samples]
  int sample_count = 0;
  // kd-tree histogram of this set as built at its creation (resample or init), pf_kdtree.cpp
  int leaf_count = 0, node_count = 0, cluster_count = 0;
  DevBuf label;          // int32[max_samples]: cluster label of every sample
  DevBuf cluster_stats;  // f64[10 * cluster_count]
  size_t cluster_cap = 0;
};

// Scan staged for the scoring kernel: valid beams only, in cell units.
struct StagedScan
{
  bool valid = false;
  bamcl_planar_model model;
  int n_beams = 0;      // beams that survive the step / max-range / NaN filter
  int n_slots = 0;      // beam_ind slots (prob model with beam skipping)
  DevBuf beams;         // double2[n_beams]: (r cos b, r sin b) / resolution
  DevBuf table;         // f64[table_n]: per-k model value
  int table_n = 0;
  DevBuf table_q;       // int32[table_n]: round(table * 2^q_shift), -1 shift = unusable
  int q_shift = -1;
  double range_max = 0.0;
  uint64_t map_version = 0;
  // raw copies for the beam model
  DevBuf ranges, angles;
  int range_count = 0, step = 1;
  // beam skipping (planar_scanner.cpp:361-529): slot (beam_ind) of every staged beam, the staged beams on the host,
  // "cached distance < beam_skip_distance" per table rank, and device scratch for the agreement counts / the
  // beams that survive the mask
  std::vector<int> beam_slot;
  std::vector<double> h_beams;
  DevBuf agree;      // u8[table_n]
  DevBuf skip_count; // int32[n_beams]
  DevBuf beams_sel;  // double2[n_beams]
};
// Point cloud staged for the 3-D scoring kernel.
struct StagedCloud
{
  bool valid = false;
  bamcl_cloud_model model;
  int n_points = 0;   // staged (finite) points
  int n_dropped = 0;  // points with a non-finite coordinate, left out of the staged cloud
  DevBuf xyz, px, py, pz;  // packed upload + SoA split (f32, sensor frame)
  DevBuf table;            // f64[257]: per-quantum model value, [256] = outside the map
  uint64_t map_version = 0;
};
}  // namespace bamcl

namespace bamcl
{
// What Node::randomFreeSpacePose (node.cpp:823-840) draws a cell from and how the cell becomes a world position.
// 2-D (map_type 2): the list of Node2D::updateFreeSpaceIndices, OccupancyMap::convertMapToWorld (occupancy_map.cpp:75-88).
// 3-D (map_type 3): Node3D::updateFreeSpaceIndices (node_3d.cpp:306-318) lists every (i, j) of the OctoMap's cropped cell
// rectangle [min, max), i outer - never materialised here - and OctoMap::convertMapToWorld is i * resolution
// (octomap.cpp:83-95).
struct PoseSource
{
  const int2* free_ij = nullptr;  // null = the rectangle
  long free_n = 0;
  int size_x = 0, size_y = 0;
  double resolution = 0.0, origin_x = 0.0, origin_y = 0.0;
  int min_i = 0, min_j = 0, ny = 1;
  __device__ __forceinline__ void world(unsigned int k, double* x, double* y) const
  {
    if (free_ij)
    {
      const int2 c = free_ij[k];
      *x = origin_x + (c.x - size_x / 2) * resolution;
      *y = origin_y + (c.y - size_y / 2) * resolution;
    }
    else
    {
      *x = (min_i + (int)(k / (unsigned int)ny)) * resolution;
      *y = (min_j + (int)(k % (unsigned int)ny)) * resolution;
    }
  }
};
}  // namespace bamcl

struct bamcl_pf
{
  bamcl_ctx* ctx = nullptr;
  int min_samples = 0, max_samples = 0;
  bamcl::SampleSet sets[2];
  int current = 0;
  bamcl::StagedScan scan;
  bamcl::StagedCloud cloud;
  bamcl::DevBuf scalars;   // small device scratch for totals etc.
  bamcl::PinBuf h_scalars; // pinned mirror
  bamcl::PinBuf h_stage;   // pinned staging for uploads
  bamcl::DevBuf d_stage;   // device staging (AoS <-> SoA)
  bamcl::DevBuf partials;  // block partial sums
  // resample scratch
  bamcl::DevBuf cdf;       // f64[max_samples + 1]
  bamcl::DevBuf src;       // int32[max_samples]
  bamcl::DevBuf scratch;   // generic scratch, grown on demand
  bamcl::DevBuf scratch2;
  bamcl::DevBuf random_poses;
  bamcl::DevBuf seg_buf;   // systematic target segments
  // multinomial resampling with random-pose injection: stream positions of the outputs + jump tables (multinomial.cu)
  bamcl::DevBuf mn_buf;
  // random_pose_fn_ = Node::uniformPoseGenerator over this map's free-space list (bamcl_pf_set_random_pose_source), or
  // over the OctoMap's cropped cell rectangle (bamcl_pf_set_random_pose_source_3d)
  bamcl_map2d* rp_map = nullptr;
  bamcl_map3d* rp_map3d = nullptr;
  double rp_threshold = 0.0, rp_multiplier = 0.0;
  // kd-tree / clustering workspace (kdtree.cu), shared by both sets
  bamcl::DevBuf t_hkeys, t_hfirst, t_hnode;                 // hash table over pose bins
  bamcl::DevBuf t_slot, t_flag;                             // per sample
  bamcl::DevBuf t_nkey, t_cur, t_child, t_first_at, t_side; // per node (creation order)
  bamcl::DevBuf t_parent, t_cid, t_nfirst, t_root;
  // the cluster chain (union-find, labels, statistics) runs on this stream next to the leaf-count chain
  cudaStream_t aux_stream = nullptr;     // same priority as the main stream
  cudaStream_t aux_stream_lo = nullptr;  // lowest priority
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  // copies of host sample buffers overlap the kernels (bamcl_pf_update_planar_host / bamcl_pf_resample_host)
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_copy[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  // the three bucket size classes of the leaf count run side by side
  cudaStream_t bucket_stream[2] = {nullptr, nullptr};
  cudaEvent_t ev_bucket_fork = nullptr, ev_bucket_join[2] = {nullptr, nullptr};
  bamcl::DevBuf t_scan, t_out, t_sums, t_fast;
  int tree_fast = 0;        // the last tree build used the bucketed shared-memory path
  bamcl::PinBuf t_hout;
  size_t t_cluster_cap = 0;
  int coop_blocks = 0;
  cudaEvent_t score_ev0 = nullptr, score_ev1 = nullptr;  // around the last scoring kernel (bamcl_pf_last_score_ms)
  bool score_timed = false;
  // temp_obs_ bookkeeping of the beam-skipping model (planar_scanner.h:160-164): rows written per slot since the last
  // clearTempData (planar_scanner.cpp:684-690); the values themselves are recomputed, never stored
  std::vector<int> skip_written;
  // sharded beam skipping: agreement counts of ALL ranks' windows (bamcl_pf_set_skip_counts), used by the next scoring
  // call instead of this window's own; skip_counts_out = where a counts-only pass leaves this window's counts
  std::vector<int> skip_global_counts;
  long skip_global_n = 0;
  bool skip_global_set = false;
  std::vector<int>* skip_counts_out = nullptr;
  int skip_max_obs = 0, skip_max_samples = 0;
  bool force_general_tree = false;  // keep every node's first foreign key (multinomial.cu needs t_first_at)
  int set_converged = 0;     // set->converged as handed to the last scoring entry point
  bool exact_terms = false;  // bamcl_pf_set_exact_terms
  int tree_levels = 0;      // levels the last tree build needed (diagnostics)
  // Sharded filter (multi-GPU): every rank holds the whole set; scoring, totals, normalisation and the
  // CDF work on this rank's window [win_off, win_off + win_cnt) of the current set. win_cnt < 0 = whole set.
  long win_off = 0;
  int win_cnt = -1;
  // Distributed filter (bamcl_pf_shard_attach): every rank OWNS its window of the global set -- only those samples are
  // valid in this rank's arrays (global indexing is kept, so sample i lives at index i on its owner).
  bamcl_comm* comm = nullptr;
  int cdf_slow_cap = 32;          // slow CDF tiles a rank can ship in its summary; grown on demand
  double* cdf_bounds = nullptr;   // device, [n_ranks + 1]: exact CDF value at every rank boundary (last pf_build_cdf)
  // Dense pose-bin grid (kdtree.cu): bins of the box [box_min, box_min + box_dim) are addressed directly instead of
  // hashed. The box is sticky between resamples and re-fitted when a key falls outside it or it has become far too big.
  bool box_valid = false, box_dense = false;
  int box_min[3] = {0, 0, 0}, box_dim[3] = {0, 0, 0};
  long box_cells = 0;
  bamcl::DevBuf g_first, g_node, t_bits, t_bpre, t_setpart, t_ints;
  bamcl::DevBuf stage_x, stage_y;  // pose exchange of the distributed resample: blocks to send / blocks received
  int tree_dense = 0;  // the last tree build addressed the bins directly
  // scoring order: particles grouped by coarse (x, y) cell along a Morton curve, so the particles a GPU works on at any
  // moment look at the same part of the map (L2 reuse); weights are written back by original index
  bamcl::DevBuf order_perm, order_hist;
};

namespace bamcl
{
// device int flags of the resampler: bit0 target beyond the CDF, bit1 unsupported value
inline int* pf_flags(bamcl_pf* pf) { return reinterpret_cast<int*>(pf->scalars.as<double>() + 32); }
// this rank's window of the current set
inline long pf_win_off(const bamcl_pf* pf) { return pf->win_cnt < 0 ? 0 : pf->win_off; }
inline int pf_win_cnt(const bamcl_pf* pf) { return pf->win_cnt < 0 ? pf->sets[pf->current].sample_count : pf->win_cnt; }
inline int pf_ranks(const bamcl_pf* pf) { return pf->comm ? pf->comm->n_ranks : 1; }
inline int pf_rank(const bamcl_pf* pf) { return pf->comm ? pf->comm->rank : 0; }
// rank-major contiguous partition of n samples: rank r owns [shard_begin(n, G, r), shard_begin(n, G, r + 1))
inline long shard_begin(long n, int n_ranks, int r) { return n * (long)r / n_ranks; }
// the samples this rank holds: its window of a distributed set, else the whole current set
inline void pf_owned(const bamcl_pf* pf, long* off, long* cnt)
{
  const long n = pf->sets[pf->current].sample_count;
  const int G = pf_ranks(pf), r = pf_rank(pf);
  *off = G > 1 ? shard_begin(n, G, r) : 0;
  *cnt = G > 1 ? shard_begin(n, G, r + 1) - *off : n;
}
// a distributed filter works on the window it owns: set it from the current set's (global) sample count
inline void pf_apply_shard_window(bamcl_pf* pf)
{
  if (pf_ranks(pf) <= 1)
    return;
  const long n = pf->sets[pf->current].sample_count;
  pf->win_off = shard_begin(n, pf_ranks(pf), pf_rank(pf));
  pf->win_cnt = (int)(shard_begin(n, pf_ranks(pf), pf_rank(pf) + 1) - pf->win_off);
}
}

namespace bamcl
{
// phase timing of the distributed resample (BAMCL_SHARD_TIMING=1): synchronises at every mark, so only for diagnosis
void pf_mark(bamcl_pf* pf, const char* phase);
void pf_mark_report(bamcl_pf* pf);
// perm_out (device, window-local indices) = this rank's window ordered by the Morton code of its coarse (x, y) cell;
// cell = edge of a coarse cell in metres, (x0, y0) = lower corner of the area. Order inside a cell is arbitrary.
// the registered pose source (rp_map's list or rp_map3d's rectangle); false = none registered / empty (init.cu)
bool pf_pose_source(const bamcl_pf* pf, PoseSource* out);
int pf_spatial_order(bamcl_pf* pf, double x0, double y0, double cell, const int** perm_out);
int pf_reduce_total(bamcl_pf* pf, double* d_total);  // deterministic sum of the current set's weights -> device scalar
int pf_read_scalar(bamcl_pf* pf, int slot, double* out);  // device scalar slot -> host (synchronises)
int pf_normalise_by_device_total(bamcl_pf* pf);           // w /= scalars[0] (or 1/n), particle_filter.cpp:237-266
int pf_ensure_streams(bamcl_pf* pf);                      // side streams + events, created on first use
int pf_download_async(bamcl_pf* pf, double* aos, long n); // current set -> host AoS on the copy stream (after the main stream's work so far)
int pf_build_cdf(bamcl_pf* pf);  // exact sequential CDF of this rank's window (global values) + the rank bounds
}
