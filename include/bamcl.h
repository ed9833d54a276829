/* include/bamcl.h -- C ABI of libbamcl_cuda.so (sm_100a), the device side of the badger_amcl hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types. The host
 * C++ classes that mirror the reference API (badger_amcl_b200/host/amcl/...) and the Python test/bench
 * binding (badger_amcl_b200/cabi.py) call ONLY these entry points. Each entry cites the reference
 * interface it replaces (paths relative to /root/reference).
 *
 * Conventions
 *  - every function returns 0 on success, a negative BAMCL_E_* code otherwise; bamcl_last_error() gives text
 *  - host buffers are caller-owned; device buffers are owned by the handles
 *  - particles on the host side are the reference's AoS PFSample {double pose[3]; double weight} = 4 doubles
 *    (particle_filter.h:41-49); on the device they are SoA (x, y, theta, w as separate f64 arrays)
 *  - all work is enqueued on the context's stream; functions that return scalars synchronise that stream
 *  - there is no CPU fallback: without a CUDA device every entry point fails with BAMCL_E_CUDA
 */
#ifndef BAMCL_H
#define BAMCL_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BAMCL_OK 0
#define BAMCL_E_CUDA -1       /* CUDA runtime error (text in bamcl_last_error) */
#define BAMCL_E_ARG -2        /* invalid argument / precondition of the reference violated */
#define BAMCL_E_UNSUPPORTED -3 /* input outside what the device path represents (e.g. non-canonical cache) */
#define BAMCL_E_STATE -4      /* call order violated (e.g. scoring before a distance cache exists) */
#define BAMCL_E_COMM -5       /* NCCL error, NCCL missing, or a peer rank failed */

typedef struct bamcl_ctx bamcl_ctx;
typedef struct bamcl_map2d bamcl_map2d;
typedef struct bamcl_map3d bamcl_map3d;
typedef struct bamcl_pf bamcl_pf;
typedef struct bamcl_comm bamcl_comm;

const char* bamcl_last_error(void);
int bamcl_version(void);

/* ---- context: one per GPU (one process per GPU in multi-GPU runs) */
int bamcl_ctx_create(int device, bamcl_ctx** out);
int bamcl_ctx_destroy(bamcl_ctx* ctx);
int bamcl_ctx_sync(bamcl_ctx* ctx);
/* cudaStream_t of the context (for event timing / NCCL on the same stream) */
void* bamcl_ctx_stream(bamcl_ctx* ctx);
/* CUDA-event stopwatch on the context's stream */
int bamcl_timer_start(bamcl_ctx* ctx);
int bamcl_timer_stop_ms(bamcl_ctx* ctx, float* ms);
/* kernels launched by this library through this context since creation (bench.py "gpu_launches") */
long bamcl_ctx_launch_count(bamcl_ctx* ctx);
/* write a buffer larger than L2 (bench hygiene between timed iterations) */
int bamcl_ctx_flush_l2(bamcl_ctx* ctx);

/* measurement aid ("L2-resident random-gather peak"): n_gathers random 2-byte loads over a table of
 * table_bytes, 16 independent loads in flight per thread -- the likelihood-field scorer's access pattern without its
 * arithmetic. ms_out = device time for n_gathers (best of three warm passes). */
int bamcl_bench_gather(bamcl_ctx* ctx, size_t table_bytes, long n_gathers, float* ms_out);

/* page-locked host memory for the caller's sample / scan buffers (the host mirror of PFSampleSet::samples,
 * particle_filter.h:72, lives in such a buffer so uploads and downloads run at full PCIe rate) */
int bamcl_host_alloc(size_t bytes, void** out);
int bamcl_host_free(void* p);
/* page-lock memory the caller already owns (the host mirror's std::vector<PFSample>, particle_filter.h:72), so copies of it
 * run at full PCIe rate and overlap kernels; the range must stay allocated until bamcl_host_unregister */
int bamcl_host_register(void* p, size_t bytes);
int bamcl_host_unregister(void* p);

/* ---- OccupancyMap (include/amcl/map/occupancy_map.h:57-81) */
/* ctor + setSize + setOrigin (occupancy_map.cpp:30-57) */
int bamcl_map2d_create(bamcl_ctx* ctx, int size_x, int size_y, double resolution, float origin_x, float origin_y,
                       bamcl_map2d** out);
int bamcl_map2d_destroy(bamcl_map2d* m);
/* setCellState for every cell (occupancy_map.cpp:112-115); cells[i + j*size_x] in {-1 free, 0 unknown, +1 occupied} */
int bamcl_map2d_set_cells(bamcl_map2d* m, const int8_t* cells, long n);
/* updateDistancesLUT (occupancy_map.cpp:138-160): builds the distance cache on the device as the EXACT capped
 * Euclidean distance transform with the reference's value formula (the reference's heap-order
 * brushfire is >= this in a few % of cells and cannot be reproduced in parallel). */
int bamcl_map2d_build_distances(bamcl_map2d* m, double max_distance);
/* install a precomputed distance cache (e.g. the reference's own brushfire result); values must be of the
 * form (float)(sqrt(k)*resolution), k integer, or (float)max_distance -- what updateDistancesLUT can produce */
int bamcl_map2d_upload_distances(bamcl_map2d* m, const float* distances, long n, double max_distance);
/* distances_lut_ readback (getDistanceToObject, occupancy_map.cpp:64-73) */
int bamcl_map2d_download_distances(bamcl_map2d* m, float* out, long n);
/* calcRange (occupancy_map.cpp:257-364) for a batch of rays; steps_total_or_null gets the number of cells tested */
int bamcl_map2d_calc_range(bamcl_map2d* m, const double* ox, const double* oy, const double* oa, long n,
                           double max_range, double* ranges_out, long* steps_total_or_null);

/* ---- OctoMap (include/amcl/map/octomap.h:61-81) */
/* ctor + initFromOctree (octomap.cpp:30-74). The octomap::OcTree argument becomes the list of its occupied leaves
 * as integer keys (leaf centre = (key + 0.5) * resolution, octomap's keyToCoord), 3 ints per leaf. */
int bamcl_map3d_create(bamcl_ctx* ctx, double resolution, double max_distance, const int* keys, long n_keys,
                       bamcl_map3d** out);
int bamcl_map3d_destroy(bamcl_map3d* m);
/* getMinMaxCells (octomap.h:75): inclusive voxel bounds after cropping */
int bamcl_map3d_get_bounds(bamcl_map3d* m, int mn[3], int mx[3]);
/* setMapBounds (octomap.cpp:128-150): crop x/y to [mn - max_distance, mx + max_distance] and rebuild the cache */
int bamcl_map3d_set_bounds(bamcl_map3d* m, const double mn[2], const double mx[2]);
/* updateDistancesLUT (octomap.cpp:174-207) as the EXACT quantised distance transform
 * floor(min(d, max) / max * 255) (the reference's FIFO brushfire differs in a few voxels per million) */
int bamcl_map3d_build_distances(bamcl_map3d* m);
/* install / read the cache as a dense u8 field over the bounds, x fastest, then y, then z
 * (getDistanceToObject(i, j, k) = field * max_distance / 255, octomap.cpp:336-350) */
int bamcl_map3d_upload_distances(bamcl_map3d* m, const uint8_t* dense, long n);
int bamcl_map3d_download_distances(bamcl_map3d* m, uint8_t* out, long n);

/* ---- PointCloudScanner (include/amcl/sensors/point_cloud_scanner.h:59-78) */
enum
{
  BAMCL_CLOUD_PLAIN = 0,   /* setPointCloudModel :61 */
  BAMCL_CLOUD_GOMPERTZ = 1 /* setPointCloudModelGompertz :62 */
};
typedef struct
{
  int model_type;
  int max_beams; /* init :59 */
  double z_hit, z_rand, sigma_hit;
  double gompertz_a, gompertz_b, gompertz_c, input_shift, input_scale, output_shift;
  double off_map_factor;       /* setMapFactors :73 (the non-free-space factors are unused in 3-D) */
  double mount_translation[3]; /* setPointCloudScannerToFootprintTF :76 (geometry_msgs::Transform) */
  double mount_rotation[4];    /* quaternion x, y, z, w */
} bamcl_cloud_model;

/* ---- PlanarScanner (include/amcl/sensors/planar_scanner.h:62-96) */
enum
{
  BAMCL_PLANAR_BEAM = 0,       /* setModelBeam :64 */
  BAMCL_PLANAR_LF = 1,         /* setModelLikelihoodField :66 */
  BAMCL_PLANAR_LF_PROB = 2,    /* setModelLikelihoodFieldProb :69 */
  BAMCL_PLANAR_LF_GOMPERTZ = 3 /* setModelLikelihoodFieldGompertz :73 */
};
typedef struct
{
  int model_type;
  int max_beams; /* init :62 */
  double z_hit, z_short, z_max, z_rand, sigma_hit, lambda_short;
  double gompertz_a, gompertz_b, gompertz_c, input_shift, input_scale, output_shift;
  int do_beamskip;
  double beam_skip_distance, beam_skip_threshold, beam_skip_error_threshold;
  double off_map_factor, non_free_space_factor, non_free_space_radius; /* setMapFactors :82 */
  double scanner_pose[3];                                              /* setPlanarScannerPose :93 */
} bamcl_planar_model;

/* ---- ParticleFilter (include/amcl/pf/particle_filter.h:96-130) */
/* ctor (particle_filter.cpp:38-98): two sample sets of max_samples, uniform weights 1/max_samples */
int bamcl_pf_create(bamcl_ctx* ctx, int min_samples, int max_samples, bamcl_pf** out);
int bamcl_pf_destroy(bamcl_pf* pf);
/* current set <- host AoS (what initWithPoseFn/initWithGaussian/Odom::updateAction write through
 * getCurrentSet()); sets sample_count = n */
int bamcl_pf_upload_samples(bamcl_pf* pf, const double* aos, long n);
int bamcl_pf_download_samples(bamcl_pf* pf, double* aos, long n);
int bamcl_pf_sample_count(bamcl_pf* pf, int* n);
/* weights only (n doubles) */
int bamcl_pf_upload_weights(bamcl_pf* pf, const double* w, long n);
int bamcl_pf_download_weights(bamcl_pf* pf, double* w, long n);

/* PlanarScanner::applyModelToSampleSet (planar_scanner.cpp:141-164) on the current set: the four models
 * (:168,:236,:325,:552), coordAdd (:693) and recalcWeight (:642) in one launch. `converged` is
 * set->converged (beam skipping, :361). total_out receives the returned total (synchronises). */
int bamcl_pf_score_planar(bamcl_pf* pf, bamcl_map2d* map, const bamcl_planar_model* model, const double* ranges,
                          const double* angles, int range_count, double range_max, int converged, double* total_out);
/* ParticleFilter::updateSensor after the sensor model returned `total` (particle_filter.cpp:237-266):
 * w /= total, or 1/n each if total <= 0. */
int bamcl_pf_normalise(bamcl_pf* pf, double total);
/* fused PlanarScanner::updateSensor (planar_scanner.cpp:125-137): score + total + normalise with no host
 * round trip in between; total_out as above (w_avg of :240-248 is total / sample_count). */
int bamcl_pf_update_planar(bamcl_pf* pf, bamcl_map2d* map, const bamcl_planar_model* model, const double* ranges,
                           const double* angles, int range_count, double range_max, int converged,
                           double* total_out);
/* The same update with the particles still in HOST memory (the state after Odom::updateAction wrote the poses through
 * getCurrentSet(), odom.cpp:107-121): `aos` (n PFSample, page-locked for full PCIe rate) is uploaded in chunks and every
 * chunk is scored while the next one is on the bus; then total + normalise as above. Sets sample_count = n. */
int bamcl_pf_update_planar_host(bamcl_pf* pf, bamcl_map2d* map, const bamcl_planar_model* model, const double* ranges,
                                const double* angles, int range_count, double range_max, int converged, const double* aos,
                                long n, double* total_out);
/* Same, with the scan already staged on the device by bamcl_pf_stage_scan (bench "inputs resident in HBM"). */
int bamcl_pf_stage_scan(bamcl_pf* pf, bamcl_map2d* map, const bamcl_planar_model* model, const double* ranges,
                        const double* angles, int range_count, double range_max);
int bamcl_pf_update_planar_staged(bamcl_pf* pf, bamcl_map2d* map, int converged, double* total_out_or_null);
/* applyModelToSampleSet only (no normalisation) with the staged scan: the sharded filter normalises by the GLOBAL total */
int bamcl_pf_score_planar_staged(bamcl_pf* pf, bamcl_map2d* map, int converged, double* total_out);

/* PointCloudScanner::applyModelToSampleSet (point_cloud_scanner.cpp:111-129) on the current set: per-particle
 * cloud transform (getMapCloud :231-248), calcPointCloudModel / Gompertz (:132-203) and recalcWeight (:205-229)
 * in one launch. points_xyz = PointCloudData::points_ (point_cloud_scanner.h:50) as packed float x, y, z in the
 * sensor frame. */
int bamcl_pf_score_cloud(bamcl_pf* pf, bamcl_map3d* map, const bamcl_cloud_model* model, const float* points_xyz,
                         long n_points, double* total_out);
/* fused PointCloudScanner::updateSensor (:92-102): score + total + normalise */
int bamcl_pf_update_cloud(bamcl_pf* pf, bamcl_map3d* map, const bamcl_cloud_model* model, const float* points_xyz,
                          long n_points, double* total_out);
int bamcl_pf_stage_cloud(bamcl_pf* pf, bamcl_map3d* map, const bamcl_cloud_model* model, const float* points_xyz,
                         long n_points);
int bamcl_pf_update_cloud_staged(bamcl_pf* pf, bamcl_map3d* map, double* total_out_or_null);

/* Likelihood-field scoring sums one table term per beam. Default: terms in 32-bit fixed point summed as integers
 * (order-independent, |error| <= 2^-28 of the largest term per beam -- far inside the 1e-5 log-weight bound).
 * exact = 1: every term is the bit-exact f64 the reference computes (planar_scanner.cpp:303-315), summed in f64. */
int bamcl_pf_set_exact_terms(bamcl_pf* pf, int exact);

/* device duration of the most recent scoring kernel launch alone (CUDA events around that launch on the
 * context's stream); bench.py's roofline figure */
int bamcl_pf_last_score_ms(bamcl_pf* pf, float* ms);

/* sequential CDF c[0..n] of the current set's weights exactly as particle_filter.cpp:280-285 rounds it */
int bamcl_pf_cdf(bamcl_pf* pf, double* c_out, long n_plus_1);

typedef struct
{
  int resample_systematic;  /* setResampleModel :100 (1 = PF_RESAMPLE_SYSTEMATIC) */
  int min_samples, max_samples;
  double pop_err, pop_z;    /* setPopulationSizeParameters :121 */
  double w_diff;            /* max(0, 1 - w_fast/w_slow), particle_filter.cpp:438-440 */
  double dist_threshold;    /* 0.5, particle_filter.cpp:60 */
  double convergence_threshold; /* global_localization_convergence_threshold_ */
  /* drand48 contract: the host hands over the generator state; the device consumes draws
   * exactly like the reference (1 draw for systematic, :309; 2 per output for multinomial, :385,:393)
   * and returns the advanced state. */
  uint64_t drand48_state;
  /* poses returned by random_pose_fn_ (:320,:387), in call order; n_random_poses available */
  const double* random_poses;
  long n_random_poses;
} bamcl_resample_params;

typedef struct
{
  int sample_count;       /* set_b->sample_count */
  int leaf_count;         /* set_b->kdtree->getLeafCount() (pf_kdtree.cpp:92) */
  int node_count;         /* number of kd-tree nodes = occupied pose bins */
  int cluster_count;      /* set->cluster_count */
  int converged;          /* updateConverged (particle_filter.cpp:170-220) */
  int random_poses_used;
  int status_flags;       /* bit0: a systematic target fell at/after c[P] (reference would hang) */
  int reserved;
  uint64_t drand48_state; /* generator state after the draws */
  double mean[3];         /* set->mean */
  double cov[5];          /* set->cov (0,0) (0,1) (1,0) (1,1) (2,2) */
} bamcl_resample_result;

/* ParticleFilter::updateResample (particle_filter.cpp:423-471): KLD-adaptive systematic (:269) or
 * multinomial (:356) resampling into the other set, kd-tree histogram (pf_kdtree.cpp:49-150), clustering
 * (:58-76,:169-194), cluster/set statistics (particle_filter.cpp:505-636), set swap, updateConverged. */
int bamcl_pf_resample(bamcl_pf* pf, const bamcl_resample_params* params, bamcl_resample_result* result);
/* updateResample that also hands the new set to the host (what a caller reads next through getCurrentSet()->samples,
 * node.cpp:337-349): the device->host copy of the resampled AoS overlaps the pose-bin tree / clustering / statistics.
 * aos_out holds aos_capacity PFSample; result->sample_count of them are written. */
int bamcl_pf_resample_host(bamcl_pf* pf, const bamcl_resample_params* params, bamcl_resample_result* result, double* aos_out,
                           long aos_capacity);
/* resampleMultinomial (particle_filter.cpp:356-421) with w_diff > 0 and an ARBITRARY host random_pose_fn_: the reference
 * interleaves, per output, drand48() < w_diff ? random_pose_fn_() : drand48() (:385-393) - the pose function draws from the
 * same stream in between, so the device cannot replay the stream on its own. The HOST runs exactly that branch loop for
 * max_samples outputs (calling its pose function, recording libc's generator state before every output) and hands over
 * draws[k] = the second uniform of output k (:393), or -(1 + i) when output k is the i-th pose of params->random_poses.
 * The device does the rest in parallel: CDF, look-ups, candidate tree, the KLD stop (:416), clusters, statistics.
 * result->sample_count = m outputs were kept: the host restores the generator state it recorded before output m, so libc's
 * stream ends where the sequential loop would have left it; result->random_poses_used = how many of the m are random
 * poses; result->drand48_state = params->drand48_state (nothing is drawn on the device). */
int bamcl_pf_resample_multinomial_drawn(bamcl_pf* pf, const bamcl_resample_params* params, const double* draws,
                                        bamcl_resample_result* result);
/* kd-tree + cluster statistics of the CURRENT set as initWithPoseFn/initWithGaussian compute them
 * (particle_filter.cpp:106-162): fills leaf_count, node_count, cluster_count, mean, cov. */
int bamcl_pf_init_stats(bamcl_pf* pf, bamcl_resample_result* result);
/* source index in the previous set of every sample of the last resample (-1 = random pose) */
int bamcl_pf_resample_sources(bamcl_pf* pf, int* src, long n);
/* cluster label of every sample of the current set (PFKDTree::getCluster, pf_kdtree.cpp:79-90) */
int bamcl_pf_sample_clusters(bamcl_pf* pf, int* labels, long n);
/* per cluster 10 doubles: count, weight, mean[3], cov(0,0),(0,1),(1,0),(1,1),(2,2) (getClusterStats :639) */
int bamcl_pf_cluster_stats(bamcl_pf* pf, double* out, int max_clusters);
/* the systematic target sequence of particle_filter.cpp:309,337-341 (target = start; target += delta; wrap above
 * 1.0) in the closed per-binade form the device uses; host arithmetic, exposed for parity tests (no GPU needed) */
int bamcl_systematic_targets(double start, double delta, long n, double* out);
/* resampleLimit (particle_filter.cpp:475-502), host arithmetic exposed for parity tests */
int bamcl_resample_limit(int k, int min_samples, int max_samples, double pop_err, double pop_z);

/* ---- distributed filter (multi-GPU, one process per GPU) ---------------------------------
 * The reference resamples ONE global set (particle_filter.cpp:269-354). With G ranks the global order is rank-major:
 * rank r OWNS the contiguous window [n*r/G, n*(r+1)/G) of the set (only its owner holds a sample; map and scan are
 * replicated). The exchanges run inside this library, on the context's stream, with NCCL over NVLink / NVSwitch:
 *   update    window scored locally; window totals all-gathered and summed in rank order (= sample order of the
 *             reference's sequential sum, particle_filter.cpp:246-252); window normalised by the global total
 *   resample  every rank summarises its window's CDF tiles as functions of an unknown start value; ONE all-gather of the
 *             summaries lets every rank resolve the exact CDF value at every rank boundary (the exclusive scan over
 *             ranks of particle_filter.cpp:280-285) without a chain; systematic targets are planned per source rank;
 *             selected poses go to the owners of the output slots in one grouped send/receive (all-to-all-v);
 *             pose bins: per-rank first occurrences on a dense bin grid, all-reduced (min) -> identical creation order
 *             everywhere; KLD leaf count with the bucket subtrees split over the ranks (sum all-reduced); cluster sums
 *             and the convergence count all-reduced. */
#define BAMCL_UNIQUE_ID_BYTES 128
/* ncclGetUniqueId on ONE rank; ship the bytes to the other ranks over any host channel */
int bamcl_comm_unique_id(void* id_out);
/* ncclCommInitRank on every rank (collective); n_ranks = 1 needs no id and no NCCL */
int bamcl_comm_create(bamcl_ctx* ctx, const void* id, int rank, int n_ranks, bamcl_comm** out);
/* the same with the id moved through a file (a C++ host without any other channel: rank 0 writes it, the others wait for
 * it; `path` must be fresh for every job and visible to all ranks of the node) */
int bamcl_comm_create_file(bamcl_ctx* ctx, const char* path, int rank, int n_ranks, int timeout_s, bamcl_comm** out);
int bamcl_comm_destroy(bamcl_comm* comm);
int bamcl_comm_rank(bamcl_comm* comm, int* rank, int* n_ranks);
/* TEST back end: n_ranks ranks = n_ranks threads of THIS process on one GPU (each with its own context); collectives are
 * stream-ordered device copies. Lets a single-GPU box run the multi-rank path; never selected implicitly. */
int bamcl_comm_local_group(int n_ranks, void** group_out);
int bamcl_comm_local_group_free(void* group);
int bamcl_comm_create_local(bamcl_ctx* ctx, void* group, int rank, bamcl_comm** out);
/* small host vectors summed over the ranks (beam-skipping agreement counts, planar_scanner.cpp:448-451, 487) */
int bamcl_comm_all_reduce_sum_i32(bamcl_comm* comm, int* host_inout, int n);
/* op: 0 sum, 1 min, 2 max (also serves as a barrier between ranks) */
int bamcl_comm_all_reduce_i32(bamcl_comm* comm, int* host_inout, int n, int op);

/* the filter becomes rank `comm->rank`'s shard of a global set; NULL detaches. */
int bamcl_pf_shard_attach(bamcl_pf* pf, bamcl_comm* comm);
/* the current set holds n_global samples (this rank owns its window of them); returns the window */
int bamcl_pf_shard_set_count(bamcl_pf* pf, long n_global, long* window_offset, long* window_count);
int bamcl_pf_shard_window(bamcl_pf* pf, long* window_offset, long* window_count);
/* kd-tree + statistics of the global current set (bamcl_pf_init_stats, collective) */
int bamcl_pf_shard_init_stats(bamcl_pf* pf, bamcl_resample_result* result);
/* PlanarScanner::updateSensor / PointCloudScanner::updateSensor over the global set with the staged scan / cloud
 * (collective): total_out = the global total */
int bamcl_pf_shard_update_planar_staged(bamcl_pf* pf, bamcl_map2d* map, int converged, double* total_out);
int bamcl_pf_shard_update_cloud_staged(bamcl_pf* pf, bamcl_map3d* map, double* total_out);
/* ParticleFilter::updateResample (systematic) over the global set (collective). result is identical on every rank.
 * params->random_poses (host, all n_random poses on every rank) or the registered device pose source feed :316-324. */
int bamcl_pf_shard_resample(bamcl_pf* pf, const bamcl_resample_params* params, bamcl_resample_result* result);
/* exact CDF value at every rank boundary of the last resample, bounds[0 .. n_ranks] (parity tests) */
int bamcl_pf_shard_bounds(bamcl_pf* pf, double* bounds, int n_plus_1);

/* window of the current set the scoring / CDF entry points work on; count < 0 = whole set. (A distributed filter
 * manages its window itself; this is for callers that score a set in pieces.) */
int bamcl_pf_set_window(bamcl_pf* pf, long offset, long count);
/* host arithmetic: for the systematic target sequence (start, delta, n) of particle_filter.cpp:309,337-341 and the
 * ranks' CDF bounds[0..n_ranks], the runs (k_begin, k_end, rank) of output slots whose source lies in each rank's
 * window; targets at/after bounds[n_ranks] go to the last rank that holds weight. */
int bamcl_systematic_plan(double start, double delta, long n, const double* bounds, int n_ranks, long* runs, int max_runs,
                          int* n_runs);
/* host arithmetic: the transfers of the pose exchange for such a plan. Output slot o = n_random + k belongs to the rank
 * that owns it in the partition of n_random + n outputs. moves[5 * i ..] = (source rank, destination rank, k_begin,
 * k_end, position of k_begin in the source rank's selected rows), in the order both sides issue them. */
int bamcl_shard_exchange_plan(const long* runs, int n_runs, long n_random, long n, int n_ranks, long* moves, int max_moves,
                              int* n_moves);
/* Beam skipping on a distributed set (planar_scanner.cpp:398-451, 483-494; SURVEY 8e collective (2)): the per-beam
 * agreement counts are sums over ALL particles and the skip ratio is against the whole set. bamcl_pf_skip_counts runs
 * the counting pass over this rank's window for the staged scan (n_out = 0 when this update does no beam skipping: model
 * not likelihood_field_prob, do_beamskip off, or the set not converged); the caller sums the vectors over the ranks
 * (bamcl_comm_all_reduce_sum_i32) and hands them back with the global sample count; the next scoring call then masks
 * and integrates with them. */
int bamcl_pf_skip_counts(bamcl_pf* pf, bamcl_map2d* map, int converged, int* counts_out, int capacity, int* n_out);
int bamcl_pf_set_skip_counts(bamcl_pf* pf, const int* counts, int n, long global_sample_count);
/* host AoS <-> the samples [offset, offset + n) of the current set (a rank's window); sample_count is unchanged */
int bamcl_pf_upload_window_samples(bamcl_pf* pf, const double* aos, long offset, long n);
int bamcl_pf_download_window_samples(bamcl_pf* pf, double* aos, long offset, long n);
/* getLeafCount of the current set's tree (pf_kdtree.cpp:92), as of its last resample / init */
int bamcl_pf_leaf_count(bamcl_pf* pf, int* leaf_count);

/* ---- Odom (include/amcl/sensors/odom.h:33-86): the motion update, the step before scoring */
#define BAMCL_ODOM_DIFF 0 /* OdomModelType, odom.h:33-40 */
#define BAMCL_ODOM_OMNI 1
#define BAMCL_ODOM_DIFF_CORRECTED 2
#define BAMCL_ODOM_OMNI_CORRECTED 3
#define BAMCL_ODOM_GAUSSIAN 4
typedef struct bamcl_odom_model /* Odom::setModel, odom.cpp:63-72 */
{
  int model_type;
  double alpha[5];
} bamcl_odom_model;
typedef struct bamcl_odom_data /* OdomData, odom.h:43-52 */
{
  double pose[3];            /* odometric pose */
  double delta[3];           /* change in odometric pose */
  double absolute_motion[3]; /* accumulated |motion| (gaussian model) */
} bamcl_odom_data;
/* Odom::updateAction (odom.cpp:74-311) on the device-resident current set: every sample is moved by the odometry
 * delta plus three PDFGaussian::draw values (pdf_gaussian.cpp:77-97). The draws are the ones the reference would take
 * from the process's drand48 stream in state `drand48_state` (48 bits): the polar rejection loop is evaluated for all
 * attempts at once (generator skip-ahead + a prefix sum over the acceptances), so sample i gets accepted attempts
 * 3i..3i+2 exactly as in the sequential loop. *drand48_state_out = the state after the last draw consumed (hand it to
 * seed48 to keep the host stream in step). A draw of exactly 0.0 (probability 2^-48 each) would shift the pairing and is
 * reported as BAMCL_E_UNSUPPORTED. On a distributed filter every rank replays the draw stream of the whole set (the
 * number of draws a sample consumes depends on every sample before it) and moves the window it owns. Synchronises. */
int bamcl_pf_update_action(bamcl_pf* pf, const bamcl_odom_model* model, const bamcl_odom_data* data,
                           uint64_t drand48_state, uint64_t* drand48_state_out);

/* ---- scan / map ingestion */
/* Node2D::convertMap (node_2d.cpp:265-295): nav_msgs/OccupancyGrid (data row-major, width x height, 0 free / 100
 * occupied / anything else unknown) -> a map of (width x scale_up) x (height x scale_up) cells at resolution / scale_up
 * with the origin moved to the map centre (:275-277, narrowed to float as pcl::PointXYZ does). `resolution` is the
 * message's float32 field and the division by scale_up is done in float (:267: float / int), as the compiled reference
 * does: a 0.05 m grid becomes a map of resolution 0.05000000074505806. The message bytes are uploaded once and
 * expanded on the device. */
int bamcl_map2d_from_occupancy_grid(bamcl_ctx* ctx, const int8_t* data, int width, int height, float resolution,
                                    double origin_x, double origin_y, int scale_up, bamcl_map2d** out);
/* cell states back to the host (-1 free, 0 unknown, +1 occupied; index = i + j*size_x), and the map's geometry */
int bamcl_map2d_download_cells(bamcl_map2d* m, int8_t* cells, long n);
int bamcl_map2d_geometry(bamcl_map2d* m, int* size_xy, double* resolution, float* origin_xy);
/* Node2D::updateLatestScanData (node_2d.cpp:534-560): sensor_msgs/LaserScan ranges (float) -> PlanarData ranges_ /
 * angles_ / range_max_ (double): min/max thresholds applied in float (:538-546), readings <= range_min mapped to
 * range_max (:551-556), bearing = angle_min + i * angle_increment (:558). Host arithmetic, no GPU work. */
int bamcl_scan_from_laserscan(const float* msg_ranges, int range_count, float msg_range_min, float msg_range_max,
                              double sensor_min_range, double sensor_max_range, double angle_min, double angle_increment,
                              double* ranges_out, double* angles_out, double* range_max_out);
/* Node3D::updateLatestScanData (node_3d.cpp:467-480): every step-th point, step = max((count-1)/(max_beams-1), 1);
 * xyz / xyz_out are packed float triples. Host arithmetic, no GPU work. */
int bamcl_cloud_subsample(const float* xyz, long data_count, int max_beams, float* xyz_out, long capacity, long* n_out);

/* ---- global localisation */
/* Node2D::updateFreeSpaceIndices (node_2d.cpp:320-338): the (i, j) of every CELL_FREE cell whose cached distance exceeds
 * non_free_space_radius, in the reference's order (i outer, j inner), kept on the device; count_out =
 * free_space_indices_.size(). Needs the cells and a distance cache. */
int bamcl_map2d_update_free_space(bamcl_map2d* m, double non_free_space_radius, long* count_out);
/* the list as n (i, j) int pairs (parity tests) */
int bamcl_map2d_download_free_space(bamcl_map2d* m, int* ij_out, long n);
/* ParticleFilter::initWithPoseFn (particle_filter.cpp:136-162) with Node::uniformPoseGenerator (node.cpp:847-868) as
 * the pose function: max_samples poses drawn from the free-space list (randomFreeSpacePose, node.cpp:823-840: cell
 * index = drand48() * size, heading = drand48() * 2 pi - pi), weights 1/max_samples, pose-bin tree, clusters and
 * statistics; w_slow/w_fast reset is the caller's (:158). The de-weighting retry loop compares against
 * Node2D::scorePose, which always returns 1.0 in the reference (it scores a copy, node_2d.cpp:310-313): with
 * starting_weight_threshold > 1 every pose throws away a fixed number of generations, reproduced here.
 * drand48_state in / out as for bamcl_pf_update_action. */
int bamcl_pf_init_uniform(bamcl_pf* pf, bamcl_map2d* map, double starting_weight_threshold, double deweight_multiplier,
                          uint64_t drand48_state, uint64_t* drand48_state_out, bamcl_resample_result* result);

/* The same with map_type 3 (Node3D): Node3D::updateFreeSpaceIndices (node_3d.cpp:306-318) lists every (i, j) of the
 * OctoMap's cropped cell rectangle [min, max) (OctoMap::getMinMaxCells, i outer), and OctoMap::convertMapToWorld
 * (octomap.cpp:83-95) places cell (i, j) at (i * resolution, j * resolution). The rectangle is never materialised. */
int bamcl_pf_init_uniform_3d(bamcl_pf* pf, bamcl_map3d* map, double starting_weight_threshold, double deweight_multiplier,
                             uint64_t drand48_state, uint64_t* drand48_state_out, bamcl_resample_result* result);

/* ParticleFilter::initWithGaussian (particle_filter.cpp:105-133) on the device: max_samples poses, each PDFGaussian::sample()
 * (pdf_gaussian.cpp:52-71) = mean + cr * (draw(cd[0]), draw(cd[1]), draw(cd[2])), the draws taken from the drand48 stream in
 * the reference's order (all rejection-loop attempts evaluated at once, as in bamcl_pf_update_action), weights
 * 1/max_samples, pose-bin tree, clusters and statistics; w_slow/w_fast reset is the caller's (:126). The caller passes the
 * DECOMPOSED covariance PDFGaussian's constructor computes (:33-43): cr9 = the rotation (row-major, eigenvectors in
 * columns), cd3 = the standard deviations along them - for the diagonal covariances the node passes (node.cpp:281-287,
 * 688-701) that is the identity and sqrt of the diagonal; Eigen::EigenSolver's eigenvector order / sign for other matrices
 * is third-party and stays with the caller. drand48_state in / out as for bamcl_pf_update_action. */
int bamcl_pf_init_gaussian(bamcl_pf* pf, const double* mean3, const double* cr9, const double* cd3, uint64_t drand48_state,
                           uint64_t* drand48_state_out, bamcl_resample_result* result);

/* random_pose_fn_ of the filter (particle_filter.h:128; the node wires in Node::uniformPoseGenerator, node.cpp:262-264):
 * registers `map`'s free-space list (bamcl_map2d_update_free_space) and the generator's two parameters, so that
 * multinomial resampling with w_diff > 0 (particle_filter.cpp:385-388) draws its random poses on the device from the
 * same drand48 stream, in the reference's order; the systematic model (:309-324) does the same when
 * bamcl_resample_params::random_poses is NULL (start draw first, then the poses). NULL map = no source (bamcl_pf_resample then
 * reports that case as BAMCL_E_UNSUPPORTED; a caller with an opaque host pose function uses
 * bamcl_pf_resample_multinomial_drawn). The map must outlive the filter or be unregistered first. */
int bamcl_pf_set_random_pose_source(bamcl_pf* pf, bamcl_map2d* map_or_null, double starting_weight_threshold,
                                    double deweight_multiplier);
/* the 3-D node's pose function (the OctoMap's cropped cell rectangle); replaces a 2-D source and vice versa */
int bamcl_pf_set_random_pose_source_3d(bamcl_pf* pf, bamcl_map3d* map_or_null, double starting_weight_threshold,
                                       double deweight_multiplier);

/* ---- pose read-out */
/* Node2D::getMaxWeightPose (node_2d.cpp:588-617): weight and mean pose of the first cluster with the greatest weight,
 * from the cluster statistics of the last resample / init, reduced on the device (no download of the cluster table).
 * Where the reference would index hyps[-1] (no cluster with positive weight) this returns BAMCL_E_STATE. */
int bamcl_pf_max_weight_pose(bamcl_pf* pf, double* max_weight, double* pose3, int* cluster_index_or_null);
/* covariance packing of Node::updatePose (node.cpp:381-395): the set's covariance (result->cov) into the row-major
 * 6x6 of geometry_msgs/PoseWithCovariance: [0],[1],[6],[7] and [35] (COVARIANCE_AA, node.h:67); host arithmetic */
int bamcl_pose_covariance(const bamcl_resample_result* result, double* cov36);

/* ---- allocation-free variant (no counterpart in the reference)
 * For callers that cannot call cudaMalloc - code under CUDA graph capture, tight serving loops, harnesses that forbid it:
 * the weight total and normalisation of ParticleFilter::updateSensor (w /= total, or 1/n each where the total is not
 * positive) over caller-owned DEVICE buffers, with the block partials of the deterministic two-level sum kept in a
 * file-scope `static __device__` array instead of a cudaMalloc'ed buffer. Same tree shape as the filter's own total, so
 * total and weights are bit-identical to bamcl_pf_score_planar's total followed by bamcl_pf_normalise.
 *  - weights_dev: n doubles (device); weights_out_dev: n doubles (device, may be weights_dev itself); total_out_dev: one
 *    double (device). Work is queued on the context's stream and NOT synchronised (bamcl_ctx_sync to wait).
 *  - n > BAMCL_NOALLOC_MAX_SAMPLES is refused with BAMCL_E_ARG. The scratch is 64 KB per GPU, held from library load to
 *    process exit on every GPU that loads the library.
 *  - calls no cudaMalloc / cudaFree / allocating routine. The first call on a device looks the scratch address up
 *    (cudaGetSymbolAddress); make that call outside stream capture.
 *  - every call on a device shares the one scratch array, which keeps the last call's contents (nothing in it needs to
 *    start at zero: each call writes what it reads). Two calls that overlap on the GPU - different contexts / streams on
 *    the same device - RACE on it: serialise them. */
#define BAMCL_NOALLOC_MAX_SAMPLES (1L << 24)
int bamcl_weights_normalise_noalloc(bamcl_ctx* ctx, const double* weights_dev, long n, double* weights_out_dev,
                                    double* total_out_dev);

#ifdef __cplusplus
}
#endif
#endif
