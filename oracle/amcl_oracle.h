/* oracle/amcl_oracle.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C CPU restatement of badger_amcl's measurement-and-resample hot path.
 * Every function cites the reference file:line it follows (paths relative to /root/reference).
 * Pinned against (i) the reference's own gtest KATs (test/test_badger_amcl.cpp) and (ii) the
 * unmodified reference compiled into oracle/_ref/libamcl_ref.so -- see tests/test_oracle_*.py.
 * Unpinned parts (third-party arithmetic absent from /root/reference): angles::normalize_angle
 * (last ulp of theta) and the tf2/Eigen-float point-cloud transform -- "parity unpinned" there.
 */
#ifndef AMCL_ORACLE_H
#define AMCL_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- drand48 (POSIX LCG; particle_filter.cpp:309,385,393 and pdf_gaussian.cpp:85,90 draw from it) */
typedef struct
{
  uint64_t x; /* 48-bit state */
} orc_rng;
void orc_srand48(orc_rng* r, long seed);
double orc_drand48(orc_rng* r);
double orc_gaussian_draw(orc_rng* r, double sigma); /* pdf_gaussian.cpp:77-97 */

/* ---- OccupancyMap (map/occupancy_map.cpp) */
typedef struct
{
  int size_x, size_y;
  double resolution;
  float origin_x, origin_y; /* pcl::PointXYZ members are float (map.h:48) */
  const int8_t* cells;      /* -1 free, 0 unknown, +1 occupied; index = i + j*size_x */
  const float* distances;   /* distance cache, same indexing; may be NULL before it is built */
  double max_distance;
} orc_map2d;

void orc_map2d_world_to_map(const orc_map2d* m, double x, double y, int* i, int* j); /* :90-98 */
void orc_map2d_map_to_world(const orc_map2d* m, int i, int j, double* x, double* y); /* :75-88 */
int orc_map2d_is_valid(const orc_map2d* m, int i, int j);                            /* :100-105 */
float orc_map2d_distance(const orc_map2d* m, int i, int j);                          /* :64-73 */
/* Bresenham ray cast (:257-364). steps_out (may be NULL) receives the number of cells tested. */
double orc_map2d_calc_range(const orc_map2d* m, double ox, double oy, double oa, double max_range, long* steps_out);
/* priority-queue brushfire (:138-242), including libstdc++'s heap tie order. out has size_x*size_y floats. */
void orc_map2d_update_distances(const orc_map2d* m, double max_distance, float* out);
/* exact capped Euclidean distance transform with the reference's value formula. */
void orc_map2d_exact_edt(const orc_map2d* m, double max_distance, float* out);

/* ---- PlanarScanner (sensors/planar_scanner.cpp) */
enum
{
  ORC_PLANAR_BEAM = 0,
  ORC_PLANAR_LF = 1,
  ORC_PLANAR_LF_PROB = 2,
  ORC_PLANAR_LF_GOMPERTZ = 3
};
typedef struct
{
  int model_type;
  int max_beams;
  double z_hit, z_short, z_max, z_rand, sigma_hit, lambda_short;
  double gompertz_a, gompertz_b, gompertz_c, input_shift, input_scale, output_shift;
  int do_beamskip;
  double beam_skip_distance, beam_skip_threshold, beam_skip_error_threshold;
  double off_map_factor, non_free_space_factor, non_free_space_radius;
  double scanner_pose[3];
} orc_planar_model;

/* beam-skip scratch that persists across calls (planar_scanner.h:134-136) */
typedef struct orc_planar_state orc_planar_state;
orc_planar_state* orc_planar_state_create(void);
void orc_planar_state_destroy(orc_planar_state* s);

double orc_normalize_angle(double a);                                           /* third-party angles, unpinned */
void orc_coord_add(const double a[3], const double b[3], double c[3]);          /* :693-701 */
double orc_apply_gompertz(const orc_planar_model* p, double v);                 /* :540-550 */
/* applyModelToSampleSet (:141-164): samples are AoS {x, y, theta, weight}; returns the total. */
double orc_planar_apply(const orc_map2d* m, const orc_planar_model* p, orc_planar_state* st, const double* ranges,
                        const double* angles, int range_count, double range_max, double* samples, long n,
                        int converged);
double orc_planar_recalc_weight(const orc_map2d* m, const orc_planar_model* p, double* samples, long n); /* :642-682 */

/* ---- PFKDTree (pf/pf_kdtree.cpp) */
typedef struct orc_kdtree orc_kdtree;
orc_kdtree* orc_kdtree_create(void);
void orc_kdtree_destroy(orc_kdtree* t);
void orc_kdtree_clear(orc_kdtree* t);                                      /* :42-47 */
void orc_kdtree_insert(orc_kdtree* t, const double pose[3], double value); /* :49-56, 97-150 */
void orc_kdtree_cluster(orc_kdtree* t);                                    /* :58-76, 169-194 */
int orc_kdtree_get_cluster(const orc_kdtree* t, const double pose[3]);     /* :79-90 */
int orc_kdtree_leaf_count(const orc_kdtree* t);                            /* :92-95 */
long orc_kdtree_node_count(const orc_kdtree* t);
void orc_kdtree_get_nodes(const orc_kdtree* t, int* keys, int* clusters, int* pivots, double* values);
void orc_pose_key(const double pose[3], int key[3]);                       /* :51-54 */

/* ---- ParticleFilter (pf/particle_filter.cpp) */
typedef struct orc_pf orc_pf;
orc_pf* orc_pf_create(int min_samples, int max_samples, double alpha_slow, double alpha_fast, double conv_thresh); /* :38-98 */
void orc_pf_destroy(orc_pf* pf);
void orc_pf_set_resample_model(orc_pf* pf, int systematic);                  /* :100-103 */
void orc_pf_set_population_params(orc_pf* pf, double pop_err, double pop_z); /* :651-655 */
void orc_pf_set_decay_rates(orc_pf* pf, double a_slow, double a_fast);       /* :657-661 */
void orc_pf_set_random_poses(orc_pf* pf, const double* poses, long n);       /* queue behind random_pose_fn_ */
long orc_pf_random_poses_used(const orc_pf* pf);
/* random_pose_fn_ = Node::uniformPoseGenerator (node.cpp:847-868) over free_ij, drawing from the rng handed to
 * orc_pf_update_resample -- what the node wires in (node.cpp:262-264); the caller keeps m and free_ij alive */
void orc_pf_set_uniform_pose_source(orc_pf* pf, const orc_map2d* m, const int* free_ij, long free_n, double threshold,
                                    double multiplier);
void orc_pf_init_with_poses(orc_pf* pf, const double* poses, long n);        /* initWithPoseFn :136-162 */
int orc_pf_sample_count(const orc_pf* pf);
void orc_pf_get_samples(const orc_pf* pf, double* aos);
void orc_pf_set_samples(orc_pf* pf, const double* aos, long n);
double* orc_pf_samples_ptr(orc_pf* pf); /* AoS view of the current set (x, y, theta, w) */
/* ParticleFilter::updateSensor (:223-267) after the sensor model produced `total` */
void orc_pf_normalise(orc_pf* pf, double total);
/* PlanarScanner::updateSensor -> ParticleFilter::updateSensor */
int orc_pf_update_sensor_planar(orc_pf* pf, const orc_map2d* m, const orc_planar_model* p, orc_planar_state* st,
                                const double* ranges, const double* angles, int range_count, double range_max);
void orc_pf_update_resample(orc_pf* pf, orc_rng* rng); /* :423-471 */
/* source index of every sample produced by the last resample (-1 = random pose); length sample_count */
void orc_pf_get_resample_sources(const orc_pf* pf, int* src);
int orc_pf_resample_limit(const orc_pf* pf, int k); /* :475-502 */
void orc_pf_get_averages(const orc_pf* pf, double* w_slow, double* w_fast);
void orc_pf_set_averages(orc_pf* pf, double w_slow, double w_fast);
int orc_pf_leaf_count(const orc_pf* pf);
long orc_pf_node_count(const orc_pf* pf);
int orc_pf_is_converged(const orc_pf* pf);
int orc_pf_cluster_count(const orc_pf* pf);
void orc_pf_get_sample_clusters(const orc_pf* pf, int* labels);
void orc_pf_get_cluster_stats(const orc_pf* pf, double* out10); /* count, weight, mean[3], cov00,01,10,11,22 */
void orc_pf_get_set_stats(const orc_pf* pf, double* out8);      /* mean[3], cov00,01,10,11,22 */
/* sequential CDF of the current set (:280-285); c has sample_count+1 entries */
void orc_pf_cdf(const orc_pf* pf, double* c);

/* ---- global localisation (node_2d.cpp:320-338, node.cpp:823-868). The node translation units are ROS nodes and cannot
 * be compiled here: PARITY UNPINNED for these three helpers (no reference test or golden vector covers them). */
/* Node2D::updateFreeSpaceIndices: free cells with cached distance > radius, i outer / j inner; ij_out = pairs, may be
 * NULL to count only */
long orc_map2d_free_space(const orc_map2d* m, double non_free_space_radius, int* ij_out);
/* Node::uniformPoseGenerator n times (randomFreeSpacePose + the de-weighting retry loop with Node2D::scorePose == 1.0,
 * which is what the reference's copy-of-the-sample bug makes it, node_2d.cpp:310-313) */
void orc_uniform_poses(const orc_map2d* m, const int* free_ij, long free_n, orc_rng* rng, double starting_weight_threshold,
                       double deweight_multiplier, long n, double* poses3);

/* ---- Odom::updateAction (sensors/odom.cpp:74-311): the motion update of the current set, drawing from rng.
 * model_type = OdomModelType (odom.h:33-40): 0 diff, 1 omni, 2 diff corrected, 3 omni corrected, 4 gaussian;
 * alpha[5]; pose / delta / absolute_motion = OdomData (odom.h:43-52). */
void orc_pf_update_action(orc_pf* pf, orc_rng* rng, int model_type, const double* alpha, const double* pose,
                          const double* delta, const double* absolute_motion);

/* ---- OctoMap (map/octomap.cpp) */
typedef struct orc_map3d orc_map3d;
/* occupied leaves as integer keys (centre = (key + 0.5) * res, octomap's keyToCoord) */
orc_map3d* orc_map3d_create(double resolution, const int* keys, long n_keys, double max_distance); /* :53-74 */
void orc_map3d_destroy(orc_map3d* m);
void orc_map3d_set_bounds(orc_map3d* m, const double mn[2], const double mx[2]); /* :128-150 (runs the update) */
void orc_map3d_update_distances(orc_map3d* m);                                   /* :174-333 FIFO brushfire */
void orc_map3d_get_bounds(const orc_map3d* m, int mn[3], int mx[3]);
double orc_map3d_distance(const orc_map3d* m, int i, int j, int k);              /* :336-350 */
void orc_map3d_get_dense(const orc_map3d* m, uint8_t* out);                      /* x fastest, then y, then z */
/* install a dense quantised field (x fastest, then y, then z) over the current bounds */
void orc_map3d_set_dense(orc_map3d* m, const uint8_t* in);
/* exact quantised EDT with the reference's value formula */
void orc_map3d_exact_edt(const orc_map3d* m, uint8_t* out);
void orc_map3d_world_to_map(const double xyz[3], double res, int ijk[3]);        /* :98-109 */
int orc_map3d_is_pose_valid(const orc_map3d* m, int i, int j);                   /* :112-116 */

/* ---- PointCloudScanner (sensors/point_cloud_scanner.cpp) */
typedef struct
{
  int model_type; /* 0 plain (:132-167), 1 gompertz (:169-203) */
  int max_beams;
  double z_hit, z_rand, sigma_hit;
  double gompertz_a, gompertz_b, gompertz_c, input_shift, input_scale, output_shift;
  double off_map_factor;
  double mount_translation[3]; /* scanner -> footprint transform (geometry_msgs::Transform) */
  double mount_rotation[4];    /* quaternion x, y, z, w */
} orc_cloud_model;
/* getMapCloud (:231-248) restated: f64 tf2 compose, f32 Eigen apply (third-party, parity unpinned).
 * points_in/out are packed float xyz triples. */
void orc_cloud_transform(const orc_cloud_model* p, const double pose[3], const float* points_in, long n,
                         float* points_out);
double orc_cloud_apply(const orc_map3d* m, const orc_cloud_model* p, const float* points, long n_points,
                       double* samples, long n); /* applyModelToSampleSet :111-129 */

#ifdef __cplusplus
}
#endif
#endif
