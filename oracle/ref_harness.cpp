// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// A flat extern "C" surface over the UNMODIFIED reference classes, so that tests/ and bench.py's
// cpu_baseline / --impl reference leg can drive the reference's own CPU implementation of the hot
// path through ctypes. The reference translation units are compiled where they lie under
// /root/reference by oracle/Makefile into oracle/_ref/libamcl_ref.so; nothing is copied.
//
// Reference entry points wrapped (paths relative to /root/reference):
//   include/amcl/map/occupancy_map.h:57-81      OccupancyMap
//   include/amcl/map/octomap.h:61-81            OctoMap
//   include/amcl/pf/pf_kdtree.h:46-51           PFKDTree
//   include/amcl/pf/particle_filter.h:96-130    ParticleFilter
//   include/amcl/sensors/planar_scanner.h:62-96 PlanarScanner
//   include/amcl/sensors/point_cloud_scanner.h:59-78 PointCloudScanner (its tf2 / pcl dependencies are the header-only
//                                               look-alikes under oracle/shim/)
//   include/amcl/node/node.h:78-262, node_2d.h:50-140, node_3d.h:57-150  Node / Node2D / Node3D (roscpp, tf2_ros,
//                                               dynamic_reconfigure, message_filters, yaml-cpp are look-alikes under
//                                               oracle/shim/ too: parameters and tf from tables, topics in process)
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <limits>
#include <memory>
#include <queue>
#include <string>
#include <vector>

// everything the reference headers pull in, once, BEFORE access specifiers are redefined below
#include <array>
#include <complex>
#include <fstream>
#include <map>
#include <mutex>
#include <sstream>

#include <badger_amcl/AMCLConfig.h>
#include <dynamic_reconfigure/server.h>
#include <message_filters/subscriber.h>
#include <octomap_msgs/conversions.h>
#include <pcl_conversions/pcl_conversions.h>
#include <ros/shim_msgs.h>
#include <tf2/utils.h>
#include <tf2_geometry_msgs/tf2_geometry_msgs.h>
#include <tf2_ros/message_filter.h>
#include <tf2_ros/transform_broadcaster.h>
#include <tf2_ros/transform_listener.h>
#include <yaml-cpp/yaml.h>

// The harness needs w_slow_/w_fast_, the inactive sample set and a few map internals that the
// reference keeps private. Access specifiers do not change the object layout with GCC.
#define private public
#define protected public
#include "map/occupancy_map.h"
#include "map/octomap.h"
#include "pf/particle_filter.h"
#include "pf/pf_kdtree.h"
#include "sensors/odom.h"
#include "sensors/planar_scanner.h"
#include "sensors/point_cloud_scanner.h"
#include "node/node.h"
#include "node/node_2d.h"
#include "node/node_3d.h"
#undef private
#undef protected

using namespace badger_amcl;

#include "pf/pdf_gaussian.h"

namespace
{
struct Map2D
{
  std::shared_ptr<OccupancyMap> m;
};
struct Map3D
{
  std::shared_ptr<OctoMap> m;
};
struct Scanner2D
{
  std::shared_ptr<PlanarScanner> s;
};
struct Scanner3D
{
  std::shared_ptr<PointCloudScanner> s;
};
struct PF
{
  std::shared_ptr<ParticleFilter> pf;
  // queue of poses handed out by random_pose_fn_ (x, y, theta triples), consumed in order
  std::vector<double> random_poses;
  size_t random_next = 0;
  // alternative source: Node::uniformPoseGenerator restated (node.cpp:823-868; kept beside the compiled node sources
  // here) over a free-space list, drawing from the process-global drand48 stream like the node's pose function does
  std::shared_ptr<OccupancyMap> uniform_map;
  std::vector<int> free_ij;
  double uniform_threshold = 0.0, uniform_multiplier = 0.0;
};
}  // namespace

extern "C" {

// ---------------------------------------------------------------- RNG (process-global drand48)
void ref_srand48(long seed) { srand48(seed); }
double ref_drand48() { return drand48(); }
void ref_seed48(const unsigned short* s)
{
  unsigned short t[3] = {s[0], s[1], s[2]};
  seed48(t);
}
void ref_get_seed48(unsigned short* out)
{
  unsigned short zero[3] = {0, 0, 0};
  unsigned short* old = seed48(zero);
  out[0] = old[0];
  out[1] = old[1];
  out[2] = old[2];
  seed48(out);
}
// PDFGaussian::draw, the reference's own object code (pf/pdf_gaussian.cpp compiles against shim/Eigen/Eigenvalues)
double ref_gaussian_draw(double sigma) { return PDFGaussian::draw(sigma); }
// PDFGaussian(mean, cov).sample() n times -> out[3n]; the decomposition inside is the shim's (see shim/Eigen/Eigenvalues)
void ref_gaussian_sample(const double* mean3, const double* cov9, long n, double* out)
{
  Eigen::Vector3d mean(mean3[0], mean3[1], mean3[2]);
  Eigen::Matrix3d cov;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      cov(i, j) = cov9[3 * i + j];
  PDFGaussian pdf(mean, cov);
  for (long k = 0; k < n; k++)
  {
    Eigen::Vector3d v = pdf.sample();
    out[3 * k] = v[0];
    out[3 * k + 1] = v[1];
    out[3 * k + 2] = v[2];
  }
}

// ---------------------------------------------------------------- OccupancyMap
void* ref_map2d_create(double resolution, int size_x, int size_y, float ox, float oy)
{
  Map2D* h = new Map2D();
  h->m = std::make_shared<OccupancyMap>(resolution);
  h->m->setSize({size_x, size_y});
  h->m->setOrigin(pcl::PointXYZ(ox, oy, 0.0f));
  return h;
}
void ref_map2d_destroy(void* h) { delete static_cast<Map2D*>(h); }
// cells[index] in {-1 free, 0 unknown, +1 occupied}, index = i + j * size_x
void ref_map2d_set_cells(void* h, const int8_t* cells, long n)
{
  OccupancyMap& m = *static_cast<Map2D*>(h)->m;
  for (long k = 0; k < n; k++)
    m.setCellState(static_cast<int>(k), static_cast<MapCellState>(cells[k]));
}
int ref_map2d_get_cell(void* h, int i, int j) { return static_cast<Map2D*>(h)->m->getCellState(i, j); }
void ref_map2d_update_distances(void* h, double max_dist) { static_cast<Map2D*>(h)->m->updateDistancesLUT(max_dist); }
void ref_map2d_get_distances(void* h, float* out, long n)
{
  OccupancyMap& m = *static_cast<Map2D*>(h)->m;
  std::memcpy(out, m.distances_lut_.data(), sizeof(float) * n);
}
// install a precomputed cache (what the reference would have built) without running the brushfire
void ref_map2d_set_distances(void* h, const float* in, long n, double max_dist)
{
  OccupancyMap& m = *static_cast<Map2D*>(h)->m;
  m.max_distance_to_object_ = max_dist;
  m.distances_lut_.assign(in, in + n);
  m.distances_lut_created_ = true;
}
float ref_map2d_distance(void* h, int i, int j) { return static_cast<Map2D*>(h)->m->getDistanceToObject(i, j); }
double ref_map2d_calc_range(void* h, double ox, double oy, double oa, double max_range)
{
  return static_cast<Map2D*>(h)->m->calcRange(ox, oy, oa, max_range);
}
void ref_map2d_calc_ranges(void* h, const double* ox, const double* oy, const double* oa, double max_range, long n,
                           double* out)
{
  OccupancyMap& m = *static_cast<Map2D*>(h)->m;
  for (long k = 0; k < n; k++)
    out[k] = m.calcRange(ox[k], oy[k], oa[k], max_range);
}
void ref_map2d_world_to_map(void* h, double x, double y, int* ij)
{
  std::vector<int> mc(2);
  static_cast<Map2D*>(h)->m->convertWorldToMap({x, y}, &mc);
  ij[0] = mc[0];
  ij[1] = mc[1];
}
void ref_map2d_map_to_world(void* h, int i, int j, double* xy)
{
  std::vector<double> wc(2);
  static_cast<Map2D*>(h)->m->convertMapToWorld({i, j}, &wc);
  xy[0] = wc[0];
  xy[1] = wc[1];
}
int ref_map2d_is_valid(void* h, int i, int j) { return static_cast<Map2D*>(h)->m->isValid({i, j}) ? 1 : 0; }

// ---------------------------------------------------------------- OctoMap
// occupied voxels are given as integer leaf keys k (leaf centre = (k + 0.5) * res, see shim/octomap)
void* ref_map3d_create(double resolution, const int* keys, long n_keys, double max_dist)
{
  Map3D* h = new Map3D();
  h->m = std::make_shared<OctoMap>(resolution, false);
  std::shared_ptr<octomap::OcTree> tree = std::make_shared<octomap::OcTree>(resolution);
  for (long k = 0; k < n_keys; k++)
    tree->addLeaf(keys[3 * k], keys[3 * k + 1], keys[3 * k + 2], true);
  h->m->initFromOctree(tree, max_dist);
  return h;
}
void ref_map3d_destroy(void* h) { delete static_cast<Map3D*>(h); }
void ref_map3d_set_bounds(void* h, const double* mn, const double* mx)
{
  static_cast<Map3D*>(h)->m->setMapBounds({mn[0], mn[1]}, {mx[0], mx[1]});
}
void ref_map3d_update_distances(void* h) { static_cast<Map3D*>(h)->m->updateDistancesLUT(); }
void ref_map3d_get_bounds(void* h, int* mn, int* mx)
{
  std::vector<int> a, b;
  static_cast<Map3D*>(h)->m->getMinMaxCells(&a, &b);
  for (int k = 0; k < 3; k++)
  {
    mn[k] = a[k];
    mx[k] = b[k];
  }
}
double ref_map3d_distance(void* h, int i, int j, int k) { return static_cast<Map3D*>(h)->m->getDistanceToObject(i, j, k); }
// dense dump of the quantised field over the cropped bounds, x fastest then y then z
void ref_map3d_get_dense(void* h, uint8_t* out)
{
  OctoMap& m = *static_cast<Map3D*>(h)->m;
  const int* mn = m.cropped_min_cells_.data();
  const int* mx = m.cropped_max_cells_.data();
  long w = mx[0] - mn[0] + 1, hh = mx[1] - mn[1] + 1, d = mx[2] - mn[2] + 1;
  for (long k = 0; k < d; k++)
    for (long j = 0; j < hh; j++)
      for (long i = 0; i < w; i++)
      {
        uint32_t col = m.pose_indices_[j * w + i];
        out[(k * hh + j) * w + i] = m.distance_ratios_[col + k];
      }
}
void ref_map3d_world_to_map(void* h, const double* xyz, int* ijk)
{
  std::vector<int> mc(3);
  static_cast<Map3D*>(h)->m->convertWorldToMap({xyz[0], xyz[1], xyz[2]}, &mc);
  ijk[0] = mc[0];
  ijk[1] = mc[1];
  ijk[2] = mc[2];
}
void ref_map3d_map_to_world(void* h, const int* ijk, double* xyz)
{
  std::vector<double> wc(3);
  static_cast<Map3D*>(h)->m->convertMapToWorld({ijk[0], ijk[1], ijk[2]}, &wc);
  xyz[0] = wc[0];
  xyz[1] = wc[1];
  xyz[2] = wc[2];
}

// ---------------------------------------------------------------- PFKDTree
void* ref_kdtree_create() { return new PFKDTree(); }
void ref_kdtree_destroy(void* t) { delete static_cast<PFKDTree*>(t); }
void ref_kdtree_clear(void* t) { static_cast<PFKDTree*>(t)->clearKDTree(); }
void ref_kdtree_insert(void* t, double x, double y, double a, double value)
{
  static_cast<PFKDTree*>(t)->insertPose(Eigen::Vector3d(x, y, a), value);
}
void ref_kdtree_insert_many(void* t, const double* poses, long n, double value, int* leaf_counts_or_null)
{
  PFKDTree* kt = static_cast<PFKDTree*>(t);
  for (long k = 0; k < n; k++)
  {
    kt->insertPose(Eigen::Vector3d(poses[3 * k], poses[3 * k + 1], poses[3 * k + 2]), value);
    if (leaf_counts_or_null)
      leaf_counts_or_null[k] = kt->getLeafCount();
  }
}
void ref_kdtree_cluster(void* t) { static_cast<PFKDTree*>(t)->cluster(); }
int ref_kdtree_get_cluster(void* t, double x, double y, double a)
{
  return static_cast<PFKDTree*>(t)->getCluster(Eigen::Vector3d(x, y, a));
}
int ref_kdtree_leaf_count(void* t) { return static_cast<PFKDTree*>(t)->getLeafCount(); }
long ref_kdtree_node_count(void* t) { return static_cast<long>(static_cast<PFKDTree*>(t)->nodes_.size()); }
// nodes in creation order: key[3], cluster, pivot_dim, value
void ref_kdtree_get_nodes(void* t, int* keys, int* clusters, int* pivots, double* values)
{
  PFKDTree* kt = static_cast<PFKDTree*>(t);
  for (size_t k = 0; k < kt->nodes_.size(); k++)
  {
    const PFKDTreeNode& n = kt->nodes_[k];
    keys[3 * k] = n.key[0];
    keys[3 * k + 1] = n.key[1];
    keys[3 * k + 2] = n.key[2];
    clusters[k] = n.cluster;
    pivots[k] = n.pivot_dim;
    values[k] = n.value;
  }
}

// ---------------------------------------------------------------- PlanarScanner
void* ref_scanner2d_create(void* map2d, int max_beams)
{
  Scanner2D* h = new Scanner2D();
  h->s = std::make_shared<PlanarScanner>();
  h->s->init(max_beams, static_cast<Map2D*>(map2d)->m);
  h->s->setPlanarScannerPose(Eigen::Vector3d(0.0, 0.0, 0.0));
  // the reference leaves these uninitialised until a model setter runs
  h->s->do_beamskip_ = false;
  return h;
}
void ref_scanner2d_destroy(void* h) { delete static_cast<Scanner2D*>(h); }
void ref_scanner2d_set_pose(void* h, double x, double y, double a)
{
  static_cast<Scanner2D*>(h)->s->setPlanarScannerPose(Eigen::Vector3d(x, y, a));
}
void ref_scanner2d_set_map_factors(void* h, double off_map, double non_free, double radius)
{
  static_cast<Scanner2D*>(h)->s->setMapFactors(off_map, non_free, radius);
}
void ref_scanner2d_model_beam(void* h, double z_hit, double z_short, double z_max, double z_rand, double sigma,
                              double lambda)
{
  static_cast<Scanner2D*>(h)->s->setModelBeam(z_hit, z_short, z_max, z_rand, sigma, lambda);
}
// NOTE: the likelihood-field setters rebuild the distance cache (planar_scanner.cpp:74,91,112).
void ref_scanner2d_model_lf(void* h, double z_hit, double z_rand, double sigma, double max_dist)
{
  static_cast<Scanner2D*>(h)->s->setModelLikelihoodField(z_hit, z_rand, sigma, max_dist);
}
void ref_scanner2d_model_lf_prob(void* h, double z_hit, double z_rand, double sigma, double max_dist, int do_beamskip,
                                 double skip_dist, double skip_thresh, double skip_err)
{
  static_cast<Scanner2D*>(h)->s->setModelLikelihoodFieldProb(z_hit, z_rand, sigma, max_dist, do_beamskip != 0,
                                                              skip_dist, skip_thresh, skip_err);
}
void ref_scanner2d_model_lf_gompertz(void* h, double z_hit, double z_rand, double sigma, double max_dist, double a,
                                     double b, double c, double in_shift, double in_scale, double out_shift)
{
  static_cast<Scanner2D*>(h)->s->setModelLikelihoodFieldGompertz(z_hit, z_rand, sigma, max_dist, a, b, c, in_shift,
                                                                  in_scale, out_shift);
}
// Setters that do NOT re-run the brushfire: they write the same members the reference setters write
// (planar_scanner.cpp:67-113) and leave the (already built or installed) cache alone.
void ref_scanner2d_model_lf_nocache(void* h, double z_hit, double z_rand, double sigma)
{
  PlanarScanner& s = *static_cast<Scanner2D*>(h)->s;
  s.model_type_ = PLANAR_MODEL_LIKELIHOOD_FIELD;
  s.z_hit_ = z_hit;
  s.z_rand_ = z_rand;
  s.sigma_hit_ = sigma;
}
void ref_scanner2d_model_lf_prob_nocache(void* h, double z_hit, double z_rand, double sigma, int do_beamskip,
                                         double skip_dist, double skip_thresh, double skip_err)
{
  PlanarScanner& s = *static_cast<Scanner2D*>(h)->s;
  s.model_type_ = PLANAR_MODEL_LIKELIHOOD_FIELD_PROB;
  s.z_hit_ = z_hit;
  s.z_rand_ = z_rand;
  s.sigma_hit_ = sigma;
  s.do_beamskip_ = do_beamskip != 0;
  s.beam_skip_distance_ = skip_dist;
  s.beam_skip_threshold_ = skip_thresh;
  s.beam_skip_error_threshold_ = skip_err;
}
void ref_scanner2d_model_lf_gompertz_nocache(void* h, double z_hit, double z_rand, double sigma, double a, double b,
                                             double c, double in_shift, double in_scale, double out_shift)
{
  PlanarScanner& s = *static_cast<Scanner2D*>(h)->s;
  s.model_type_ = PLANAR_MODEL_LIKELIHOOD_FIELD_GOMPERTZ;
  s.z_hit_ = z_hit;
  s.z_rand_ = z_rand;
  s.sigma_hit_ = sigma;
  s.gompertz_a_ = a;
  s.gompertz_b_ = b;
  s.gompertz_c_ = c;
  s.input_shift_ = in_shift;
  s.input_scale_ = in_scale;
  s.output_shift_ = out_shift;
}

static std::shared_ptr<PlanarData> make_scan(const double* ranges, const double* angles, int n, double range_max)
{
  std::shared_ptr<PlanarData> d = std::make_shared<PlanarData>();
  d->range_count_ = n;
  d->range_max_ = range_max;
  d->ranges_.assign(ranges, ranges + n);
  d->angles_.assign(angles, angles + n);
  return d;
}

// Apply the sensor model to a free-standing sample set (AoS x,y,theta,w), no normalisation:
// PlanarScanner::applyModelToSampleSet (planar_scanner.cpp:141-164). Returns the model's total.
double ref_scanner2d_apply(void* h, const double* ranges, const double* angles, int n_ranges, double range_max,
                           double* samples_aos, long n_samples, int converged)
{
  std::shared_ptr<PFSampleSet> set = std::make_shared<PFSampleSet>();
  set->sample_count = static_cast<int>(n_samples);
  set->samples.resize(n_samples);
  set->converged = converged;
  for (long k = 0; k < n_samples; k++)
  {
    set->samples[k].pose = Eigen::Vector3d(samples_aos[4 * k], samples_aos[4 * k + 1], samples_aos[4 * k + 2]);
    set->samples[k].weight = samples_aos[4 * k + 3];
  }
  double total = static_cast<Scanner2D*>(h)->s->applyModelToSampleSet(make_scan(ranges, angles, n_ranges, range_max), set);
  for (long k = 0; k < n_samples; k++)
    samples_aos[4 * k + 3] = set->samples[k].weight;
  return total;
}

// ---------------------------------------------------------------- PointCloudScanner
void* ref_scanner3d_create(void* map3d, int max_beams)
{
  Scanner3D* h = new Scanner3D();
  h->s = std::make_shared<PointCloudScanner>();
  h->s->init(max_beams, static_cast<Map3D*>(map3d)->m);
  return h;
}
void ref_scanner3d_destroy(void* h) { delete static_cast<Scanner3D*>(h); }
void ref_scanner3d_model(void* h, int gompertz, double z_hit, double z_rand, double sigma, double a, double b, double c,
                         double in_shift, double in_scale, double out_shift)
{
  PointCloudScanner& s = *static_cast<Scanner3D*>(h)->s;
  if (gompertz)
    s.setPointCloudModelGompertz(z_hit, z_rand, sigma, a, b, c, in_shift, in_scale, out_shift);
  else
    s.setPointCloudModel(z_hit, z_rand, sigma);
}
void ref_scanner3d_set_map_factors(void* h, double off_map, double non_free, double radius)
{
  static_cast<Scanner3D*>(h)->s->setMapFactors(off_map, non_free, radius);
}
void ref_scanner3d_set_mount(void* h, const double* translation3, const double* quaternion_xyzw)
{
  geometry_msgs::Transform t;
  t.translation.x = translation3[0];
  t.translation.y = translation3[1];
  t.translation.z = translation3[2];
  t.rotation.x = quaternion_xyzw[0];
  t.rotation.y = quaternion_xyzw[1];
  t.rotation.z = quaternion_xyzw[2];
  t.rotation.w = quaternion_xyzw[3];
  static_cast<Scanner3D*>(h)->s->setPointCloudScannerToFootprintTF(t);
}
// PointCloudScanner::applyModelToSampleSet (point_cloud_scanner.cpp:111-129) on a free-standing sample set
// (AoS x, y, theta, w), no normalisation. points = packed float x, y, z in the sensor frame. Returns the model's total.
double ref_scanner3d_apply(void* h, const float* points, long n_points, double* samples_aos, long n_samples)
{
  std::shared_ptr<PointCloudData> data = std::make_shared<PointCloudData>();
  data->frame_id_ = "scanner";
  data->points_.points.resize(n_points);
  for (long k = 0; k < n_points; k++)
    data->points_.points[k] = pcl::PointXYZ(points[3 * k], points[3 * k + 1], points[3 * k + 2]);
  data->points_.width = static_cast<uint32_t>(n_points);
  data->points_.height = 1;
  std::shared_ptr<PFSampleSet> set = std::make_shared<PFSampleSet>();
  set->sample_count = static_cast<int>(n_samples);
  set->samples.resize(n_samples);
  for (long k = 0; k < n_samples; k++)
  {
    set->samples[k].pose = Eigen::Vector3d(samples_aos[4 * k], samples_aos[4 * k + 1], samples_aos[4 * k + 2]);
    set->samples[k].weight = samples_aos[4 * k + 3];
  }
  const double total = static_cast<Scanner3D*>(h)->s->applyModelToSampleSet(data, set);
  for (long k = 0; k < n_samples; k++)
    samples_aos[4 * k + 3] = set->samples[k].weight;
  return total;
}

// ---------------------------------------------------------------- ParticleFilter
void* ref_pf_create(int min_samples, int max_samples, double alpha_slow, double alpha_fast, double conv_thresh)
{
  PF* h = new PF();
  h->pf = std::make_shared<ParticleFilter>(min_samples, max_samples, alpha_slow, alpha_fast, conv_thresh, [h]() {
    Eigen::Vector3d p(0.0, 0.0, 0.0);
    if (h->uniform_map)
    {
      auto generate = [h, &p]() {
        const unsigned int k = drand48() * (h->free_ij.size() / 2);
        std::vector<double> w(2);
        h->uniform_map->convertMapToWorld({h->free_ij[2 * k], h->free_ij[2 * k + 1]}, &w);
        p[0] = w[0];
        p[1] = w[1];
        p[2] = drand48() * 2 * M_PI - M_PI;
      };
      double good = h->uniform_threshold;
      generate();
      if (good > 0.0 && h->uniform_multiplier < 1.0 && h->uniform_multiplier >= 0.0)
        while (1.0 < good)  // scorePose(p) is always 1.0 in the reference (node_2d.cpp:310-313)
        {
          generate();
          good *= h->uniform_multiplier;
        }
      h->random_next += 3;
      return p;
    }
    if (h->random_next + 3 <= h->random_poses.size())
    {
      p = Eigen::Vector3d(h->random_poses[h->random_next], h->random_poses[h->random_next + 1],
                          h->random_poses[h->random_next + 2]);
      h->random_next += 3;
    }
    return p;
  });
  return h;
}
void ref_pf_destroy(void* h) { delete static_cast<PF*>(h); }
void ref_pf_set_resample_model(void* h, int systematic)
{
  static_cast<PF*>(h)->pf->setResampleModel(systematic ? PF_RESAMPLE_SYSTEMATIC : PF_RESAMPLE_MULTINOMIAL);
}
void ref_pf_set_population_params(void* h, double err, double z) { static_cast<PF*>(h)->pf->setPopulationSizeParameters(err, z); }
void ref_pf_set_decay_rates(void* h, double a_slow, double a_fast) { static_cast<PF*>(h)->pf->setDecayRates(a_slow, a_fast); }
void ref_pf_set_random_poses(void* h, const double* poses, long n)
{
  PF* p = static_cast<PF*>(h);
  p->random_poses.assign(poses, poses + 3 * n);
  p->random_next = 0;
}
void ref_pf_set_uniform_pose_source(void* h, void* map2d, const int* free_ij, long n, double threshold, double multiplier)
{
  PF* p = static_cast<PF*>(h);
  p->uniform_map = static_cast<Map2D*>(map2d)->m;
  p->free_ij.assign(free_ij, free_ij + 2 * n);
  p->uniform_threshold = threshold;
  p->uniform_multiplier = multiplier;
  p->random_next = 0;
}
long ref_pf_random_poses_used(void* h) { return static_cast<long>(static_cast<PF*>(h)->random_next / 3); }
// ParticleFilter::initWithPoseFn (particle_filter.cpp:136-162) with poses handed over in order
void ref_pf_init_with_poses(void* h, const double* poses, long n)
{
  long next = 0;
  static_cast<PF*>(h)->pf->initWithPoseFn([&]() {
    Eigen::Vector3d p(poses[3 * (next % n)], poses[3 * (next % n) + 1], poses[3 * (next % n) + 2]);
    next++;
    return p;
  });
}
// ParticleFilter::initWithGaussian (particle_filter.cpp:105-133), drawing from the process-global drand48 stream
void ref_pf_init_gaussian(void* h, const double* mean3, const double* cov9)
{
  Eigen::Vector3d mean(mean3[0], mean3[1], mean3[2]);
  Eigen::Matrix3d cov;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      cov(i, j) = cov9[3 * i + j];
  static_cast<PF*>(h)->pf->initWithGaussian(mean, cov);
}
int ref_pf_sample_count(void* h) { return static_cast<PF*>(h)->pf->getCurrentSet()->sample_count; }
void ref_pf_get_samples(void* h, double* aos)
{
  std::shared_ptr<PFSampleSet> set = static_cast<PF*>(h)->pf->getCurrentSet();
  for (int k = 0; k < set->sample_count; k++)
  {
    aos[4 * k] = set->samples[k].pose[0];
    aos[4 * k + 1] = set->samples[k].pose[1];
    aos[4 * k + 2] = set->samples[k].pose[2];
    aos[4 * k + 3] = set->samples[k].weight;
  }
}
// overwrite poses and/or weights of the current set (what Odom::updateAction does through getCurrentSet())
void ref_pf_set_samples(void* h, const double* aos, long n)
{
  std::shared_ptr<PFSampleSet> set = static_cast<PF*>(h)->pf->getCurrentSet();
  for (long k = 0; k < n && k < set->sample_count; k++)
  {
    set->samples[k].pose = Eigen::Vector3d(aos[4 * k], aos[4 * k + 1], aos[4 * k + 2]);
    set->samples[k].weight = aos[4 * k + 3];
  }
}
// PlanarScanner::updateSensor -> ParticleFilter::updateSensor (planar_scanner.cpp:125, particle_filter.cpp:223)
int ref_pf_update_sensor_planar(void* h, void* scanner, const double* ranges, const double* angles, int n_ranges,
                                double range_max)
{
  return static_cast<Scanner2D*>(scanner)->s->updateSensor(static_cast<PF*>(h)->pf,
                                                           make_scan(ranges, angles, n_ranges, range_max))
             ? 1
             : 0;
}
void ref_pf_update_resample(void* h) { static_cast<PF*>(h)->pf->updateResample(); }
void ref_pf_get_averages(void* h, double* w_slow, double* w_fast)
{
  *w_slow = static_cast<PF*>(h)->pf->w_slow_;
  *w_fast = static_cast<PF*>(h)->pf->w_fast_;
}
void ref_pf_set_averages(void* h, double w_slow, double w_fast)
{
  static_cast<PF*>(h)->pf->w_slow_ = w_slow;
  static_cast<PF*>(h)->pf->w_fast_ = w_fast;
}
int ref_pf_leaf_count(void* h) { return static_cast<PF*>(h)->pf->getCurrentSet()->kdtree->getLeafCount(); }
long ref_pf_node_count(void* h) { return static_cast<long>(static_cast<PF*>(h)->pf->getCurrentSet()->kdtree->nodes_.size()); }
int ref_pf_is_converged(void* h) { return static_cast<PF*>(h)->pf->isConverged() ? 1 : 0; }
int ref_pf_cluster_count(void* h) { return static_cast<PF*>(h)->pf->getCurrentSet()->cluster_count; }
int ref_pf_resample_limit(void* h, int k) { return static_cast<PF*>(h)->pf->resampleLimit(k); }
// cluster label of every sample of the current set (PFKDTree::getCluster, pf_kdtree.cpp:79-90)
void ref_pf_get_sample_clusters(void* h, int* labels)
{
  std::shared_ptr<PFSampleSet> set = static_cast<PF*>(h)->pf->getCurrentSet();
  for (int k = 0; k < set->sample_count; k++)
    labels[k] = set->kdtree->getCluster(set->samples[k].pose);
}
// per cluster: count, weight, mean[3], cov(0,0),(0,1),(1,0),(1,1),(2,2)  (10 doubles; Q14: only these are written)
void ref_pf_get_cluster_stats(void* h, double* out)
{
  std::shared_ptr<PFSampleSet> set = static_cast<PF*>(h)->pf->getCurrentSet();
  for (int k = 0; k < set->cluster_count; k++)
  {
    const PFCluster& c = set->clusters[k];
    double* o = out + 10 * k;
    o[0] = c.count;
    o[1] = c.weight;
    o[2] = c.mean[0];
    o[3] = c.mean[1];
    o[4] = c.mean[2];
    o[5] = c.cov(0, 0);
    o[6] = c.cov(0, 1);
    o[7] = c.cov(1, 0);
    o[8] = c.cov(1, 1);
    o[9] = c.cov(2, 2);
  }
}
// set mean[3], cov(0,0),(0,1),(1,0),(1,1),(2,2)
void ref_pf_get_set_stats(void* h, double* out)
{
  std::shared_ptr<PFSampleSet> set = static_cast<PF*>(h)->pf->getCurrentSet();
  out[0] = set->mean[0];
  out[1] = set->mean[1];
  out[2] = set->mean[2];
  out[3] = set->cov(0, 0);
  out[4] = set->cov(0, 1);
  out[5] = set->cov(1, 0);
  out[6] = set->cov(1, 1);
  out[7] = set->cov(2, 2);
}

// ---------------------------------------------------------------- Odom::updateAction (sensors/odom.cpp:74-311)
// model_type = OdomModelType (odom.h:33-40); the draws come from the process-global drand48 stream.
void ref_pf_update_action(void* h, int model_type, const double* alpha, const double* pose, const double* delta,
                          const double* absolute_motion)
{
  Odom odom;
  odom.setModel(static_cast<OdomModelType>(model_type), alpha[0], alpha[1], alpha[2], alpha[3], alpha[4]);
  std::shared_ptr<OdomData> data = std::make_shared<OdomData>();
  for (int i = 0; i < 3; i++)
  {
    data->pose[i] = pose[i];
    data->delta[i] = delta[i];
    data->absolute_motion[i] = absolute_motion[i];
  }
  odom.updateAction(static_cast<PF*>(h)->pf, data);
}

// ---------------------------------------------------------------- the ROS node: see node_harness.inc
#include "node_harness.inc"  // (inside the extern "C" block)

// handles onto what the node owns, usable with the ref_pf_* / ref_map2d_* / ref_map3d_* readers above (NULL if absent)
void* ref_node_pf(void* h)
{
  std::shared_ptr<ParticleFilter> pf = static_cast<Node*>(h)->getPfPtr();
  if (!pf)
    return nullptr;
  PF* p = new PF();
  p->pf = pf;
  return p;
}
void* ref_node_map2d(void* h)
{
  Node2D* n = node2d(h);
  if (!n || !n->map_)
    return nullptr;
  Map2D* m = new Map2D();
  m->m = n->map_;
  return m;
}
void* ref_node_map3d(void* h)
{
  Node3D* n = node3d(h);
  if (!n || !n->map_)
    return nullptr;
  Map3D* m = new Map3D();
  m->m = n->map_;
  return m;
}
void ref_map2d_geometry(void* h, int* size_xy, double* resolution, double* origin_xy)
{
  OccupancyMap* m = static_cast<Map2D*>(h)->m.get();
  std::vector<int> sz = m->getSize();
  size_xy[0] = sz[0];
  size_xy[1] = sz[1];
  *resolution = m->resolution_;
  pcl::PointXYZ o = m->getOrigin();
  origin_xy[0] = o.x;
  origin_xy[1] = o.y;
}
}  // extern "C"
