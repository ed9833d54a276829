// Host mirror of include/amcl/pf/particle_filter.h:35-188 of the reference: same types, same method names and
// argument meaning. The particle store is device-resident SoA (x, y, theta, w as separate f64 arrays); the public
// AoS vector PFSampleSet::samples is a lazily synchronised mirror:
//   * getCurrentSet() brings the mirror up to date and -- because callers mutate it (odom.cpp:107-121) -- assumes
//     it was modified: the next device operation uploads it first;
//   * peekCurrentSet() brings it up to date for reading only (no upload afterwards).
// Scalars the reference keeps in the class (w_slow_/w_fast_, converged_, current_set_) stay on the host.
#ifndef BAMCL_AMCL_PF_PARTICLE_FILTER_H
#define BAMCL_AMCL_PF_PARTICLE_FILTER_H
#include <functional>
#include <memory>
#include <vector>

#include "amcl/compat.h"
#include "amcl/pf/pf_kdtree.h"

struct bamcl_ctx;
struct bamcl_pf;
struct bamcl_map2d;

namespace badger_amcl
{
enum PFResampleModelType
{
  PF_RESAMPLE_MULTINOMIAL,
  PF_RESAMPLE_SYSTEMATIC,
};

struct PFSample
{
  Eigen::Vector3d pose;
  double weight;
};

struct PFCluster
{
  int count;
  double weight;
  Eigen::Vector3d mean;
  Eigen::Matrix3d cov;
  double m[4], c[2][2];
};

class ParticleFilter;

struct PFSampleSet
{
  int sample_count;
  std::vector<PFSample> samples;
  std::shared_ptr<PFKDTree> kdtree;
  int cluster_count, cluster_max_count;
  std::vector<PFCluster> clusters;  // cluster_count entries (the reference keeps max_samples of them)
  Eigen::Vector3d mean;
  Eigen::Matrix3d cov;  // only (0..1, 0..1) and (2, 2) are written, as in the reference (particle_filter.cpp:624-635)
  int converged;
  ParticleFilter* owner;  // device port: the filter whose device set this mirrors
  bool device_scored;     // device port: a device sensor model has updated the weights since the last sync
  bool device_failed = false;  // device port: the last sensor model call failed on the device (see lastDeviceError())
};

class SensorData;

// shared per-process device context (one GPU per process; BAMCL_DEVICE or LOCAL_RANK selects it)
bamcl_ctx* deviceContext();

class ParticleFilter
{
public:
  ParticleFilter(int min_samples, int max_samples, double alpha_slow, double alpha_fast,
                 double global_localization_convergence_threshold, std::function<Eigen::Vector3d()> random_pose_fn);
  ~ParticleFilter();
  ParticleFilter(const ParticleFilter&) = delete;
  ParticleFilter& operator=(const ParticleFilter&) = delete;
  void setResampleModel(PFResampleModelType resample_model);
  void initWithGaussian(const Eigen::Vector3d& mean, const Eigen::Matrix3d& cov);
  void initWithPoseFn(std::function<Eigen::Vector3d()> pose_fn);
  void updateSensor(std::function<double(std::shared_ptr<SensorData>, std::shared_ptr<PFSampleSet>)> sensor_fn_ptr,
                    std::shared_ptr<SensorData> sensor_data);
  void updateResample();
  bool getClusterStats(int cluster, double* weight, Eigen::Vector3d* mean);
  void setPopulationSizeParameters(double pop_err, double pop_z);
  void setDecayRates(double alpha_slow, double alpha_fast);
  std::shared_ptr<PFSampleSet> getCurrentSet();
  bool isConverged();

  // ---- additions of the device port
  std::shared_ptr<PFSampleSet> peekCurrentSet();  // up-to-date mirror, read-only contract
  bamcl_pf* deviceHandle() { return dev_; }
  void pushIfHostDirty();                         // upload the mirror if the caller may have modified it
  void markDeviceModified()                       // a device operation changed the set: the host mirror is stale
  {
    host_valid_ = false;
    host_dirty_ = false;
  }
  int resampleLimit(int k);                       // particle_filter.cpp:475-502
  // The node's pose function (Node::uniformPoseGenerator over the free-space list of `map`, node.cpp:262-264, 823-868)
  // replayed on the device from libc's drand48 stream: random poses of updateResample (both models; the only way to
  // run multinomial resampling with w_diff > 0 on the device) are then not requested from random_pose_fn_.
  // map = a bamcl_map2d with a free-space list (OccupancyMap::deviceHandle + bamcl_map2d_update_free_space).
  void setDevicePoseSource(struct ::bamcl_map2d* map, double starting_weight_threshold, double deweight_multiplier);
  double slowAverage() const { return w_slow_; }
  double fastAverage() const { return w_fast_; }
  // sources of the last resample (index into the previous set, -1 = random pose): test / diagnostics aid
  std::vector<int> lastResampleSources();

private:
  void installStats(const void* resample_result);
  void pullToHost();
  bamcl_pf* dev_;
  void* pinned_samples_ = nullptr;  // the samples vector's storage while it is page-locked
  bool clusters_pending_ = false;   // set->clusters not fetched from the device yet (syncClusters)
  void syncClusters();
  int min_samples_, max_samples_;
  double pop_err_, pop_z_;
  double w_slow_, w_fast_, alpha_slow_, alpha_fast_;
  PFResampleModelType resample_model_;
  std::function<Eigen::Vector3d()> random_pose_fn_;
  double dist_threshold_;
  bool converged_;
  double global_localization_convergence_threshold_;
  std::shared_ptr<PFSampleSet> set_;  // mirror of the CURRENT device set (the device double-buffers internally)
  bool host_valid_, host_dirty_;
  bool device_pose_source_ = false;
};
}  // namespace badger_amcl
#endif
