// Host mirror of ParticleFilter / PFKDTree / PDFGaussian over the C ABI.
// Reference: src/amcl/pf/particle_filter.cpp:38-98 (ctor), 106-168 (init), 223-267 (updateSensor), 423-471
// (updateResample), 475-502 (resampleLimit), 639-675 (accessors); src/amcl/pf/pf_kdtree.cpp:42-95;
// src/amcl/pf/pdf_gaussian.cpp:77-97 (draw).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

#include "amcl/compat.h"
#include "amcl/pf/particle_filter.h"
#include "amcl/pf/pdf_gaussian.h"
#include "amcl/sensors/sensor.h"
#include "bamcl.h"

namespace badger_amcl
{
static void check(int rc, const char* what)
{
  if (rc != BAMCL_OK)
    throw std::runtime_error(std::string(what) + ": " + bamcl_last_error());
}

// current state of the process-global POSIX drand48 generator, without disturbing it
static uint64_t peekDrand48State()
{
  unsigned short zero[3] = {0, 0, 0};
  unsigned short* prev = seed48(zero);  // returns a pointer to the previous state
  unsigned short saved[3] = {prev[0], prev[1], prev[2]};
  seed48(saved);
  return (uint64_t)saved[0] | ((uint64_t)saved[1] << 16) | ((uint64_t)saved[2] << 32);
}

// ------------------------------------------------------------------------------------------ PDFGaussian
double PDFGaussian::draw(double sigma)  // pdf_gaussian.cpp:77-97: polar Box-Muller with rejection
{
  double x1, x2, w, r;
  do
  {
    do
    {
      r = drand48();
    } while (r == 0.0);
    x1 = 2.0 * r - 1.0;
    do
    {
      r = drand48();
    } while (r == 0.0);
    x2 = 2.0 * r - 1.0;
    w = x1 * x1 + x2 * x2;
  } while (w > 1.0 || w == 0.0);
  return sigma * x2 * std::sqrt(-2.0 * std::log(w) / w);
}

PDFGaussian::PDFGaussian(const Eigen::Vector3d& x, const Eigen::Matrix3d& cx) : x_(x)
{
  // Jacobi eigen-decomposition of the symmetric covariance (see the header: unpinned against Eigen::EigenSolver)
  double a[3][3], v[3][3];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++)
    {
      a[r][c] = 0.5 * (cx(r, c) + cx(c, r));
      v[r][c] = r == c ? 1.0 : 0.0;
    }
  for (int sweep = 0; sweep < 50; sweep++)
  {
    const double off = std::fabs(a[0][1]) + std::fabs(a[0][2]) + std::fabs(a[1][2]);
    if (off < 1e-300)
      break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++)
      {
        if (std::fabs(a[p][q]) < 1e-300)
          continue;
        const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
        const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
        const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < 3; k++)
        {
          const double akp = a[k][p], akq = a[k][q];
          a[k][p] = c * akp - s * akq;
          a[k][q] = s * akp + c * akq;
        }
        for (int k = 0; k < 3; k++)
        {
          const double apk = a[p][k], aqk = a[q][k];
          a[p][k] = c * apk - s * aqk;
          a[q][k] = s * apk + c * aqk;
        }
        for (int k = 0; k < 3; k++)
        {
          const double vkp = v[k][p], vkq = v[k][q];
          v[k][p] = c * vkp - s * vkq;
          v[k][q] = s * vkp + c * vkq;
        }
      }
  }
  for (int r = 0; r < 3; r++)
  {
    cd_[r] = std::sqrt(a[r][r] > 0.0 ? a[r][r] : 0.0);
    for (int c = 0; c < 3; c++)
      cr_(r, c) = v[r][c];
  }
}

Eigen::Vector3d PDFGaussian::sample()  // pdf_gaussian.cpp:53-75: x = mean + R * (sigma_i * N(0,1))
{
  double r[3];
  for (int i = 0; i < 3; i++)
    r[i] = draw(cd_[i]);
  Eigen::Vector3d x;
  for (int i = 0; i < 3; i++)
  {
    x[i] = x_[i];
    for (int j = 0; j < 3; j++)
      x[i] += cr_(i, j) * r[j];
  }
  return x;
}

// ------------------------------------------------------------------------------------------ PFKDTree
static const double kCell[3] = {0.50, 0.50, (10 * M_PI / 180)};  // pf_kdtree.cpp:35-37

PFKDTree::PFKDTree() : flushed_(true), clustered_(false), leaf_count_(0), node_count_(0) {}

void PFKDTree::clearKDTree()
{
  fetch_labels_ = nullptr;
  pending_.clear();
  labels_.clear();
  flushed_ = true;
  clustered_ = false;
  leaf_count_ = node_count_ = 0;
}

PFKDTree::Key PFKDTree::keyOf(const Eigen::Vector3d& pose)
{
  return Key((int)std::floor(pose[0] / kCell[0]), (int)std::floor(pose[1] / kCell[1]), (int)std::floor(pose[2] / kCell[2]));
}

void PFKDTree::insertPose(const Eigen::Vector3d& pose, double value)
{
  pending_.push_back(pose[0]);
  pending_.push_back(pose[1]);
  pending_.push_back(pose[2]);
  pending_.push_back(value);
  flushed_ = false;
}

void PFKDTree::install(int leaf_count, int node_count, const std::vector<Eigen::Vector3d>& poses, const std::vector<int>& labels)
{
  pending_.clear();
  labels_.clear();
  for (size_t i = 0; i < poses.size(); i++)
    labels_[keyOf(poses[i])] = labels[i];
  leaf_count_ = leaf_count;
  node_count_ = node_count;
  flushed_ = true;
  clustered_ = true;
}

// A stand-alone tree (the reference's unit tests build one directly): the buffered poses go through a private
// device sample set and the same kernels that bin a real particle set.
void PFKDTree::flush()
{
  if (flushed_)
    return;
  const long n = (long)(pending_.size() / 4);
  bamcl_pf* pf = nullptr;
  check(bamcl_pf_create(deviceContext(), 0, (int)n, &pf), "bamcl_pf_create");
  bamcl_resample_result res;
  std::vector<int> labels((size_t)n);
  int rc = bamcl_pf_upload_samples(pf, pending_.data(), n);
  if (rc == BAMCL_OK)
    rc = bamcl_pf_init_stats(pf, &res);
  if (rc == BAMCL_OK)
    rc = bamcl_pf_sample_clusters(pf, labels.data(), n);
  bamcl_pf_destroy(pf);
  check(rc, "PFKDTree::flush");
  labels_.clear();
  for (long i = 0; i < n; i++)
    labels_[keyOf(Eigen::Vector3d(pending_[4 * i], pending_[4 * i + 1], pending_[4 * i + 2]))] = labels[(size_t)i];
  leaf_count_ = res.leaf_count;
  node_count_ = res.node_count;
  flushed_ = true;
}

void PFKDTree::cluster()
{
  flush();
  clustered_ = true;
}

int PFKDTree::getCluster(const Eigen::Vector3d& pose)
{
  flush();
  if (labels_.empty() && fetch_labels_)
    fetch_labels_();
  std::map<Key, int>::const_iterator it = labels_.find(keyOf(pose));
  if (it == labels_.end())
    return -1;
  return clustered_ ? it->second : -1;  // nodes carry cluster = -1 until cluster() runs (pf_kdtree.cpp:126)
}

int PFKDTree::getLeafCount()
{
  flush();
  return leaf_count_;
}

int PFKDTree::getNodeCount()
{
  flush();
  return node_count_;
}

// ------------------------------------------------------------------------------------------ ParticleFilter
ParticleFilter::ParticleFilter(int min_samples, int max_samples, double alpha_slow, double alpha_fast,
                               double global_localization_convergence_threshold,
                               std::function<Eigen::Vector3d()> random_pose_fn)
    : dev_(nullptr), min_samples_(min_samples), max_samples_(max_samples), pop_err_(0.01), pop_z_(3), w_slow_(0.0),
      w_fast_(0.0), alpha_slow_(alpha_slow), alpha_fast_(alpha_fast), resample_model_(PF_RESAMPLE_MULTINOMIAL),
      random_pose_fn_(random_pose_fn), dist_threshold_(0.5), converged_(false),
      global_localization_convergence_threshold_(global_localization_convergence_threshold), host_valid_(true),
      host_dirty_(false)
{
  check(bamcl_pf_create(deviceContext(), min_samples, max_samples, &dev_), "bamcl_pf_create");
  set_ = std::make_shared<PFSampleSet>();
  set_->sample_count = max_samples_;
  set_->samples.resize((size_t)max_samples_);
  // the public AoS vector is the reference's type; page-locking it in place makes its lazy copies run at PCIe rate
  // (it never reallocates: every later resize stays within max_samples)
  if (bamcl_host_register(set_->samples.data(), sizeof(PFSample) * (size_t)max_samples_) == BAMCL_OK)
    pinned_samples_ = set_->samples.data();
  for (int i = 0; i < max_samples_; i++)  // particle_filter.cpp:69-79 (the device set is created in the same state)
  {
    set_->samples[i].pose = Eigen::Vector3d(0.0, 0.0, 0.0);
    set_->samples[i].weight = 1.0 / max_samples_;
  }
  set_->kdtree = std::make_shared<PFKDTree>();
  set_->cluster_count = 0;
  set_->cluster_max_count = max_samples_;
  set_->converged = 0;
  set_->owner = this;
  set_->device_scored = false;
}

ParticleFilter::~ParticleFilter()
{
  if (pinned_samples_)
    bamcl_host_unregister(pinned_samples_);
  if (dev_)
    bamcl_pf_destroy(dev_);
}

void ParticleFilter::setResampleModel(PFResampleModelType resample_model) { resample_model_ = resample_model; }

void ParticleFilter::setPopulationSizeParameters(double pop_err, double pop_z)
{
  pop_err_ = pop_err;
  pop_z_ = pop_z;
}

void ParticleFilter::setDecayRates(double alpha_slow, double alpha_fast)
{
  alpha_slow_ = alpha_slow;
  alpha_fast_ = alpha_fast;
}

bool ParticleFilter::isConverged() { return converged_; }

int ParticleFilter::resampleLimit(int k) { return bamcl_resample_limit(k, min_samples_, max_samples_, pop_err_, pop_z_); }

// PFSample {Vector3d pose; double weight} is 4 packed doubles with either Eigen: the C ABI's AoS layout
static_assert(sizeof(PFSample) == 4 * sizeof(double), "PFSample must be 4 packed doubles");

void ParticleFilter::pushIfHostDirty()
{
  if (!host_dirty_)
    return;
  if (!deviceOk(bamcl_pf_upload_samples(dev_, reinterpret_cast<const double*>(set_->samples.data()), set_->sample_count),
                "bamcl_pf_upload_samples"))
    return;  // the host copy stays the valid one
  host_dirty_ = false;
  host_valid_ = true;
}

void ParticleFilter::pullToHost()
{
  if (host_valid_)
    return;
  int n = 0;
  if (!deviceOk(bamcl_pf_sample_count(dev_, &n), "bamcl_pf_sample_count"))
    return;
  set_->sample_count = n;
  if ((int)set_->samples.size() < n)
    set_->samples.resize((size_t)n);
  if (n > 0 && !deviceOk(bamcl_pf_download_samples(dev_, reinterpret_cast<double*>(set_->samples.data()), n),
                         "bamcl_pf_download_samples"))
    return;
  host_valid_ = true;
  set_->device_scored = false;
}

// set->clusters of the mirror <- the device's cluster table (computeClusterStatsForSet, particle_filter.cpp:505-636)
void ParticleFilter::syncClusters()
{
  if (!clusters_pending_)
    return;
  const int count = set_->cluster_count;
  set_->clusters.assign((size_t)count, PFCluster());
  if (count > 0)
  {
    std::vector<double> rows((size_t)count * 10);
    if (!deviceOk(bamcl_pf_cluster_stats(dev_, rows.data(), count), "bamcl_pf_cluster_stats"))
      return;
    for (int c = 0; c < count; c++)
    {
      const double* o = &rows[(size_t)c * 10];
      PFCluster& cl = set_->clusters[(size_t)c];
      cl.count = (int)o[0];
      cl.weight = o[1];
      cl.mean = Eigen::Vector3d(o[2], o[3], o[4]);
      cl.cov = Eigen::Matrix3d::Zero();
      cl.cov(0, 0) = o[5];
      cl.cov(0, 1) = o[6];
      cl.cov(1, 0) = o[7];
      cl.cov(1, 1) = o[8];
      cl.cov(2, 2) = o[9];
    }
  }
  clusters_pending_ = false;
}

std::shared_ptr<PFSampleSet> ParticleFilter::getCurrentSet()
{
  syncClusters();
  pullToHost();
  host_dirty_ = true;  // the caller holds a mutable set (odom.cpp:107-121 writes the poses through it)
  return set_;
}

std::shared_ptr<PFSampleSet> ParticleFilter::peekCurrentSet()
{
  syncClusters();
  pullToHost();
  return set_;
}

// install kd-tree / cluster / set statistics of the current device set into the mirror
void ParticleFilter::installStats(const void* result)
{
  const bamcl_resample_result& r = *static_cast<const bamcl_resample_result*>(result);
  set_->sample_count = r.sample_count;
  set_->cluster_count = r.cluster_count;
  clusters_pending_ = true;  // the cluster table (176 B per cluster in the reference's type) is fetched when someone looks
  set_->mean = Eigen::Vector3d(r.mean[0], r.mean[1], r.mean[2]);
  set_->cov = Eigen::Matrix3d::Zero();
  set_->cov(0, 0) = r.cov[0];
  set_->cov(0, 1) = r.cov[1];
  set_->cov(1, 0) = r.cov[2];
  set_->cov(1, 1) = r.cov[3];
  set_->cov(2, 2) = r.cov[4];
  // the set's kd-tree facade: counts now, labels on demand would need the poses -- fetch both lazily via peek
  set_->kdtree->clearKDTree();
  set_->kdtree->leaf_count_ = r.leaf_count;
  set_->kdtree->node_count_ = r.node_count;
  set_->kdtree->flushed_ = true;
  set_->kdtree->clustered_ = true;
  PFKDTree* tree = set_->kdtree.get();
  tree->fetch_labels_ = [this, tree]() {
    // PFKDTree::getCluster on a set's own tree (particle_filter.cpp:573 does this per sample): one download
    pullToHost();
    const int n = set_->sample_count;
    std::vector<int> labels((size_t)n);
    if (n > 0 && !deviceOk(bamcl_pf_sample_clusters(dev_, labels.data(), n), "bamcl_pf_sample_clusters"))
      return;
    for (int i = 0; i < n; i++)
      tree->labels_[PFKDTree::keyOf(set_->samples[(size_t)i].pose)] = labels[(size_t)i];
  };
}

void ParticleFilter::initWithPoseFn(std::function<Eigen::Vector3d()> pose_fn)
{
  set_->sample_count = max_samples_;
  set_->samples.resize((size_t)max_samples_);
  for (int i = 0; i < max_samples_; i++)
  {
    set_->samples[i].weight = 1.0 / max_samples_;
    set_->samples[i].pose = pose_fn();
  }
  host_dirty_ = true;
  host_valid_ = true;
  pushIfHostDirty();
  w_slow_ = w_fast_ = 0.0;
  bamcl_resample_result res;
  check(bamcl_pf_init_stats(dev_, &res), "bamcl_pf_init_stats");  // computeClusterStatsForSet + kd-tree
  installStats(&res);
  set_->converged = false;  // initConverged
  converged_ = false;
}

void ParticleFilter::initWithGaussian(const Eigen::Vector3d& mean, const Eigen::Matrix3d& cov)
{
  // particle_filter.cpp:105-133 on the device: the covariance is decomposed here (PDFGaussian's constructor), the
  // 3 * max_samples Gaussian draws and the poses are made on the device from libc's drand48 state, which is then advanced
  // to where the reference's loop would have left it
  PDFGaussian pdf(mean, cov);
  double m[3], cr[9], cd[3];
  for (int i = 0; i < 3; i++)
  {
    m[i] = pdf.mean()[i];
    cd[i] = pdf.sigmas()[i];
    for (int j = 0; j < 3; j++)
      cr[3 * i + j] = pdf.rotation()(i, j);
  }
  uint64_t state_out = 0;
  bamcl_resample_result res;
  if (!deviceOk(bamcl_pf_init_gaussian(dev_, m, cr, cd, peekDrand48State(), &state_out, &res), "bamcl_pf_init_gaussian"))
  {
    // e.g. a drand48 output of exactly 0.0 inside the draw range (2^-48 per draw): the sequential loop handles it
    clearDeviceError();
    initWithPoseFn([&pdf]() { return pdf.sample(); });
    return;
  }
  unsigned short st[3] = {(unsigned short)(state_out & 0xFFFF), (unsigned short)((state_out >> 16) & 0xFFFF),
                          (unsigned short)((state_out >> 32) & 0xFFFF)};
  seed48(st);
  w_slow_ = w_fast_ = 0.0;
  host_valid_ = false;
  host_dirty_ = false;
  installStats(&res);
  set_->converged = false;  // initConverged
  converged_ = false;
}

void ParticleFilter::updateSensor(std::function<double(std::shared_ptr<SensorData>, std::shared_ptr<PFSampleSet>)> sensor_fn,
                                  std::shared_ptr<SensorData> sensor_data)
{
  // The device scanners (PlanarScanner / PointCloudScanner::applyModelToSampleSet) score the device set in place
  // and flag it; any other functor gets the host mirror, as in the reference, and its weights are uploaded.
  pushIfHostDirty();
  set_->device_scored = false;
  set_->device_failed = false;
  const double total = sensor_fn(sensor_data, set_);
  if (set_->device_failed)
    return;  // the device could not score (lastDeviceError()): weights untouched, no exception
  if (!set_->device_scored)
  {
    if (!host_valid_)
    {
      deviceFail("ParticleFilter::updateSensor: functor neither scored on the device nor had a valid host set");
      return;
    }
    host_dirty_ = true;
    pushIfHostDirty();
  }
  if (!deviceOk(bamcl_pf_normalise(dev_, total), "bamcl_pf_normalise"))  // w /= total, or 1/n each (:237-266)
    return;
  host_valid_ = false;
  if (total > 0.0)
  {
    int n = 0;
    if (!deviceOk(bamcl_pf_sample_count(dev_, &n), "bamcl_pf_sample_count") || n <= 0)
      return;
    const double w_avg = total / n;  // the reference's w_avg is the same sum of the un-normalised weights
    if (w_slow_ == 0.0)
      w_slow_ = w_avg;
    else
      w_slow_ += alpha_slow_ * (w_avg - w_slow_);
    if (w_fast_ == 0.0)
      w_fast_ = w_avg;
    else
      w_fast_ += alpha_fast_ * (w_avg - w_fast_);
  }
}

void ParticleFilter::updateResample()
{
  pushIfHostDirty();
  double w_diff = 1.0 - w_fast_ / w_slow_;
  if (w_diff < 0.0)
    w_diff = 0.0;
  bamcl_resample_params prm;
  std::memset(&prm, 0, sizeof(prm));
  prm.resample_systematic = resample_model_ == PF_RESAMPLE_SYSTEMATIC ? 1 : 0;
  prm.min_samples = min_samples_;
  prm.max_samples = max_samples_;
  prm.pop_err = pop_err_;
  prm.pop_z = pop_z_;
  prm.w_diff = w_diff;
  prm.dist_threshold = dist_threshold_;
  prm.convergence_threshold = global_localization_convergence_threshold_;
  // drand48 contract: the device replays the draws from the generator state; the host advances libc's generator
  // by the same draws so that later consumers (random_pose_fn_, the motion model) see the reference's stream.
  prm.drand48_state = peekDrand48State();
  std::vector<double> random_poses;
  if (device_pose_source_ && w_diff > 0.0)
  {
    // both models: the device draws the start / branch values AND the random poses from the stream (bamcl.h)
  }
  else if (prm.resample_systematic)
  {
    // particle_filter.cpp:292-309: set size from the CURRENT set's leaf count, one draw, then the random poses
    int new_count = resampleLimit(set_->kdtree->getLeafCount());
    if (w_diff > 0.0)
    {
      new_count *= (1.0 + w_diff);
      if (new_count > max_samples_)
        new_count = max_samples_;
    }
    const int n_random = w_diff * new_count;
    (void)drand48();
    random_poses.resize((size_t)3 * n_random);
    for (int i = 0; i < n_random; i++)
    {
      const Eigen::Vector3d p = random_pose_fn_();
      random_poses[3 * i] = p[0];
      random_poses[3 * i + 1] = p[1];
      random_poses[3 * i + 2] = p[2];
    }
  }
  else if (w_diff > 0.0)
  {
    // multinomial with an arbitrary host pose function (particle_filter.cpp:385-393): the branch loop runs here - it is the
    // only part that cannot be replayed from the stream, the pose function draws from it too - and everything else on the
    // device (bamcl_pf_resample_multinomial_drawn). libc's generator state is recorded before every output so that it can
    // be put back to where the reference's loop stops (the KLD bound, :416).
    std::vector<double> draws((size_t)max_samples_);
    std::vector<uint64_t> state_before((size_t)max_samples_ + 1);
    for (int k = 0; k < max_samples_; k++)
    {
      state_before[(size_t)k] = peekDrand48State();
      if (drand48() < w_diff)
      {
        const Eigen::Vector3d p = random_pose_fn_();
        random_poses.push_back(p[0]);
        random_poses.push_back(p[1]);
        random_poses.push_back(p[2]);
        draws[(size_t)k] = -(double)(random_poses.size() / 3);
      }
      else
        draws[(size_t)k] = drand48();
    }
    state_before[(size_t)max_samples_] = peekDrand48State();
    prm.random_poses = random_poses.empty() ? nullptr : random_poses.data();
    prm.n_random_poses = (long)(random_poses.size() / 3);
    bamcl_resample_result res;
    const bool ok = deviceOk(bamcl_pf_resample_multinomial_drawn(dev_, &prm, draws.data(), &res), "bamcl_pf_resample_multinomial_drawn");
    const uint64_t st64 = state_before[ok ? (size_t)res.sample_count : 0];  // on failure: as if nothing had been drawn
    unsigned short st[3] = {(unsigned short)(st64 & 0xFFFF), (unsigned short)((st64 >> 16) & 0xFFFF),
                            (unsigned short)((st64 >> 32) & 0xFFFF)};
    seed48(st);
    if (!ok)
      return;  // the set is left as it was
    w_slow_ = w_fast_ = 0.0;
    host_valid_ = false;
    host_dirty_ = false;
    installStats(&res);
    set_->converged = res.converged;
    converged_ = res.converged != 0;
    return;
  }
  prm.random_poses = random_poses.empty() ? nullptr : random_poses.data();
  prm.n_random_poses = (long)(random_poses.size() / 3);
  bamcl_resample_result res;
  if (!deviceOk(bamcl_pf_resample(dev_, &prm, &res), "bamcl_pf_resample"))
    return;  // the set is left as it was
  if (!prm.resample_systematic || (device_pose_source_ && w_diff > 0.0))
  {
    // draws were consumed on the device: advance libc's generator to the returned state
    unsigned short st[3] = {(unsigned short)(res.drand48_state & 0xFFFF), (unsigned short)((res.drand48_state >> 16) & 0xFFFF),
                            (unsigned short)((res.drand48_state >> 32) & 0xFFFF)};
    seed48(st);
  }
  if (w_diff > 0.0)
    w_slow_ = w_fast_ = 0.0;
  host_valid_ = false;
  host_dirty_ = false;
  installStats(&res);
  set_->converged = res.converged;
  converged_ = res.converged != 0;
}

void ParticleFilter::setDevicePoseSource(bamcl_map2d* map, double starting_weight_threshold, double deweight_multiplier)
{
  check(bamcl_pf_set_random_pose_source(dev_, map, starting_weight_threshold, deweight_multiplier),
        "bamcl_pf_set_random_pose_source");
  device_pose_source_ = map != nullptr;
}

bool ParticleFilter::getClusterStats(int cidx, double* weight, Eigen::Vector3d* mean)
{
  if (cidx >= set_->cluster_count)
    return false;
  syncClusters();
  if (cidx >= (int)set_->clusters.size())
    return false;
  *weight = set_->clusters[(size_t)cidx].weight;
  *mean = set_->clusters[(size_t)cidx].mean;
  return true;
}

std::vector<int> ParticleFilter::lastResampleSources()
{
  int n = 0;
  check(bamcl_pf_sample_count(dev_, &n), "bamcl_pf_sample_count");
  std::vector<int> src((size_t)n);
  if (n > 0)
    check(bamcl_pf_resample_sources(dev_, src.data(), n), "bamcl_pf_resample_sources");
  return src;
}
}  // namespace badger_amcl
