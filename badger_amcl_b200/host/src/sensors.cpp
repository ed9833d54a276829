// Host mirror of PlanarScanner / PointCloudScanner over the C ABI.
// Reference: src/amcl/sensors/planar_scanner.cpp:33-164, 540-550, 684-701;
// src/amcl/sensors/point_cloud_scanner.cpp:36-129, 250-260. The per-particle x beam / point loops
// (:168-640 and :132-248) are the device kernels K-a, K-b, K-c behind bamcl_pf_score_planar / bamcl_pf_score_cloud.
#include <cmath>
#include <stdexcept>
#include <string>

#include "amcl/sensors/planar_scanner.h"
#include "amcl/sensors/point_cloud_scanner.h"
#include "amcl/compat.h"
#include "bamcl.h"

namespace badger_amcl
{
// ------------------------------------------------------------------------------------------ PlanarScanner
PlanarScanner::PlanarScanner()
    : model_type_(PLANAR_MODEL_BEAM), planar_scanner_pose_(0.0, 0.0, 0.0), max_beams_(0), do_beamskip_(false),
      beam_skip_distance_(0.0), beam_skip_threshold_(0.0), beam_skip_error_threshold_(0.0), z_hit_(0.0), z_short_(0.0),
      z_max_(0.0), z_rand_(0.0), sigma_hit_(0.0), lambda_short_(0.0), gompertz_a_(0.0), gompertz_b_(0.0), gompertz_c_(0.0),
      input_shift_(0.0), input_scale_(0.0), output_shift_(0.0), off_map_factor_(1.0), non_free_space_factor_(1.0),
      non_free_space_radius_(0.0)
{
}

void PlanarScanner::init(int max_beams, std::shared_ptr<OccupancyMap> map)
{
  max_beams_ = max_beams;
  map_ = map;
}

void PlanarScanner::setModelBeam(double z_hit, double z_short, double z_max, double z_rand, double sigma_hit,
                                 double lambda_short)
{
  model_type_ = PLANAR_MODEL_BEAM;
  z_hit_ = z_hit;
  z_short_ = z_short;
  z_max_ = z_max;
  z_rand_ = z_rand;
  sigma_hit_ = sigma_hit;
  lambda_short_ = lambda_short;
}

void PlanarScanner::setModelLikelihoodField(double z_hit, double z_rand, double sigma_hit, double max_distance_to_object)
{
  model_type_ = PLANAR_MODEL_LIKELIHOOD_FIELD;
  z_hit_ = z_hit;
  z_rand_ = z_rand;
  sigma_hit_ = sigma_hit;
  map_->updateDistancesLUT(max_distance_to_object);
}

void PlanarScanner::setModelLikelihoodFieldProb(double z_hit, double z_rand, double sigma_hit, double max_distance_to_object,
                                                bool do_beamskip, double beam_skip_distance, double beam_skip_threshold,
                                                double beam_skip_error_threshold)
{
  model_type_ = PLANAR_MODEL_LIKELIHOOD_FIELD_PROB;
  z_hit_ = z_hit;
  z_rand_ = z_rand;
  sigma_hit_ = sigma_hit;
  do_beamskip_ = do_beamskip;
  beam_skip_distance_ = beam_skip_distance;
  beam_skip_threshold_ = beam_skip_threshold;
  beam_skip_error_threshold_ = beam_skip_error_threshold;
  map_->updateDistancesLUT(max_distance_to_object);
}

void PlanarScanner::setModelLikelihoodFieldGompertz(double z_hit, double z_rand, double sigma_hit,
                                                    double max_distance_to_object, double gompertz_a, double gompertz_b,
                                                    double gompertz_c, double input_shift, double input_scale,
                                                    double output_shift)
{
  model_type_ = PLANAR_MODEL_LIKELIHOOD_FIELD_GOMPERTZ;
  z_hit_ = z_hit;
  z_rand_ = z_rand;
  sigma_hit_ = sigma_hit;
  gompertz_a_ = gompertz_a;
  gompertz_b_ = gompertz_b;
  gompertz_c_ = gompertz_c;
  input_shift_ = input_shift;
  input_scale_ = input_scale;
  output_shift_ = output_shift;
  map_->updateDistancesLUT(max_distance_to_object);
}

void PlanarScanner::setMapFactors(double off_map_factor, double non_free_space_factor, double non_free_space_radius)
{
  off_map_factor_ = off_map_factor;
  non_free_space_factor_ = non_free_space_factor;
  non_free_space_radius_ = non_free_space_radius;
}

void PlanarScanner::setPlanarScannerPose(const Eigen::Vector3d& scanner_pose) { planar_scanner_pose_ = scanner_pose; }

double PlanarScanner::applyGompertz(double p)  // planar_scanner.cpp:540-550
{
  p = p * input_scale_ + input_shift_;
  p = gompertz_a_ * std::exp(-1.0 * gompertz_b_ * std::exp(-1.0 * gompertz_c_ * p));
  p += output_shift_;
  return p;
}

bool PlanarScanner::updateSensor(std::shared_ptr<ParticleFilter> pf, std::shared_ptr<SensorData> data)
{
  if (max_beams_ < 2)
    return false;
  std::function<double(std::shared_ptr<SensorData>, std::shared_ptr<PFSampleSet>)> sensor_fn =
      (std::bind(&PlanarScanner::applyModelToSampleSet, this, std::placeholders::_1, std::placeholders::_2));
  pf->updateSensor(sensor_fn, data);
  return true;
}

double PlanarScanner::applyModelToSampleSet(std::shared_ptr<SensorData> data, std::shared_ptr<PFSampleSet> set)
{
  if (max_beams_ < 2)
    return 0.0;
  std::shared_ptr<PlanarData> ndata = std::dynamic_pointer_cast<PlanarData>(data);
  if (!ndata || !set || !set->owner)
  {
    deviceFail("PlanarScanner::applyModelToSampleSet: need PlanarData and a set owned by a ParticleFilter");
    return 0.0;
  }
  ParticleFilter* pf = set->owner;
  pf->pushIfHostDirty();
  bamcl_planar_model m;
  m.model_type = (int)model_type_;  // the enums agree with BAMCL_PLANAR_* by construction
  m.max_beams = max_beams_;
  m.z_hit = z_hit_;
  m.z_short = z_short_;
  m.z_max = z_max_;
  m.z_rand = z_rand_;
  m.sigma_hit = sigma_hit_;
  m.lambda_short = lambda_short_;
  m.gompertz_a = gompertz_a_;
  m.gompertz_b = gompertz_b_;
  m.gompertz_c = gompertz_c_;
  m.input_shift = input_shift_;
  m.input_scale = input_scale_;
  m.output_shift = output_shift_;
  m.do_beamskip = do_beamskip_ ? 1 : 0;
  m.beam_skip_distance = beam_skip_distance_;
  m.beam_skip_threshold = beam_skip_threshold_;
  m.beam_skip_error_threshold = beam_skip_error_threshold_;
  m.off_map_factor = off_map_factor_;
  m.non_free_space_factor = non_free_space_factor_;
  m.non_free_space_radius = non_free_space_radius_;
  for (int i = 0; i < 3; i++)
    m.scanner_pose[i] = planar_scanner_pose_[i];
  double total = 0.0;
  if (!deviceOk(bamcl_pf_score_planar(pf->deviceHandle(), map_->deviceHandle(), &m, ndata->ranges_.data(), ndata->angles_.data(),
                                      ndata->range_count_, ndata->range_max_, set->converged, &total),
                "bamcl_pf_score_planar"))
  {
    set->device_failed = true;  // ParticleFilter::updateSensor leaves the weights as they were
    return 0.0;
  }
  set->device_scored = true;
  return total;
}

// ------------------------------------------------------------------------------------------ PointCloudScanner
PointCloudScanner::PointCloudScanner()
    : model_type_(POINT_CLOUD_MODEL), gompertz_a_(0.0), gompertz_b_(0.0), gompertz_c_(0.0), input_shift_(0.0), input_scale_(0.0),
      output_shift_(0.0), off_map_factor_(1.0), non_free_space_factor_(1.0), non_free_space_radius_(0.0), z_hit_(0.0),
      z_rand_(0.0), sigma_hit_(0.0), max_beams_(0)
{
}

void PointCloudScanner::init(int max_beams, std::shared_ptr<OctoMap> map)
{
  max_beams_ = max_beams;
  map_ = map;
}

void PointCloudScanner::setPointCloudModel(double z_hit, double z_rand, double sigma_hit)
{
  model_type_ = POINT_CLOUD_MODEL;
  z_hit_ = z_hit;
  z_rand_ = z_rand;
  sigma_hit_ = sigma_hit;
}

void PointCloudScanner::setPointCloudModelGompertz(double z_hit, double z_rand, double sigma_hit, double gompertz_a,
                                                   double gompertz_b, double gompertz_c, double input_shift,
                                                   double input_scale, double output_shift)
{
  model_type_ = POINT_CLOUD_MODEL_GOMPERTZ;
  z_hit_ = z_hit;
  z_rand_ = z_rand;
  sigma_hit_ = sigma_hit;
  gompertz_a_ = gompertz_a;
  gompertz_b_ = gompertz_b;
  gompertz_c_ = gompertz_c;
  input_shift_ = input_shift;
  input_scale_ = input_scale;
  output_shift_ = output_shift;
}

void PointCloudScanner::setMapFactors(double off_map_factor, double non_free_space_factor, double non_free_space_radius)
{
  off_map_factor_ = off_map_factor;
  non_free_space_factor_ = non_free_space_factor;
  non_free_space_radius_ = non_free_space_radius;
}

void PointCloudScanner::setPointCloudScannerToFootprintTF(geometry_msgs::Transform tf_msg)
{
  point_cloud_scanner_to_footprint_tf_ = tf_msg;
}

int PointCloudScanner::getMaxBeams() { return max_beams_; }

double PointCloudScanner::applyGompertz(double p)  // point_cloud_scanner.cpp:250-260
{
  p = p * input_scale_ + input_shift_;
  p = gompertz_a_ * std::exp(-1.0 * gompertz_b_ * std::exp(-1.0 * gompertz_c_ * p));
  p += output_shift_;
  return p;
}

bool PointCloudScanner::updateSensor(std::shared_ptr<ParticleFilter> pf, std::shared_ptr<SensorData> data)
{
  if (max_beams_ < 2)
    return false;
  std::function<double(std::shared_ptr<SensorData>, std::shared_ptr<PFSampleSet>)> sensor_fn =
      (std::bind(&PointCloudScanner::applyModelToSampleSet, this, std::placeholders::_1, std::placeholders::_2));
  pf->updateSensor(sensor_fn, data);
  return true;
}

double PointCloudScanner::applyModelToSampleSet(std::shared_ptr<SensorData> data, std::shared_ptr<PFSampleSet> set)
{
  if (max_beams_ < 2)
    return 0.0;
  std::shared_ptr<PointCloudData> ndata = std::dynamic_pointer_cast<PointCloudData>(data);
  if (!ndata || !set || !set->owner)
  {
    deviceFail("PointCloudScanner::applyModelToSampleSet: need PointCloudData and a set owned by a ParticleFilter");
    return 0.0;
  }
  ParticleFilter* pf = set->owner;
  pf->pushIfHostDirty();
  bamcl_cloud_model m;
  m.model_type = (int)model_type_;
  m.max_beams = max_beams_;
  m.z_hit = z_hit_;
  m.z_rand = z_rand_;
  m.sigma_hit = sigma_hit_;
  m.gompertz_a = gompertz_a_;
  m.gompertz_b = gompertz_b_;
  m.gompertz_c = gompertz_c_;
  m.input_shift = input_shift_;
  m.input_scale = input_scale_;
  m.output_shift = output_shift_;
  m.off_map_factor = off_map_factor_;
  const geometry_msgs::Transform& t = point_cloud_scanner_to_footprint_tf_;
  m.mount_translation[0] = t.translation.x;
  m.mount_translation[1] = t.translation.y;
  m.mount_translation[2] = t.translation.z;
  m.mount_rotation[0] = t.rotation.x;
  m.mount_rotation[1] = t.rotation.y;
  m.mount_rotation[2] = t.rotation.z;
  m.mount_rotation[3] = t.rotation.w;
  // pcl::PointXYZ is padded to 16 bytes in PCL and 12 here: repack to the ABI's packed float triples
  const size_t n = ndata->points_.points.size();
  std::vector<float> xyz(3 * n);
  for (size_t i = 0; i < n; i++)
  {
    xyz[3 * i] = ndata->points_.points[i].x;
    xyz[3 * i + 1] = ndata->points_.points[i].y;
    xyz[3 * i + 2] = ndata->points_.points[i].z;
  }
  double total = 0.0;
  if (!deviceOk(bamcl_pf_score_cloud(pf->deviceHandle(), map_->deviceHandle(), &m, xyz.data(), (long)n, &total), "bamcl_pf_score_cloud"))
  {
    set->device_failed = true;
    return 0.0;
  }
  set->device_scored = true;
  return total;
}
}  // namespace badger_amcl
