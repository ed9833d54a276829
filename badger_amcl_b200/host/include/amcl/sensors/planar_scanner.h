// Host mirror of include/amcl/sensors/planar_scanner.h:35-172 of the reference.
#ifndef BAMCL_AMCL_SENSORS_PLANAR_SCANNER_H
#define BAMCL_AMCL_SENSORS_PLANAR_SCANNER_H
#include <memory>
#include <vector>

#include "amcl/map/occupancy_map.h"
#include "amcl/pf/particle_filter.h"
#include "amcl/sensors/sensor.h"

namespace badger_amcl
{
enum PlanarModelType
{
  PLANAR_MODEL_BEAM,
  PLANAR_MODEL_LIKELIHOOD_FIELD,
  PLANAR_MODEL_LIKELIHOOD_FIELD_PROB,
  PLANAR_MODEL_LIKELIHOOD_FIELD_GOMPERTZ,
};

class PlanarData : public SensorData
{
public:
  virtual ~PlanarData() = default;
  int range_count_;
  double range_max_;
  std::vector<double> ranges_;
  std::vector<double> angles_;
};

// Cheaply copyable (the node copies a prototype per sensor frame, node_2d.cpp:454); the map is shared.
class PlanarScanner : public Sensor
{
public:
  PlanarScanner();
  void init(int max_beams, std::shared_ptr<OccupancyMap> map);
  void setModelBeam(double z_hit, double z_short, double z_max, double z_rand, double sigma_hit, double lambda_short);
  void setModelLikelihoodField(double z_hit, double z_rand, double sigma_hit, double max_distance_to_object);
  void setModelLikelihoodFieldProb(double z_hit, double z_rand, double sigma_hit, double max_distance_to_object,
                                   bool do_beamskip, double beam_skip_distance, double beam_skip_threshold,
                                   double beam_skip_error_threshold);
  void setModelLikelihoodFieldGompertz(double z_hit, double z_rand, double sigma_hit, double max_distance_to_object,
                                       double gompertz_a, double gompertz_b, double gompertz_c, double input_shift,
                                       double input_scale, double output_shift);
  void setMapFactors(double off_map_factor, double non_free_space_factor, double non_free_space_radius);
  bool updateSensor(std::shared_ptr<ParticleFilter> pf, std::shared_ptr<SensorData> data);
  double applyModelToSampleSet(std::shared_ptr<SensorData> data, std::shared_ptr<PFSampleSet> set);
  void setPlanarScannerPose(const Eigen::Vector3d& scanner_pose);
  double applyGompertz(double p);

private:
  PlanarModelType model_type_;
  std::shared_ptr<OccupancyMap> map_;
  Eigen::Vector3d planar_scanner_pose_;
  int max_beams_;
  bool do_beamskip_;
  double beam_skip_distance_, beam_skip_threshold_, beam_skip_error_threshold_;
  double z_hit_, z_short_, z_max_, z_rand_, sigma_hit_, lambda_short_;
  double gompertz_a_, gompertz_b_, gompertz_c_, input_shift_, input_scale_, output_shift_;
  double off_map_factor_, non_free_space_factor_, non_free_space_radius_;
};
}  // namespace badger_amcl
#endif
