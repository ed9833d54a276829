// Host mirror of include/amcl/sensors/point_cloud_scanner.h:37-124 of the reference.
#ifndef BAMCL_AMCL_SENSORS_POINT_CLOUD_SCANNER_H
#define BAMCL_AMCL_SENSORS_POINT_CLOUD_SCANNER_H
#include <memory>
#include <string>

#include "amcl/map/octomap.h"
#include "amcl/pf/particle_filter.h"
#include "amcl/sensors/sensor.h"

namespace badger_amcl
{
enum PointCloudModelType
{
  POINT_CLOUD_MODEL,
  POINT_CLOUD_MODEL_GOMPERTZ,
};

class PointCloudData : public SensorData
{
public:
  virtual ~PointCloudData() = default;
  std::string frame_id_;
  pcl::PointCloud<pcl::PointXYZ> points_;
};

class PointCloudScanner : public Sensor
{
public:
  PointCloudScanner();
  ~PointCloudScanner() = default;
  void init(int max_beams, std::shared_ptr<OctoMap> map);
  void setPointCloudModel(double z_hit, double z_rand, double sigma_hit);
  void setPointCloudModelGompertz(double z_hit, double z_rand, double sigma_hit, double gompertz_a, double gompertz_b,
                                  double gompertz_c, double input_shift, double input_scale, double output_shift);
  bool updateSensor(std::shared_ptr<ParticleFilter> pf, std::shared_ptr<SensorData> data);
  double applyModelToSampleSet(std::shared_ptr<SensorData> data, std::shared_ptr<PFSampleSet> set);
  void setMapFactors(double off_map_factor, double non_free_space_factor, double non_free_space_radius);
  void setPointCloudScannerToFootprintTF(geometry_msgs::Transform tf_msg);
  int getMaxBeams();
  double applyGompertz(double p);

private:
  std::shared_ptr<OctoMap> map_;
  PointCloudModelType model_type_;
  double gompertz_a_, gompertz_b_, gompertz_c_, input_shift_, input_scale_, output_shift_;
  double off_map_factor_, non_free_space_factor_, non_free_space_radius_;
  double z_hit_, z_rand_, sigma_hit_;
  int max_beams_;
  geometry_msgs::Transform point_cloud_scanner_to_footprint_tf_;
};
}  // namespace badger_amcl
#endif
