// Host mirror of include/amcl/sensors/odom.h:33-86 of the reference: same types, same method names and argument
// meaning. updateAction runs on the device-resident sample set (bamcl_pf_update_action): the poses never visit the
// host, and libc's drand48 generator is advanced to the state the reference's loop would have left it in.
#ifndef BAMCL_AMCL_SENSORS_ODOM_H
#define BAMCL_AMCL_SENSORS_ODOM_H
#include <memory>

#include "amcl/compat.h"
#include "amcl/pf/particle_filter.h"
#include "amcl/sensors/sensor.h"

namespace badger_amcl
{
enum OdomModelType
{
  ODOM_MODEL_DIFF,
  ODOM_MODEL_OMNI,
  ODOM_MODEL_DIFF_CORRECTED,
  ODOM_MODEL_OMNI_CORRECTED,
  ODOM_MODEL_GAUSSIAN,
};

class OdomData : public SensorData
{
public:
  Eigen::Vector3d pose;             // odometric pose
  Eigen::Vector3d delta;            // change in odometric pose
  Eigen::Vector3d absolute_motion;  // accumulated absolute motion (gaussian model)
};

class Odom : public Sensor
{
public:
  Odom() = default;
  void setModelDiff(double alpha1, double alpha2, double alpha3, double alpha4);
  void setModelOmni(double alpha1, double alpha2, double alpha3, double alpha4, double alpha5);
  void setModelGaussian(double alpha1, double alpha2, double alpha3, double alpha4, double alpha5);
  void setModel(OdomModelType type, double alpha1, double alpha2, double alpha3, double alpha4, double alpha5 = 0);
  bool updateAction(std::shared_ptr<ParticleFilter> pf, std::shared_ptr<SensorData> data) override;

private:
  OdomModelType model_type_ = ODOM_MODEL_DIFF;
  double alpha_[5] = {0, 0, 0, 0, 0};
};
}  // namespace badger_amcl
#endif
