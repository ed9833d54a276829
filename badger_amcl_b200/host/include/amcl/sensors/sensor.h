// Host mirror of include/amcl/sensors/sensor.h:33-56 of the reference.
#ifndef BAMCL_AMCL_SENSORS_SENSOR_H
#define BAMCL_AMCL_SENSORS_SENSOR_H
#include <memory>

#include "amcl/pf/particle_filter.h"

namespace badger_amcl
{
class SensorData
{
public:
  virtual ~SensorData() = default;
};

class Sensor
{
public:
  virtual ~Sensor() = default;
  virtual bool updateAction(std::shared_ptr<ParticleFilter>, std::shared_ptr<SensorData>) { return false; }
  virtual bool updateSensor(std::shared_ptr<ParticleFilter>, std::shared_ptr<SensorData>) { return false; }
};
}  // namespace badger_amcl
#endif
