// Host mirror of src/amcl/sensors/odom.cpp: the class keeps the model parameters, the motion update itself is
// bamcl_pf_update_action on the device-resident set.
#include "amcl/sensors/odom.h"

#include <stdlib.h>

#include <stdexcept>
#include <string>

#include "amcl/compat.h"
#include "bamcl.h"

namespace badger_amcl
{
void Odom::setModel(OdomModelType type, double alpha1, double alpha2, double alpha3, double alpha4, double alpha5)
{
  model_type_ = type;  // odom.cpp:63-72
  alpha_[0] = alpha1;
  alpha_[1] = alpha2;
  alpha_[2] = alpha3;
  alpha_[3] = alpha4;
  alpha_[4] = alpha5;
}

// odom.cpp:34-61: the named setters leave the parameters they do not take as they were
void Odom::setModelDiff(double alpha1, double alpha2, double alpha3, double alpha4)
{
  setModel(ODOM_MODEL_DIFF, alpha1, alpha2, alpha3, alpha4, alpha_[4]);
}
void Odom::setModelOmni(double alpha1, double alpha2, double alpha3, double alpha4, double alpha5)
{
  setModel(ODOM_MODEL_OMNI, alpha1, alpha2, alpha3, alpha4, alpha5);
}
void Odom::setModelGaussian(double alpha1, double alpha2, double alpha3, double alpha4, double alpha5)
{
  setModel(ODOM_MODEL_GAUSSIAN, alpha1, alpha2, alpha3, alpha4, alpha5);
}

bool Odom::updateAction(std::shared_ptr<ParticleFilter> pf, std::shared_ptr<SensorData> data)
{
  std::shared_ptr<OdomData> ndata = std::dynamic_pointer_cast<OdomData>(data);
  if (!pf || !ndata)
    return false;
  pf->pushIfHostDirty();  // poses a caller edited through getCurrentSet() go to the device first
  bamcl_odom_model model;
  model.model_type = static_cast<int>(model_type_);
  bamcl_odom_data od;
  for (int i = 0; i < 5; i++)
    model.alpha[i] = alpha_[i];
  for (int i = 0; i < 3; i++)
  {
    od.pose[i] = ndata->pose[i];
    od.delta[i] = ndata->delta[i];
    od.absolute_motion[i] = ndata->absolute_motion[i];
  }
  // drand48 contract: read libc's generator state, let the device replay the draws, move libc's generator on
  unsigned short zero[3] = {0, 0, 0};
  unsigned short* prev = seed48(zero);
  unsigned short saved[3] = {prev[0], prev[1], prev[2]};
  seed48(saved);
  const uint64_t state = (uint64_t)saved[0] | ((uint64_t)saved[1] << 16) | ((uint64_t)saved[2] << 32);
  uint64_t next = state;
  if (!deviceOk(bamcl_pf_update_action(pf->deviceHandle(), &model, &od, state, &next), "bamcl_pf_update_action"))
    return false;  // nothing moved, libc's generator untouched (lastDeviceError() says why)
  unsigned short st[3] = {(unsigned short)(next & 0xFFFF), (unsigned short)((next >> 16) & 0xFFFF),
                          (unsigned short)((next >> 32) & 0xFFFF)};
  seed48(st);
  pf->markDeviceModified();
  return true;
}
}  // namespace badger_amcl
