// Multinomial KLD resampling (particle_filter.cpp:356-420) -- placeholder until the device version lands.
#include "pf.cuh"
namespace bamcl
{
int resample_multinomial(bamcl_pf*, const bamcl_resample_params*, bamcl_resample_result*)
{
  return set_error(BAMCL_E_UNSUPPORTED, "multinomial resampling is not implemented on the device yet");
}
}  // namespace bamcl
