// placeholder: replaced by the KLD resampler
#include "pf.cuh"
using namespace bamcl;
extern "C" {
int bamcl_pf_cdf(bamcl_pf*, double*, long) { return set_error(BAMCL_E_UNSUPPORTED, "not built yet"); }
int bamcl_pf_resample(bamcl_pf*, const bamcl_resample_params*, bamcl_resample_result*) { return set_error(BAMCL_E_UNSUPPORTED, "not built yet"); }
int bamcl_pf_init_stats(bamcl_pf*, bamcl_resample_result*) { return set_error(BAMCL_E_UNSUPPORTED, "not built yet"); }
int bamcl_pf_resample_sources(bamcl_pf*, int*, long) { return set_error(BAMCL_E_UNSUPPORTED, "not built yet"); }
int bamcl_pf_sample_clusters(bamcl_pf*, int*, long) { return set_error(BAMCL_E_UNSUPPORTED, "not built yet"); }
int bamcl_pf_cluster_stats(bamcl_pf*, double*, int) { return set_error(BAMCL_E_UNSUPPORTED, "not built yet"); }
int bamcl_resample_limit(int, int, int, double, double) { return -1; }
}
