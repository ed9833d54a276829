// Scan ingestion: the on-wire -> hot-path conversions of Node2D::updateLatestScanData
// (src/amcl/node/node_2d.cpp:534-560) and Node3D::updateLatestScanData (src/amcl/node/node_3d.cpp:467-480).
// A scan is ~1e3 beams / ~1e4 points: host arithmetic, restated with the reference's types (float message fields, double
// scan data) so that the staged scan is bit-identical. Map ingestion (convertMap) is in map2d.cu, on the device.
#include <algorithm>

#include "common.cuh"

using namespace bamcl;

extern "C" int bamcl_scan_from_laserscan(const float* msg_ranges, int range_count, float msg_range_min, float msg_range_max,
                                         double sensor_min_range, double sensor_max_range, double angle_min,
                                         double angle_increment, double* ranges_out, double* angles_out,
                                         double* range_max_out)
{
  if (range_count < 0 || (range_count > 0 && (!msg_ranges || !ranges_out || !angles_out)) || !range_max_out)
    return set_error(BAMCL_E_ARG, "bamcl_scan_from_laserscan: bad arguments");
  // :538-546: thresholds are applied in float, the results widen to double
  double range_max, range_min;
  if (sensor_max_range > 0.0)
    range_max = std::min(msg_range_max, static_cast<float>(sensor_max_range));
  else
    range_max = msg_range_max;
  if (sensor_min_range > 0.0)
    range_min = std::max(msg_range_min, static_cast<float>(sensor_min_range));
  else
    range_min = msg_range_min;
  for (int i = 0; i < range_count; i++)
  {
    ranges_out[i] = msg_ranges[i] <= range_min ? range_max : msg_ranges[i];  // :551-556 short readings -> max range
    angles_out[i] = angle_min + (i * angle_increment);                       // :558
  }
  *range_max_out = range_max;
  return BAMCL_OK;
}

extern "C" int bamcl_cloud_subsample(const float* xyz, long data_count, int max_beams, float* xyz_out, long capacity,
                                     long* n_out)
{
  if (data_count < 0 || (data_count > 0 && (!xyz || !xyz_out)) || !n_out)
    return set_error(BAMCL_E_ARG, "bamcl_cloud_subsample: bad arguments");
  if (max_beams < 2)
    return set_error(BAMCL_E_ARG, "bamcl_cloud_subsample: max_beams < 2 divides by zero in the reference (node_3d.cpp:471)");
  if (data_count > 0x7fffffffl)
    return set_error(BAMCL_E_ARG, "bamcl_cloud_subsample: the reference holds the point count in an int");
  const int count = (int)data_count;
  int step = (count - 1) / (max_beams - 1);  // :471-472
  step = std::max(step, 1);
  long n = 0;
  for (int i = 0; i < count; i += step)  // :474-478
  {
    if (n >= capacity)
      return set_error(BAMCL_E_ARG, "bamcl_cloud_subsample: output capacity %ld too small", capacity);
    xyz_out[3 * n] = xyz[3 * (long)i];
    xyz_out[3 * n + 1] = xyz[3 * (long)i + 1];
    xyz_out[3 * n + 2] = xyz[3 * (long)i + 2];
    n++;
  }
  *n_out = n;
  return BAMCL_OK;
}
