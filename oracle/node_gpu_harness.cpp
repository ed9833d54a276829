// oracle/node_gpu_harness.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// The drop-in claim of INTEGRATION.md made executable: the reference's ROS node sources (node/node.cpp, node_2d.cpp,
// node_3d.cpp), UNMODIFIED, compiled against the host mirror's headers (badger_amcl_b200/host/include/amcl in front of the
// reference's include directory, so that "pf/particle_filter.h", "map/occupancy_map.h", "sensors/planar_scanner.h" ... resolve
// to the mirror and only "node/*.h" to the reference) and linked with the mirror's sources + libbamcl_cuda.so instead of the
// reference's pf/, map/ and sensors/ translation units. roscpp / tf2 / pcl / octomap / Eigen are the same look-alikes under
// oracle/shim that the all-reference build uses (the mirror is built with BAMCL_HAVE_EIGEN / _PCL / _ROS / _OCTOMAP, i.e.
// "the real third-party headers are present"). The extern "C" surface is node_harness.inc, shared with ref_harness.cpp, so
// the same Python driver (oracle.RefNode) runs the same scenario through both and tests compare what the node publishes.
#include <array>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <vector>

#include <badger_amcl/AMCLConfig.h>
#include <dynamic_reconfigure/server.h>
#include <message_filters/subscriber.h>
#include <octomap_msgs/conversions.h>
#include <pcl_conversions/pcl_conversions.h>
#include <ros/shim_msgs.h>
#include <tf2/utils.h>
#include <tf2_geometry_msgs/tf2_geometry_msgs.h>
#include <tf2_ros/message_filter.h>
#include <tf2_ros/transform_broadcaster.h>
#include <tf2_ros/transform_listener.h>
#include <yaml-cpp/yaml.h>

// the harness reads a few members the node keeps private; access specifiers do not change the object layout with GCC
#define private public
#define protected public
#include "map/occupancy_map.h"
#include "map/octomap.h"
#include "pf/particle_filter.h"
#include "pf/pf_kdtree.h"
#include "sensors/odom.h"
#include "sensors/planar_scanner.h"
#include "sensors/point_cloud_scanner.h"
#include "node/node.h"
#include "node/node_2d.h"
#include "node/node_3d.h"
#undef private
#undef protected

using namespace badger_amcl;

#define NODE_HARNESS_MIRROR 1
extern "C" {
void ref_srand48(long seed) { srand48(seed); }
const char* ref_node_last_device_error() { return badger_amcl::lastDeviceError(); }
#include "node_harness.inc"
// install a distance cache computed elsewhere (the reference's brushfire from the golden run) into the node's map, so that
// the filter underneath scores against the same field as the reference did (the device builds the exact EDT)
int ref_node_install_cache_2d(void* h, const float* dist, long n, double max_dist)
{
  Node2D* nd = node2d(h);
  if (!nd || !nd->map_)
    return -1;
  nd->map_->setDistancesLUT(std::vector<float>(dist, dist + n), max_dist);
  return 0;
}
int ref_node_install_cache_3d(void* h, const uint8_t* dense, long n)
{
  Node3D* nd = node3d(h);
  if (!nd || !nd->map_)
    return -1;
  nd->map_->setDistancesLUT(std::vector<uint8_t>(dense, dense + n));
  return 0;
}
}  // extern "C"
