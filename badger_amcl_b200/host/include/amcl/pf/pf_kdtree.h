// Host mirror of include/amcl/pf/pf_kdtree.h:33-69 of the reference. The tree itself is never materialised on
// the host: poses are buffered and the observable results (leaf count, node count, cluster labels) are computed
// by the device kernels (csrc/kdtree.cu) the first time they are asked for.
#ifndef BAMCL_AMCL_PF_KDTREE_H
#define BAMCL_AMCL_PF_KDTREE_H
#include <functional>
#include <map>
#include <tuple>
#include <vector>

#include "amcl/compat.h"

namespace badger_amcl
{
class ParticleFilter;

class PFKDTree
{
public:
  PFKDTree();
  void clearKDTree();
  void insertPose(const Eigen::Vector3d& pose, double value);  // pf_kdtree.cpp:49-56
  void cluster();                                              // :58-76
  int getCluster(const Eigen::Vector3d& pose);                 // :79-90 (-1 if the pose's bin is not in the tree)
  int getLeafCount();                                          // :92

  int getNodeCount();  // number of occupied pose bins (nodes_.size() in the reference)

private:
  friend class ParticleFilter;
  void flush();
  // results installed by ParticleFilter for the tree of a sample set it owns
  void install(int leaf_count, int node_count, const std::vector<Eigen::Vector3d>& poses, const std::vector<int>& labels);
  typedef std::tuple<int, int, int> Key;
  static Key keyOf(const Eigen::Vector3d& pose);
  std::vector<double> pending_;  // x, y, theta, value per inserted pose
  std::function<void()> fetch_labels_;  // set by ParticleFilter: fills labels_ of a device-built tree on demand
  bool flushed_, clustered_;
  int leaf_count_, node_count_;
  std::map<Key, int> labels_;
};
}  // namespace badger_amcl
#endif
