// oracle/shim: TEST INFRASTRUCTURE ONLY. Stand-in for octomap::OcTree (third-party, not vendored,
// version unpinned): a flat list of occupied leaf keys at the finest resolution. Leaf centre
// coordinates follow octomap's published keyToCoord: (key + 0.5) * resolution with key relative to
// the tree centre, i.e. centre = (k + 0.5) * res for signed integer k. getMetricMin/Max return the
// bounding box of the leaf volumes (centre -/+ res/2), like octomap's calcMinMax over leaves.
#pragma once
#include <algorithm>
#include <limits>
#include <vector>
namespace octomap
{
struct ShimLeaf
{
  int k[3];
  bool occupied;
};
class AbstractOcTree
{
public:
  virtual ~AbstractOcTree() {}
};
class OcTree : public AbstractOcTree
{
public:
  explicit OcTree(double res) : res_(res) {}
  void addLeaf(int i, int j, int k, bool occ = true) { leaves_.push_back({{i, j, k}, occ}); }
  void expand() {}
  double getResolution() const { return res_; }
  void getMetricMin(double& x, double& y, double& z) const { metric(x, y, z, true); }
  void getMetricMax(double& x, double& y, double& z) const { metric(x, y, z, false); }
  class leaf_iterator
  {
  public:
    leaf_iterator(const OcTree* t, std::size_t i) : t_(t), i_(i) {}
    double getX() const { return (t_->leaves_[i_].k[0] + 0.5) * t_->res_; }
    double getY() const { return (t_->leaves_[i_].k[1] + 0.5) * t_->res_; }
    double getZ() const { return (t_->leaves_[i_].k[2] + 0.5) * t_->res_; }
    const ShimLeaf& operator*() const { return t_->leaves_[i_]; }
    leaf_iterator& operator++() { ++i_; return *this; }
    bool operator!=(const leaf_iterator& o) const { return i_ != o.i_; }
  private:
    const OcTree* t_;
    std::size_t i_;
  };
  leaf_iterator begin_leafs() const { return leaf_iterator(this, 0); }
  leaf_iterator end_leafs() const { return leaf_iterator(this, leaves_.size()); }
  bool isNodeOccupied(const ShimLeaf& l) const { return l.occupied; }
private:
  void metric(double& x, double& y, double& z, bool mn) const
  {
    double v[3];
    for (int a = 0; a < 3; a++)
    {
      v[a] = mn ? std::numeric_limits<double>::max() : -std::numeric_limits<double>::max();
      for (const ShimLeaf& l : leaves_)
      {
        // octomap's calcMinMax: x = centre - size/2 feeds min; x += size feeds max
        double c = (l.k[a] + 0.5) * res_ - res_ / 2.0;
        if (mn)
          v[a] = std::min(v[a], c);
        else
          v[a] = std::max(v[a], c + res_);
      }
    }
    x = v[0]; y = v[1]; z = v[2];
  }
  double res_;
  std::vector<ShimLeaf> leaves_;
  friend class leaf_iterator;
};
}  // namespace octomap
