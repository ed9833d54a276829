// oracle/shim: TEST INFRASTRUCTURE ONLY. The slice of yaml-cpp's Node interface Node::loadPoseFromFile / savePoseToFile
// use (node.cpp:493-668), enough to compile them. LoadFile always throws BadFile: the saved-pose file is host I/O outside
// the hot path, and a missing file is the reference's own "use the default initial pose" branch (node.cpp:467-478).
#pragma once
#include <memory>
#include <ostream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>
namespace YAML
{
class BadFile : public std::runtime_error
{
public:
  BadFile() : std::runtime_error("yaml-cpp shim: bad file") {}
};
class Node
{
  struct Impl;
public:
  Node() : p_(std::make_shared<Impl>()) {}
  template <typename K>
  Node operator[](const K& key) const
  {
    std::ostringstream ks;
    ks << key;
    for (std::pair<Node, Node>& kv : p_->map)
      if (kv.first.p_->scalar == ks.str())
        return kv.second;
    Node k, v;
    k.p_->scalar = ks.str();
    k.p_->defined = true;
    p_->map.push_back(std::make_pair(k, v));
    p_->defined = true;
    return v;
  }
  template <typename T>
  Node& operator=(const T& value)
  {
    std::ostringstream vs;
    vs.precision(17);
    vs << value;
    p_->scalar = vs.str();
    p_->defined = true;
    return *this;
  }
  Node& operator=(const Node& o)
  {
    *p_ = *o.p_;  // assignment through a handle replaces the content it refers to
    return *this;
  }
  Node(const Node& o) : p_(o.p_) {}
  template <typename T>
  T as() const
  {
    if (!p_->defined)
      throw std::runtime_error("yaml-cpp shim: bad conversion");
    std::istringstream is(p_->scalar);
    T v;
    is >> v;
    return v;
  }
  explicit operator bool() const { return p_->defined; }
  typedef std::vector<std::pair<Node, Node>>::const_iterator const_iterator;
  const_iterator begin() const { return p_->map.begin(); }
  const_iterator end() const { return p_->map.end(); }
  void write(std::ostream& os, int indent) const
  {
    if (p_->map.empty())
    {
      os << p_->scalar << "\n";
      return;
    }
    if (indent)
      os << "\n";
    for (const std::pair<Node, Node>& kv : p_->map)
    {
      os << std::string(indent, ' ') << kv.first.p_->scalar << ": ";
      kv.second.write(os, indent + 2);
    }
  }
private:
  struct Impl
  {
    std::string scalar;
    bool defined = false;
    std::vector<std::pair<Node, Node>> map;
  };
  std::shared_ptr<Impl> p_;
};
template <>
inline std::string Node::as<std::string>() const
{
  if (!p_->defined)
    throw std::runtime_error("yaml-cpp shim: bad conversion");
  return p_->scalar;
}
inline std::ostream& operator<<(std::ostream& os, const Node& n)
{
  n.write(os, 0);
  return os;
}
inline Node LoadFile(const std::string&) { throw BadFile(); }
}  // namespace YAML
