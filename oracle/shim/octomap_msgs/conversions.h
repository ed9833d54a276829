// oracle/shim: TEST INFRASTRUCTURE ONLY. octomap_msgs::binaryMsgToMap / fullMsgToMap: the message shim carries the decoded
// occupied leaf keys (ros/shim_msgs.h), the tree shim is a flat list of them (octomap/OcTree.h).
#pragma once
#include <octomap/OcTree.h>
#include <ros/shim_msgs.h>
namespace octomap_msgs
{
inline octomap::AbstractOcTree* fullMsgToMap(const Octomap& msg)
{
  octomap::OcTree* tree = new octomap::OcTree(msg.resolution);
  for (std::size_t k = 0; k + 2 < msg.occupied_keys.size(); k += 3)
    tree->addLeaf(msg.occupied_keys[k], msg.occupied_keys[k + 1], msg.occupied_keys[k + 2], true);
  return tree;
}
inline octomap::AbstractOcTree* binaryMsgToMap(const Octomap& msg) { return fullMsgToMap(msg); }
}  // namespace octomap_msgs
