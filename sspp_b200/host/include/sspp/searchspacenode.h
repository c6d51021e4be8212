// SearchSpaceNode — one way-point of the search space (reference include/sspp/searchspacenode.h:35-45):
// pose, the poses of its sensors, the singly linked `next`, and the adjacency as raw non-owning
// pointers.  `index` is new: the node's position in list order, maintained by SearchSpace::flatten.
// `uid` is new: a process-wide serial number, never reused, so that caches keyed by way-point (the
// coverage heuristic's visible sets) cannot mistake a node allocated at a freed node's address for it.
#pragma once
#include <atomic>
#include <cstdint>
#include <vector>
#include "geometry_msgs.h"

enum { BridgeNode, RegGridNode };

namespace SSPP {
class SearchSpaceNode {
 public:
  geometry_msgs::Pose location;
  geometry_msgs::PoseArray sensorLocation;
  SearchSpaceNode* next = nullptr;
  int type = RegGridNode, id = 0;
  int index = -1;  // list position after SearchSpace::flatten()
  long id_order_ = 0;  // insertion stamp (the list is ordered newest first)
  uint64_t uid = nextUid();
  static uint64_t nextUid() {
    static std::atomic<uint64_t> counter{1};
    return counter.fetch_add(1, std::memory_order_relaxed);
  }
  std::vector<SearchSpaceNode*> children;
};
}  // namespace SSPP
