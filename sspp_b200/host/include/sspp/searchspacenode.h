// SearchSpaceNode — one way-point of the search space (reference include/sspp/searchspacenode.h:35-45):
// pose, the poses of its sensors, the singly linked `next`, and the adjacency as raw non-owning
// pointers.  `index` is new: the node's position in list order, maintained by SearchSpace::flatten.
#pragma once
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
  std::vector<SearchSpaceNode*> children;
};
}  // namespace SSPP
