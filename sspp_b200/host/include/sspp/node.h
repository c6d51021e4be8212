// Node — one node of the A* search tree (reference include/sspp/node.h:49-72): pose, sensor poses,
// g/h/f, tree and list links, and the coverage payload a coverage heuristic accumulates along a
// path.  The PCL/CGAL/octomap payload members of the reference (surfaceTriangles, octree, clouds)
// are replaced by the index sets the GPU visibility pass returns.
#pragma once
#include <vector>
#include "utils.h"

namespace SSPP {
class SearchSpaceNode;
class Node {
 public:
  int id = 0, depth = 0, direction = 0, coveredVoxelsNum = 0;
  double nearest_obstacle = 0, g_value = 0, h_value = 0, f_value = 0, totalEntroby = 0, distance = 0, coverage = 0,
         coveredVolume = 0;
  Node *parent = nullptr, *next = nullptr, *prev = nullptr;
  Pose pose;
  std::vector<Pose> senPoses;
  std::vector<int32_t> coveredPoints;      // model-point indices seen so far along the path
  const SearchSpaceNode* ssnode = nullptr; // the way-point this tree node stands on
  Node() {}
  explicit Node(const Node* n) { *this = *n; }
  // pose equality within 1e-12 (reference src/node.cpp:75-83)
  bool operator==(const Node& a) const {
    return isPositionEqual(pose.p.position, a.pose.p.position) && isOrientationEqual(pose.p.orientation, a.pose.p.orientation);
  }
  bool operator!=(const Node& a) const { return !(*this == a); }
};
}  // namespace SSPP
