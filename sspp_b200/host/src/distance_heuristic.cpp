// DistanceHeuristic (reference src/distance_heuristic.cpp).
#include "sspp/distance_heuristic.h"

namespace SSPP {
DistanceHeuristic::DistanceHeuristic(bool d, GpuWorld* w) : debug(d), world(w) {}

// g = parent's g + the step length; the root costs nothing (:41-56)
double DistanceHeuristic::gCost(Node* node) {
  if (!node || !node->parent) return 0.0;
  return node->parent->g_value + Dist(node->pose.p, node->parent->pose.p);
}
// h = straight-line distance to the goal (:58-65)
double DistanceHeuristic::hCost(Node* node) {
  if (!node) return 0;
  return Dist(endPose, node->pose.p);
}
void DistanceHeuristic::calculateHeuristic(Node* node) {
  if (!node) return;
  node->g_value = gCost(node);
  node->h_value = hCost(node);
  node->f_value = node->g_value + node->h_value;  // in this order, in double, unfused (:209-216)
}
bool DistanceHeuristic::terminateConditionReached(Node* node) { return Dist(node->pose.p, endPose) <= tolerance2Goal; }

// the per-pair form: one segment through the batched entry point (:99-127)
bool DistanceHeuristic::isConnectionConditionSatisfied(geometry_msgs::Pose temp, geometry_msgs::Pose S) {
  if (!world) return true;
  const double a[3] = {temp.position.x, temp.position.y, temp.position.z};
  const double b[3] = {S.position.x, S.position.y, S.position.z};
  return world->getLineStatusBoundingBox(a, b, box) == GpuWorld::kFree;
}
bool DistanceHeuristic::isConnectionConditionSatisfied(SearchSpaceNode* temp, SearchSpaceNode* S) {
  return isConnectionConditionSatisfied(temp->location, S->location);
}

// the reference leaves this hook as `return true` (:134-138); with setFilterByValidity(true) it
// becomes the node-validity query (DESIGN.md validity spec)
bool DistanceHeuristic::isFilteringConditionSatisfied(geometry_msgs::Pose pose, geometry_msgs::PoseArray&, double, double,
                                                      pcl::PointCloud<pcl::PointXYZ>&,
                                                      std::vector<pcl::PointCloud<pcl::PointXYZ> >&, double) {
  if (!filterByValidity || !world) return true;
  const double p[3] = {pose.position.x, pose.position.y, pose.position.z};
  return world->getCellStatusBoundingBox(p, box) == GpuWorld::kFree;
}

bool DistanceHeuristic::connectAll(const double* xyz, int64_t n, double radius, int position_mode, std::vector<int64_t>& row_ptr,
                                   std::vector<int32_t>& col) {
  if (!world) throw SSPPException("DistanceHeuristic::connectAll needs a GpuWorld (a map-free graph still uses its radius search)");
  GpuWorld::ConnectResult r = world->connect(xyz, n, radius, box, world->hasMap() && useMap, position_mode, row_ptr, col);
  lastTestedPairs = r.tested_pairs;
  return true;
}

void DistanceHeuristic::connectionConditionAll(const double* starts, const double* ends, int64_t n, std::vector<uint8_t>& ok) {
  ok.assign((size_t)n, 1);
  if (!world || !world->hasMap() || !useMap || n == 0) return;  // manager == NULL: every pair passes
  std::vector<uint8_t> st((size_t)n);
  world->checkSegments(starts, ends, n, box, st.data());
  for (int64_t i = 0; i < n; ++i) ok[(size_t)i] = st[(size_t)i] == GpuWorld::kFree;
}

bool DistanceHeuristic::filterAll(const double* xyz, int64_t n, std::vector<uint8_t>& valid) {
  valid.assign((size_t)n, 1);
  if (!filterByValidity || !world || n == 0) return true;
  std::vector<uint8_t> st((size_t)n);
  world->validity(xyz, n, box, st.data());
  for (int64_t i = 0; i < n; ++i) valid[(size_t)i] = st[(size_t)i] == GpuWorld::kFree;
  return true;
}
}  // namespace SSPP
