// CoverageHeuristic (authored; see the header for what survives of it in the reference).
#include "sspp/coverage_heuristic.h"

#include <algorithm>

namespace SSPP {
CoverageHeuristic::CoverageHeuristic(GpuWorld* world, bool debug) : DistanceHeuristic(debug, world), world_(world) {
  if (!world) throw SSPPException("CoverageHeuristic needs a GpuWorld (visibility runs on the engine)");
}

void CoverageHeuristic::setModelPoints(const float* xyz, int64_t n) {
  world_->setModelPoints(xyz, n);
  modelSize = n;
  seen_.clear();
}

void CoverageHeuristic::setFrustum(double h, double v, double n, double f) {
  hfov = h; vfov = v; nearD = n; farD = f;
  seen_.clear();
}

// visible sets of the way-points not seen yet: all their sensor poses in ONE engine call
void CoverageHeuristic::ensureVisible(const std::vector<const SearchSpaceNode*>& wps) {
  std::vector<const SearchSpaceNode*> todo;
  for (const SearchSpaceNode* w : wps)
    if (w && !seen_.count(w) && std::find(todo.begin(), todo.end(), w) == todo.end()) todo.push_back(w);
  if (todo.empty()) return;
  std::vector<double> poses;
  std::vector<size_t> first(todo.size() + 1, 0);
  for (size_t i = 0; i < todo.size(); ++i) {
    for (const geometry_msgs::Pose& p : todo[i]->sensorLocation.poses) {
      const double v[7] = {p.position.x, p.position.y, p.position.z, p.orientation.x, p.orientation.y, p.orientation.z, p.orientation.w};
      poses.insert(poses.end(), v, v + 7);
    }
    first[i + 1] = poses.size() / 7;
  }
  std::vector<int64_t> view_ptr;
  std::vector<int32_t> idx;
  int64_t rays = 0;
  world_->visibility(poses.data(), (int64_t)(poses.size() / 7), hfov, vfov, nearD, farD, view_ptr, idx, &rays);
  raysCast += rays;
  ++visibilityCalls;
  for (size_t i = 0; i < todo.size(); ++i) {
    std::vector<int32_t> u;
    for (size_t s = first[i]; s < first[i + 1]; ++s) {  // union over the way-point's sensors (each list is sorted)
      std::vector<int32_t> m;
      std::set_union(u.begin(), u.end(), idx.begin() + view_ptr[s], idx.begin() + view_ptr[s + 1], std::back_inserter(m));
      u.swap(m);
    }
    seen_[todo[i]].swap(u);
  }
}

const std::vector<int32_t>& CoverageHeuristic::visibleFrom(const SearchSpaceNode* wp) {
  ensureVisible(std::vector<const SearchSpaceNode*>(1, wp));
  return seen_[wp];
}

void CoverageHeuristic::score(Node* n) {
  const std::vector<int32_t>& mine = seen_[n->ssnode];
  n->coveredPoints.clear();
  if (n->parent)
    std::set_union(n->parent->coveredPoints.begin(), n->parent->coveredPoints.end(), mine.begin(), mine.end(),
                   std::back_inserter(n->coveredPoints));
  else
    n->coveredPoints = mine;
  n->coveredVoxelsNum = (int)n->coveredPoints.size();
  n->coverage = modelSize ? 100.0 * (double)n->coveredPoints.size() / (double)modelSize : 0.0;
  if (n->parent) {
    const double d = Dist(n->pose.p, n->parent->pose.p);
    const double c = n->coverage - n->parent->coverage;
    n->distance = n->parent->distance + d;
    n->g_value = n->distance;
    n->f_value = (d != 0.0) ? n->parent->f_value + (1.0 / d) * c : n->parent->f_value + c;
  } else {
    n->distance = 0;
    n->g_value = 0;
    n->f_value = n->coverage;
  }
  n->h_value = n->coverage;
}

void CoverageHeuristic::calculateHeuristic(Node* n) {
  if (!n) return;
  ensureVisible(std::vector<const SearchSpaceNode*>(1, n->ssnode));
  score(n);
}

void CoverageHeuristic::calculateHeuristicAll(std::vector<Node*>& nodes) {
  std::vector<const SearchSpaceNode*> wps;
  for (Node* n : nodes) wps.push_back(n->ssnode);
  ensureVisible(wps);
  for (Node* n : nodes) score(n);
}

bool CoverageHeuristic::terminateConditionReached(Node* node) { return coverageTarget - node->coverage <= coverageTolerance; }

bool CoverageHeuristic::isFilteringConditionSatisfied(geometry_msgs::Pose pose, geometry_msgs::PoseArray& sensorPoses, double, double,
                                                      pcl::PointCloud<pcl::PointXYZ>&, std::vector<pcl::PointCloud<pcl::PointXYZ> >&,
                                                      double) {
  const double p[3] = {pose.position.x, pose.position.y, pose.position.z};
  const double box[3] = {0.5, 0.5, 0.3};
  if (world_->hasMap() && world_->getCellStatusBoundingBox(p, box) != GpuWorld::kFree) return false;
  std::vector<double> poses;
  for (const geometry_msgs::Pose& s : sensorPoses.poses) {
    const double v[7] = {s.position.x, s.position.y, s.position.z, s.orientation.x, s.orientation.y, s.orientation.z, s.orientation.w};
    poses.insert(poses.end(), v, v + 7);
  }
  if (poses.empty()) return false;
  std::vector<int64_t> view_ptr;
  std::vector<int32_t> idx;
  int64_t rays = 0;
  world_->visibility(poses.data(), (int64_t)(poses.size() / 7), hfov, vfov, nearD, farD, view_ptr, idx, &rays);
  raysCast += rays;
  ++visibilityCalls;
  return !idx.empty();
}

double CoverageHeuristic::uncoveredShare(const std::vector<const SearchSpaceNode*>& waypoints) {
  ensureVisible(waypoints);
  std::vector<int32_t> u;
  for (const SearchSpaceNode* w : waypoints) {
    std::vector<int32_t> m;
    const std::vector<int32_t>& s = seen_[w];
    std::set_union(u.begin(), u.end(), s.begin(), s.end(), std::back_inserter(m));
    u.swap(m);
  }
  return modelSize ? 1.0 - (double)u.size() / (double)modelSize : 0.0;
}
}  // namespace SSPP
