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
  seenGen_ = world_->mapGeneration();
  model_.assign(xyz, xyz + 3 * n);
  covered_.assign((size_t)n, 0);
  coveredCount = 0;
}

void CoverageHeuristic::setFrustum(double h, double v, double n, double f) {
  hfov = h; vfov = v; nearD = n; farD = f;
  seen_.clear();
}

// occlusion answers belong to one state of the map: after setFree / insertPointcloud / a new upload
// the cached visible sets and the coverage accumulated from them are stale
void CoverageHeuristic::dropStale() {
  const uint64_t g = world_->mapGeneration();
  if (g == seenGen_) return;
  seenGen_ = g;
  seen_.clear();
  resetCoverage();
}

// visible sets of the way-points not seen yet: all their sensor poses in ONE engine call
void CoverageHeuristic::ensureVisible(const std::vector<const SearchSpaceNode*>& wps) {
  dropStale();
  std::vector<const SearchSpaceNode*> todo;
  for (const SearchSpaceNode* w : wps)
    if (w && !seen_.count(w->uid) && std::find(todo.begin(), todo.end(), w) == todo.end()) todo.push_back(w);
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
    seen_[todo[i]->uid].swap(u);
  }
}

const std::vector<int32_t>& CoverageHeuristic::visibleFrom(const SearchSpaceNode* wp) {
  ensureVisible(std::vector<const SearchSpaceNode*>(1, wp));
  return seen_[wp->uid];
}

void CoverageHeuristic::score(Node* n) {
  const std::vector<int32_t>& mine = seen_[n->ssnode->uid];
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
    const std::vector<int32_t>& s = seen_[w->uid];
    std::set_union(u.begin(), u.end(), s.begin(), s.end(), std::back_inserter(m));
    u.swap(m);
  }
  return modelSize ? 1.0 - (double)u.size() / (double)modelSize : 0.0;
}
bool CoverageHeuristic::filterAllPoses(const std::vector<geometry_msgs::Pose>& poses, const std::vector<geometry_msgs::PoseArray>& sensorPoses,
                                       std::vector<uint8_t>& valid) {
  const size_t n = poses.size();
  valid.assign(n, 0);
  if (n == 0) return true;
  dropStale();
  std::vector<uint8_t> st(n, GpuWorld::kFree);
  if (world_->hasMap()) {
    std::vector<double> xyz(n * 3);
    for (size_t i = 0; i < n; ++i) { xyz[3 * i] = poses[i].position.x; xyz[3 * i + 1] = poses[i].position.y; xyz[3 * i + 2] = poses[i].position.z; }
    const double box[3] = {0.5, 0.5, 0.3};
    world_->validity(xyz.data(), (int64_t)n, box, st.data());
  }
  std::vector<double> sp;
  std::vector<size_t> owner;
  for (size_t i = 0; i < n; ++i) {
    if (st[i] != GpuWorld::kFree) continue;
    for (const geometry_msgs::Pose& s : sensorPoses[i].poses) {
      const double v[7] = {s.position.x, s.position.y, s.position.z, s.orientation.x, s.orientation.y, s.orientation.z, s.orientation.w};
      sp.insert(sp.end(), v, v + 7);
      owner.push_back(i);
    }
  }
  if (owner.empty()) return true;
  std::vector<int64_t> view_ptr;
  std::vector<int32_t> idx;
  int64_t rays = 0;
  world_->visibility(sp.data(), (int64_t)owner.size(), hfov, vfov, nearD, farD, view_ptr, idx, &rays);
  raysCast += rays;
  ++visibilityCalls;
  for (size_t v = 0; v < owner.size(); ++v) {
    if (view_ptr[v + 1] == view_ptr[v]) continue;
    valid[owner[v]] = 1;
    for (int64_t e = view_ptr[v]; e < view_ptr[v + 1]; ++e)
      if (!covered_[(size_t)idx[(size_t)e]]) { covered_[(size_t)idx[(size_t)e]] = 1; ++coveredCount; }
  }
  return true;
}

double CoverageHeuristic::pointCloudDiff(pcl::PointCloud<pcl::PointXYZ>::Ptr, pcl::PointCloud<pcl::PointXYZ>::Ptr& diff) {
  if (!diff) diff.reset(new pcl::PointCloud<pcl::PointXYZ>);
  dropStale();
  diff->points.clear();
  for (int64_t i = 0; i < modelSize; ++i)
    if (!covered_[(size_t)i]) diff->points.push_back(pcl::PointXYZ(model_[3 * i], model_[3 * i + 1], model_[3 * i + 2]));
  return modelSize ? (double)diff->points.size() / (double)modelSize : 0.0;
}

// connected components of the radius graph, built by the engine's own radius search (no map)
std::vector<int32_t> CoverageHeuristic::clusterLabels(const float* xyz, int64_t n) {
  std::vector<int32_t> label((size_t)n, -1);
  if (n == 0) return label;
  std::vector<double> d((size_t)n * 3);
  for (int64_t i = 0; i < 3 * n; ++i) d[(size_t)i] = xyz[i];
  std::vector<int64_t> rp;
  std::vector<int32_t> col;
  const double box[3] = {0, 0, 0};
  world_->connect(d.data(), n, clusterTolerance, box, false, 0, rp, col);
  std::vector<int32_t> stack;
  int32_t next = 0;
  for (int64_t s = 0; s < n; ++s) {
    if (label[(size_t)s] != -1) continue;
    std::vector<int32_t> members;
    stack.assign(1, (int32_t)s);
    label[(size_t)s] = -2;
    while (!stack.empty()) {
      int32_t u = stack.back();
      stack.pop_back();
      members.push_back(u);
      for (int64_t e = rp[(size_t)u]; e < rp[(size_t)u + 1]; ++e) {
        int32_t v = col[(size_t)e];
        if (label[(size_t)v] == -1) { label[(size_t)v] = -2; stack.push_back(v); }
      }
    }
    const int32_t id = ((int)members.size() >= minClusterSize) ? next++ : -3;
    for (int32_t m : members) label[(size_t)m] = id;
  }
  for (int32_t& l : label) if (l == -3) l = -1;
  return label;
}

void CoverageHeuristic::clusteringPointCloud(std::vector<pcl::PointCloud<pcl::PointXYZ> >& clusters, pcl::PointCloud<pcl::PointXYZ>::Ptr diff) {
  clusters.clear();
  if (!diff || diff->points.empty()) return;
  const int64_t n = (int64_t)diff->points.size();
  std::vector<float> xyz((size_t)n * 3);
  for (int64_t i = 0; i < n; ++i) { xyz[3 * i] = diff->points[(size_t)i].x; xyz[3 * i + 1] = diff->points[(size_t)i].y; xyz[3 * i + 2] = diff->points[(size_t)i].z; }
  std::vector<int32_t> label = clusterLabels(xyz.data(), n);
  int32_t k = -1;
  for (int32_t l : label) k = std::max(k, l);
  clusters.resize((size_t)(k + 1));
  for (int64_t i = 0; i < n; ++i)
    if (label[(size_t)i] >= 0) clusters[(size_t)label[(size_t)i]].push_back(diff->points[(size_t)i]);
}

void CoverageHeuristic::findClusterBB(pcl::PointCloud<pcl::PointXYZ> cluster, geometry_msgs::Vector3& gridSize, geometry_msgs::Pose& gridStart) {
  if (cluster.points.empty()) { gridSize = geometry_msgs::Vector3(); return; }
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (const pcl::PointXYZ& p : cluster.points) {
    const double v[3] = {p.x, p.y, p.z};
    for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], v[a]); hi[a] = std::max(hi[a], v[a]); }
  }
  gridStart.position.x = 0.5 * (lo[0] + hi[0]);
  gridStart.position.y = 0.5 * (lo[1] + hi[1]);
  gridStart.position.z = 0.5 * (lo[2] + hi[2]);
  gridStart.orientation.w = 1;
  gridSize.x = (hi[0] - lo[0]) + 2 * standoff;
  gridSize.y = (hi[1] - lo[1]) + 2 * standoff;
  gridSize.z = (hi[2] - lo[2]) + 2 * standoff;
}
}  // namespace SSPP
