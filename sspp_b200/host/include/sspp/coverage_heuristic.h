// CoverageHeuristic — the coverage-driven heuristic the planner API names but whose source is not in
// the reference snapshot (only the interface include/sspp/heuristic_interface.h:35-57,
// the per-node coverage payload include/sspp/node.h:51-63, the sensor model include/sspp/sensors.h:57-68
// and the commented-out call site src/distance_heuristic_test.cpp:79-86 survive).  AUTHORED here on top
// of the engine's visibility pass (DESIGN.md "visibility spec v1"):
//   * a way-point sees the model points visible from any of its sensor poses;
//   * a tree node covers what its parent covered plus what its way-point sees; coverage is the covered
//     share of the model in percent;
//   * reward f = parent's f + (coverage gained) / (distance travelled), as the surface-coverage variant
//     of the original planner rewards a step (coverage gained alone when the step has zero length);
//   * the search ends when coverage is within `tolerance` of the target;
//   * edges are the swept-robot-box test of DistanceHeuristic against the same occupancy map.
// Visible sets depend on the way-point and the map only, so they are computed once per way-point and
// map generation (GpuWorld::mapGeneration), for all children of an expansion in one batched call.
#pragma once
#include <unordered_map>
#include "distance_heuristic.h"
#include "sensors.h"

namespace SSPP {
class CoverageHeuristic : public DistanceHeuristic {
 public:
  CoverageHeuristic(GpuWorld* world, bool debug = false);
  void setModelPoints(const float* xyz, int64_t n);
  void setFrustum(double hfov_deg, double vfov_deg, double near_d, double far_d);
  void setFrustum(const Sensors& s) { setFrustum(s.hFOV, s.vFOV, s.near, s.far); }
  void setCoverageTarget(double pct) { coverageTarget = pct; }
  void setCoverageTolerance(double pct) { coverageTolerance = pct; }
  bool isCost() override { return false; }  // a reward: the open list is kept in descending f
  void calculateHeuristic(Node* n) override;
  void calculateHeuristicAll(std::vector<Node*>& nodes) override;
  bool terminateConditionReached(Node* node) override;
  // sample filter: the robot box is free there AND at least one model point is visible
  bool isFilteringConditionSatisfied(geometry_msgs::Pose pose, geometry_msgs::PoseArray& correspondingSensorPoses, double minDist,
                                     double maxDist, pcl::PointCloud<pcl::PointXYZ>& globalCloud,
                                     std::vector<pcl::PointCloud<pcl::PointXYZ> >& accuracyClusters,
                                     double accuracyThreshhold) override;
  // share of the model not yet covered by the union of `waypoints`' visible sets
  double uncoveredShare(const std::vector<const SearchSpaceNode*>& waypoints);
  // batched sample filter: validity of every sample (one call), visibility of the valid ones' sensor poses
  // (one call); kept samples add what they see to the running coverage of the sampling loop
  bool filterAllPoses(const std::vector<geometry_msgs::Pose>& poses, const std::vector<geometry_msgs::PoseArray>& sensorPoses,
                      std::vector<uint8_t>& valid) override;
  // uncovered share of the model after the samples kept so far; diff = the uncovered model points
  double pointCloudDiff(pcl::PointCloud<pcl::PointXYZ>::Ptr globalCloudPtr, pcl::PointCloud<pcl::PointXYZ>::Ptr& pointCloudDiffPtr) override;
  // Euclidean clusters (connected components of the radius graph at `clusterTolerance`) of at least minClusterSize points
  void clusteringPointCloud(std::vector<pcl::PointCloud<pcl::PointXYZ> >& clusters, pcl::PointCloud<pcl::PointXYZ>::Ptr diff) override;
  // sampling box of a cluster: its bounding box grown by `standoff` on every side
  void findClusterBB(pcl::PointCloud<pcl::PointXYZ> cluster, geometry_msgs::Vector3& gridSize, geometry_msgs::Pose& gridStart) override;
  void setClustering(double tolerance, int minSize) { clusterTolerance = tolerance; minClusterSize = minSize; }
  void setStandoff(double d) { standoff = d; }
  void resetCoverage() { std::fill(covered_.begin(), covered_.end(), 0); coveredCount = 0; }
  int64_t coveredCount = 0;   // model points seen by the samples kept so far
  std::vector<int32_t> clusterLabels(const float* xyz, int64_t n);  // component label per point (-1 = dropped as too small)
  const std::vector<int32_t>& visibleFrom(const SearchSpaceNode* wp);
  int64_t raysCast = 0;        // occlusion rays cast so far
  int64_t visibilityCalls = 0; // batched engine calls so far

 private:
  GpuWorld* world_;
  int64_t modelSize = 0;
  double hfov = 58.0, vfov = 45.0, nearD = 0.7, farD = 6.0;
  double coverageTarget = 100.0, coverageTolerance = 0.0;
  // way-point (by its never-reused uid, not its address) -> sorted visible indices; valid for the map
  // generation seenGen_ only: any change of the occupancy map drops the cache and the running coverage
  std::unordered_map<uint64_t, std::vector<int32_t> > seen_;
  uint64_t seenGen_ = 0;
  void dropStale();
  std::vector<float> model_;        // host copy of the model points (pointCloudDiff hands the uncovered ones back)
  std::vector<uint8_t> covered_;    // per model point: seen by a kept sample
  double clusterTolerance = 0.6, standoff = 3.0;
  int minClusterSize = 10;
  void ensureVisible(const std::vector<const SearchSpaceNode*>& wps);
  void score(Node* n);
};
}  // namespace SSPP
