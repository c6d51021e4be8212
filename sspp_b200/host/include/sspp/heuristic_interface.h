// Heuristic — the plug-in interface of the planner, THE drop-in boundary (reference
// include/sspp/heuristic_interface.h:35-57): same eleven virtuals, same names, same argument meaning.
// Added, with defaults so existing heuristics keep compiling: the batched forms the GPU path uses
// (one call per connectNodes() / per expansion instead of one per pair / per child).
#pragma once
#include <cstdint>
#include <vector>
#include "node.h"
#include "pointcloud.h"
#include "searchspacenode.h"
#include "ssppexception.h"
#include "utils.h"

namespace SSPP {
class Heuristic {
 public:
  Heuristic() {}
  virtual ~Heuristic() {}
  virtual void calculateHeuristic(Node*) = 0;
  virtual bool terminateConditionReached(Node*) = 0;
  virtual bool isConnectionConditionSatisfied(SearchSpaceNode*, SearchSpaceNode*) = 0;
  virtual bool isConnectionConditionSatisfied(geometry_msgs::Pose, geometry_msgs::Pose) = 0;
  virtual bool isFilteringConditionSatisfied(geometry_msgs::Pose pose, geometry_msgs::PoseArray& correspondingSensorPoses,
                                             double minDist, double maxDist, pcl::PointCloud<pcl::PointXYZ>& globalCloud,
                                             std::vector<pcl::PointCloud<pcl::PointXYZ> >& accuracyClusters,
                                             double accuracyThreshhold) = 0;
  virtual void displayProgress(std::vector<Tree> tree) = 0;
  virtual double pointCloudDiff(pcl::PointCloud<pcl::PointXYZ>::Ptr globalCloudPtr,
                                pcl::PointCloud<pcl::PointXYZ>::Ptr& pointCloudDiffPtr) = 0;
  virtual void clusteringPointCloud(std::vector<pcl::PointCloud<pcl::PointXYZ> >& clustersPointCloudVec,
                                    pcl::PointCloud<pcl::PointXYZ>::Ptr pointCloudDiffPtr) = 0;
  virtual void findClusterOuterPoints(geometry_msgs::PoseArray waypoints, pcl::PointCloud<pcl::PointXYZ>& cloudHull) = 0;
  virtual void findClusterBB(pcl::PointCloud<pcl::PointXYZ> clusterPoints, geometry_msgs::Vector3& gridSize,
                             geometry_msgs::Pose& gridStart) = 0;
  // whether this is a cost (lower is better) or a reward function
  virtual bool isCost() = 0;

  // ---- batched forms (new) -------------------------------------------------------------------
  // Adjacency of the whole list in one call.  Returns false when the heuristic has no batched
  // implementation (the planner then raises: there is no per-pair CPU loop behind connectNodes()).
  // xyz: n x 3 positions in list order; row_ptr: n+1; col: filled with list indices.
  virtual bool connectAll(const double* /*xyz*/, int64_t /*n*/, double /*radius*/, int /*position_mode*/,
                          std::vector<int64_t>& /*row_ptr*/, std::vector<int32_t>& /*col*/) { return false; }
  // isConnectionConditionSatisfied for n segments (positions only) in one call.  Default: the per-pair virtual.
  virtual void connectionConditionAll(const double* starts, const double* ends, int64_t n, std::vector<uint8_t>& ok) {
    ok.assign((size_t)n, 0);
    for (int64_t i = 0; i < n; ++i) {
      geometry_msgs::Pose a, b;
      a.position.x = starts[3 * i]; a.position.y = starts[3 * i + 1]; a.position.z = starts[3 * i + 2];
      b.position.x = ends[3 * i]; b.position.y = ends[3 * i + 1]; b.position.z = ends[3 * i + 2];
      ok[(size_t)i] = isConnectionConditionSatisfied(a, b) ? 1 : 0;
    }
  }
  // calculateHeuristic for all children of one expansion, in the order the search will consume them.
  // Default: one call per node, exactly what the reference's loop does (src/astar.cpp:318).
  virtual void calculateHeuristicAll(std::vector<Node*>& nodes) {
    for (Node* n : nodes) calculateHeuristic(n);
  }
  // isFilteringConditionSatisfied for many samples with their sensor poses in one call; false = not offered.
  virtual bool filterAllPoses(const std::vector<geometry_msgs::Pose>& /*poses*/,
                              const std::vector<geometry_msgs::PoseArray>& /*sensorPoses*/, std::vector<uint8_t>& /*valid*/) {
    return false;
  }
  // Validity of many poses in one call (isFilteringConditionSatisfied batched); false = not offered.
  virtual bool filterAll(const double* /*xyz*/, int64_t /*n*/, std::vector<uint8_t>& /*valid*/) { return false; }
};
}  // namespace SSPP
