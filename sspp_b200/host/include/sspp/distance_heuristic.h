// DistanceHeuristic — Euclidean g/h, goal test, and the edge condition "the swept robot box is free
// along the segment" (reference include/sspp/distance_heuristic.h, src/distance_heuristic.cpp).
// The world model is a GpuWorld* where the reference takes an OctomapManager*; a null world keeps the
// reference's behaviour of accepting every in-radius pair (src/distance_heuristic.cpp:101,126).
#pragma once
#include "gpu_world.h"
#include "heuristic_interface.h"

namespace SSPP {
class DistanceHeuristic : public Heuristic {
 public:
  explicit DistanceHeuristic(bool d = false, GpuWorld* world = nullptr);
  ~DistanceHeuristic() {}
  double gCost(Node* node);
  double hCost(Node* node);
  bool isCost() override { return true; }
  void setDebug(bool d) { debug = d; }
  void setEndPose(geometry_msgs::Pose pose) { endPose = pose; }
  void setTolerance2Goal(double t) { tolerance2Goal = t; }
  void setRobotBox(double x, double y, double z) { box[0] = x; box[1] = y; box[2] = z; }
  void calculateHeuristic(Node* n) override;
  bool terminateConditionReached(Node* node) override;
  bool isConnectionConditionSatisfied(SearchSpaceNode* temp, SearchSpaceNode* S) override;
  bool isConnectionConditionSatisfied(geometry_msgs::Pose temp, geometry_msgs::Pose S) override;
  bool isFilteringConditionSatisfied(geometry_msgs::Pose pose, geometry_msgs::PoseArray& correspondingSensorPoses, double minDist,
                                     double maxDist, pcl::PointCloud<pcl::PointXYZ>& globalCloud,
                                     std::vector<pcl::PointCloud<pcl::PointXYZ> >& accuracyClusters,
                                     double accuracyThreshhold) override;
  double pointCloudDiff(pcl::PointCloud<pcl::PointXYZ>::Ptr, pcl::PointCloud<pcl::PointXYZ>::Ptr&) override { return 1; }
  void clusteringPointCloud(std::vector<pcl::PointCloud<pcl::PointXYZ> >&, pcl::PointCloud<pcl::PointXYZ>::Ptr) override {}
  void findClusterOuterPoints(geometry_msgs::PoseArray, pcl::PointCloud<pcl::PointXYZ>&) override {}
  void findClusterBB(pcl::PointCloud<pcl::PointXYZ>, geometry_msgs::Vector3&, geometry_msgs::Pose&) override {}
  void displayProgress(std::vector<Tree>) override {}
  bool connectAll(const double* xyz, int64_t n, double radius, int position_mode, std::vector<int64_t>& row_ptr,
                  std::vector<int32_t>& col) override;
  bool filterAll(const double* xyz, int64_t n, std::vector<uint8_t>& valid) override;
  void connectionConditionAll(const double* starts, const double* ends, int64_t n, std::vector<uint8_t>& ok) override;
  // node validity is an extension (the reference answers true): off unless asked for
  void setFilterByValidity(bool on) { filterByValidity = on; }
  // false = the manager==NULL behaviour of the reference (every in-radius pair accepted) while still
  // using the engine's radius search
  void setUseMap(bool on) { useMap = on; }

  int64_t lastTestedPairs = 0;

 protected:
  bool debug;
  bool filterByValidity = false;
  bool useMap = true;
  geometry_msgs::Pose endPose;
  double tolerance2Goal = 0;
  double box[3] = {0.5, 0.5, 0.3};  // the reference hard-codes this (src/distance_heuristic.cpp:107)
  GpuWorld* world;
};
}  // namespace SSPP
