// PathPlanner — the graph-build facade (reference include/sspp/pathplanner.h:57-114,
// src/pathplanner.cpp): regular-grid sampling, node connection, search.  Same entry points and member
// names; connectNodes() / connectClustersInternalNodes() are batched through the heuristic's
// connectAll() (one C-ABI call) instead of looping over all ordered pairs on the host.
#pragma once
#include <string>
#include <vector>
#include "astar.h"
#include "sensors.h"

namespace SSPP {
class PathPlanner : public Astar {
 public:
  double regGridConRadius;
  PathPlanner(Robot* rob, double regGridConRadius, int progressDisplayFrequency);
  PathPlanner(Robot* rob, double regGridConRadius, int progressDisplayFrequency, std::vector<Sensors>& rSensors);
  ~PathPlanner();
  unsigned int getPlanningSteps();
  void setConRad(double r) { regGridConRadius = r; }
  void setMultiAgentSupport(bool on) { multiAgentSupport = on; }
  void freeResources();
  void freePath();
  void generateRegularGrid(geometry_msgs::Pose gridStartPose, geometry_msgs::Vector3 gridSize);
  void generateRegularGrid(geometry_msgs::Pose gridStartPose, geometry_msgs::Vector3 gridSize, float gridRes);
  void generateRegularGrid(geometry_msgs::Pose gridStartPose, geometry_msgs::Vector3 gridSize, float gridRes,
                           bool sampleOrientations = false, float orientationRes = 360, bool samplesFiltering = false,
                           bool insertSearchSpace = true);  // defaults as include/sspp/pathplanner.h:73 (a three-argument
                                                            // call is ambiguous with the overload above, there as here)
  void connectNodes();
  void connectClustersInternalNodes(SearchSpaceNode* space, double connRadius);
  // links every point of cloudHull2 to its nearest point of cloudHull1, in all yaw combinations
  void connectToNN(pcl::PointCloud<pcl::PointXYZ> cloudHull1, pcl::PointCloud<pcl::PointXYZ> cloudHull2);
  // coverage-driven refinement: sample, connect, find what is still uncovered, cluster it, re-sample the
  // clusters at a finer resolution, connect each stage to the previous one; until <= 1 % is uncovered
  void dynamicNodesGenerationAndConnection(geometry_msgs::Pose gridStartPose, geometry_msgs::Vector3 gridSize, double startRes,
                                           double resDecrement);
  long checkSearchSpaceDuplications();
  void disconnectNodes();
  void blockPath(Node* path);
  Node* startSearch(Pose startPose);
  std::vector<geometry_msgs::Point> getConnections();
  std::vector<geometry_msgs::Point> getSearchSpace();
  void getRobotSensorPoses(geometry_msgs::PoseArray& robotPoses, std::vector<geometry_msgs::PoseArray>& sensorsPoses);
  void getRobotSensorPoses(geometry_msgs::PoseArray& robotPoses, geometry_msgs::PoseArray& sensorPoses);
  void loadRegularGrid(const char* robotFile, const char* sensor1File, const char* sensor2File);
  void saveSearchSpace(const char* filename);
  bool loadSearchSpace(const char* filename);
  // lossless binary form of the whole graph: poses and sensor poses as doubles, ids,
  // adjacency as CSR over list indices.  The text format above keeps three decimals and drops ids.
  void saveSearchSpaceBinary(const char* filename);
  bool loadSearchSpaceBinary(const char* filename);

  std::vector<Sensors> robotSensors;
  bool sampleOrientations = false, samplesFiltering = false, multiAgentSupport = false, insertSearchSpace = true;
  float gridResolution = 1.0f;          // float, as in the reference: the double loop variable is stepped by it
  float orientationResolution = 360.0f;
  pcl::PointCloud<pcl::PointXYZ> globalCloud;
  std::vector<pcl::PointCloud<pcl::PointXYZ> > accuracyClusters;
  geometry_msgs::PoseArray robotFilteredPoses;
  std::vector<geometry_msgs::PoseArray> sensorsFilteredPoses;
  std::vector<Node*> paths;
};
}  // namespace SSPP
