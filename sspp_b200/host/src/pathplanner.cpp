// PathPlanner (reference src/pathplanner.cpp).  Cited lines are the reference's.
#include "sspp/pathplanner.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iomanip>

namespace SSPP {
namespace {
// tf::createQuaternionFromYaw = Quaternion::setRPY(0, 0, yaw), every product written out so the
// bits equal tf's
geometry_msgs::Quaternion quaternion_from_yaw(double yaw) {
  const double hy = yaw * 0.5, hp = 0.0 * 0.5, hr = 0.0 * 0.5;
  const double cy = std::cos(hy), sy = std::sin(hy), cp = std::cos(hp), sp = std::sin(hp), cr = std::cos(hr), sr = std::sin(hr);
  geometry_msgs::Quaternion q;
  q.x = sr * cp * cy - cr * sp * sy;
  q.y = cr * sp * cy + sr * cp * sy;
  q.z = cr * cp * sy - sr * sp * cy;
  q.w = cr * cp * cy + sr * sp * sy;
  return q;
}
}  // namespace

PathPlanner::PathPlanner(Robot* rob, double radius, int freq) : Astar(rob, freq), regGridConRadius(radius) {}
PathPlanner::PathPlanner(Robot* rob, double radius, int freq, std::vector<Sensors>& rSensors)
    : Astar(rob, freq), regGridConRadius(radius), robotSensors(rSensors) { MAXNODES = 0; }
PathPlanner::~PathPlanner() { freeResources(); }

void PathPlanner::freeResources() {
  freeSearchSpace();
  freePath();
  root = nullptr;
}

void PathPlanner::freePath() {
  for (Node* path : paths)
    while (path) {
      Node* nx = path->next;
      delete path;
      path = nx;
    }
  paths.clear();
}

void PathPlanner::generateRegularGrid(geometry_msgs::Pose s, geometry_msgs::Vector3 size, float res, bool sampleOr, float orRes,
                                      bool filtering, bool insert) {
  gridResolution = res;
  sampleOrientations = sampleOr;
  orientationResolution = orRes;
  samplesFiltering = filtering;
  insertSearchSpace = insert;
  generateRegularGrid(s, size);
}
void PathPlanner::generateRegularGrid(geometry_msgs::Pose s, geometry_msgs::Vector3 size, float res) {
  gridResolution = res;
  generateRegularGrid(s, size);
}

// z outer, x middle, y inner; the loop variable is a double stepped by the FLOAT resolution, so the
// number of samples for a resolution that is not exactly representable follows the accumulated
// rounding, as in the reference (:122-126).  Samples are collected first and filtered in one batched
// call when the heuristic offers it; insertion order is the generation order either way.
void PathPlanner::generateRegularGrid(geometry_msgs::Pose gridStartPose, geometry_msgs::Vector3 gridSize) {
  for (size_t j = sensorsFilteredPoses.size(); j < robotSensors.size(); ++j) sensorsFilteredPoses.push_back(geometry_msgs::PoseArray());
  int orientationsNum = 1;
  if (sampleOrientations) orientationsNum = 360.0f / orientationResolution;  // float quotient truncated (:117)
  std::vector<geometry_msgs::Pose> samples;
  for (double z = (gridStartPose.position.z - gridSize.z / 2.0); z <= (gridStartPose.position.z + gridSize.z / 2.0); z += gridResolution)
    for (double x = (gridStartPose.position.x - gridSize.x / 2.0); x <= (gridStartPose.position.x + gridSize.x / 2.0); x += gridResolution)
      for (double y = (gridStartPose.position.y - gridSize.y / 2.0); y <= (gridStartPose.position.y + gridSize.y / 2.0); y += gridResolution) {
        geometry_msgs::Pose pose;
        pose.position.x = x;
        pose.position.y = y;
        pose.position.z = z;
        if (sampleOrientations) {
          double yaw = 0.0;
          for (int i = 0; i < orientationsNum; ++i) {
            pose.orientation = quaternion_from_yaw(yaw);
            yaw += (orientationResolution * M_PI / 180.0f);
            samples.push_back(pose);
          }
        } else {
          pose.orientation.x = 0; pose.orientation.y = 0; pose.orientation.z = 0; pose.orientation.w = 1;
          samples.push_back(pose);
        }
      }
  std::vector<geometry_msgs::PoseArray> sensorPosesAll(samples.size());
  for (size_t i = 0; i < samples.size(); ++i)
    for (size_t j = 0; j < robotSensors.size(); ++j) sensorPosesAll[i].poses.push_back(robotSensors[j].robot2sensorTransformation(samples[i]));
  std::vector<uint8_t> keep(samples.size(), 1);
  const bool filter_here = samplesFiltering && sampleOrientations;  // the reference filters only in the orientation branch (:151)
  bool batched = false;
  if (filter_here && heuristic) {
    batched = heuristic->filterAllPoses(samples, sensorPosesAll, keep);
    if (!batched) {
      std::vector<double> xyz(samples.size() * 3);
      for (size_t i = 0; i < samples.size(); ++i) {
        xyz[3 * i] = samples[i].position.x; xyz[3 * i + 1] = samples[i].position.y; xyz[3 * i + 2] = samples[i].position.z;
      }
      batched = heuristic->filterAll(xyz.data(), (int64_t)samples.size(), keep);
    }
  }
  for (size_t i = 0; i < samples.size(); ++i) {
    geometry_msgs::PoseArray& sensorPoses = sensorPosesAll[i];
    if (filter_here) {
      const double desiredMaxDepth = 6.9;
      const double maxErrorThresh = 0.0000285 * desiredMaxDepth * desiredMaxDepth;
      bool ok = batched ? keep[i] != 0
                        : heuristic->isFilteringConditionSatisfied(samples[i], sensorPoses, 1, 4, globalCloud, accuracyClusters, maxErrorThresh);
      if (!ok) continue;
      if (insertSearchSpace) {
        insertNode(samples[i], sensorPoses);
        robotFilteredPoses.poses.push_back(samples[i]);
        for (size_t j = 0; j < robotSensors.size(); ++j) sensorsFilteredPoses[j].poses.push_back(sensorPoses.poses[j]);
      }
    } else if (insertSearchSpace) {
      insertNode(samples[i], sensorPoses);
    }
  }
}

// All ordered in-radius pairs of the list, each through the heuristic's connection test, children
// pushed in list order (:527-567) — done by one batched call instead of the N x N host loop.
void PathPlanner::connectNodes() {
  if (!searchspace) return;
  if (!heuristic) throw SSPPException("connectNodes: no heuristic set");
  const std::vector<SearchSpaceNode*>& list = flatten();
  std::vector<double> xyz(list.size() * 3);
  for (size_t i = 0; i < list.size(); ++i) {
    xyz[3 * i] = list[i]->location.position.x; xyz[3 * i + 1] = list[i]->location.position.y; xyz[3 * i + 2] = list[i]->location.position.z;
  }
  std::vector<int64_t> row_ptr;
  std::vector<int32_t> col;
  if (!heuristic->connectAll(xyz.data(), (int64_t)list.size(), regGridConRadius, 0, row_ptr, col))
    throw SSPPException("connectNodes: the heuristic offers no batched connection test (there is no per-pair host loop)");
  for (size_t i = 0; i < list.size(); ++i) {
    std::vector<SearchSpaceNode*>& ch = list[i]->children;
    ch.reserve(ch.size() + (size_t)(row_ptr[i + 1] - row_ptr[i]));
    for (int64_t e = row_ptr[i]; e < row_ptr[i + 1]; ++e) ch.push_back(list[(size_t)col[(size_t)e]]);
  }
  MAXNODES = (long)col.size();  // numConnections (:566)
}

// Same, over a temporary list, with the 1e-12 position test, edges recorded on the search-space
// nodes of the same poses (:612-656)
void PathPlanner::connectClustersInternalNodes(SearchSpaceNode* space, double connRadius) {
  if (!space) return;
  if (!heuristic) throw SSPPException("connectClustersInternalNodes: no heuristic set");
  std::vector<SearchSpaceNode*> list;
  for (SearchSpaceNode* t = space; t; t = t->next) list.push_back(t);
  std::vector<double> xyz(list.size() * 3);
  for (size_t i = 0; i < list.size(); ++i) {
    xyz[3 * i] = list[i]->location.position.x; xyz[3 * i + 1] = list[i]->location.position.y; xyz[3 * i + 2] = list[i]->location.position.z;
  }
  std::vector<int64_t> row_ptr;
  std::vector<int32_t> col;
  if (!heuristic->connectAll(xyz.data(), (int64_t)list.size(), connRadius, 1, row_ptr, col))
    throw SSPPException("connectClustersInternalNodes: the heuristic offers no batched connection test");
  std::vector<SearchSpaceNode*> real(list.size());
  for (size_t i = 0; i < list.size(); ++i) real[i] = nodeExists(list[i]->location);
  long added = 0;
  for (size_t i = 0; i < list.size(); ++i)
    for (int64_t e = row_ptr[i]; e < row_ptr[i + 1]; ++e) {
      SearchSpaceNode* a = real[i];
      SearchSpaceNode* b = real[(size_t)col[(size_t)e]];
      if (!a || !b) continue;
      a->children.push_back(b);
      ++added;
    }
  MAXNODES += added;
}

// number of repeated (parent pose, child pose) connections (:569-610; the reference prints it, O(E^2);
// here a sort of the index pairs)
long PathPlanner::checkSearchSpaceDuplications() {
  const std::vector<SearchSpaceNode*>& list = flatten();
  std::vector<std::pair<int, int> > e;
  for (SearchSpaceNode* t : list)
    for (SearchSpaceNode* c : t->children) e.push_back(std::make_pair(t->index, c->index));
  std::sort(e.begin(), e.end());
  long dup = 0;
  for (size_t i = 1; i < e.size(); ++i) dup += (e[i] == e[i - 1]);
  return dup;
}

// For every point of hull 2 its nearest point of hull 1 (the reference asks a kd-tree for the 1-NN;
// squared float distances, first minimum wins), then every yaw of the one against every yaw of the
// other: nodes that exist in the search space are linked in both directions when the straight segment
// passes the heuristic's connection test — one batched test for all segments instead of one per yaw
// pair, the test only depends on the positions (:658-763).
void PathPlanner::connectToNN(pcl::PointCloud<pcl::PointXYZ> hull1, pcl::PointCloud<pcl::PointXYZ> hull2) {
  if (hull1.points.empty() || hull2.points.empty()) return;
  if (!heuristic) throw SSPPException("connectToNN: no heuristic set");
  const size_t n2 = hull2.points.size();
  std::vector<int> nn(n2);
  for (size_t i = 0; i < n2; ++i) {
    const pcl::PointXYZ& q = hull2.points[i];
    int best = 0;
    float bd = -1;
    for (size_t j = 0; j < hull1.points.size(); ++j) {
      const pcl::PointXYZ& p = hull1.points[j];
      float d = (p.x - q.x) * (p.x - q.x) + (p.y - q.y) * (p.y - q.y) + (p.z - q.z) * (p.z - q.z);
      if (bd < 0 || d < bd) { bd = d; best = (int)j; }
    }
    nn[i] = best;
  }
  std::vector<double> starts(n2 * 3), ends(n2 * 3);
  for (size_t i = 0; i < n2; ++i) {
    const pcl::PointXYZ& q = hull2.points[i];
    const pcl::PointXYZ& p = hull1.points[(size_t)nn[i]];
    starts[3 * i] = q.x; starts[3 * i + 1] = q.y; starts[3 * i + 2] = q.z;   // node -> nearest, as the reference calls it
    ends[3 * i] = p.x; ends[3 * i + 1] = p.y; ends[3 * i + 2] = p.z;
  }
  std::vector<uint8_t> ok;
  heuristic->connectionConditionAll(starts.data(), ends.data(), (int64_t)n2, ok);
  const int orientationsNum = 360.0f / orientationResolution;
  long numConnections = 0;
  for (size_t i = 0; i < n2; ++i) {
    geometry_msgs::Pose node, nearest;
    node.position.x = starts[3 * i]; node.position.y = starts[3 * i + 1]; node.position.z = starts[3 * i + 2];
    nearest.position.x = ends[3 * i]; nearest.position.y = ends[3 * i + 1]; nearest.position.z = ends[3 * i + 2];
    const bool same_place = isPositionEqual(node.position, nearest.position);
    double yaw1 = 0.0;
    for (int k = 0; k < orientationsNum; ++k) {
      node.orientation = quaternion_from_yaw(yaw1);
      double yaw2 = 0.0;
      for (int m = 0; m < orientationsNum; ++m) {
        nearest.orientation = quaternion_from_yaw(yaw2);
        yaw2 += (orientationResolution * M_PI / 180.0f);
        if (samePosition(node, nearest)) continue;
        if (!same_place && !ok[i]) continue;
        SearchSpaceNode* a = nodeExists(node);
        SearchSpaceNode* b = nodeExists(nearest);
        if (a && b) {
          a->children.push_back(b);
          b->children.push_back(a);
          numConnections += 2;
        }
      }
      yaw1 += (orientationResolution * M_PI / 180.0f);
    }
  }
  MAXNODES += numConnections;
}

// (:765-891) with the coverage heuristic's pointCloudDiff / clusteringPointCloud / findClusterBB
void PathPlanner::dynamicNodesGenerationAndConnection(geometry_msgs::Pose gridStartPose, geometry_msgs::Vector3 gridSize, double startRes,
                                                      double resDecrement) {
  if (!heuristic) throw SSPPException("dynamicNodesGenerationAndConnection: no heuristic set");
  double res = startRes;
  generateRegularGrid(gridStartPose, gridSize, (float)startRes, true, 45, true, true);
  SearchSpaceNode* S = insertTempSearchSpace(robotFilteredPoses, sensorsFilteredPoses);
  double connRadius = std::sqrt((res * res) + (res * res)) + 0.01;
  connectClustersInternalNodes(S, connRadius);
  freeTempSearchSpace(S);
  pcl::PointCloud<pcl::PointXYZ> hullCloud1;
  for (const geometry_msgs::Pose& p : robotFilteredPoses.poses) hullCloud1.push_back(pcl::PointXYZ((float)p.position.x, (float)p.position.y, (float)p.position.z));
  robotFilteredPoses.poses.clear();
  sensorsFilteredPoses.clear();
  for (;;) {
    pcl::PointCloud<pcl::PointXYZ>::Ptr globalCloudPtr(new pcl::PointCloud<pcl::PointXYZ>);
    globalCloudPtr->points = globalCloud.points;
    pcl::PointCloud<pcl::PointXYZ>::Ptr diffPtr(new pcl::PointCloud<pcl::PointXYZ>);
    const double uncovered = heuristic->pointCloudDiff(globalCloudPtr, diffPtr);
    if (uncovered <= 0.01) break;
    std::vector<pcl::PointCloud<pcl::PointXYZ> > clusters;
    heuristic->clusteringPointCloud(clusters, diffPtr);
    for (size_t i = 0; i < accuracyClusters.size(); ++i) clusters.push_back(accuracyClusters[i]);
    accuracyClusters.clear();
    res -= resDecrement;
    connRadius = std::sqrt((res * res) + (res * res)) + 0.01;
    if (!(res > 0)) break;
    pcl::PointCloud<pcl::PointXYZ> clustersCloudHull;
    for (size_t i = 0; i < clusters.size(); ++i) {
      geometry_msgs::Pose clusterGridStart;
      geometry_msgs::Vector3 clusterGridSize;
      heuristic->findClusterBB(clusters[i], clusterGridSize, clusterGridStart);
      for (size_t j = sensorsFilteredPoses.size(); j < robotSensors.size(); ++j) sensorsFilteredPoses.push_back(geometry_msgs::PoseArray());
      generateRegularGrid(clusterGridStart, clusterGridSize, (float)res, true, 45, true, true);
      if (robotFilteredPoses.poses.size() > 1) {
        SearchSpaceNode* T = insertTempSearchSpace(robotFilteredPoses, sensorsFilteredPoses);
        connectClustersInternalNodes(T, connRadius);
        freeTempSearchSpace(T);
      }
      for (const geometry_msgs::Pose& p : robotFilteredPoses.poses)
        clustersCloudHull.push_back(pcl::PointXYZ((float)p.position.x, (float)p.position.y, (float)p.position.z));
      robotFilteredPoses.poses.clear();
      sensorsFilteredPoses.clear();
    }
    connectToNN(hullCloud1, clustersCloudHull);
    hullCloud1.points = clustersCloudHull.points;
  }
}

void PathPlanner::disconnectNodes() {
  for (SearchSpaceNode* t = searchspace; t; t = t->next) t->children.clear();
}

void PathPlanner::blockPath(Node* path) {
  for (; path; path = path->next) removeNode(path->pose.p);
}

Node* PathPlanner::startSearch(Pose startPose) {
  if (multiAgentSupport) {
    disconnectNodes();
    for (Node* p : paths) blockPath(p);
    connectNodes();
  }
  Node* p = astarSearch(startPose);
  if (!multiAgentSupport) freePath();  // only one path kept at any time
  paths.push_back(p);
  return p;
}

std::vector<geometry_msgs::Point> PathPlanner::getSearchSpace() {
  std::vector<geometry_msgs::Point> pts;
  for (SearchSpaceNode* t = searchspace; t; t = t->next) pts.push_back(t->location.position);
  return pts;
}

std::vector<geometry_msgs::Point> PathPlanner::getConnections() {
  std::vector<geometry_msgs::Point> seg;
  for (SearchSpaceNode* t = searchspace; t; t = t->next)
    for (SearchSpaceNode* c : t->children) {
      seg.push_back(t->location.position);
      seg.push_back(c->location.position);
    }
  return seg;
}

// way-points and sensor viewpoints, one PoseArray per sensor (:378-404).  As in the reference the number of
// sensor arrays follows the head node's sensor poses; an empty search space (which the reference
// dereferences) yields nothing.
void PathPlanner::getRobotSensorPoses(geometry_msgs::PoseArray& robotPoses, std::vector<geometry_msgs::PoseArray>& sensorsPoses) {
  if (!searchspace) return;
  for (size_t i = 0; i < searchspace->sensorLocation.poses.size(); ++i) sensorsPoses.push_back(geometry_msgs::PoseArray());
  for (SearchSpaceNode* t = searchspace; t; t = t->next) {
    robotPoses.poses.push_back(t->location);
    for (size_t i = 0; i < sensorsPoses.size() && i < t->sensorLocation.poses.size(); ++i)
      sensorsPoses[i].poses.push_back(t->sensorLocation.poses[i]);
  }
}

// way-points and ALL sensor viewpoints in one array, node by node (:407-427)
void PathPlanner::getRobotSensorPoses(geometry_msgs::PoseArray& robotPoses, geometry_msgs::PoseArray& sensorPoses) {
  for (SearchSpaceNode* t = searchspace; t; t = t->next) {
    robotPoses.poses.push_back(t->location);
    for (const geometry_msgs::Pose& sp : t->sensorLocation.poses) sensorPoses.poses.push_back(sp);
  }
}

// declared at include/sspp/pathplanner.h:63 of the reference and defined nowhere in it; here: the number of
// way-points of the path the last startSearch returned (0 when it found none)
unsigned int PathPlanner::getPlanningSteps() {
  unsigned int steps = 0;
  for (Node* q = paths.empty() ? nullptr : paths.back(); q; q = q->next) ++steps;
  return steps;
}

// three pose files read in lock-step: robot pose + two sensor poses per way-point (:204-269)
void PathPlanner::loadRegularGrid(const char* f1, const char* f2, const char* f3) {
  FILE* a = fopen(f1, "r");
  FILE* b = fopen(f2, "r");
  FILE* c = fopen(f3, "r");
  if (!a || !b || !c) {
    if (a) fclose(a);
    if (b) fclose(b);
    if (c) fclose(c);
    throw SSPPException("loadRegularGrid: cannot open search space file");
  }
  auto read_pose = [](FILE* f, geometry_msgs::Pose& p) {
    return fscanf(f, "%lf %lf %lf %lf %lf %lf %lf\n", &p.position.x, &p.position.y, &p.position.z, &p.orientation.x,
                  &p.orientation.y, &p.orientation.z, &p.orientation.w) == 7;
  };
  while (!feof(a) && !feof(b) && !feof(c)) {
    geometry_msgs::Pose pose, s1, s2;
    if (!read_pose(a, pose) || !read_pose(b, s1) || !read_pose(c, s2)) break;
    geometry_msgs::PoseArray sensors;
    sensors.poses.push_back(s1);
    sensors.poses.push_back(s2);
    insertNode(pose, sensors);
  }
  fclose(a);
  fclose(b);
  fclose(c);
}

// edge list, 14 columns, 3 decimals (:271-301)
void PathPlanner::saveSearchSpace(const char* filename) {
  std::ofstream out(filename);
  out << std::fixed << std::setprecision(3);
  for (SearchSpaceNode* t = searchspace; t; t = t->next)
    for (SearchSpaceNode* c : t->children) {
      const geometry_msgs::Pose* two[2] = {&t->location, &c->location};
      for (int k = 0; k < 2; ++k) {
        const geometry_msgs::Pose& p = *two[k];
        out << p.position.x << " " << p.position.y << " " << p.position.z << " " << p.orientation.x << " " << p.orientation.y
            << " " << p.orientation.z << " " << p.orientation.w << (k == 0 ? " " : "\n");
      }
    }
}

// reads the edge list back.  The reference's eof()/getline/sscanf pattern parses the empty string
// after a trailing newline with the previous line's values still in place, so the last edge is
// pushed twice (:319-356); kept, since graphs saved and reloaded by the reference look like that.
bool PathPlanner::loadSearchSpace(const char* fileName) {
  std::ifstream in(fileName);
  if (!in) throw SSPPException(std::string("loadSearchSpace: cannot open ") + fileName);
  double v[14] = {0};
  char buffer[1024];
  while (!in.eof()) {
    in.getline(buffer, sizeof(buffer));
    if (buffer[0] == '#') continue;
    sscanf(buffer, "%lf %lf %lf %lf %lf %lf %lf %lf %lf %lf %lf %lf %lf %lf", &v[0], &v[1], &v[2], &v[3], &v[4], &v[5], &v[6], &v[7],
           &v[8], &v[9], &v[10], &v[11], &v[12], &v[13]);
    geometry_msgs::Pose p1, p2;
    p1.position.x = v[0]; p1.position.y = v[1]; p1.position.z = v[2];
    p1.orientation.x = v[3]; p1.orientation.y = v[4]; p1.orientation.z = v[5]; p1.orientation.w = v[6];
    p2.position.x = v[7]; p2.position.y = v[8]; p2.position.z = v[9];
    p2.orientation.x = v[10]; p2.orientation.y = v[11]; p2.orientation.z = v[12]; p2.orientation.w = v[13];
    geometry_msgs::PoseArray s1, s2;
    for (size_t j = 0; j < robotSensors.size(); ++j) {
      s1.poses.push_back(robotSensors[j].robot2sensorTransformation(p1));
      s2.poses.push_back(robotSensors[j].robot2sensorTransformation(p2));
    }
    SearchSpaceNode* n1 = insertNode(p1, s1);
    SearchSpaceNode* n2 = insertNode(p2, s2);
    n1->children.push_back(n2);
  }
  return true;
}
// ---- lossless binary graph file ------------------------------------------------------------------
// layout (little endian): magic "SSPPCSR1", u64 n, u64 E, u64 sensors_per_node, then per node in LIST
// order: i32 id, 7 doubles pose, sensors_per_node x 7 doubles; then (n+1) i64 row_ptr, E i32 col_idx.
namespace {
const char kMagic[8] = {'S', 'S', 'P', 'P', 'C', 'S', 'R', '1'};
void put_pose(std::ofstream& out, const geometry_msgs::Pose& p) {
  const double v[7] = {p.position.x, p.position.y, p.position.z, p.orientation.x, p.orientation.y, p.orientation.z, p.orientation.w};
  out.write((const char*)v, sizeof(v));
}
bool get_pose(std::ifstream& in, geometry_msgs::Pose& p) {
  double v[7];
  if (!in.read((char*)v, sizeof(v))) return false;
  p.position.x = v[0]; p.position.y = v[1]; p.position.z = v[2];
  p.orientation.x = v[3]; p.orientation.y = v[4]; p.orientation.z = v[5]; p.orientation.w = v[6];
  return true;
}
}  // namespace

void PathPlanner::saveSearchSpaceBinary(const char* filename) {
  const std::vector<SearchSpaceNode*>& list = flatten();
  std::ofstream out(filename, std::ios::binary);
  if (!out) throw SSPPException(std::string("saveSearchSpaceBinary: cannot open ") + filename);
  uint64_t n = list.size(), E = 0, ns = list.empty() ? 0 : list[0]->sensorLocation.poses.size();
  for (SearchSpaceNode* t : list) {
    E += t->children.size();
    if (t->sensorLocation.poses.size() != ns) throw SSPPException("saveSearchSpaceBinary: nodes carry different numbers of sensor poses");
  }
  out.write(kMagic, 8);
  out.write((const char*)&n, 8);
  out.write((const char*)&E, 8);
  out.write((const char*)&ns, 8);
  for (SearchSpaceNode* t : list) {
    int32_t id = t->id;
    out.write((const char*)&id, 4);
    put_pose(out, t->location);
    for (const geometry_msgs::Pose& sp : t->sensorLocation.poses) put_pose(out, sp);
  }
  int64_t off = 0;
  for (SearchSpaceNode* t : list) {
    out.write((const char*)&off, 8);
    off += (int64_t)t->children.size();
  }
  out.write((const char*)&off, 8);
  for (SearchSpaceNode* t : list)
    for (SearchSpaceNode* c : t->children) {
      int32_t j = c->index;
      out.write((const char*)&j, 4);
    }
}

// replaces the current search space; the list order, the ids and the children order come back exactly
bool PathPlanner::loadSearchSpaceBinary(const char* filename) {
  std::ifstream in(filename, std::ios::binary);
  if (!in) throw SSPPException(std::string("loadSearchSpaceBinary: cannot open ") + filename);
  char magic[8];
  uint64_t n = 0, E = 0, ns = 0;
  if (!in.read(magic, 8) || memcmp(magic, kMagic, 8) != 0) throw SSPPException("loadSearchSpaceBinary: not an SSPPCSR1 file");
  in.read((char*)&n, 8);
  in.read((char*)&E, 8);
  in.read((char*)&ns, 8);
  std::vector<geometry_msgs::Pose> poses(n);
  std::vector<geometry_msgs::PoseArray> sens(n);
  std::vector<int32_t> ids(n);
  for (uint64_t i = 0; i < n; ++i) {
    if (!in.read((char*)&ids[i], 4) || !get_pose(in, poses[i])) throw SSPPException("loadSearchSpaceBinary: truncated node table");
    sens[i].poses.resize(ns);
    for (uint64_t k = 0; k < ns; ++k)
      if (!get_pose(in, sens[i].poses[k])) throw SSPPException("loadSearchSpaceBinary: truncated node table");
  }
  std::vector<int64_t> rp(n + 1);
  std::vector<int32_t> col(E);
  if (!in.read((char*)rp.data(), (std::streamsize)((n + 1) * 8)) || (E && !in.read((char*)col.data(), (std::streamsize)(E * 4))))
    throw SSPPException("loadSearchSpaceBinary: truncated adjacency");
  if (rp[n] != (int64_t)E) throw SSPPException("loadSearchSpaceBinary: row_ptr and edge count disagree");
  freeSearchSpace();
  // head insertion reverses: insert from the tail of the stored list so the list order comes back
  for (uint64_t i = n; i-- > 0;) {
    SearchSpaceNode* t = insertNode(poses[i], sens[i], ids[i]);
    if (t->id != ids[i]) throw SSPPException("loadSearchSpaceBinary: duplicate pose in file");
  }
  const std::vector<SearchSpaceNode*>& list = flatten();
  for (uint64_t i = 0; i < n; ++i)
    for (int64_t e = rp[i]; e < rp[i + 1]; ++e) {
      if (col[e] < 0 || (uint64_t)col[e] >= n) throw SSPPException("loadSearchSpaceBinary: column index out of range");
      list[i]->children.push_back(list[(size_t)col[e]]);
    }
  MAXNODES = (long)E;
  return true;
}
}  // namespace SSPP
