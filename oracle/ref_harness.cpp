// ORACLE — TEST INFRASTRUCTURE ONLY.
// C harness around the REFERENCE'S OWN planner classes, compiled unmodified from
// /root/reference/src/*.cpp against oracle/stubs (see oracle/Makefile).  It drives
//   SSPP::PathPlanner::generateRegularGrid   src/pathplanner.cpp:90-201
//   SSPP::SearchSpace::insertNode            src/searchspace.cpp:64-105
//   SSPP::PathPlanner::connectNodes          src/pathplanner.cpp:527-567
//   SSPP::PathPlanner::startSearch/Astar     src/pathplanner.cpp:502-525, src/astar.cpp:220-373
//   SSPP::DistanceHeuristic                  src/distance_heuristic.cpp
// exactly as src/distance_heuristic_test.cpp:40-117 and src/reactive_planner_server.cpp:83-129 do,
// and returns nodes (list order), adjacency (as list indices) and the A* path.  The octomap behind
// DistanceHeuristic is the restated one (octomap_world/octomap_manager.h stub).
#include <cstdint>
#include <cstdlib>
#include <iostream>
#include <map>
#include <sstream>
#include <vector>

#include "sspp/pathplanner.h"
#include "sspp/distance_heuristic.h"

// src/rviz_drawing_tools.cpp builds RViz markers; headless stand-ins (no graphics on this box)
visualization_msgs::Marker drawLines(std::vector<geometry_msgs::Point>, int, int, int, double) {
  return visualization_msgs::Marker();
}
visualization_msgs::Marker drawPoints(std::vector<geometry_msgs::Point>, int, int) {
  return visualization_msgs::Marker();
}

// src/sensors.cpp needs real Eigen 4x4 products, so it is not compiled; Sensors is restated here:
// sensorPose = T_robot * T_mount * Rx(+90deg)   (src/sensors.cpp:47-92), products accumulated
// left to right over k as Eigen's coefficient-wise evaluation does.
Sensors::Sensors(double hfov, double vfov, double fl, double nearR, double farR, double width, double height,
                 Eigen::Vector3f sP, Eigen::Vector3f sRPY)
    : hFOV(hfov), vFOV(vfov), focalLength(fl), near(nearR), far(farR), resWidth(width), resHeight(height),
      sensorPose(sP), sensorRPY(sRPY) {}
Sensors::Sensors() {}
Sensors::~Sensors() {}
static void mat4mul(const double A[4][4], const double B[4][4], double C[4][4]) {
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      double acc = A[i][0] * B[0][j];
      for (int k = 1; k < 4; ++k) acc = acc + A[i][k] * B[k][j];
      C[i][j] = acc;
    }
}
geometry_msgs::Pose Sensors::robot2sensorTransformation(geometry_msgs::Pose pose) {
  double A[4][4] = {{0}}, B[4][4] = {{0}}, C[4][4], D[4][4];
  tf::Matrix3x3 R1(tf::Quaternion(pose.orientation.x, pose.orientation.y, pose.orientation.z, pose.orientation.w));
  tf::Matrix3x3 R2(tf::createQuaternionFromRPY(sensorRPY[0], sensorRPY[1], sensorRPY[2]));
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { A[i][j] = R1.m[i][j]; B[i][j] = R2.m[i][j]; }
  A[0][3] = pose.position.x; A[1][3] = pose.position.y; A[2][3] = pose.position.z; A[3][3] = 1;
  B[0][3] = sensorPose[0]; B[1][3] = sensorPose[1]; B[2][3] = sensorPose[2]; B[3][3] = 1;
  mat4mul(A, B, C);
  const double S[4][4] = {{1, 0, 0, 0}, {0, 0, -1, 0}, {0, 1, 0, 0}, {0, 0, 0, 1}};
  mat4mul(C, S, D);
  tf::Matrix3x3 R3;
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) R3.m[i][j] = D[i][j];
  tf::Quaternion qt;
  R3.getRotation(qt);
  geometry_msgs::Pose p;
  p.position.x = D[0][3]; p.position.y = D[1][3]; p.position.z = D[2][3];
  p.orientation.x = qt.getX(); p.orientation.y = qt.getY(); p.orientation.z = qt.getZ(); p.orientation.w = qt.getW();
  return p;
}

namespace {
struct Quiet {  // the reference narrates on std::cout; keep test logs clean
  std::streambuf* old;
  std::ostringstream sink;
  Quiet() : old(std::cout.rdbuf(getenv("SSPP_REF_VERBOSE") ? std::cout.rdbuf() : sink.rdbuf())) {  // SSPP_REF_VERBOSE: let it talk
    if (getenv("SSPP_REF_VERBOSE")) std::cout.setf(std::ios::unitbuf);
  }
  ~Quiet() { std::cout.rdbuf(old); }
};

struct RefPlanner {
  ros::NodeHandle nh;
  Robot* robot;
  SSPP::PathPlanner* planner;
  SSPP::DistanceHeuristic* heuristic;
  volumetric_mapping::OctomapManager* manager;
  std::string last_log;
  SSPP::Heuristic* coverage = NULL;  // set by ref_coverage_planner_create (heuristic stays NULL then)
};

geometry_msgs::Pose mkpose(const double p[7]) {
  geometry_msgs::Pose q;
  q.position.x = p[0]; q.position.y = p[1]; q.position.z = p[2];
  q.orientation.x = p[3]; q.orientation.y = p[4]; q.orientation.z = p[5]; q.orientation.w = p[6];
  return q;
}
}  // namespace

extern "C" {

// world: an orc::OctomapWorld* made by orc_world_create in THIS library, or NULL (manager == NULL)
void* ref_planner_create(double con_radius, void* world) {
  Quiet q;
  RefPlanner* R = new RefPlanner;
  R->robot = NULL;  // PathPlanner only stores the pointer (src/astar.cpp:29)
  R->planner = new SSPP::PathPlanner(R->nh, R->robot, con_radius, -1);
  R->planner->setDebugDelay(0.0);
  R->manager = world ? new volumetric_mapping::OctomapManager((orc::OctomapWorld*)world) : NULL;
  R->heuristic = new SSPP::DistanceHeuristic(R->nh, false, R->manager);
  R->planner->setHeuristicFucntion(R->heuristic);
  return R;
}

void ref_planner_destroy(void* h) {
  Quiet q;
  RefPlanner* R = (RefPlanner*)h;
  delete R->planner;
  delete R->heuristic;
  delete R->coverage;
  delete R->manager;
  delete R;
}

void ref_set_goal(void* h, const double end_pose[7], double tolerance) {
  RefPlanner* R = (RefPlanner*)h;
  R->heuristic->setEndPose(mkpose(end_pose));
  R->heuristic->setTolerance2Goal(tolerance);
}

void ref_generate_grid(void* h, const double start[3], const double size[3], float res, int sample_orient,
                       float orient_res) {
  Quiet q;
  RefPlanner* R = (RefPlanner*)h;
  geometry_msgs::Pose gs;
  gs.position.x = start[0]; gs.position.y = start[1]; gs.position.z = start[2];
  geometry_msgs::Vector3 sz;
  sz.x = size[0]; sz.y = size[1]; sz.z = size[2];
  R->planner->generateRegularGrid(gs, sz, res, sample_orient != 0, orient_res, false, true);
}

void ref_insert_node(void* h, const double pose[7]) {
  RefPlanner* R = (RefPlanner*)h;
  R->planner->insertNode(mkpose(pose));
}

// id of the node SearchSpace::insertNode returns: the new node's own id, or the stored node's when
// the pose is refused (ids count insertNode calls, src/searchspace.cpp:64-67)
long ref_insert_node_id(void* h, const double pose[7]) {
  RefPlanner* R = (RefPlanner*)h;
  SSPP::SearchSpaceNode* t = R->planner->insertNode(mkpose(pose));
  return t ? (long)t->id : -1;
}

long ref_num_nodes(void* h) {
  RefPlanner* R = (RefPlanner*)h;
  long n = 0;
  for (SSPP::SearchSpaceNode* t = R->planner->searchspace; t; t = t->next) ++n;
  return n;
}

// poses: n x 7 doubles in LIST order (head first); ids: SearchSpaceNode::id
void ref_get_nodes(void* h, double* poses, int32_t* ids) {
  RefPlanner* R = (RefPlanner*)h;
  long i = 0;
  for (SSPP::SearchSpaceNode* t = R->planner->searchspace; t; t = t->next, ++i) {
    poses[7 * i + 0] = t->location.position.x; poses[7 * i + 1] = t->location.position.y;
    poses[7 * i + 2] = t->location.position.z; poses[7 * i + 3] = t->location.orientation.x;
    poses[7 * i + 4] = t->location.orientation.y; poses[7 * i + 5] = t->location.orientation.z;
    poses[7 * i + 6] = t->location.orientation.w;
    if (ids) ids[i] = t->id;
  }
}

// file formats of the reference (src/pathplanner.cpp:204-359), driven through its own code
void ref_load_regular_grid(void* h, const char* f1, const char* f2, const char* f3) {
  Quiet q;
  ((RefPlanner*)h)->planner->loadRegularGrid(f1, f2, f3);
}
void ref_save_search_space(void* h, const char* fn) { ((RefPlanner*)h)->planner->saveSearchSpace(fn); }
void ref_load_search_space(void* h, const char* fn) {
  Quiet q;
  ((RefPlanner*)h)->planner->loadSearchSpace(fn);
}

// PathPlanner::connectToNN (src/pathplanner.cpp:658-763) with the kd-tree stand-in (exact brute-force 1-NN)
void ref_connect_to_nn(void* h, const float* hull1, long n1, const float* hull2, long n2, float orient_res) {
  Quiet q;
  RefPlanner* R = (RefPlanner*)h;
  R->planner->orientationResolution = orient_res;
  pcl::PointCloud<pcl::PointXYZ> a, b;
  for (long i = 0; i < n1; ++i) a.push_back(pcl::PointXYZ(hull1[3 * i], hull1[3 * i + 1], hull1[3 * i + 2]));
  for (long i = 0; i < n2; ++i) b.push_back(pcl::PointXYZ(hull2[3 * i], hull2[3 * i + 1], hull2[3 * i + 2]));
  R->planner->connectToNN(a, b);
}
long ref_maxnodes(void* h) { return ((RefPlanner*)h)->planner->MAXNODES; }

void ref_disconnect(void* h) { ((RefPlanner*)h)->planner->disconnectNodes(); }

// runs the reference's O(N^2) connectNodes; returns MAXNODES (= connection count it printed)
long ref_connect(void* h) {
  Quiet q;
  RefPlanner* R = (RefPlanner*)h;
  R->planner->connectNodes();
  return R->planner->MAXNODES;
}

long ref_collision_calls(void* h) {
  RefPlanner* R = (RefPlanner*)h;
  return R->manager ? R->manager->calls : 0;
}

// children vectors -> CSR over list indices; returns E (col filled up to cap)
long ref_get_csr(void* h, int64_t* row_ptr, int32_t* col, long cap) {
  RefPlanner* R = (RefPlanner*)h;
  std::map<SSPP::SearchSpaceNode*, int32_t> index;
  int32_t i = 0;
  for (SSPP::SearchSpaceNode* t = R->planner->searchspace; t; t = t->next) index[t] = i++;
  long E = 0;
  i = 0;
  for (SSPP::SearchSpaceNode* t = R->planner->searchspace; t; t = t->next, ++i) {
    row_ptr[i] = E;
    for (size_t c = 0; c < t->children.size(); ++c) {
      if (E < cap) col[E] = index[t->children[c]];
      ++E;
    }
  }
  row_ptr[i] = E;
  return E;
}

// PathPlanner::startSearch(start).  path: cap x 10 doubles (pose7, g, h, f) from root to goal.
// Returns the number of path nodes (0 = no path).  stats[0] = nodes created (Astar::globalID),
// stats[1] = nodes expanded (parsed from the reference's own log line, -1 if absent).
long ref_astar(void* h, const double start_pose[7], double* path, long cap, long* stats) {
  RefPlanner* R = (RefPlanner*)h;
  std::string log;
  SSPP::Node* p;
  {
    Quiet q;
    Pose s;
    s.p = mkpose(start_pose);
    s.phi = 0;
    p = R->planner->startSearch(s);
    log = q.sink.str();
  }
  if (stats) {
    stats[0] = R->planner->globalID;
    stats[1] = -1;
    size_t k = log.find("nodes created and :");
    if (k != std::string::npos) stats[1] = atol(log.c_str() + k + 19);
  }
  long n = 0;
  for (SSPP::Node* t = p; t; t = t->next, ++n) {
    if (n < cap) {
      double* o = path + 10 * n;
      o[0] = t->pose.p.position.x; o[1] = t->pose.p.position.y; o[2] = t->pose.p.position.z;
      o[3] = t->pose.p.orientation.x; o[4] = t->pose.p.orientation.y; o[5] = t->pose.p.orientation.z;
      o[6] = t->pose.p.orientation.w; o[7] = t->g_value; o[8] = t->h_value; o[9] = t->f_value;
    }
  }
  return n;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// The reference's dynamic sampling loop (PathPlanner::dynamicNodesGenerationAndConnection,
// src/pathplanner.cpp:765-891) needs a heuristic whose coverage virtuals do something; the snapshot
// holds no CoverageHeuristic source.  This harness heuristic implements the AUTHORED
// semantics of sspp_b200/host/include/sspp/coverage_heuristic.h on the ORACLE (restated octomap +
// the oracle's visibility), so that the reference's own loop — its grid sampling, temp spaces,
// connectClustersInternalNodes, connectToNN, termination rule — can be run end to end and compared
// with the mirror's GPU-backed loop node by node and edge by edge:
//   * a sample is kept iff the robot box is free there (getCellStatusBoundingBox == kFree) and at least
//     one model point is visible from one of its sensor poses; what kept samples see is marked covered;
//   * pointCloudDiff = the model points not covered yet, in index order, and their share;
//   * clusteringPointCloud = connected components of the radius graph (Dist <= tolerance) of at least
//     minClusterSize points, numbered by their smallest member, members in index order;
//   * findClusterBB = centre of the bounding box, size = extent + 2 * standoff;
//   * connection condition = DistanceHeuristic's: swept robot box (0.5, 0.5, 0.3) free (:99-127).
// ------------------------------------------------------------------------------------------------
extern "C" {
void orc_cell_status_bbox(void* w, const double* pts, long n, const double box[3], uint8_t* status);
void orc_line_status_bbox(void* w, const double* starts, const double* ends, long n, const double box[3], uint8_t* status,
                          long* steps, long* rays);
void orc_frustum_consts(double hfov_deg, double vfov_deg, double near_d, double far_d, float out[4]);
long orc_visibility(void* w, const double* poses, long nv, const float fr[4], const float* pts, long np, int64_t* view_ptr,
                    int32_t* idx, long cap, int64_t* counters);
}

namespace {
class HarnessCoverage : public SSPP::Heuristic {
 public:
  void* world;
  float fr[4];
  std::vector<float> model;
  std::vector<uint8_t> covered;
  long coveredCount, raysCast;
  double clusterTolerance, standoff;
  int minClusterSize;
  HarnessCoverage(void* w, double hfov, double vfov, double nearD, double farD)
      : world(w), coveredCount(0), raysCast(0), clusterTolerance(0.6), standoff(3.0), minClusterSize(10) {
    orc_frustum_consts(hfov, vfov, nearD, farD, fr);
  }
  void calculateHeuristic(SSPP::Node*) {}
  bool terminateConditionReached(SSPP::Node*) { return true; }
  bool isCost() { return false; }
  void displayProgress(vector<Tree>) {}
  void findClusterOuterPoints(geometry_msgs::PoseArray, pcl::PointCloud<pcl::PointXYZ>&) {}
  bool isConnectionConditionSatisfied(geometry_msgs::Pose a, geometry_msgs::Pose b) {
    if (!world) return true;
    const double s[3] = {a.position.x, a.position.y, a.position.z}, e[3] = {b.position.x, b.position.y, b.position.z};
    const double box[3] = {0.5, 0.5, 0.3};
    uint8_t st = 0;
    long steps = 0, rays = 0;
    orc_line_status_bbox(world, s, e, 1, box, &st, &steps, &rays);
    return st == 0;
  }
  bool isConnectionConditionSatisfied(SSPP::SearchSpaceNode* t, SSPP::SearchSpaceNode* s) {
    return isConnectionConditionSatisfied(t->location, s->location);
  }
  bool isFilteringConditionSatisfied(geometry_msgs::Pose pose, geometry_msgs::PoseArray& sensorPoses, double, double,
                                     pcl::PointCloud<pcl::PointXYZ>&, std::vector<pcl::PointCloud<pcl::PointXYZ> >&, double) {
    const double p[3] = {pose.position.x, pose.position.y, pose.position.z};
    const double box[3] = {0.5, 0.5, 0.3};
    uint8_t st = 0;
    if (world) orc_cell_status_bbox(world, p, 1, box, &st);
    if (st != 0) return false;
    bool any = false;
    for (size_t k = 0; k < sensorPoses.poses.size(); ++k) {
      const geometry_msgs::Pose& q = sensorPoses.poses[k];
      const double sp[7] = {q.position.x, q.position.y, q.position.z, q.orientation.x, q.orientation.y, q.orientation.z, q.orientation.w};
      int64_t vp[2] = {0, 0}, cnt[2] = {0, 0};
      const long np = (long)(model.size() / 3);
      long nvis = orc_visibility(world, sp, 1, fr, model.data(), np, vp, NULL, 0, cnt);
      raysCast += cnt[0];
      if (nvis <= 0) continue;
      std::vector<int32_t> idx((size_t)nvis);
      orc_visibility(world, sp, 1, fr, model.data(), np, vp, idx.data(), nvis, cnt);
      any = true;
      for (long e = 0; e < nvis; ++e)
        if (!covered[(size_t)idx[(size_t)e]]) { covered[(size_t)idx[(size_t)e]] = 1; ++coveredCount; }
    }
    return any;
  }
  double pointCloudDiff(pcl::PointCloud<pcl::PointXYZ>::Ptr, pcl::PointCloud<pcl::PointXYZ>::Ptr& diff) {
    if (!diff) diff.reset(new pcl::PointCloud<pcl::PointXYZ>);
    diff->points.clear();
    const size_t np = model.size() / 3;
    for (size_t i = 0; i < np; ++i)
      if (!covered[i]) diff->points.push_back(pcl::PointXYZ(model[3 * i], model[3 * i + 1], model[3 * i + 2]));
    return np ? (double)diff->points.size() / (double)np : 0.0;
  }
  // connected components of the radius graph: Dist (double, left to right, as include/sspp/utils.h:356-359) <= tolerance
  void clusteringPointCloud(std::vector<pcl::PointCloud<pcl::PointXYZ> >& clusters, pcl::PointCloud<pcl::PointXYZ>::Ptr diff) {
    clusters.clear();
    if (!diff || diff->points.empty()) return;
    const long n = (long)diff->points.size();
    std::vector<double> d((size_t)n * 3);
    for (long i = 0; i < n; ++i) { d[3 * i] = diff->points[i].x; d[3 * i + 1] = diff->points[i].y; d[3 * i + 2] = diff->points[i].z; }
    std::vector<int32_t> label((size_t)n, -1), stack;
    int32_t next = 0;
    for (long s = 0; s < n; ++s) {
      if (label[s] != -1) continue;
      std::vector<int32_t> members;
      stack.assign(1, (int32_t)s);
      label[s] = -2;
      while (!stack.empty()) {
        const int32_t u = stack.back();
        stack.pop_back();
        members.push_back(u);
        for (long v = 0; v < n; ++v) {
          if (label[v] != -1) continue;
          const double dx = d[3 * v] - d[3 * u], dy = d[3 * v + 1] - d[3 * u + 1], dz = d[3 * v + 2] - d[3 * u + 2];
          if (sqrt(dx * dx + dy * dy + dz * dz) <= clusterTolerance) { label[v] = -2; stack.push_back((int32_t)v); }
        }
      }
      const int32_t id = ((int)members.size() >= minClusterSize) ? next++ : -3;
      for (size_t k = 0; k < members.size(); ++k) label[members[k]] = id;
    }
    clusters.resize((size_t)next);
    for (long i = 0; i < n; ++i)
      if (label[i] >= 0) clusters[(size_t)label[i]].push_back(diff->points[i]);
  }
  void findClusterBB(pcl::PointCloud<pcl::PointXYZ> cluster, geometry_msgs::Vector3& gridSize, geometry_msgs::Pose& gridStart) {
    if (cluster.points.empty()) { gridSize = geometry_msgs::Vector3(); return; }
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (size_t i = 0; i < cluster.points.size(); ++i) {
      const double v[3] = {cluster.points[i].x, cluster.points[i].y, cluster.points[i].z};
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
};
}  // namespace

extern "C" {
// a reference PathPlanner with one sensor (mount pose as Sensors takes it) and the harness coverage heuristic on `world`
void* ref_coverage_planner_create(double con_radius, void* world, const double mount_xyz[3], const double mount_rpy[3],
                                  double hfov, double vfov, double near_d, double far_d) {
  Quiet q;
  RefPlanner* R = new RefPlanner;
  R->robot = NULL;
  std::vector<Sensors> sensors;
  sensors.push_back(Sensors(hfov, vfov, 0.0, near_d, far_d, 640, 480,
                            Eigen::Vector3f((float)mount_xyz[0], (float)mount_xyz[1], (float)mount_xyz[2]),
                            Eigen::Vector3f((float)mount_rpy[0], (float)mount_rpy[1], (float)mount_rpy[2])));
  R->planner = new SSPP::PathPlanner(R->nh, R->robot, con_radius, -1, sensors);
  R->planner->setDebugDelay(0.0);
  R->manager = NULL;
  R->heuristic = NULL;
  R->coverage = new HarnessCoverage(world, hfov, vfov, near_d, far_d);
  R->planner->setHeuristicFucntion(R->coverage);
  return R;
}
void ref_coverage_set_model(void* h, const float* xyz, long n) {
  HarnessCoverage* c = (HarnessCoverage*)((RefPlanner*)h)->coverage;
  c->model.assign(xyz, xyz + 3 * n);
  c->covered.assign((size_t)n, 0);
  c->coveredCount = 0;
}
void ref_coverage_set_clustering(void* h, double tolerance, int min_size, double standoff) {
  HarnessCoverage* c = (HarnessCoverage*)((RefPlanner*)h)->coverage;
  c->clusterTolerance = tolerance;
  c->minClusterSize = min_size;
  c->standoff = standoff;
}
// PathPlanner::dynamicNodesGenerationAndConnection, unmodified; out = nodes, MAXNODES, covered model points
void ref_dynamic_sampling(void* h, const double start[3], const double size[3], double start_res, double res_decrement, int64_t out[3]) {
  Quiet q;
  RefPlanner* R = (RefPlanner*)h;
  geometry_msgs::Pose gs;
  gs.position.x = start[0]; gs.position.y = start[1]; gs.position.z = start[2];
  geometry_msgs::Vector3 sz;
  sz.x = size[0]; sz.y = size[1]; sz.z = size[2];
  R->planner->dynamicNodesGenerationAndConnection(gs, sz, start_res, res_decrement);
  out[0] = ref_num_nodes(h);
  out[1] = R->planner->MAXNODES;
  out[2] = ((HarnessCoverage*)R->coverage)->coveredCount;
}
}  // extern "C"
