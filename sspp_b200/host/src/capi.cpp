// C hooks over the host mirror so the Python test-suite can drive it the way the reference's own
// test programs drive the reference (src/distance_heuristic_test.cpp, src/reactive_planner_server.cpp).
// Not part of the drop-in boundary (that is include/sspp_gpu.h below, and the C++ classes above).
#include <cstdint>
#include <cstring>
#include <string>

#include "sspp/coverage_heuristic.h"
#include "sspp/distance_heuristic.h"
#include "sspp/pathplanner.h"

using namespace SSPP;

namespace {
struct Handle {
  GpuWorld* world = nullptr;
  bool own_world = false;
  DistanceHeuristic* heuristic = nullptr;
  CoverageHeuristic* coverage = nullptr;  // == heuristic when the handle was made by sspph_create_coverage
  std::vector<Sensors> sensors;
  PathPlanner* planner = nullptr;
  std::string err;
};
geometry_msgs::Pose mkpose(const double p[7]) {
  geometry_msgs::Pose q;
  q.position.x = p[0]; q.position.y = p[1]; q.position.z = p[2];
  q.orientation.x = p[3]; q.orientation.y = p[4]; q.orientation.z = p[5]; q.orientation.w = p[6];
  return q;
}
// a heuristic whose adjacency is injected (CPU tests of A* and the list logic: no engine needed)
class InjectedHeuristic : public DistanceHeuristic {
 public:
  InjectedHeuristic() : DistanceHeuristic(false, nullptr) {}
  std::vector<int64_t> rp;
  std::vector<int32_t> cl;
  bool connectAll(const double*, int64_t n, double, int, std::vector<int64_t>& row_ptr, std::vector<int32_t>& col) override {
    if ((int64_t)rp.size() != n + 1) throw SSPPException("injected adjacency does not match the node count");
    row_ptr = rp;
    col = cl;
    return true;
  }
};
}  // namespace

#define GUARD(h, ...)                          \
  try { __VA_ARGS__; return 0; }               \
  catch (const std::exception& e) { (h)->err = e.what(); return 1; }

extern "C" {

// device < 0: no engine (adjacency must be injected with sspph_set_adjacency)
void* sspph_create(double radius, int device) {
  Handle* h = new Handle;
  try {
    if (device >= 0) {
      h->world = new GpuWorld(device);
      h->own_world = true;
      h->heuristic = new DistanceHeuristic(false, h->world);
    } else {
      h->heuristic = new InjectedHeuristic();
    }
    h->planner = new PathPlanner(nullptr, radius, -1);
    h->planner->setHeuristicFucntion(h->heuristic);
  } catch (const std::exception& e) {
    h->err = e.what();
  }
  return h;
}
// planner with a CoverageHeuristic and one camera (mount pose relative to the robot, frustum)
void* sspph_create_coverage(double radius, int device, const double mount_xyz[3], const double mount_rpy[3], double hfov,
                            double vfov, double near_d, double far_d) {
  Handle* h = new Handle;
  try {
    h->world = new GpuWorld(device);
    h->own_world = true;
    h->coverage = new CoverageHeuristic(h->world);
    h->heuristic = h->coverage;
    float sp[3] = {(float)mount_xyz[0], (float)mount_xyz[1], (float)mount_xyz[2]};
    float rp[3] = {(float)mount_rpy[0], (float)mount_rpy[1], (float)mount_rpy[2]};
    h->sensors.push_back(Sensors(hfov, vfov, 0.0, near_d, far_d, 640, 480, sp, rp));
    h->coverage->setFrustum(h->sensors[0]);
    h->planner = new PathPlanner(nullptr, radius, -1, h->sensors);
    h->planner->setHeuristicFucntion(h->heuristic);
  } catch (const std::exception& e) {
    h->err = e.what();
  }
  return h;
}
int sspph_set_model_points(void* p, const float* xyz, int64_t n) {
  Handle* h = (Handle*)p;
  if (!h->coverage) { h->err = "not a coverage planner"; return 1; }
  GUARD(h, h->coverage->setModelPoints(xyz, n));
}
void sspph_set_coverage_goal(void* p, double target_pct, double tolerance_pct) {
  Handle* h = (Handle*)p;
  if (!h->coverage) return;
  h->coverage->setCoverageTarget(target_pct);
  h->coverage->setCoverageTolerance(tolerance_pct);
}
// visible model points of way-point `index` (list order); returns the count (idx filled up to cap), -1 on error
int64_t sspph_visible_from(void* p, int64_t index, int32_t* idx, int64_t cap) {
  Handle* h = (Handle*)p;
  try {
    const std::vector<SearchSpaceNode*>& l = h->planner->flatten();
    if (!h->coverage || index < 0 || index >= (int64_t)l.size()) return -1;
    const std::vector<int32_t>& v = h->coverage->visibleFrom(l[(size_t)index]);
    for (size_t i = 0; i < v.size() && (int64_t)i < cap; ++i) idx[i] = v[i];
    return (int64_t)v.size();
  } catch (const std::exception& e) {
    h->err = e.what();
    return -1;
  }
}
void sspph_get_sensor_poses(void* p, double* poses) {  // first sensor pose of every way-point, list order
  Handle* h = (Handle*)p;
  const std::vector<SearchSpaceNode*>& l = h->planner->flatten();
  for (size_t i = 0; i < l.size(); ++i) {
    geometry_msgs::Pose q = l[i]->sensorLocation.poses.empty() ? geometry_msgs::Pose() : l[i]->sensorLocation.poses[0];
    double* o = poses + 7 * i;
    o[0] = q.position.x; o[1] = q.position.y; o[2] = q.position.z;
    o[3] = q.orientation.x; o[4] = q.orientation.y; o[5] = q.orientation.z; o[6] = q.orientation.w;
  }
}
void sspph_coverage_counters(void* p, int64_t out[2]) {
  Handle* h = (Handle*)p;
  out[0] = h->coverage ? h->coverage->raysCast : 0;
  out[1] = h->coverage ? h->coverage->visibilityCalls : 0;
}

void sspph_destroy(void* p) {
  Handle* h = (Handle*)p;
  delete h->planner;
  delete h->heuristic;
  if (h->own_world) delete h->world;
  delete h;
}
const char* sspph_error(void* p) { return ((Handle*)p)->err.c_str(); }
int sspph_ok(void* p) { return ((Handle*)p)->planner != nullptr; }

int sspph_map_create(void* p, double res, const int32_t kmin[3], const int32_t dims[3], int treat_unknown) {
  Handle* h = (Handle*)p;
  GUARD(h, h->world->createMap(res, kmin, dims, treat_unknown != 0));
}
int sspph_map_set_box(void* p, const double pos[3], const double box[3], int occupied) {
  Handle* h = (Handle*)p;
  GUARD(h, occupied ? h->world->setOccupied(pos, box) : h->world->setFree(pos, box));
}
int sspph_map_insert_boxes(void* p, const double* centres, const double* sizes, const uint8_t* occ, int64_t n) {
  Handle* h = (Handle*)p;
  GUARD(h, h->world->insertBoxes(centres, sizes, occ, n));
}
int sspph_map_upload_dense(void* p, const uint32_t* occ, const uint32_t* known) {
  Handle* h = (Handle*)p;
  GUARD(h, h->world->uploadDense(occ, known));
}

void sspph_set_goal(void* p, const double end_pose[7], double tol) {
  Handle* h = (Handle*)p;
  h->heuristic->setEndPose(mkpose(end_pose));
  h->heuristic->setTolerance2Goal(tol);
}
void sspph_set_filter_by_validity(void* p, int on) { ((Handle*)p)->heuristic->setFilterByValidity(on != 0); }

int sspph_generate_grid(void* p, const double start[3], const double size[3], float res, int sample_orient, float orient_res,
                        int filtering) {
  Handle* h = (Handle*)p;
  geometry_msgs::Pose s;
  s.position.x = start[0]; s.position.y = start[1]; s.position.z = start[2];
  geometry_msgs::Vector3 z;
  z.x = size[0]; z.y = size[1]; z.z = size[2];
  GUARD(h, h->planner->generateRegularGrid(s, z, res, sample_orient != 0, orient_res, filtering != 0, true));
}
void sspph_insert_nodes(void* p, const double* poses, int64_t n) {
  Handle* h = (Handle*)p;
  for (int64_t i = 0; i < n; ++i) {
    const geometry_msgs::Pose q = mkpose(poses + 7 * i);
    if (h->sensors.empty()) { h->planner->insertNode(q); continue; }
    geometry_msgs::PoseArray sp;  // a planner with sensors stores every way-point with its viewpoints
    for (Sensors& s : h->sensors) sp.poses.push_back(s.robot2sensorTransformation(q));
    h->planner->insertNode(q, sp);
  }
}
int sspph_load_regular_grid(void* p, const char* f1, const char* f2, const char* f3) {
  Handle* h = (Handle*)p;
  GUARD(h, h->planner->loadRegularGrid(f1, f2, f3));
}
int sspph_save_search_space(void* p, const char* fn) {
  Handle* h = (Handle*)p;
  GUARD(h, h->planner->saveSearchSpace(fn));
}
int sspph_load_search_space(void* p, const char* fn) {
  Handle* h = (Handle*)p;
  GUARD(h, h->planner->loadSearchSpace(fn));
}
int sspph_save_binary(void* p, const char* fn) {
  Handle* h = (Handle*)p;
  GUARD(h, h->planner->saveSearchSpaceBinary(fn));
}
int sspph_load_binary(void* p, const char* fn) {
  Handle* h = (Handle*)p;
  GUARD(h, h->planner->loadSearchSpaceBinary(fn));
}
int64_t sspph_num_nodes(void* p) { return (int64_t)((Handle*)p)->planner->flatten().size(); }
void sspph_get_nodes(void* p, double* poses, int32_t* ids) {
  Handle* h = (Handle*)p;
  const std::vector<SearchSpaceNode*>& l = h->planner->flatten();
  for (size_t i = 0; i < l.size(); ++i) {
    const geometry_msgs::Pose& q = l[i]->location;
    double* o = poses + 7 * i;
    o[0] = q.position.x; o[1] = q.position.y; o[2] = q.position.z;
    o[3] = q.orientation.x; o[4] = q.orientation.y; o[5] = q.orientation.z; o[6] = q.orientation.w;
    if (ids) ids[i] = l[i]->id;
  }
}
int sspph_remove_node(void* p, const double pose[7]) { return ((Handle*)p)->planner->removeNode(mkpose(pose)) ? 1 : 0; }

int sspph_set_adjacency(void* p, const int64_t* row_ptr, const int32_t* col, int64_t n) {
  Handle* h = (Handle*)p;
  InjectedHeuristic* ih = dynamic_cast<InjectedHeuristic*>(h->heuristic);
  if (!ih) { h->err = "set_adjacency needs a handle created with device < 0"; return 1; }
  ih->rp.assign(row_ptr, row_ptr + n + 1);
  ih->cl.assign(col, col + row_ptr[n]);
  return 0;
}
int sspph_connect_to_nn(void* p, const float* hull1, int64_t n1, const float* hull2, int64_t n2, float orient_res) {
  Handle* h = (Handle*)p;
  GUARD(h, {
    h->planner->orientationResolution = orient_res;
    pcl::PointCloud<pcl::PointXYZ> a, b;
    for (int64_t i = 0; i < n1; ++i) a.push_back(pcl::PointXYZ(hull1[3 * i], hull1[3 * i + 1], hull1[3 * i + 2]));
    for (int64_t i = 0; i < n2; ++i) b.push_back(pcl::PointXYZ(hull2[3 * i], hull2[3 * i + 1], hull2[3 * i + 2]));
    h->planner->connectToNN(a, b);
  });
}
int64_t sspph_maxnodes(void* p) { return ((Handle*)p)->planner->MAXNODES; }
int64_t sspph_duplicate_edges(void* p) { return ((Handle*)p)->planner->checkSearchSpaceDuplications(); }

// coverage-driven dynamic sampling (PathPlanner::dynamicNodesGenerationAndConnection); out: [0] nodes, [1] MAXNODES,
// [2] model points covered by the kept samples
int sspph_dynamic_sampling(void* p, const double start[3], const double size[3], double start_res, double res_decrement,
                           int64_t out[3]) {
  Handle* h = (Handle*)p;
  if (!h->coverage) { h->err = "not a coverage planner"; return 1; }
  geometry_msgs::Pose s;
  s.position.x = start[0]; s.position.y = start[1]; s.position.z = start[2];
  geometry_msgs::Vector3 z;
  z.x = size[0]; z.y = size[1]; z.z = size[2];
  GUARD(h, {
    h->planner->dynamicNodesGenerationAndConnection(s, z, start_res, res_decrement);
    out[0] = (int64_t)h->planner->flatten().size();
    out[1] = h->planner->MAXNODES;
    out[2] = h->coverage->coveredCount;
  });
}
int sspph_set_clustering(void* p, double tolerance, int min_size, double standoff) {
  Handle* h = (Handle*)p;
  if (!h->coverage) { h->err = "not a coverage planner"; return 1; }
  h->coverage->setClustering(tolerance, min_size);
  h->coverage->setStandoff(standoff);
  return 0;
}
int sspph_cluster_labels(void* p, const float* xyz, int64_t n, int32_t* labels) {
  Handle* h = (Handle*)p;
  if (!h->coverage) { h->err = "not a coverage planner"; return 1; }
  GUARD(h, {
    std::vector<int32_t> l = h->coverage->clusterLabels(xyz, n);
    for (int64_t i = 0; i < n; ++i) labels[i] = l[(size_t)i];
  });
}

void sspph_disconnect(void* p) { ((Handle*)p)->planner->disconnectNodes(); }
int sspph_connect(void* p, int64_t* n_edges, int64_t* tested) {
  Handle* h = (Handle*)p;
  GUARD(h, {
    h->planner->connectNodes();
    if (n_edges) *n_edges = h->planner->MAXNODES;
    if (tested) *tested = h->heuristic->lastTestedPairs;
  });
}
int64_t sspph_get_csr(void* p, int64_t* row_ptr, int32_t* col, int64_t cap) {
  Handle* h = (Handle*)p;
  const std::vector<SearchSpaceNode*>& l = h->planner->flatten();
  int64_t E = 0;
  for (size_t i = 0; i < l.size(); ++i) {
    row_ptr[i] = E;
    for (SearchSpaceNode* c : l[i]->children) {
      if (E < cap) col[E] = c->index;
      ++E;
    }
  }
  row_ptr[l.size()] = E;
  return E;
}

// path: cap x 10 doubles (pose7, g, h, f); stats: [0] created, [1] expanded.  Returns the path
// length (0 = none), -1 on error.  order (optional, cap_order): list index of every expanded node.
int64_t sspph_astar(void* p, const double start_pose[7], double* path, int64_t cap, int64_t* stats, int32_t* order,
                    int64_t cap_order) {
  Handle* h = (Handle*)p;
  try {
    Pose s;
    s.p = mkpose(start_pose);
    Node* head = h->planner->startSearch(s);
    if (stats) { stats[0] = h->planner->globalID; stats[1] = h->planner->nodesExpanded; }
    if (order)
      for (size_t i = 0; i < h->planner->expansionOrder.size() && (int64_t)i < cap_order; ++i) order[i] = h->planner->expansionOrder[i];
    int64_t n = 0;
    for (Node* t = head; t; t = t->next, ++n)
      if (n < cap) {
        double* o = path + 10 * n;
        const geometry_msgs::Pose& q = t->pose.p;
        o[0] = q.position.x; o[1] = q.position.y; o[2] = q.position.z;
        o[3] = q.orientation.x; o[4] = q.orientation.y; o[5] = q.orientation.z; o[6] = q.orientation.w;
        o[7] = t->g_value; o[8] = t->h_value; o[9] = t->f_value;
      }
    return n;
  } catch (const std::exception& e) {
    h->err = e.what();
    return -1;
  }
}

// Sensors::robot2sensorTransformation for n robot poses
void sspph_robot2sensor(const double mount_xyz[3], const double mount_rpy[3], const double* robot_poses, int64_t n, double* out) {
  float sp[3] = {(float)mount_xyz[0], (float)mount_xyz[1], (float)mount_xyz[2]};
  float rp[3] = {(float)mount_rpy[0], (float)mount_rpy[1], (float)mount_rpy[2]};
  Sensors s(58, 45, 0, 0.7, 6.0, 640, 480, sp, rp);
  for (int64_t i = 0; i < n; ++i) {
    geometry_msgs::Pose q = s.robot2sensorTransformation(mkpose(robot_poses + 7 * i));
    double* o = out + 7 * i;
    o[0] = q.position.x; o[1] = q.position.y; o[2] = q.position.z;
    o[3] = q.orientation.x; o[4] = q.orientation.y; o[5] = q.orientation.z; o[6] = q.orientation.w;
  }
}

}  // extern "C"
