// GpuWorld: C++ face of the C ABI (include/sspp_gpu.h).  Every method is one or two entry-point calls.
#include "sspp/gpu_world.h"

#include <string>

#include "sspp_gpu.h"

namespace SSPP {
GpuWorld::GpuWorld(int device) {
  int rc = sspp_gpu_create(&ctx_, device);
  if (rc != SSPP_OK) throw SSPPException("sspp_gpu_create failed: no usable CUDA device (the engine has no CPU fallback)");
}
GpuWorld::~GpuWorld() { sspp_gpu_destroy(ctx_); }

void GpuWorld::check(int rc, const char* where) {
  if (rc != SSPP_OK) throw SSPPException(std::string(where) + ": " + sspp_gpu_last_error(ctx_));
}

void GpuWorld::createMap(double resolution, const int32_t key_min[3], const int32_t dims[3], bool treat_unknown) {
  check(sspp_gpu_map_create(ctx_, resolution, key_min, dims, treat_unknown ? 1 : 0), "map_create");
  resolution_ = resolution;
  has_map_ = true;
  ++map_generation_;
}
void GpuWorld::setTreatUnknownAsOccupied(bool flag) {
  ++map_generation_;
  check(sspp_gpu_map_set_treat_unknown(ctx_, flag ? 1 : 0), "map_set_treat_unknown");
}
void GpuWorld::setFree(const double position[3], const double box[3]) {
  ++map_generation_;
  check(sspp_gpu_map_set_box(ctx_, position, box, 0), "setFree");
}
void GpuWorld::setOccupied(const double position[3], const double box[3]) {
  ++map_generation_;
  check(sspp_gpu_map_set_box(ctx_, position, box, 1), "setOccupied");
}

void GpuWorld::insertBoxes(const std::vector<std::pair<std::vector<double>, double> >& boxes, bool occupied) {
  std::vector<double> c, s;
  std::vector<uint8_t> o(boxes.size(), occupied ? 1 : 0);
  c.reserve(boxes.size() * 3);
  s.reserve(boxes.size());
  for (const auto& b : boxes) {
    c.insert(c.end(), b.first.begin(), b.first.begin() + 3);
    s.push_back(b.second);
  }
  insertBoxes(c.data(), s.data(), o.data(), (int64_t)s.size());
}
void GpuWorld::insertBoxes(const double* centres, const double* sizes, const uint8_t* occupied, int64_t n) {
  ++map_generation_;
  check(sspp_gpu_map_upload_boxes(ctx_, centres, sizes, occupied, n), "map_upload_boxes");
}
void GpuWorld::uploadDense(const uint32_t* occ_bits, const uint32_t* known_bits) {
  ++map_generation_;
  check(sspp_gpu_map_upload_dense(ctx_, occ_bits, known_bits), "map_upload_dense");
}

void GpuWorld::insertPointcloud(const float* pts, int64_t n, const double origin[3], double max_range) {
  ++map_generation_;
  check(sspp_gpu_map_insert_pointcloud(ctx_, pts, n, origin, max_range), "map_insert_pointcloud");
}

GpuWorld::CellStatus GpuWorld::getLineStatusBoundingBox(const double start[3], const double end[3], const double box[3]) {
  uint8_t st = 0;
  check(sspp_gpu_check_segments(ctx_, start, end, 1, box, &st), "check_segments");
  return (CellStatus)st;
}
GpuWorld::CellStatus GpuWorld::getCellStatusBoundingBox(const double point[3], const double box[3]) {
  uint8_t st = 0;
  check(sspp_gpu_validity(ctx_, point, 1, box, &st), "validity");
  return (CellStatus)st;
}
void GpuWorld::checkSegments(const double* starts, const double* ends, int64_t n, const double box[3], uint8_t* status) {
  check(sspp_gpu_check_segments(ctx_, starts, ends, n, box, status), "check_segments");
}
void GpuWorld::validity(const double* points, int64_t n, const double box[3], uint8_t* status) {
  check(sspp_gpu_validity(ctx_, points, n, box, status), "validity");
}

GpuWorld::ConnectResult GpuWorld::connect(const double* xyz, int64_t n, double radius, const double box[3], bool use_map,
                                          int position_mode, std::vector<int64_t>& row_ptr, std::vector<int32_t>& col) {
  sspp_connect_params p = {};
  p.radius = radius;
  p.box[0] = box[0]; p.box[1] = box[1]; p.box[2] = box[2];
  p.use_map = use_map ? 1 : 0;
  p.position_mode = position_mode;
  p.use_pyramid = 1;
  sspp_connect_stats st;
  check(sspp_gpu_set_nodes(ctx_, xyz, n), "set_nodes");
  check(sspp_gpu_connect(ctx_, &p, &st), "connect");
  row_ptr.assign((size_t)n + 1, 0);
  col.assign((size_t)st.edges, 0);
  check(sspp_gpu_connect_fetch(ctx_, row_ptr.data(), col.empty() ? nullptr : col.data()), "connect_fetch");
  ConnectResult r;
  r.edges = st.edges;
  r.tested_pairs = st.tested_pairs;
  r.in_radius_pairs = st.in_radius_pairs;
  return r;
}

void GpuWorld::setModelPoints(const float* xyz, int64_t n) { check(sspp_gpu_set_model_points(ctx_, xyz, n), "set_model_points"); }

void GpuWorld::visibility(const double* sensor_poses, int64_t nv, double hfov_deg, double vfov_deg, double near_d, double far_d,
                          std::vector<int64_t>& view_ptr, std::vector<int32_t>& point_idx, int64_t* rays) {
  sspp_frustum fr = {hfov_deg, vfov_deg, near_d, far_d};
  sspp_visibility_stats st;
  check(sspp_gpu_visibility(ctx_, sensor_poses, nv, &fr, &st), "visibility");
  view_ptr.assign((size_t)nv + 1, 0);
  point_idx.assign((size_t)st.visible_pairs, 0);
  check(sspp_gpu_visibility_fetch(ctx_, view_ptr.data(), point_idx.empty() ? nullptr : point_idx.data()), "visibility_fetch");
  if (rays) *rays = st.rays;
}
}  // namespace SSPP
