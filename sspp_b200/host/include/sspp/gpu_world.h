// GpuWorld — the world model behind the heuristics: plays the part volumetric_mapping::OctomapManager
// plays in the reference (include/sspp/distance_heuristic.h:37, src/reactive_planner_server.cpp:42,60)
// with the same method names, but every query is a call into the CUDA engine through the C ABI of
// include/sspp_gpu.h.  A non-zero status becomes an SSPPException.
#pragma once
#include <cstdint>
#include <utility>
#include <vector>
#include "ssppexception.h"

struct sspp_gpu_ctx;

namespace SSPP {
class GpuWorld {
 public:
  enum CellStatus { kFree = 0, kOccupied = 1, kUnknown = 2 };  // as WorldBase::CellStatus upstream

  explicit GpuWorld(int device = 0);
  ~GpuWorld();
  GpuWorld(const GpuWorld&) = delete;
  GpuWorld& operator=(const GpuWorld&) = delete;

  // map ingest
  void createMap(double resolution, const int32_t key_min[3], const int32_t dims[3], bool treat_unknown_as_occupied);
  void setTreatUnknownAsOccupied(bool flag);
  void setFree(const double position[3], const double box[3]);
  void setOccupied(const double position[3], const double box[3]);
  // (centre, edge length) leaf lists as getAllOccupiedBoxes / getAllFreeBoxes return them upstream
  void insertBoxes(const std::vector<std::pair<std::vector<double>, double> >& boxes, bool occupied);
  void insertBoxes(const double* centres, const double* sizes, const uint8_t* occupied, int64_t n);
  void uploadDense(const uint32_t* occ_bits, const uint32_t* known_bits);
  // OctomapManager::insertPointcloudWithTf after its TF step: world-frame points, sensor origin, range
  void insertPointcloud(const float* points_world_xyz, int64_t n, const double sensor_origin[3], double sensor_max_range);
  double getResolution() const { return resolution_; }
  bool hasMap() const { return has_map_; }
  // bumped by every call that can change an occupancy answer: results cached from queries (the coverage
  // heuristic's visible sets) are only valid for the generation they were computed in
  uint64_t mapGeneration() const { return map_generation_; }

  // queries
  CellStatus getLineStatusBoundingBox(const double start[3], const double end[3], const double box[3]);
  CellStatus getCellStatusBoundingBox(const double point[3], const double box[3]);
  void checkSegments(const double* starts, const double* ends, int64_t n, const double box[3], uint8_t* status);
  void validity(const double* points, int64_t n, const double box[3], uint8_t* status);

  // graph build: CSR over list indices
  struct ConnectResult { int64_t edges = 0, tested_pairs = 0, in_radius_pairs = 0; };
  ConnectResult connect(const double* xyz, int64_t n, double radius, const double box[3], bool use_map, int position_mode,
                        std::vector<int64_t>& row_ptr, std::vector<int32_t>& col);

  // coverage visibility
  void setModelPoints(const float* xyz, int64_t n);
  void visibility(const double* sensor_poses, int64_t nv, double hfov_deg, double vfov_deg, double near_d, double far_d,
                  std::vector<int64_t>& view_ptr, std::vector<int32_t>& point_idx, int64_t* rays = nullptr);

  sspp_gpu_ctx* handle() { return ctx_; }

 private:
  sspp_gpu_ctx* ctx_ = nullptr;
  double resolution_ = 0;
  bool has_map_ = false;
  uint64_t map_generation_ = 0;
  void check(int rc, const char* where);
};
}  // namespace SSPP
