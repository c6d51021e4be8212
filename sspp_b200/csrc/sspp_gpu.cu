// sspp_gpu.cu — CUDA engine (sm_100a) and C ABI of include/sspp_gpu.h.
//
// Hot path (reference file:line in include/sspp_gpu.h):
//   map ingest      k_upload_boxes / k_set_box / k_dense_in  -> bit planes, k_blocked, k_pyramid
//   neighbour search k_bbox, k_cell_hist, scan, k_cell_scatter  (uniform grid, cell edge >= radius)
//                    k_count                                    (in-radius pairs per source row, written
//                                                                straight into fixed per-row slots)
//   edge test        k_edges   persistent, one work row per warp: pairs ordered by list index, exact
//                              key-box shortcuts per row / pair / ray, queued exact DDA walks
//   CSR              scan + k_compact; in row chunks with the copies to host / peers overlapped
//                    (sspp_gpu_connect_to_host / _to_peers); k_push_csr = compaction result + all-gather
//   node store       k_dedup_*  batched SearchSpace::insertNode
//   queries          k_check_segments, k_validity, k_cast_rays; visibility (k_view_*)
// There is no CPU fallback: every entry point runs these kernels or fails with a status.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/sspp_gpu.h"
#include "engine.cuh"

using namespace sspp;


// one translation unit, split by concern (every part lives in the same anonymous namespace):
#include "context.cuh"
#include "scan.cuh"
#include "map_kernels.cuh"
#include "connect_kernels.cuh"
#include "query_kernels.cuh"
#include "visibility_kernels.cuh"
#include "dedup_kernels.cuh"

namespace {
// ------------------------------------------------------------------------------------------------
// host helpers
// ------------------------------------------------------------------------------------------------
// largest double t with sqrt(t) <= radius (IEEE sqrt on the host = __dsqrt_rn on the device)
double radius_threshold(double radius) {
  if (!(radius >= 0.0)) return -1.0;  // sqrt(d2) >= 0 > radius: no pair connects
  if (std::isinf(radius)) return radius;
  double t = radius * radius;
  while (t > 0.0 && sqrt(t) > radius) t = nextafter(t, 0.0);
  for (;;) {
    double u = nextafter(t, INFINITY);
    if (std::isinf(u) || sqrt(u) > radius) break;
    t = u;
  }
  return t;
}

int ensure_planes(sspp_gpu_ctx* ctx) {
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map: call sspp_gpu_map_create first");
  if (!ctx->planes_dirty) return SSPP_OK;
  k_blocked<<<148 * 8, 256, 0, ctx->stream>>>(ctx->occ.as<uint32_t>(), ctx->known.as<uint32_t>(), ctx->blocked.as<uint32_t>(),
                                              ctx->map_words, ctx->map.treat_unknown);
  LAUNCHED(ctx);
  k_pyramid<<<148 * 4, 256, 0, ctx->stream>>>(ctx->map, ctx->blocked.as<uint32_t>(), ctx->cblocked.as<uint32_t>());
  LAUNCHED(ctx);
  ctx->planes_dirty = false;
  return SSPP_OK;
}

int build_lattice(sspp_gpu_ctx* ctx, const double box[3], RayLattice& lat) {
  for (int a = 0; a < 3; ++a) {
    if (!(box[a] >= 0.0) || !std::isfinite(box[a])) return fail(ctx, SSPP_ERR_INVALID, "box size must be finite and >= 0");
    int n = ray_lattice_axis(box[a], ctx->map.ks.res, lat.off[a], SSPP_MAX_LATTICE);
    if (n < 0) return fail(ctx, SSPP_ERR_RANGE, "ray lattice exceeds 64 offsets per axis (box / resolution too large)");
    lat.n[a] = n;
    for (int i = n; i < SSPP_MAX_LATTICE; ++i) lat.off[a][i] = 0.0;
  }
  lat.rays = lat.n[0] * lat.n[1] * lat.n[2];
  return SSPP_OK;
}

int upload_lattice(sspp_gpu_ctx* ctx, const RayLattice& lat) {
  // the table only changes with the box or the resolution: a repeated pass finds it on the device already
  if (ctx->lattice_valid && ctx->lattice.p && memcmp(&ctx->lattice_host, &lat, sizeof(RayLattice)) == 0) return SSPP_OK;
  ctx->lattice_valid = false;
  CK(ctx->lattice.reserve(sizeof(RayLattice)));
  // `lat` is pageable host memory: the call returns once it has been copied to the driver's staging
  // buffer, so the caller's stack object may go away afterwards
  CK(cudaMemcpyAsync(ctx->lattice.p, &lat, sizeof(RayLattice), cudaMemcpyHostToDevice, ctx->stream));
  memcpy(&ctx->lattice_host, &lat, sizeof(RayLattice));
  ctx->lattice_valid = true;
  return SSPP_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

int sspp_gpu_create(sspp_gpu_ctx** out, int device) {
  if (!out) return SSPP_ERR_INVALID;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || device < 0 || device >= count) return SSPP_ERR_CUDA;
  if (cudaSetDevice(device) != cudaSuccess) return SSPP_ERR_CUDA;
  sspp_gpu_ctx* ctx = new sspp_gpu_ctx();
  ctx->device = device;
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return SSPP_ERR_CUDA; }
  for (int i = 0; i < 8; ++i)
    if (cudaEventCreate(&ctx->ev[i]) != cudaSuccess) { delete ctx; return SSPP_ERR_CUDA; }
  memset(&ctx->map, 0, sizeof(ctx->map));
  if (cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || ctx->sm_count < 1)
    ctx->sm_count = 148;
  *out = ctx;
  return SSPP_OK;
}

void sspp_gpu_destroy(sspp_gpu_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  DevBuf* bufs[] = {&ctx->occ, &ctx->known, &ctx->blocked, &ctx->cblocked, &ctx->map_stage, &ctx->logodds, &ctx->scan_free,
                    &ctx->scan_occ, &ctx->pc_pts, &ctx->pc_keys, &ctx->pc_vals, &ctx->xyz, &ctx->bbox, &ctx->vis_bbox,
                    &ctx->cell_of, &ctx->cell_start, &ctx->cell_cursor, &ctx->sorted_idx, &ctx->sx, &ctx->sy, &ctx->sz,
                    &ctx->rows, &ctx->row_flag, &ctx->row_free, &ctx->row_class, &ctx->cand_count, &ctx->cand_off, &ctx->candA, &ctx->candB,
                    &ctx->acc, &ctx->deg, &ctx->row_ptr, &ctx->col_idx, &ctx->counters, &ctx->scan_tmp, &ctx->lattice, &ctx->work_rows,
                    &ctx->q0, &ctx->q1, &ctx->q2, &ctx->q3, &ctx->q4, &ctx->pts, &ctx->vis_cnt, &ctx->vis_ptr,
                    &ctx->vis_idx, &ctx->vis_poses, &ctx->vis_counters, &ctx->vis_words, &ctx->vis_mask, &ctx->pts_cell_start, &ctx->pts_sorted,
                    &ctx->vis_views, &ctx->vis_rayoff, &ctx->vis_rayview, &ctx->vis_raypt, &ctx->vis_flag, &ctx->dd_poses, &ctx->dd_xyz, &ctx->dd_state,
                    &ctx->dd_pend, &ctx->dd_dup, &ctx->dd_keep, &ctx->dd_misc};
  for (DevBuf* b : bufs) b->release();
  ctx->chunk_base.release();
  for (int c = 0; c < 8; ++c) if (ctx->chunk_ev[c]) cudaEventDestroy(ctx->chunk_ev[c]);
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  if (ctx->push_stream) cudaStreamDestroy(ctx->push_stream);
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  for (void* q : ctx->peer_opened) cudaIpcCloseMemHandle(q);
  for (void* q : ctx->peer_created) cudaFree(q);
  for (int i = 0; i < 8; ++i) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
  for (int i = 0; i < 6; ++i) if (ctx->vev[i]) cudaEventDestroy(ctx->vev[i]);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

const char* sspp_gpu_last_error(const sspp_gpu_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }
void* sspp_gpu_stream(sspp_gpu_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int64_t sspp_gpu_launch_count(const sspp_gpu_ctx* ctx) { return ctx ? ctx->launches : 0; }

int sspp_gpu_synchronize(sspp_gpu_ctx* ctx) {
  if (!ctx) return SSPP_ERR_INVALID;
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

// ---- map -------------------------------------------------------------------------------------
int sspp_gpu_map_create(sspp_gpu_ctx* ctx, double resolution, const int32_t key_min[3], const int32_t dims[3],
                        int treat_unknown_as_occupied) {
  if (!ctx || !key_min || !dims) return fail(ctx, SSPP_ERR_INVALID, "map_create: null argument");
  if (!(resolution > 0.0) || !std::isfinite(resolution)) return fail(ctx, SSPP_ERR_INVALID, "map_create: resolution must be > 0");
  for (int a = 0; a < 3; ++a) {
    if (dims[a] <= 0) return fail(ctx, SSPP_ERR_INVALID, "map_create: dims must be positive");
    if (key_min[a] < 0 || (int64_t)key_min[a] + dims[a] > 65536)
      return fail(ctx, SSPP_ERR_RANGE, "map_create: window must lie inside the 16-bit key space [0,65536)");
  }
  CK(cudaSetDevice(ctx->device));
  DeviceMap& m = ctx->map;
  m.ks.res = resolution;
  m.ks.res_factor = 1.0 / resolution;
  m.ks.half_res_f = (float)(resolution * 0.5);
  m.ks.half_res_d = resolution * 0.5;
  for (int a = 0; a < 3; ++a) { m.kmin[a] = key_min[a]; ctx->user_dims[a] = dims[a]; }
  m.dims[0] = (dims[0] + 127) / 128 * 128;
  m.dims[1] = (dims[1] + 3) / 4 * 4;
  m.dims[2] = (dims[2] + 3) / 4 * 4;
  m.nxw = m.dims[0] / 32;
  for (int a = 0; a < 3; ++a) m.cdims[a] = m.dims[a] / 4;
  m.cnxw = m.cdims[0] / 32;
  m.treat_unknown = treat_unknown_as_occupied ? 1 : 0;
  ctx->map_words = (size_t)m.nxw * m.dims[1] * m.dims[2];
  ctx->cmap_words = (size_t)m.cnxw * m.cdims[1] * m.cdims[2];
  CK(ctx->occ.reserve(ctx->map_words * 4));
  CK(ctx->known.reserve(ctx->map_words * 4));
  CK(ctx->blocked.reserve(ctx->map_words * 4));
  CK(ctx->cblocked.reserve(ctx->cmap_words * 4));
  CK(cudaMemsetAsync(ctx->occ.p, 0, ctx->map_words * 4, ctx->stream));
  CK(cudaMemsetAsync(ctx->known.p, 0, ctx->map_words * 4, ctx->stream));
  m.occ = ctx->occ.as<uint32_t>();
  m.known = ctx->known.as<uint32_t>();
  m.blocked = ctx->blocked.as<uint32_t>();
  m.cblocked = ctx->cblocked.as<uint32_t>();
  ctx->has_map = true;
  ctx->planes_dirty = true;
  ctx->has_logodds = false;  // a new map starts without a log-odds grid (allocated by the first point cloud)
  return SSPP_OK;
}

int sspp_gpu_map_set_treat_unknown(sspp_gpu_ctx* ctx, int flag) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  int v = flag ? 1 : 0;
  if (v != ctx->map.treat_unknown) { ctx->map.treat_unknown = v; ctx->planes_dirty = true; }
  return SSPP_OK;
}

int sspp_gpu_map_upload_boxes(sspp_gpu_ctx* ctx, const double* centres, const double* sizes, const uint8_t* occupied,
                              int64_t n) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  if (n < 0 || (n > 0 && (!centres || !sizes || !occupied))) return fail(ctx, SSPP_ERR_INVALID, "upload_boxes: null argument");
  if (n == 0) return SSPP_OK;
  CK(cudaSetDevice(ctx->device));
  CK(ctx->q0.reserve((size_t)n * 24));
  CK(ctx->q1.reserve((size_t)n * 8));
  CK(ctx->q2.reserve((size_t)n));
  CK(cudaMemcpyAsync(ctx->q0.p, centres, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->q1.p, sizes, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->q2.p, occupied, (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
  k_upload_boxes<<<ceil_div(n * 32, 256), 256, 0, ctx->stream>>>(ctx->map, ctx->occ.as<uint32_t>(), ctx->known.as<uint32_t>(),
                                                               ctx->q0.as<double>(), ctx->q1.as<double>(),
                                                               ctx->q2.as<uint8_t>(), n,
                                                               ctx->has_logodds ? ctx->logodds.as<float>() : nullptr,
                                                               ctx->clamp_min_log, ctx->clamp_max_log);
  LAUNCHED(ctx);
  CK(cudaStreamSynchronize(ctx->stream));
  ctx->planes_dirty = true;
  return SSPP_OK;
}

int sspp_gpu_map_set_box(sspp_gpu_ctx* ctx, const double position[3], const double box[3], int occupied) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  if (!position || !box) return fail(ctx, SSPP_ERR_INVALID, "set_box: null argument");
  CK(cudaSetDevice(ctx->device));
  // upstream OctomapWorld::setLogOddsBoundingBox: lattice from min-eps while <= max+eps step res,
  // each point narrowed to float (octomap::point3d) before coordToKey
  const double epsilon = 0.001;
  const double res = ctx->map.ks.res;
  std::vector<int> keys[3];
  for (int a = 0; a < 3; ++a) {
    if (!std::isfinite(position[a]) || !std::isfinite(box[a])) return fail(ctx, SSPP_ERR_INVALID, "set_box: non-finite value");
    double lo = position[a] - box[a] / 2 - epsilon, hi = position[a] + box[a] / 2 + epsilon;
    if ((hi - lo) / res > 70000.0) return fail(ctx, SSPP_ERR_RANGE, "set_box: box spans more than the key space");
    for (double v = lo; v <= hi; v += res) {
      int k;
      if (coord_to_key(ctx->map.ks, (double)(float)v, k)) keys[a].push_back(k - ctx->map.kmin[a]);
    }
    std::sort(keys[a].begin(), keys[a].end());
    keys[a].erase(std::unique(keys[a].begin(), keys[a].end()), keys[a].end());
    if (keys[a].empty()) return SSPP_OK;
  }
  size_t nx = keys[0].size(), ny = keys[1].size(), nz = keys[2].size();
  CK(ctx->q0.reserve((nx + ny + nz) * sizeof(int)));
  int* d = ctx->q0.as<int>();
  CK(cudaMemcpyAsync(d, keys[0].data(), nx * 4, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d + nx, keys[1].data(), ny * 4, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d + nx + ny, keys[2].data(), nz * 4, cudaMemcpyHostToDevice, ctx->stream));
  int64_t total = (int64_t)nx * ny * nz;
  int blocks = (int)std::min<int64_t>(ceil_div(total, 256), 148 * 16);
  k_set_box<<<blocks, 256, 0, ctx->stream>>>(ctx->map, ctx->occ.as<uint32_t>(), ctx->known.as<uint32_t>(), d, (int)nx, d + nx,
                                             (int)ny, d + nx + ny, (int)nz, occupied ? 1 : 0,
                                             ctx->has_logodds ? ctx->logodds.as<float>() : nullptr, ctx->clamp_min_log,
                                             ctx->clamp_max_log);
  LAUNCHED(ctx);
  CK(cudaStreamSynchronize(ctx->stream));
  ctx->planes_dirty = true;
  return SSPP_OK;
}

int sspp_gpu_map_upload_dense(sspp_gpu_ctx* ctx, const uint32_t* occ_bits, const uint32_t* known_bits) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  if (!occ_bits || !known_bits) return fail(ctx, SSPP_ERR_INVALID, "upload_dense: null argument");
  CK(cudaSetDevice(ctx->device));
  const int* ud = ctx->user_dims;
  int wpr = (ud[0] + 31) / 32;
  size_t words = (size_t)wpr * ud[1] * ud[2];
  CK(ctx->map_stage.reserve(words * 4));
  const DeviceMap& m = ctx->map;
  const uint32_t* src[2] = {occ_bits, known_bits};
  uint32_t* dst[2] = {ctx->occ.as<uint32_t>(), ctx->known.as<uint32_t>()};
  for (int p = 0; p < 2; ++p) {
    CK(cudaMemcpyAsync(ctx->map_stage.p, src[p], words * 4, cudaMemcpyHostToDevice, ctx->stream));
    k_dense_in<<<148 * 8, 256, 0, ctx->stream>>>(ctx->map_stage.as<uint32_t>(), wpr, ud[1], ud[2], dst[p], m.nxw, m.dims[1],
                                                 m.dims[2], ud[0]);
    LAUNCHED(ctx);
  }
  if (ctx->has_logodds) {  // a full upload replaces the map: the log-odds grid restarts from the new planes
    k_logodds_init<<<148 * 8, 256, 0, ctx->stream>>>(ctx->occ.as<uint32_t>(), ctx->known.as<uint32_t>(), ctx->logodds.as<float>(),
                                                     ctx->map_words, ctx->clamp_min_log, ctx->clamp_max_log);
    LAUNCHED(ctx);
  }
  CK(cudaStreamSynchronize(ctx->stream));
  ctx->planes_dirty = true;
  return SSPP_OK;
}

int sspp_gpu_map_download_dense(sspp_gpu_ctx* ctx, uint32_t* occ_bits, uint32_t* known_bits) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  if (!occ_bits || !known_bits) return fail(ctx, SSPP_ERR_INVALID, "download_dense: null argument");
  CK(cudaSetDevice(ctx->device));
  const int* ud = ctx->user_dims;
  int wpr = (ud[0] + 31) / 32;
  size_t words = (size_t)wpr * ud[1] * ud[2];
  CK(ctx->map_stage.reserve(words * 4));
  const DeviceMap& m = ctx->map;
  uint32_t* dst[2] = {occ_bits, known_bits};
  const uint32_t* src[2] = {ctx->occ.as<uint32_t>(), ctx->known.as<uint32_t>()};
  for (int p = 0; p < 2; ++p) {
    k_dense_out<<<148 * 8, 256, 0, ctx->stream>>>(src[p], m.nxw, m.dims[1], m.dims[2], ctx->map_stage.as<uint32_t>(), wpr,
                                                  ud[1], ud[2], ud[0]);
    LAUNCHED(ctx);
    CK(cudaMemcpyAsync(dst[p], ctx->map_stage.p, words * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  return SSPP_OK;
}

// ---- incremental update ----------------------------------------------------------------------
static float logodds_f(double p) { return (float)log(p / (1 - p)); }  // octomap::logodds

static void default_sensor_model(sspp_gpu_ctx* ctx) {
  ctx->prob_hit_log = logodds_f(0.65);
  ctx->prob_miss_log = logodds_f(0.4);
  ctx->clamp_min_log = logodds_f(0.12);
  ctx->clamp_max_log = logodds_f(0.97);
  ctx->occ_thres_log = logodds_f(0.7);
}

int sspp_gpu_map_set_sensor_model(sspp_gpu_ctx* ctx, const sspp_sensor_model* model) {
  if (!ctx || !model) return fail(ctx, SSPP_ERR_INVALID, "set_sensor_model: null argument");
  const double v[5] = {model->probability_hit, model->probability_miss, model->threshold_min, model->threshold_max,
                       model->threshold_occupancy};
  for (int i = 0; i < 5; ++i)
    if (!(v[i] > 0.0 && v[i] < 1.0)) return fail(ctx, SSPP_ERR_INVALID, "set_sensor_model: probabilities must lie in (0,1)");
  if (ctx->has_logodds) return fail(ctx, SSPP_ERR_STATE, "set_sensor_model: call before the first point cloud is inserted");
  ctx->prob_hit_log = logodds_f(v[0]);
  ctx->prob_miss_log = logodds_f(v[1]);
  ctx->clamp_min_log = logodds_f(v[2]);
  ctx->clamp_max_log = logodds_f(v[3]);
  ctx->occ_thres_log = logodds_f(v[4]);
  return SSPP_OK;
}

// the log-odds grid appears with the first cloud, seeded from the planes (known-free / occupied voxels
// sit at the clamping values, as setFree / setOccupied leave them)
static int ensure_logodds(sspp_gpu_ctx* ctx) {
  if (ctx->has_logodds) return SSPP_OK;
  if (ctx->prob_hit_log == 0 && ctx->prob_miss_log == 0) default_sensor_model(ctx);
  CK(ctx->logodds.reserve(ctx->map_words * 32 * sizeof(float)));
  CK(ctx->scan_free.reserve(ctx->map_words * 4));
  CK(ctx->scan_occ.reserve(ctx->map_words * 4));
  CK(cudaMemsetAsync(ctx->scan_free.p, 0, ctx->map_words * 4, ctx->stream));
  CK(cudaMemsetAsync(ctx->scan_occ.p, 0, ctx->map_words * 4, ctx->stream));
  k_logodds_init<<<148 * 8, 256, 0, ctx->stream>>>(ctx->occ.as<uint32_t>(), ctx->known.as<uint32_t>(), ctx->logodds.as<float>(),
                                                   ctx->map_words, ctx->clamp_min_log, ctx->clamp_max_log);
  LAUNCHED(ctx);
  ctx->has_logodds = true;
  return SSPP_OK;
}

int sspp_gpu_map_insert_pointcloud(sspp_gpu_ctx* ctx, const float* pts, int64_t n, const double origin[3], double max_range) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  if (n < 0 || !origin || (n > 0 && !pts)) return fail(ctx, SSPP_ERR_INVALID, "insert_pointcloud: null argument");
  if (n >= (int64_t)0x7fffffff) return fail(ctx, SSPP_ERR_RANGE, "insert_pointcloud: too many points");
  if (!std::isfinite(origin[0]) || !std::isfinite(origin[1]) || !std::isfinite(origin[2]) || std::isnan(max_range))
    return fail(ctx, SSPP_ERR_INVALID, "insert_pointcloud: non-finite sensor origin or range");
  CK(cudaSetDevice(ctx->device));
  int rc = ensure_logodds(ctx);
  if (rc) return rc;
  if (n == 0) return SSPP_OK;
  cudaStream_t st = ctx->stream;
  unsigned long long slots = 1024;
  while (slots < (unsigned long long)n * 2 + 1) slots <<= 1;
  CK(ctx->pc_pts.reserve((size_t)n * 12));
  CK(ctx->pc_keys.reserve((size_t)slots * 8));
  CK(ctx->pc_vals.reserve((size_t)slots * 4));
  CK(cudaMemcpyAsync(ctx->pc_pts.p, pts, (size_t)n * 12, cudaMemcpyHostToDevice, st));
  CK(cudaMemsetAsync(ctx->pc_keys.p, 0xff, (size_t)slots * 8, st));
  CK(cudaMemsetAsync(ctx->pc_vals.p, 0x7f, (size_t)slots * 4, st));
  const float ox = (float)origin[0], oy = (float)origin[1], oz = (float)origin[2];
  k_pc_first<<<ceil_div(n, 256), 256, 0, st>>>(ctx->map.ks, ctx->pc_pts.as<float>(), n, ox, oy, oz, max_range,
                                              ctx->pc_keys.as<unsigned long long>(), ctx->pc_vals.as<int>(), slots - 1);
  LAUNCHED(ctx);
  k_pc_cast<<<ceil_div(n, 128), 128, 0, st>>>(ctx->map, ctx->pc_pts.as<float>(), n, ox, oy, oz, max_range,
                                             ctx->pc_keys.as<unsigned long long>(), ctx->pc_vals.as<int>(), slots - 1,
                                             ctx->scan_free.as<uint32_t>(), ctx->scan_occ.as<uint32_t>());
  LAUNCHED(ctx);
  k_pc_apply<<<148 * 8, 256, 0, st>>>(ctx->occ.as<uint32_t>(), ctx->known.as<uint32_t>(), ctx->scan_free.as<uint32_t>(),
                                      ctx->scan_occ.as<uint32_t>(), ctx->logodds.as<float>(), ctx->map_words, ctx->prob_hit_log,
                                      ctx->prob_miss_log, ctx->clamp_min_log, ctx->clamp_max_log, ctx->occ_thres_log);
  LAUNCHED(ctx);
  CK(cudaStreamSynchronize(st));  // the host cloud may be reused on return
  ctx->planes_dirty = true;
  return SSPP_OK;
}

int sspp_gpu_map_download_logodds(sspp_gpu_ctx* ctx, float* out) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  if (!out) return fail(ctx, SSPP_ERR_INVALID, "download_logodds: null argument");
  CK(cudaSetDevice(ctx->device));
  int rc = ensure_logodds(ctx);
  if (rc) return rc;
  const int* ud = ctx->user_dims;
  size_t total = (size_t)ud[0] * ud[1] * ud[2];
  CK(ctx->map_stage.reserve(total * 4));
  k_logodds_out<<<148 * 8, 256, 0, ctx->stream>>>(ctx->logodds.as<float>(), ctx->known.as<uint32_t>(), ctx->map.nxw, ctx->map.dims[1],
                                                  ctx->map.dims[2], ctx->map_stage.as<float>(), ud[0], ud[1], ud[2]);
  LAUNCHED(ctx);
  CK(cudaMemcpyAsync(out, ctx->map_stage.p, total * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

// ---- nodes -----------------------------------------------------------------------------------
static int set_nodes_impl(sspp_gpu_ctx* ctx, const double* xyz, int64_t n, cudaMemcpyKind kind) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (n < 0 || (n > 0 && !xyz)) return fail(ctx, SSPP_ERR_INVALID, "set_nodes: null argument");
  if (n >= (int64_t)0x7fffffff) return fail(ctx, SSPP_ERR_RANGE, "set_nodes: node indices must fit in int32");
  CK(cudaSetDevice(ctx->device));
  CK(ctx->xyz.reserve((size_t)std::max<int64_t>(n, 1) * 24));
  if (n) CK(cudaMemcpyAsync(ctx->xyz.p, xyz, (size_t)n * 24, kind, ctx->stream));
  ctx->n_nodes = n;
  ctx->has_result = false;
  // the bounding box (and the finiteness check) of the node set belongs to the node store: queued
  // here, read by the first connect that needs it
  ctx->bbox_valid = false;
  ctx->bbox_pending = false;
  ctx->shard_lo = ctx->shard_hi = -1;
  if (n) {
    CK(ctx->bbox.reserve(14 * 8));
    static const unsigned long long init[7] = {~0ull, ~0ull, ~0ull, 0, 0, 0, 0};
    CK(cudaMemcpyAsync(ctx->bbox.p, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
    k_bbox<<<std::min(ceil_div(n, 256), 148 * 8), 256, 0, ctx->stream>>>(ctx->xyz.as<double>(), n, ctx->bbox.as<unsigned long long>());
    LAUNCHED(ctx);
    ctx->bbox_pending = true;
  }
  return SSPP_OK;
}
int sspp_gpu_set_nodes(sspp_gpu_ctx* ctx, const double* xyz, int64_t n) {
  int rc = set_nodes_impl(ctx, xyz, n, cudaMemcpyHostToDevice);
  if (rc == SSPP_OK) CK(cudaStreamSynchronize(ctx->stream));  // the host buffer may be reused on return
  return rc;
}
int sspp_gpu_set_nodes_device(sspp_gpu_ctx* ctx, const double* xyz_device, int64_t n) {
  return set_nodes_impl(ctx, xyz_device, n, cudaMemcpyDeviceToDevice);
}

// ---- connect ---------------------------------------------------------------------------------
namespace {
// Host buffers of sspp_gpu_connect_nodes_host: the connection pass then runs the edge test and the
// compaction in row chunks and sends every finished chunk home on a second stream while the next
// one is computed.
struct HostSink {
  int64_t* row_ptr;
  int32_t* col_idx;
  int64_t col_capacity;
  bool delivered;  // out: the CSR is already in the host buffers / in the peers' block slots
  // multi-GPU variant (sspp_gpu_connect_to_peers): the chunks go to slot [rank] of every rank's block
  // buffers over NVLink (copy engines) instead of to the host; col_capacity = the slot size then
  const sspp_peer_blocks* peers;
  // exchange fused into the pass (sspp_gpu_connect_fused): rows go to all ranks' slot-format buffers from the edge kernel
  const sspp_fused_peers* fused;
};

// Order in which the edge kernel takes the rays of a pair (ConnectArgs::ray_stride): position t of a turn is ray
// (t * stride) % rays.  The stride is about a third of the lattice and coprime with its size, so every ray is taken
// once and the first few already sample the whole box.  SSPP_RAY_STRIDE=1 keeps the lattice order.
static int ray_order_stride(int rays) {
  if (const char* e = getenv("SSPP_RAY_STRIDE")) {
    int v = atoi(e);
    if (v == 1) return 1;
  }
  if (rays < 3 || rays > 1024) return 1;
  int s = (rays * 3 + 9) / 10;
  if (s < 2) s = 2;
  auto gcd = [](int x, int y) { while (y) { int t = x % y; x = y; y = t; } return x; };
  while (gcd(s, rays) != 1) ++s;
  return s < rays ? s : 1;
}

int connect_impl(sspp_gpu_ctx* ctx, const sspp_connect_params* prm, sspp_connect_stats* stats, HostSink* sink);
}  // namespace

int sspp_gpu_connect(sspp_gpu_ctx* ctx, const sspp_connect_params* prm, sspp_connect_stats* stats) {
  return connect_impl(ctx, prm, stats, nullptr);
}

namespace {
int connect_impl(sspp_gpu_ctx* ctx, const sspp_connect_params* prm, sspp_connect_stats* stats, HostSink* sink) {
  if (!ctx || !prm) return fail(ctx, SSPP_ERR_INVALID, "connect: null argument");
  if (!std::isfinite(prm->radius)) return fail(ctx, SSPP_ERR_INVALID, "connect: radius must be finite");
  if (prm->position_mode != 0 && prm->position_mode != 1) return fail(ctx, SSPP_ERR_INVALID, "connect: position_mode must be 0 or 1");
  const int64_t n = ctx->n_nodes;
  int64_t lo = prm->row_begin, hi = prm->row_end;
  if (lo == 0 && hi == 0) hi = n;
  if (lo < 0 || hi < lo || hi > n) return fail(ctx, SSPP_ERR_INVALID, "connect: row range outside [0, n]");
  const int64_t nrows = hi - lo;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  ctx->has_result = false;
  if (stats) memset(stats, 0, sizeof(*stats));

  RayLattice lat;
  memset(&lat, 0, sizeof(lat));
  lat.rays = 1;
  RowReach ts;
  ts.reach[0] = ts.reach[1] = ts.reach[2] = 0;
  const int use_map = prm->use_map ? 1 : 0;
  if (use_map) {
    int rc = ensure_planes(ctx);
    if (rc) return rc;
    rc = build_lattice(ctx, prm->box, lat);
    if (rc) return rc;
    rc = upload_lattice(ctx, lat);
    if (rc) return rc;
    const double rad = std::max(prm->radius, 0.0);
    for (int a = 0; a < 3; ++a) {
      ts.reach[a] = rad + prm->box[a] * 0.5;
    }
  }
  if (stats) stats->rays_per_pair = use_map ? lat.rays : 0;

  CK(cudaEventRecord(ctx->ev[0], st));
  // --- empty input: an empty CSR
  CK(ctx->row_ptr.reserve((size_t)(nrows + 1) * 8));
  if (n == 0 || nrows == 0) {
    CK(cudaMemsetAsync(ctx->row_ptr.p, 0, (size_t)(nrows + 1) * 8, st));
    CK(ctx->col_idx.reserve(4));
    CK(cudaStreamSynchronize(st));
    ctx->has_result = true;
    ctx->res_rows = nrows;
    ctx->res_edges = 0;
    if (stats) stats->rows = nrows;
    for (int i = 0; i < 8; ++i) ctx->timings[i] = 0;
    return SSPP_OK;
  }

  // --- 1. bounding box: of all nodes (queued by set_nodes, with the finiteness check) and, for a
  // shard, of its own rows; both are cached until the node set changes, so a repeated pass over the
  // same resident nodes does not wait for the host here
  const bool sharded = nrows != n;
  const bool need_shard_bbox = sharded && !(ctx->shard_lo == lo && ctx->shard_hi == hi);
  if (!ctx->bbox_valid || need_shard_bbox) {
    if (!ctx->bbox_valid && !ctx->bbox_pending) return fail(ctx, SSPP_ERR_STATE, "connect: no nodes set");
    if (need_shard_bbox) {
      static const unsigned long long init[7] = {~0ull, ~0ull, ~0ull, 0, 0, 0, 0};
      CK(cudaMemcpyAsync(ctx->bbox.as<unsigned long long>() + 7, init, sizeof(init), cudaMemcpyHostToDevice, st));
      k_bbox<<<std::min(ceil_div(nrows, 256), 148 * 8), 256, 0, st>>>(ctx->xyz.as<double>() + 3 * lo, nrows,
                                                                     ctx->bbox.as<unsigned long long>() + 7);
      LAUNCHED(ctx);
    }
    unsigned long long tmp[14];
    CK(cudaMemcpyAsync(tmp, ctx->bbox.p, sizeof(tmp), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (!ctx->bbox_valid) {
      memcpy(ctx->bbox_host, tmp, sizeof(ctx->bbox_host));
      ctx->bbox_valid = true;
      ctx->bbox_pending = false;
    }
    if (need_shard_bbox) {
      memcpy(ctx->shard_bbox_host, tmp + 7, sizeof(ctx->shard_bbox_host));
      ctx->shard_lo = lo;
      ctx->shard_hi = hi;
    }
  }
  unsigned long long hb[14];
  memcpy(hb, ctx->bbox_host, sizeof(ctx->bbox_host));
  memcpy(hb + 7, ctx->shard_bbox_host, sizeof(ctx->shard_bbox_host));
  if (hb[6]) return fail(ctx, SSPP_ERR_INVALID, "connect: node coordinates must be finite");

  // cell edge slightly above the radius: two nodes within `radius` of each other are then never
  // more than one cell apart, whatever the rounding of the cell index
  NodeGrid g;
  double ext[3];
  double cell = std::max(prm->radius, 0.0) * (1.0 + 1e-6);
  {
    double full_ext = 0;
    for (int a = 0; a < 3; ++a) full_ext = std::max(full_ext, dec_f64(hb[3 + a]) - dec_f64(hb[a]));
    if (!(cell > 0.0)) cell = std::max(full_ext, 1.0);
    cell = std::max(cell, full_ext * 1e-9);
  }
  for (int a = 0; a < 3; ++a) {
    if (sharded) {
      // a shard bins only what its rows can reach: its own bounding box grown by one cell (>= radius)
      const double grow = cell * 1.001;
      g.min[a] = std::max(dec_f64(hb[a]), dec_f64(hb[7 + a]) - grow);
      g.max[a] = std::min(dec_f64(hb[3 + a]), dec_f64(hb[10 + a]) + grow);
    } else {
      g.min[a] = dec_f64(hb[a]);
      g.max[a] = dec_f64(hb[3 + a]);
    }
    ext[a] = g.max[a] - g.min[a];
  }
  for (;;) {
    double total = 1;
    for (int a = 0; a < 3; ++a) total *= floor(ext[a] / cell) + 2.0;
    if (total <= (double)(1 << 25)) break;
    cell *= 1.26;
  }
  g.inv_cell = 1.0 / cell;
  int64_t ncells = 1;
  for (int a = 0; a < 3; ++a) { g.n[a] = (int)floor(ext[a] * g.inv_cell) + 1; ncells *= g.n[a]; }

  // --- 2. bin nodes into cells
  CK(ctx->cell_of.reserve((size_t)n * 4));
  CK(ctx->cell_start.reserve((size_t)(ncells + 1) * 4));
  CK(ctx->cell_cursor.reserve((size_t)(ncells + 1) * 4));
  CK(ctx->sx.reserve((size_t)n * sizeof(NodeRec)));
  CK(cudaMemsetAsync(ctx->cell_cursor.p, 0, (size_t)(ncells + 1) * 4, st));
  k_cell_hist<<<ceil_div(n, 256), 256, 0, st>>>(ctx->xyz.as<double>(), n, g, ctx->cell_of.as<int>(), ctx->cell_cursor.as<int>());
  LAUNCHED(ctx);
  {
    int rc = scan_exclusive<int>(ctx, ctx->cell_cursor.as<int>(), ncells, ctx->cell_start.as<int>());
    if (rc) return rc;
  }
  CK(cudaMemcpyAsync(ctx->cell_cursor.p, ctx->cell_start.p, (size_t)ncells * 4, cudaMemcpyDeviceToDevice, st));
  k_cell_scatter<<<ceil_div(n, 256), 256, 0, st>>>(ctx->xyz.as<double>(), n, ctx->cell_of.as<int>(), ctx->cell_cursor.as<int>(),
                                                   ctx->sx.as<NodeRec>());
  LAUNCHED(ctx);
  k_cell_sort<<<ceil_div(ncells, 256), 256, 0, st>>>(ctx->cell_start.as<int>(), (int)ncells, ctx->sx.as<NodeRec>());
  LAUNCHED(ctx);

  CK(cudaEventRecord(ctx->ev[1], st));

  // --- 3. count in-radius pairs per row
  ConnectArgs a;
  a.g = g;
  a.cell_start = ctx->cell_start.as<int>();
  a.sn = ctx->sx.as<NodeRec>();
  a.n_sorted = sharded ? ctx->cell_start.as<int>() + ncells : nullptr;
  a.nscan = (int)n;
  a.nrows = (int)nrows;
  a.lo = lo;
  a.hi = hi;
  a.r2max = radius_threshold(prm->radius);
  a.mode = prm->position_mode;
  a.use_map = use_map;
  const bool count_steps = use_map && prm->count_steps;
  a.use_pyramid = (prm->use_pyramid && !count_steps) ? 1 : 0;
  a.ts = ts;
  a.ray_stride = ray_order_stride(use_map ? lat.rays : 1);
  a.ray_magic = a.ray_stride > 1 ? 0xffffffffu / (unsigned)lat.rays + 1u : 0u;
  CK(ctx->cand_count.reserve((size_t)nrows * 4));
  CK(ctx->row_free.reserve((size_t)nrows));
  CK(ctx->row_class.reserve((size_t)nrows));
  a.row_class = ctx->row_class.as<uint8_t>();
  CK(ctx->cand_off.reserve((size_t)(nrows + 1) * 8));
  CK(ctx->counters.reserve(48 * 8));
  CK(cudaMemsetAsync(ctx->counters.p, 0, 48 * 8, st));
  // row chunks: only when the CSR goes straight to host buffers and the pass is large enough to pipeline
  const sspp_fused_peers* fz = sink ? sink->fused : nullptr;
  a.fused = fz ? 1 : 0;
  a.fworld = fz ? fz->world : 0;
  a.frow0 = fz ? (long long)fz->first_row : 0;
  a.row_free_dev = ctx->row_free.as<uint8_t>();
  for (int k = 0; k < SSPP_MAX_PEERS; ++k) {
    a.fslot[k] = (fz && k < fz->world) ? fz->slot_peers[k] : nullptr;
    a.fdeg[k] = (fz && k < fz->world) ? fz->deg_peers[k] : nullptr;
  }
  int K = 1;
  if (sink && !fz && nrows >= (1 << 16)) {
    // host sink: with the edge test at ~1.2 ms the call is bound by the D2H copy of the CSR (71 MB at S4a), so the sooner
    // the first chunk is final the better: six chunks (6:5:4:3:2:1) measured best (1 / 2 / 4 / 6 / 8: 3.61 / 3.17 / 2.90 /
    // 2.82 / 2.89 ms end to end); the peer variant keeps four
    K = sink->peers ? 4 : 6;
    if (const char* env = getenv("SSPP_HOST_CHUNKS")) K = std::min(std::max(atoi(env), 1), (int)SSPP_MAX_CHUNKS);
  }
  const bool chunked = K > 1 || (sink && sink->peers);  // the peer variant always delivers through the chunk path
  a.nchunks = K;
  // chunk sizes fall off linearly (4:3:2:1 for four chunks): the copy of the last chunk is the only one
  // nothing overlaps, so it is the smallest; the earlier copies have the later chunks' kernels to hide behind
  const char* shape_env = getenv("SSPP_HOST_CHUNK_SHAPE");  // tuning: "equal" = chunks of equal size
  const bool equal_chunks = shape_env && shape_env[0] == 'e';
  for (int c = 0; c <= SSPP_MAX_CHUNKS; ++c) {
    const int64_t done = (int64_t)c * (2 * K - c + 1) / 2, all = (int64_t)K * (K + 1) / 2;  // sum of (K - j), j < c
    a.chunk_lo[c] = (int)(c >= K ? nrows : (equal_chunks ? (nrows * c) / K : (nrows * done) / all));
  }
  const bool to_peers = sink && sink->peers;
  if (to_peers && !ctx->push_stream) {
    // the chunk pushes run beside the next chunk's edge kernel: highest priority, so that their few CTAs are
    // placed as soon as an SM has room
    int lo_prio = 0, hi_prio = 0;
    CK(cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
    CK(cudaStreamCreateWithPriority(&ctx->push_stream, cudaStreamNonBlocking, hi_prio));
  }
  if (chunked) {
    if (!ctx->copy_stream) CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (int c = 0; c < K; ++c)
      if (!ctx->chunk_ev[c]) CK(cudaEventCreateWithFlags(&ctx->chunk_ev[c], cudaEventDisableTiming));
    if (!ctx->pinned) CK(cudaMallocHost((void**)&ctx->pinned, 32 * 8));
    CK(ctx->chunk_base.reserve((SSPP_MAX_CHUNKS + 1) * 8));
    CK(cudaMemsetAsync(ctx->chunk_base.p, 0, (SSPP_MAX_CHUNKS + 1) * 8, st));
  }
  if (sink) sink->delivered = false;
  CK(ctx->work_rows.reserve((size_t)nrows * 4));
  CK(ctx->deg.reserve((size_t)nrows * 4));
  a.cand_count = ctx->cand_count.as<int>();

  // Pair storage: with few pairs per row (the expected count follows from the node density: a ball of
  // the radius holds 4.19 / 27 of the nodes of its 27 cells) every row gets a fixed slot and the
  // counting pass writes the pairs itself; otherwise — or when a row overflows its slot — the counts
  // are scanned and a second neighbour pass fills the pairs.
  auto reserve_pairs = [&](long long c) -> cudaError_t {
    const size_t want = (size_t)std::max<long long>(c, 1);
    cudaError_t e;
    if ((e = ctx->candA.reserve(want * 8)) != cudaSuccess) return e;
    if ((e = ctx->candB.reserve(want * 8)) != cudaSuccess) return e;
    if ((e = ctx->acc.reserve(want * 4)) != cudaSuccess) return e;
    return ctx->col_idx.reserve(want * 4);
  };
  {
    const double lambda = 4.19 * (sharded ? 1.1 * (double)nrows : (double)n) / (double)ncells;
    a.slot = lambda <= 10.0 ? 32 : (lambda <= 28.0 ? 64 : 0);
    if (const char* env = getenv("SSPP_CONNECT_SLOT")) a.slot = std::max(0, atoi(env));
    if (a.slot && (double)nrows * a.slot > 3.0e9) a.slot = 0;  // slots beyond ~36 GB of pair buffers: scan instead
    if (fz) a.slot = fz->slot;  // the fused exchange works on the slot layout the ranks agreed on
    if (a.slot) CK(reserve_pairs((long long)nrows * a.slot));
  }
  k_count<false><<<ceil_div(n, 256), 256, 0, st>>>(a, ctx->map, ctx->cand_count.as<int>(), ctx->row_free.as<uint8_t>(),
                                                      ctx->counters.as<unsigned long long>(), nullptr,
                                                      a.slot ? ctx->candA.as<int2>() : nullptr, ctx->work_rows.as<int>(),
                                                      ctx->deg.as<int>(), 0);
  LAUNCHED(ctx);
  if (!a.slot) {
    int rc = scan_exclusive<long long>(ctx, ctx->cand_count.as<int>(), nrows, ctx->cand_off.as<long long>());
    if (rc) return rc;
  }
  CK(cudaEventRecord(ctx->ev[2], st));
  if (prm->count_only) {
    unsigned long long hc3[3];
    CK(cudaMemcpyAsync(hc3, ctx->counters.p, sizeof(hc3), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (stats) {
      stats->rows = nrows;
      stats->in_radius_pairs = (int64_t)hc3[0];
      stats->tested_pairs = (int64_t)hc3[1];
      stats->pyramid_rows = (int64_t)hc3[2];
    }
    return SSPP_OK;
  }

  // --- 4/5. edge test and CSR.  The pair buffers keep their size from call to call; the kernels
  // refuse to run when the pairs do not fit (a slot overflowed / the scanned total exceeds the
  // buffers), which the host sees at the single synchronisation at the end; it then switches to
  // scanned offsets or grows the buffers and repeats this phase.  Only a scanned pass with no buffers
  // yet has to wait for the count here.
  auto pair_capacity = [&]() -> long long {
    return (long long)std::min(std::min(ctx->candA.cap / 8, ctx->candB.cap / 8), std::min(ctx->acc.cap / 4, ctx->col_idx.cap / 4));
  };
  long long C = -1;
  if (!a.slot && pair_capacity() == 0) {
    CK(cudaMemcpyAsync(&C, ctx->cand_off.as<long long>() + nrows, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(reserve_pairs(C + C / 16 + 1024));
  }
  unsigned long long hc[10];
  long long E = 0;
  bool range_error = false;
  for (int attempt = 0;; ++attempt) {
    const long long cap = pair_capacity();
    if (!a.slot) {
      k_count<true><<<ceil_div(n, 256), 256, 0, st>>>(a, ctx->map, nullptr, nullptr, nullptr, ctx->cand_off.as<long long>(),
                                                         ctx->candA.as<int2>(), nullptr, nullptr, cap);
      LAUNCHED(ctx);
      k_rows_accept<<<ceil_div(n, EDGE_WARPS), EDGE_WARPS * 32, 0, st>>>(a, ctx->cand_off.as<long long>(),
                                                                          ctx->row_free.as<uint8_t>(), ctx->candA.as<int2>(),
                                                                          ctx->candB.as<int2>(), ctx->acc.as<int>(),
                                                                          ctx->deg.as<int>(), ctx->counters.as<unsigned long long>(),
                                                                          cap);
      LAUNCHED(ctx);
    }
    CK(cudaEventRecord(ctx->ev[5], st));
    bool copies_off = false;    // chunked: this attempt's pairs turned out unusable, nothing is sent home
    bool range_fail = false;    // chunked: the host column buffer is too small
    volatile long long* pin = (volatile long long*)ctx->pinned;  // [0..K] edge prefix per chunk, [16..24] counters
    // what chunk c needs once its kernels are done: its piece of the CSR goes to the host buffers
    auto drain = [&](int c) -> int {
      CK(cudaEventSynchronize(ctx->chunk_ev[c]));
      if (c == 0) copies_off = a.slot ? (pin[16 + 8] != 0) : (pin[16 + 0] > cap);
      if (copies_off || range_fail) return SSPP_OK;
      const long long e0 = pin[c], e1 = pin[c + 1];
      if (e1 > sink->col_capacity) { range_fail = true; return SSPP_OK; }
      const int r0 = a.chunk_lo[c], r1 = a.chunk_lo[c + 1];
      if (e1 > e0)
        CK(cudaMemcpyAsync(sink->col_idx + e0, ctx->col_idx.as<int>() + e0, (size_t)(e1 - e0) * 4, cudaMemcpyDeviceToHost,
                           ctx->copy_stream));
      CK(cudaMemcpyAsync(sink->row_ptr + r0, ctx->row_ptr.as<long long>() + r0, (size_t)(r1 - r0) * 8, cudaMemcpyDeviceToHost,
                         ctx->copy_stream));
      return SSPP_OK;
    };
    if (chunked) pin[0] = 0;
    for (int c = 0; c < K; ++c) {
      if (use_map) {
        // exchange overlapped with the edge test: half the SMs keep one CTA slot free for the push kernel of the
        // chunk before (SSPP_PUSH_ROOM = CTAs left out)
        int room = 0;
        if (to_peers && K > 1) {
          room = ctx->sm_count / 2;
          if (const char* env = getenv("SSPP_PUSH_ROOM")) room = std::max(0, atoi(env));
        }
#define SSPP_LAUNCH_EDGES(COUNT, NW, FUSED)                                                                                  \
  k_edges<COUNT, NW, FUSED><<<std::max(ctx->sm_count, ctx->sm_count * edge_min_blocks(COUNT, NW, FUSED) - room), EDGE_WARPS * 32, \
            edge_okey_bytes(lat.rays), st>>>(                                    \
      a, ctx->map, ctx->lattice.as<RayLattice>(), ctx->cand_off.as<long long>(), ctx->work_rows.as<int>(), ctx->candA.as<int2>(), \
      ctx->candB.as<int2>(), ctx->acc.as<int>(), ctx->deg.as<int>(), ctx->counters.as<unsigned long long>(), cap, c)
        if (count_steps) SSPP_LAUNCH_EDGES(true, 1, false);  // the instrumented variant has no start stage
        else if (fz && lat.rays >= 2 && lat.rays <= 32) SSPP_LAUNCH_EDGES(false, 1, true);
        else if (fz) SSPP_LAUNCH_EDGES(false, START_RAYS_MAX / 32, true);
        else if (lat.rays >= 2 && lat.rays <= 32) SSPP_LAUNCH_EDGES(false, 1, false);
        else SSPP_LAUNCH_EDGES(false, START_RAYS_MAX / 32, false);
#undef SSPP_LAUNCH_EDGES
        LAUNCHED(ctx);
      }
      if (c == K - 1) CK(cudaEventRecord(ctx->ev[3], st));
      const int r0 = a.chunk_lo[c], r1 = a.chunk_lo[c + 1];
      if (fz) continue;  // fused exchange: the rows are already on every rank, in slot format; no local CSR
      if (!chunked) {
        int rc = scan_exclusive<long long>(ctx, ctx->deg.as<int>(), nrows, ctx->row_ptr.as<long long>());
        if (rc) return rc;
      } else {
        // the chunk's piece of row_ptr, continued from the chunks before it (chunk_base[c] -> [c + 1])
        int rc = scan_exclusive<long long>(ctx, ctx->deg.as<int>() + r0, r1 - r0, ctx->row_ptr.as<long long>() + r0,
                                           ctx->chunk_base.as<long long>() + c, ctx->chunk_base.as<long long>() + c + 1);
        if (rc) return rc;
      }
      if (r1 > r0) {
        k_compact<<<ceil_div((int64_t)(r1 - r0) * COMPACT_LANES, 256), 256, 0, st>>>(
            a, ctx->cand_off.as<long long>(), ctx->deg.as<int>(), ctx->row_ptr.as<long long>(), ctx->acc.as<int>(),
            ctx->candA.as<int2>(), ctx->row_free.as<uint8_t>(), ctx->col_idx.as<int>(), ctx->counters.as<unsigned long long>(),
            cap, r0, r1);
        LAUNCHED(ctx);
      }
      if (to_peers) {
        // device-driven: the push of chunk c is queued behind its compaction and reads the chunk's edge range from
        // chunk_base in device memory; the host waits for nothing between the chunks
        const sspp_peer_blocks* pb = sink->peers;
        PushPeers pp;
        pp.world = pb->world;
        pp.rank = pb->rank;
        for (int k = 0; k < SSPP_MAX_PEERS; ++k) { pp.rp[k] = nullptr; pp.col[k] = nullptr; }
        for (int k = 0; k < pb->world; ++k) {
          pp.rp[k] = (long long*)pb->row_ptr_blocks[k] + (size_t)pb->rank * pb->row_stride;
          pp.col[k] = pb->col_blocks[k] + (size_t)pb->rank * pb->col_stride;
        }
        CK(cudaEventRecord(ctx->chunk_ev[c], st));
        CK(cudaStreamWaitEvent(ctx->push_stream, ctx->chunk_ev[c], 0));
        int pblocks = ctx->sm_count / 2;
        if (const char* env = getenv("SSPP_PUSH_BLOCKS")) pblocks = std::max(1, atoi(env));
        k_push_chunk<<<pblocks, 256, 0, ctx->push_stream>>>(pp, ctx->chunk_base.as<long long>(), c, c == K - 1 ? 1 : 0,
                                                            ctx->row_ptr.as<long long>(), ctx->col_idx.as<int>(), r0, r1,
                                                            (long long)pb->col_stride);
        LAUNCHED(ctx);
      } else if (chunked) {
        CK(cudaMemcpyAsync((void*)(pin + c + 1), ctx->chunk_base.as<long long>() + c + 1, 8, cudaMemcpyDeviceToHost, st));
        if (c == 0) CK(cudaMemcpyAsync((void*)(pin + 16), ctx->counters.p, 9 * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(ctx->chunk_ev[c], st));
        if (c >= 1) {
          int rc = drain(c - 1);  // chunk c is queued: now wait for chunk c-1 and send it home
          if (rc) return rc;
        }
      }
    }
    CK(cudaEventRecord(ctx->ev[4], st));
    if (chunked && !to_peers) {
      pin[0] = 0;
      int rc = drain(K - 1);
      if (rc) return rc;
    }
    if (!fz) CK(cudaMemcpyAsync(&E, ctx->row_ptr.as<long long>() + nrows, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(hc, ctx->counters.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (fz) {
      E = (long long)hc[9];
      if (hc[8] != 0) return fail(ctx, SSPP_ERR_RANGE, "connect_fused: a row holds more pairs than the exchange's slot");
      sink->delivered = true;
      break;
    }
    if (to_peers) {
      CK(cudaStreamSynchronize(ctx->push_stream));  // this rank's stores have been issued and retired
      const bool usable = a.slot ? (hc[8] == 0) : ((long long)hc[0] <= cap);
      if (usable) {
        if (E > sink->peers->col_stride) range_error = true;  // the chunks beyond the slot were not sent
        else sink->delivered = true;
      }
    } else if (chunked) {
      CK(cudaStreamSynchronize(ctx->copy_stream));
      if (range_fail) range_error = true;  // reported below, once the resident result is registered (connect_fetch still works)
      if (!copies_off && !range_fail) {
        sink->row_ptr[nrows] = E;
        sink->delivered = true;
      }
    }
    C = (long long)hc[0];
    if (a.slot) {
      if (hc[8] == 0) break;
      // some row holds more pairs than a slot: redo the phase with scanned offsets (the counts, the
      // pyramid flags and the work list of the counting pass stay valid)
      a.slot = 0;
      int rc = scan_exclusive<long long>(ctx, ctx->cand_count.as<int>(), nrows, ctx->cand_off.as<long long>());
      if (rc) return rc;
      if (C > pair_capacity()) CK(reserve_pairs(C + C / 16 + 1024));
      continue;
    }
    if (C <= cap) break;
    if (attempt > 1) return fail(ctx, SSPP_ERR_CUDA, "connect: pair buffers still too small after growing them");
    CK(reserve_pairs(C + C / 16 + 1024));
  }
  if (stats) {
    stats->rows = nrows;
    stats->in_radius_pairs = (int64_t)hc[0];
    stats->tested_pairs = (int64_t)hc[1];
    stats->pyramid_rows = (int64_t)hc[2];
    stats->voxel_steps = count_steps ? (int64_t)hc[3] : 0;
    stats->tile_rows = 0;
    stats->box_pairs = (int64_t)hc[5];
    stats->lattice_pairs = (int64_t)hc[6];
    stats->edges = E;
  }
  cudaEventElapsedTime(&ctx->timings[0], ctx->ev[0], ctx->ev[4]);
  cudaEventElapsedTime(&ctx->timings[1], ctx->ev[0], ctx->ev[1]);
  cudaEventElapsedTime(&ctx->timings[2], ctx->ev[1], ctx->ev[2]);
  cudaEventElapsedTime(&ctx->timings[3], ctx->ev[2], ctx->ev[3]);
  cudaEventElapsedTime(&ctx->timings[4], ctx->ev[3], ctx->ev[4]);
  cudaEventElapsedTime(&ctx->timings[5], ctx->ev[5], ctx->ev[3]);
  ctx->has_result = !fz;  // the fused pass leaves no local CSR behind
  ctx->res_rows = nrows;
  ctx->res_edges = E;
  if (range_error)
    return fail(ctx, SSPP_ERR_RANGE, sink->peers ? "connect_to_peers: the block does not fit its column slot"
                                                 : "connect_nodes_host: col_idx capacity too small");
  return SSPP_OK;
}
}  // namespace

int sspp_gpu_connect_fetch(sspp_gpu_ctx* ctx, int64_t* row_ptr, int32_t* col_idx) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_result) return fail(ctx, SSPP_ERR_STATE, "connect_fetch: no connect result");
  if (!row_ptr || (ctx->res_edges > 0 && !col_idx)) return fail(ctx, SSPP_ERR_INVALID, "connect_fetch: null argument");
  CK(cudaSetDevice(ctx->device));
  CK(cudaMemcpyAsync(row_ptr, ctx->row_ptr.p, (size_t)(ctx->res_rows + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (ctx->res_edges)
    CK(cudaMemcpyAsync(col_idx, ctx->col_idx.p, (size_t)ctx->res_edges * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

int sspp_gpu_connect_row_classes(sspp_gpu_ctx* ctx, uint8_t* row_class) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_result) return fail(ctx, SSPP_ERR_STATE, "connect_row_classes: no connect result");
  if (!row_class) return fail(ctx, SSPP_ERR_INVALID, "connect_row_classes: null argument");
  if (ctx->res_rows == 0) return SSPP_OK;
  CK(cudaSetDevice(ctx->device));
  CK(cudaMemcpyAsync(row_class, ctx->row_class.p, (size_t)ctx->res_rows, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

int sspp_gpu_connect_fetch_device(sspp_gpu_ctx* ctx, int64_t* row_ptr_device, int32_t* col_idx_device) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_result) return fail(ctx, SSPP_ERR_STATE, "connect_fetch_device: no connect result");
  if (!row_ptr_device || (ctx->res_edges > 0 && !col_idx_device)) return fail(ctx, SSPP_ERR_INVALID, "connect_fetch_device: null argument");
  CK(cudaSetDevice(ctx->device));
  CK(cudaMemcpyAsync(row_ptr_device, ctx->row_ptr.p, (size_t)(ctx->res_rows + 1) * 8, cudaMemcpyDeviceToDevice, ctx->stream));
  if (ctx->res_edges)
    CK(cudaMemcpyAsync(col_idx_device, ctx->col_idx.p, (size_t)ctx->res_edges * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

int sspp_gpu_assemble_csr_blocks(sspp_gpu_ctx* ctx, const int64_t* row_ptr_blocks_device, const int32_t* col_blocks_device,
                                 int world, int64_t row_stride, int64_t col_stride, const int64_t* row_base_device,
                                 int64_t* row_ptr_out_device, int32_t* col_out_device, int64_t* overflow_device) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (world < 1 || !row_ptr_blocks_device || !col_blocks_device || !row_base_device || !row_ptr_out_device || !col_out_device ||
      !overflow_device)
    return fail(ctx, SSPP_ERR_INVALID, "assemble_csr_blocks: null argument");
  CK(cudaSetDevice(ctx->device));
  dim3 grid((unsigned)(ctx->sm_count * 2 / world + 1), (unsigned)world);
  k_assemble_blocks<<<grid, 256, 0, ctx->stream>>>((const long long*)row_ptr_blocks_device, col_blocks_device, world, row_stride,
                                                   col_stride, (const long long*)row_base_device, (long long*)row_ptr_out_device,
                                                   col_out_device, (long long*)overflow_device);
  LAUNCHED(ctx);
  return SSPP_OK;
}

// ---- multi-GPU assembly through peer memory ---------------------------------------------------
int sspp_gpu_peer_buffer_create(sspp_gpu_ctx* ctx, int64_t bytes, void** device_ptr, uint8_t handle[SSPP_PEER_HANDLE_BYTES]) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (bytes <= 0 || !device_ptr || !handle) return fail(ctx, SSPP_ERR_INVALID, "peer_buffer_create: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == SSPP_PEER_HANDLE_BYTES, "IPC handle size");
  CK(cudaSetDevice(ctx->device));
  void* p = nullptr;
  CK(cudaMalloc(&p, (size_t)bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return cuda_fail(ctx, e, "cudaIpcGetMemHandle");
  }
  memcpy(handle, &h, SSPP_PEER_HANDLE_BYTES);
  ctx->peer_created.push_back(p);
  *device_ptr = p;
  return SSPP_OK;
}

int sspp_gpu_peer_buffer_open(sspp_gpu_ctx* ctx, const uint8_t handle[SSPP_PEER_HANDLE_BYTES], void** device_ptr) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!handle || !device_ptr) return fail(ctx, SSPP_ERR_INVALID, "peer_buffer_open: null argument");
  CK(cudaSetDevice(ctx->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, SSPP_PEER_HANDLE_BYTES);
  void* p = nullptr;
  CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  ctx->peer_opened.push_back(p);
  *device_ptr = p;
  return SSPP_OK;
}

int sspp_gpu_peer_buffer_release(sspp_gpu_ctx* ctx, void* device_ptr) {
  if (!ctx) return SSPP_ERR_INVALID;
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->stream));
  auto c = std::find(ctx->peer_created.begin(), ctx->peer_created.end(), device_ptr);
  if (c != ctx->peer_created.end()) {
    ctx->peer_created.erase(c);
    CK(cudaFree(device_ptr));
    return SSPP_OK;
  }
  auto o = std::find(ctx->peer_opened.begin(), ctx->peer_opened.end(), device_ptr);
  if (o != ctx->peer_opened.end()) {
    ctx->peer_opened.erase(o);
    CK(cudaIpcCloseMemHandle(device_ptr));
    return SSPP_OK;
  }
  return fail(ctx, SSPP_ERR_INVALID, "peer_buffer_release: not a peer buffer of this context");
}

namespace {
// one block of a CSR-shaped result (row pointers relative to the block + items) into the whole-result buffers of all ranks
int push_block(sspp_gpu_ctx* ctx, const char* who, const long long* rp_local, const int* items_local, int world, int rank,
               const int64_t* counts_device, const int64_t* row_base_device, int64_t* const* row_ptr_peers,
               int32_t* const* col_idx_peers, int64_t col_capacity, int64_t* overflow_device) {
  if (world < 1 || world > SSPP_MAX_PEERS || rank < 0 || rank >= world)
    return fail(ctx, SSPP_ERR_INVALID, std::string(who) + ": world must be 1..SSPP_MAX_PEERS and rank inside it");
  if (!counts_device || !row_base_device || !row_ptr_peers || !col_idx_peers || !overflow_device || col_capacity < 0)
    return fail(ctx, SSPP_ERR_INVALID, std::string(who) + ": null argument");
  PushPeers p;
  p.world = world;
  p.rank = rank;
  for (int k = 0; k < SSPP_MAX_PEERS; ++k) {
    p.rp[k] = nullptr;
    p.col[k] = nullptr;
  }
  for (int k = 0; k < world; ++k) {
    if (!row_ptr_peers[k] || !col_idx_peers[k]) return fail(ctx, SSPP_ERR_INVALID, std::string(who) + ": null peer buffer");
    if (((uintptr_t)col_idx_peers[k] & 15) != 0) return fail(ctx, SSPP_ERR_INVALID, std::string(who) + ": column buffers must be 16-byte aligned");
    p.rp[k] = (long long*)row_ptr_peers[k];
    p.col[k] = col_idx_peers[k];
  }
  CK(cudaSetDevice(ctx->device));
  k_push_csr<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(p, (const long long*)counts_device, (const long long*)row_base_device,
                                                         rp_local, items_local, (long long)col_capacity, (long long*)overflow_device);
  LAUNCHED(ctx);
  return SSPP_OK;
}
}  // namespace

int sspp_gpu_connect_push_csr(sspp_gpu_ctx* ctx, int world, int rank, const int64_t* edge_counts_device,
                              const int64_t* row_base_device, int64_t* const* row_ptr_peers, int32_t* const* col_idx_peers,
                              int64_t col_capacity, int64_t* overflow_device) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_result) return fail(ctx, SSPP_ERR_STATE, "connect_push_csr: no connect result");
  return push_block(ctx, "connect_push_csr", ctx->row_ptr.as<long long>(), ctx->col_idx.as<int>(), world, rank, edge_counts_device,
                    row_base_device, row_ptr_peers, col_idx_peers, col_capacity, overflow_device);
}

int sspp_gpu_visibility_push(sspp_gpu_ctx* ctx, int world, int rank, const int64_t* pair_counts_device,
                             const int64_t* view_base_device, int64_t* const* view_ptr_peers, int32_t* const* point_idx_peers,
                             int64_t idx_capacity, int64_t* overflow_device) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_vis) return fail(ctx, SSPP_ERR_STATE, "visibility_push: no visibility result");
  return push_block(ctx, "visibility_push", ctx->vis_ptr.as<long long>(), ctx->vis_idx.as<int>(), world, rank, pair_counts_device,
                    view_base_device, view_ptr_peers, point_idx_peers, idx_capacity, overflow_device);
}

int sspp_gpu_balanced_cuts(sspp_gpu_ctx* ctx, const sspp_connect_params* params, int world, int work_row_weight, int64_t* cuts) {
  if (!ctx || !params || !cuts) return fail(ctx, SSPP_ERR_INVALID, "balanced_cuts: null argument");
  if (world < 1 || world > 4096 || work_row_weight < 1 || work_row_weight > 1024)
    return fail(ctx, SSPP_ERR_INVALID, "balanced_cuts: world must be 1..4096 and work_row_weight 1..1024");
  const int64_t n = ctx->n_nodes;
  if (n == 0 || world == 1) {
    for (int g = 0; g <= world; ++g) cuts[g] = g == world ? n : 0;
    return SSPP_OK;
  }
  sspp_connect_params p = *params;
  p.row_begin = p.row_end = 0;  // the counting pass runs over all rows
  p.count_only = 1;
  p.count_steps = 0;
  p.use_pyramid = 1;
  sspp_connect_stats st;
  int rc = connect_impl(ctx, &p, &st, nullptr);
  if (rc) return rc;
  CK(ctx->q0.reserve((size_t)n * 4));
  CK(ctx->q1.reserve((size_t)(n + 1) * 8));
  CK(ctx->q2.reserve((size_t)(world + 1) * 8));
  k_row_weights<<<ceil_div(n, 256), 256, 0, ctx->stream>>>(ctx->cand_count.as<int>(), ctx->row_free.as<uint8_t>(), p.use_map ? 1 : 0,
                                                          work_row_weight, n, ctx->q0.as<int>());
  LAUNCHED(ctx);
  rc = scan_exclusive<long long>(ctx, ctx->q0.as<int>(), n, ctx->q1.as<long long>());
  if (rc) return rc;
  k_find_cuts<<<ceil_div(world + 1, 128), 128, 0, ctx->stream>>>(ctx->q1.as<long long>(), n, world, ctx->q2.as<long long>());
  LAUNCHED(ctx);
  CK(cudaMemcpyAsync(cuts, ctx->q2.p, (size_t)(world + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  for (int g = 1; g <= world; ++g) cuts[g] = std::max(cuts[g], cuts[g - 1]);
  return SSPP_OK;
}

int sspp_gpu_connect_result_device(sspp_gpu_ctx* ctx, const int64_t** row_ptr_device, const int32_t** col_idx_device) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_result) return fail(ctx, SSPP_ERR_STATE, "no connect result");
  if (row_ptr_device) *row_ptr_device = ctx->row_ptr.as<int64_t>();
  if (col_idx_device) *col_idx_device = ctx->col_idx.as<int32_t>();
  return SSPP_OK;
}

int sspp_gpu_connect_to_host(sspp_gpu_ctx* ctx, const sspp_connect_params* params, int64_t* row_ptr, int32_t* col_idx,
                             int64_t col_capacity, sspp_connect_stats* stats) {
  sspp_connect_stats local;
  if (!stats) stats = &local;
  if (!ctx) return SSPP_ERR_INVALID;
  if (!row_ptr || (col_capacity > 0 && !col_idx) || col_capacity < 0)
    return fail(ctx, SSPP_ERR_INVALID, "connect_to_host: null argument");
  HostSink sink = {row_ptr, col_idx, col_capacity, false, nullptr, nullptr};
  int rc = connect_impl(ctx, params, stats, &sink);
  if (rc) return rc;
  if (sink.delivered) return SSPP_OK;  // the chunks were sent home while the pass was still running
  if (stats->edges > col_capacity) return fail(ctx, SSPP_ERR_RANGE, "connect_nodes_host: col_idx capacity too small");
  return sspp_gpu_connect_fetch(ctx, row_ptr, col_idx);
}

int sspp_gpu_connect_nodes_host(sspp_gpu_ctx* ctx, const double* xyz, int64_t n, const sspp_connect_params* params,
                                int64_t* row_ptr, int32_t* col_idx, int64_t col_capacity, sspp_connect_stats* stats) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!row_ptr || (col_capacity > 0 && !col_idx) || col_capacity < 0)
    return fail(ctx, SSPP_ERR_INVALID, "connect_nodes_host: null argument");
  int rc = set_nodes_impl(ctx, xyz, n, cudaMemcpyHostToDevice);
  if (rc) return rc;
  return sspp_gpu_connect_to_host(ctx, params, row_ptr, col_idx, col_capacity, stats);
}

int sspp_gpu_connect_to_peers(sspp_gpu_ctx* ctx, const sspp_connect_params* params, const sspp_peer_blocks* peers,
                              sspp_connect_stats* stats) {
  sspp_connect_stats local;
  if (!stats) stats = &local;
  if (!ctx) return SSPP_ERR_INVALID;
  if (!peers || !peers->row_ptr_blocks || !peers->col_blocks) return fail(ctx, SSPP_ERR_INVALID, "connect_to_peers: null argument");
  if (peers->world < 1 || peers->world > SSPP_MAX_PEERS || peers->rank < 0 || peers->rank >= peers->world)
    return fail(ctx, SSPP_ERR_INVALID, "connect_to_peers: world must be 1..SSPP_MAX_PEERS and rank inside it");
  for (int p = 0; p < peers->world; ++p)
    if (!peers->row_ptr_blocks[p] || !peers->col_blocks[p]) return fail(ctx, SSPP_ERR_INVALID, "connect_to_peers: null peer buffer");
  int64_t lo = params ? params->row_begin : 0, hi = params ? params->row_end : 0;
  if (lo == 0 && hi == 0) hi = ctx->n_nodes;
  if (peers->row_stride < hi - lo + 1 || peers->col_stride < 0)
    return fail(ctx, SSPP_ERR_INVALID, "connect_to_peers: row_stride must hold rows + 1 entries");
  HostSink sink = {nullptr, nullptr, peers->col_stride, false, peers, nullptr};
  int rc = connect_impl(ctx, params, stats, &sink);
  if (rc) return rc;
  if (sink.delivered) return SSPP_OK;
  // nothing went through the chunk path (an empty row range): deliver the block in one piece
  if (stats->edges > peers->col_stride) return fail(ctx, SSPP_ERR_RANGE, "connect_to_peers: the block does not fit its column slot");
  CK(cudaSetDevice(ctx->device));
  for (int p = 0; p < peers->world; ++p) {
    if (stats->edges)
      CK(cudaMemcpyAsync(peers->col_blocks[p] + (size_t)peers->rank * peers->col_stride, ctx->col_idx.p, (size_t)stats->edges * 4,
                         cudaMemcpyDefault, ctx->stream));
    CK(cudaMemcpyAsync(peers->row_ptr_blocks[p] + (size_t)peers->rank * peers->row_stride, ctx->row_ptr.p,
                       (size_t)(stats->rows + 1) * 8, cudaMemcpyDefault, ctx->stream));
  }
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

int sspp_gpu_connect_fused(sspp_gpu_ctx* ctx, const sspp_connect_params* params, const sspp_fused_peers* fp, sspp_connect_stats* stats) {
  sspp_connect_stats local;
  if (!stats) stats = &local;
  if (!ctx) return SSPP_ERR_INVALID;
  if (!params || !fp || !fp->slot_peers || !fp->deg_peers) return fail(ctx, SSPP_ERR_INVALID, "connect_fused: null argument");
  if (fp->world < 1 || fp->world > SSPP_MAX_PEERS || fp->world > 32 || fp->rank < 0 || fp->rank >= fp->world)
    return fail(ctx, SSPP_ERR_INVALID, "connect_fused: world must be 1..SSPP_MAX_PEERS and rank inside it");
  if (fp->slot != 32 && fp->slot != 64) return fail(ctx, SSPP_ERR_INVALID, "connect_fused: slot must be 32 or 64");
  if (!params->use_map) return fail(ctx, SSPP_ERR_INVALID, "connect_fused: needs use_map (without a map there is no edge test to fuse with)");
  if (params->count_steps || params->count_only) return fail(ctx, SSPP_ERR_INVALID, "connect_fused: not for instrumented passes");
  if (fp->first_row < 0) return fail(ctx, SSPP_ERR_INVALID, "connect_fused: first_row must be >= 0");
  for (int p = 0; p < fp->world; ++p)
    if (!fp->slot_peers[p] || !fp->deg_peers[p]) return fail(ctx, SSPP_ERR_INVALID, "connect_fused: null peer buffer");
  HostSink sink = {nullptr, nullptr, 0, false, nullptr, fp};
  return connect_impl(ctx, params, stats, &sink);
}

int sspp_gpu_pack_slots(sspp_gpu_ctx* ctx, const int32_t* slots_device, const int32_t* deg_device, int64_t n_rows, int32_t slot,
                        int64_t* row_ptr_out_device, int32_t* col_out_device, int64_t col_capacity, int64_t* overflow_device) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!slots_device || !deg_device || !row_ptr_out_device || !col_out_device || !overflow_device || n_rows < 0 || col_capacity < 0 ||
      (slot != 32 && slot != 64))
    return fail(ctx, SSPP_ERR_INVALID, "pack_slots: bad argument");
  CK(cudaSetDevice(ctx->device));
  int rc = scan_exclusive<long long>(ctx, (const int*)deg_device, n_rows, (long long*)row_ptr_out_device);
  if (rc) return rc;
  if (n_rows) {
    k_pack_slots<<<ceil_div(n_rows * COMPACT_LANES, 256), 256, 0, ctx->stream>>>((const int*)slots_device, (const int*)deg_device,
                                                                              (const long long*)row_ptr_out_device, (long long)n_rows,
                                                                              slot, col_out_device, (long long)col_capacity,
                                                                              (long long*)overflow_device);
    LAUNCHED(ctx);
  }
  return SSPP_OK;
}

int sspp_gpu_connect_timings(sspp_gpu_ctx* ctx, float ms[8]) {
  if (!ctx || !ms) return SSPP_ERR_INVALID;
  for (int i = 0; i < 8; ++i) ms[i] = ctx->timings[i];
  return SSPP_OK;
}

// ---- batched SearchSpace::insertNode -----------------------------------------------------------
int sspp_gpu_dedup_poses(sspp_gpu_ctx* ctx, const double* poses, int64_t n, double tolerance, uint8_t* keep, int64_t* dup_of,
                         int64_t* stored) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (n < 0 || (n > 0 && (!poses || !keep))) return fail(ctx, SSPP_ERR_INVALID, "dedup_poses: null argument");
  if (!(tolerance > 0.0) || !std::isfinite(tolerance)) return fail(ctx, SSPP_ERR_INVALID, "dedup_poses: tolerance must be > 0");
  if (n >= (int64_t)0x7fffffff) return fail(ctx, SSPP_ERR_RANGE, "dedup_poses: pose indices must fit in int32");
  if (stored) *stored = 0;
  if (n == 0) return SSPP_OK;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  CK(ctx->dd_poses.reserve((size_t)n * 56));
  CK(ctx->dd_xyz.reserve((size_t)n * 24));
  CK(ctx->dd_state.reserve((size_t)n));
  CK(ctx->dd_keep.reserve((size_t)n));
  CK(ctx->dd_pend.reserve((size_t)n * 4));
  CK(ctx->dd_misc.reserve(16 * 8));
  unsigned long long* misc = ctx->dd_misc.as<unsigned long long>();  // [0..6] bounding box, [8..10] counters
  static const unsigned long long init[16] = {~0ull, ~0ull, ~0ull, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  CK(cudaMemcpyAsync(misc, init, sizeof(init), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->dd_poses.p, poses, (size_t)n * 56, cudaMemcpyHostToDevice, st));
  const double* dposes = ctx->dd_poses.as<double>();
  k_pose_xyz<<<ceil_div(n, 256), 256, 0, st>>>(dposes, n, ctx->dd_xyz.as<double>());
  LAUNCHED(ctx);
  k_bbox<<<std::min(ceil_div(n, 256), 148 * 8), 256, 0, st>>>(ctx->dd_xyz.as<double>(), n, misc);
  LAUNCHED(ctx);
  unsigned long long hb[7];
  CK(cudaMemcpyAsync(hb, misc, sizeof(hb), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (hb[6]) return fail(ctx, SSPP_ERR_INVALID, "dedup_poses: positions must be finite");

  // grid: about one pose per cell, cell edge never below 4 x tolerance (so that every match of a pose
  // lies in its own or an adjacent cell whatever the rounding of the cell index)
  NodeGrid g;
  double ext[3], volume = 1.0, max_ext = 0.0;
  int axes = 0;
  for (int a = 0; a < 3; ++a) {
    g.min[a] = dec_f64(hb[a]);
    g.max[a] = dec_f64(hb[3 + a]);
    ext[a] = g.max[a] - g.min[a];
    max_ext = std::max(max_ext, ext[a]);
    if (ext[a] > 0) { volume *= ext[a]; ++axes; }
  }
  double cell = axes ? pow(volume / (double)n, 1.0 / axes) : 1.0;
  cell = std::max(cell, std::max(4.0 * tolerance, max_ext * 1e-9));
  if (!(cell > 0.0) || !std::isfinite(cell)) cell = 1.0;
  for (;;) {
    double total = 1;
    for (int a = 0; a < 3; ++a) total *= floor(ext[a] / cell) + 2.0;
    if (total <= (double)(1 << 24)) break;
    cell *= 1.26;
  }
  g.inv_cell = 1.0 / cell;
  int64_t ncells = 1;
  for (int a = 0; a < 3; ++a) { g.n[a] = (int)floor(ext[a] * g.inv_cell) + 1; ncells *= g.n[a]; }

  CK(ctx->cell_of.reserve((size_t)n * 4));
  CK(ctx->cell_start.reserve((size_t)(ncells + 1) * 4));
  CK(ctx->cell_cursor.reserve((size_t)(ncells + 1) * 4));
  CK(ctx->sx.reserve((size_t)n * sizeof(NodeRec)));
  CK(cudaMemsetAsync(ctx->cell_cursor.p, 0, (size_t)(ncells + 1) * 4, st));
  k_cell_hist<<<ceil_div(n, 256), 256, 0, st>>>(ctx->dd_xyz.as<double>(), n, g, ctx->cell_of.as<int>(), ctx->cell_cursor.as<int>());
  LAUNCHED(ctx);
  {
    int rc = scan_exclusive<int>(ctx, ctx->cell_cursor.as<int>(), ncells, ctx->cell_start.as<int>());
    if (rc) return rc;
  }
  CK(cudaMemcpyAsync(ctx->cell_cursor.p, ctx->cell_start.p, (size_t)ncells * 4, cudaMemcpyDeviceToDevice, st));
  k_cell_scatter<<<ceil_div(n, 256), 256, 0, st>>>(ctx->dd_xyz.as<double>(), n, ctx->cell_of.as<int>(), ctx->cell_cursor.as<int>(),
                                                   ctx->sx.as<NodeRec>());
  LAUNCHED(ctx);

  uint8_t* state = ctx->dd_state.as<uint8_t>();
  int* pend = ctx->dd_pend.as<int>();
  unsigned long long* counters = misc + 8;
  k_dedup_first<<<ceil_div(n, 256), 256, 0, st>>>(g, ctx->cell_start.as<int>(), ctx->sx.as<NodeRec>(), n, dposes, tolerance, state,
                                                  pend, counters);
  LAUNCHED(ctx);
  unsigned long long hc[3] = {0, 0, 0};
  CK(cudaMemcpyAsync(hc, counters, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  const int64_t npend = (int64_t)hc[0];
  unsigned long long waiting = npend;
  while (waiting > 0) {
    CK(cudaMemsetAsync(counters + 1, 0, 8, st));
    k_dedup_round<<<ceil_div(npend, 256), 256, 0, st>>>(g, ctx->cell_start.as<int>(), ctx->sx.as<NodeRec>(), dposes, tolerance,
                                                       state, pend, npend, counters);
    LAUNCHED(ctx);
    unsigned long long left = 0;
    CK(cudaMemcpyAsync(&left, counters + 1, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    // the pending pose with the smallest index always resolves, so every round makes progress
    if (left >= waiting) return fail(ctx, SSPP_ERR_CUDA, "dedup_poses: a round resolved nothing");
    waiting = left;
  }
  if (dup_of) {
    CK(ctx->dd_dup.reserve((size_t)n * 8));
    k_dedup_init_dup<<<ceil_div(n, 256), 256, 0, st>>>(n, ctx->dd_dup.as<long long>());
    LAUNCHED(ctx);
    if (npend) {
      k_dedup_final<<<ceil_div(npend, 256), 256, 0, st>>>(g, ctx->cell_start.as<int>(), ctx->sx.as<NodeRec>(), dposes, tolerance,
                                                         state, pend, npend, ctx->dd_dup.as<long long>());
      LAUNCHED(ctx);
    }
    CK(cudaMemcpyAsync(dup_of, ctx->dd_dup.p, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
  }
  k_dedup_keep<<<ceil_div(n, 256), 256, 0, st>>>(state, n, ctx->dd_keep.as<uint8_t>(), counters);
  LAUNCHED(ctx);
  CK(cudaMemcpyAsync(keep, ctx->dd_keep.p, (size_t)n, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(hc, counters, sizeof(hc), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (stored) *stored = (int64_t)hc[2];
  return SSPP_OK;
}

// ---- batched queries -------------------------------------------------------------------------
int sspp_gpu_check_segments(sspp_gpu_ctx* ctx, const double* starts, const double* ends, int64_t n, const double box[3],
                            uint8_t* status) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  if (n < 0 || !box || (n > 0 && (!starts || !ends || !status))) return fail(ctx, SSPP_ERR_INVALID, "check_segments: null argument");
  if (n == 0) return SSPP_OK;
  CK(cudaSetDevice(ctx->device));
  RayLattice lat;
  int rc = build_lattice(ctx, box, lat);
  if (rc) return rc;
  rc = upload_lattice(ctx, lat);
  if (rc) return rc;
  CK(ctx->q0.reserve((size_t)n * 24));
  CK(ctx->q1.reserve((size_t)n * 24));
  CK(ctx->q2.reserve((size_t)n));
  CK(cudaMemcpyAsync(ctx->q0.p, starts, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->q1.p, ends, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
  int G = lat.rays >= 32 ? 32 : lat.rays;
  int EPI = 32 / G;
  int64_t warps = (n + EPI - 1) / EPI;
  k_check_segments<<<ceil_div(warps * 32, 256), 256, 0, ctx->stream>>>(ctx->map, ctx->lattice.as<RayLattice>(),
                                                                      ctx->q0.as<double>(), ctx->q1.as<double>(), n,
                                                                      ctx->q2.as<uint8_t>());
  LAUNCHED(ctx);
  CK(cudaMemcpyAsync(status, ctx->q2.p, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

// ---- allocation-free segment check ------------------------------------------------------------
// Working memory for sspp_gpu_check_segments_noalloc: one fixed device array instead of the
// cudaMalloc-ed staging buffers the regular entry point grows on demand.  Layout (bytes):
//   [0, 2048)                      ray-offset table (RayLattice, rewritten in full by every call)
//   [2048, +24*MAX)                segment starts   (first n entries rewritten before they are read)
//   [.., +24*MAX)                  segment ends     (same)
//   [.., +MAX)                     status bytes     (first n written by the kernel, then copied out)
// Nothing in it has to start at zero, so no clearing is needed; every GPU that loads the library
// holds the array (~0.8 MB) until the process exits.
namespace {
constexpr size_t NA_MAX = SSPP_NOALLOC_MAX_SEGMENTS;
constexpr size_t NA_LATTICE_BYTES = 2048;
constexpr size_t NA_STARTS_OFF = NA_LATTICE_BYTES;
constexpr size_t NA_ENDS_OFF = NA_STARTS_OFF + NA_MAX * 24;
constexpr size_t NA_STATUS_OFF = NA_ENDS_OFF + NA_MAX * 24;
constexpr size_t NA_BYTES = NA_STATUS_OFF + NA_MAX;
static_assert(sizeof(RayLattice) <= NA_LATTICE_BYTES, "ray-offset table must fit its slot");
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address: the
// address a kernel can use is asked for once per device, with that device current.
constexpr int NA_MAX_DEVICES = 64;
unsigned char* g_noalloc_addr[NA_MAX_DEVICES] = {nullptr};

int noalloc_scratch_address(sspp_gpu_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= NA_MAX_DEVICES) return fail(ctx, SSPP_ERR_RANGE, "check_segments_noalloc: device index too large");
  if (!__atomic_load_n(&g_noalloc_addr[ctx->device], __ATOMIC_ACQUIRE)) {
    unsigned char* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess) return cuda_fail(ctx, e, "cudaGetSymbolAddress(g_noalloc_scratch)");
    if (!ptr) return fail(ctx, SSPP_ERR_CUDA, "check_segments_noalloc: the runtime returned a null address for the scratch array");
    // every caller on this device computes the same value, so a racing first call stores it twice at worst
    __atomic_store_n(&g_noalloc_addr[ctx->device], ptr, __ATOMIC_RELEASE);
  }
  *out = g_noalloc_addr[ctx->device];
  return SSPP_OK;
}
}  // namespace

// Batched getLineStatusBoundingBox without any device allocation (see include/sspp_gpu.h).  The
// first call on a device resolves the scratch array's address — do it before any stream capture.
// All contexts of a device share the array: calls whose GPU work overlaps race on it.
int sspp_gpu_check_segments_noalloc(sspp_gpu_ctx* ctx, const double* starts, const double* ends, int64_t n,
                                    const double box[3], uint8_t* status) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  if (n < 0 || !box || (n > 0 && (!starts || !ends || !status))) return fail(ctx, SSPP_ERR_INVALID, "check_segments_noalloc: null argument");
  if (n > (int64_t)NA_MAX)
    return fail(ctx, SSPP_ERR_RANGE, "check_segments_noalloc: more than SSPP_NOALLOC_MAX_SEGMENTS (16384) segments in one call; "
                                     "split the batch or use sspp_gpu_check_segments");
  if (n == 0) return SSPP_OK;
  CK(cudaSetDevice(ctx->device));
  unsigned char* base = nullptr;
  int rc = noalloc_scratch_address(ctx, &base);
  if (rc) return rc;
  RayLattice lat;
  rc = build_lattice(ctx, box, lat);
  if (rc) return rc;
  RayLattice* d_lat = (RayLattice*)base;
  double* d_starts = (double*)(base + NA_STARTS_OFF);
  double* d_ends = (double*)(base + NA_ENDS_OFF);
  uint8_t* d_status = base + NA_STATUS_OFF;
  CK(cudaMemcpyAsync(d_lat, &lat, sizeof(RayLattice), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d_starts, starts, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d_ends, ends, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
  int G = lat.rays >= 32 ? 32 : lat.rays;
  int EPI = 32 / G;
  int64_t warps = (n + EPI - 1) / EPI;
  k_check_segments<<<ceil_div(warps * 32, 256), 256, 0, ctx->stream>>>(ctx->map, d_lat, d_starts, d_ends, n, d_status);
  LAUNCHED(ctx);
  CK(cudaMemcpyAsync(status, d_status, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

int sspp_gpu_validity(sspp_gpu_ctx* ctx, const double* pts, int64_t n, const double box[3], uint8_t* status) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  if (n < 0 || !box || (n > 0 && (!pts || !status))) return fail(ctx, SSPP_ERR_INVALID, "validity: null argument");
  if (n == 0) return SSPP_OK;
  for (int a = 0; a < 3; ++a)
    if (!(box[a] >= 0.0) || !std::isfinite(box[a])) return fail(ctx, SSPP_ERR_INVALID, "validity: box must be finite and >= 0");
  CK(cudaSetDevice(ctx->device));
  CK(ctx->q0.reserve((size_t)n * 24));
  CK(ctx->q2.reserve((size_t)n));
  CK(cudaMemcpyAsync(ctx->q0.p, pts, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
  k_validity<<<ceil_div(n, 128), 128, 0, ctx->stream>>>(ctx->map, ctx->q0.as<double>(), n, box[0], box[1], box[2],
                                                       ctx->q2.as<uint8_t>());
  LAUNCHED(ctx);
  CK(cudaMemcpyAsync(status, ctx->q2.p, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

int sspp_gpu_cast_rays(sspp_gpu_ctx* ctx, const float* org, const float* dir, int64_t n, int ignore_unknown,
                       const double* max_range, uint8_t* hit, int32_t* hit_keys) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "no map");
  if (n < 0 || (n > 0 && (!org || !dir || !hit))) return fail(ctx, SSPP_ERR_INVALID, "cast_rays: null argument");
  if (n == 0) return SSPP_OK;
  CK(cudaSetDevice(ctx->device));
  CK(ctx->q0.reserve((size_t)n * 12));
  CK(ctx->q1.reserve((size_t)n * 12));
  CK(ctx->q2.reserve((size_t)n));
  CK(ctx->q3.reserve((size_t)n * 12));
  if (max_range) CK(ctx->q4.reserve((size_t)n * 8));
  CK(cudaMemcpyAsync(ctx->q0.p, org, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->q1.p, dir, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
  if (max_range) CK(cudaMemcpyAsync(ctx->q4.p, max_range, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
  k_cast_rays<<<ceil_div(n, 128), 128, 0, ctx->stream>>>(ctx->map, ctx->q0.as<float>(), ctx->q1.as<float>(), n, ignore_unknown,
                                                        max_range ? ctx->q4.as<double>() : nullptr, ctx->q2.as<uint8_t>(),
                                                        ctx->q3.as<int>());
  LAUNCHED(ctx);
  CK(cudaMemcpyAsync(hit, ctx->q2.p, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  if (hit_keys) CK(cudaMemcpyAsync(hit_keys, ctx->q3.p, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

// ---- coverage visibility ---------------------------------------------------------------------
int sspp_gpu_set_model_points(sspp_gpu_ctx* ctx, const float* xyz, int64_t n) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (n < 0 || (n > 0 && !xyz)) return fail(ctx, SSPP_ERR_INVALID, "set_model_points: null argument");
  if (n >= (int64_t)0x7fffffff) return fail(ctx, SSPP_ERR_RANGE, "set_model_points: point indices must fit in int32");
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  CK(ctx->pts.reserve((size_t)std::max<int64_t>(n, 1) * 12));
  if (n) CK(cudaMemcpyAsync(ctx->pts.p, xyz, (size_t)n * 12, cudaMemcpyHostToDevice, st));
  ctx->n_points = n;
  ctx->has_vis = false;
  ctx->pts_binned = false;
  ctx->pts_bbox_valid = false;
  // the bounding box (and the finiteness check) belongs to the model: computed here, once, so that
  // a visibility pass over resident points never waits for the host before its last kernel
  if (n) {
    CK(ctx->vis_bbox.reserve(7 * 8));
    static const unsigned long long init[7] = {~0ull, ~0ull, ~0ull, 0, 0, 0, 0};
    CK(cudaMemcpyAsync(ctx->vis_bbox.p, init, sizeof(init), cudaMemcpyHostToDevice, st));
    k_pts_bbox<<<std::min(ceil_div(n, 256), 148 * 8), 256, 0, st>>>(ctx->pts.as<float>(), n, ctx->vis_bbox.as<unsigned long long>());
    LAUNCHED(ctx);
    CK(cudaMemcpyAsync(ctx->pts_bbox_host, ctx->vis_bbox.p, sizeof(ctx->pts_bbox_host), cudaMemcpyDeviceToHost, st));
  }
  CK(cudaStreamSynchronize(st));  // the host array may be reused on return
  ctx->pts_bbox_valid = n > 0;
  return SSPP_OK;
}

int sspp_gpu_visibility(sspp_gpu_ctx* ctx, const double* poses, int64_t nv, const sspp_frustum* fr,
                        sspp_visibility_stats* stats) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_map) return fail(ctx, SSPP_ERR_STATE, "visibility: no map");
  if (nv < 0 || !fr || (nv > 0 && !poses)) return fail(ctx, SSPP_ERR_INVALID, "visibility: null argument");
  if (nv >= (int64_t)0x7fffffff) return fail(ctx, SSPP_ERR_RANGE, "visibility: too many viewpoints");
  if (!(fr->far_dist > 0.0) || !(fr->near_dist >= 0.0) || !std::isfinite(fr->far_dist) || !std::isfinite(fr->hfov_deg) ||
      !std::isfinite(fr->vfov_deg))
    return fail(ctx, SSPP_ERR_INVALID, "visibility: bad frustum");
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  ctx->has_vis = false;
  if (stats) memset(stats, 0, sizeof(*stats));
  for (int i = 0; i < 8; ++i) ctx->vis_timings[i] = 0;
  const int64_t np = ctx->n_points;
  CK(ctx->vis_ptr.reserve((size_t)(nv + 1) * 8));
  if (nv == 0 || np == 0) {
    CK(cudaMemsetAsync(ctx->vis_ptr.p, 0, (size_t)(nv + 1) * 8, st));
    CK(ctx->vis_idx.reserve(4));
    CK(cudaStreamSynchronize(st));
    ctx->has_vis = true; ctx->vis_nv = nv; ctx->vis_total = 0;
    return SSPP_OK;
  }
  for (int i = 0; i < 6; ++i)
    if (!ctx->vev[i]) CK(cudaEventCreate(&ctx->vev[i]));
  // host half of the spec: tan() of the half angles, narrowed exactly as the oracle states
  float vfov_rad = float(fr->vfov_deg * M_PI / 180);
  float hfov_rad = float(fr->hfov_deg * M_PI / 180);
  float vh = float(2 * tan(vfov_rad / 2));
  float hw = float(2 * tan(hfov_rad / 2));
  float np_dist = (float)fr->near_dist, fp_dist = (float)fr->far_dist;
  const bool count_steps = fr->count_steps != 0;

  // --- the model points binned into cells of edge far/4: kept while the points and the cell size stay
  if (!ctx->pts_bbox_valid) return fail(ctx, SSPP_ERR_STATE, "visibility: no model points");
  const unsigned long long* hb = ctx->pts_bbox_host;
  if (hb[6]) return fail(ctx, SSPP_ERR_INVALID, "visibility: model points must be finite");
  PointGrid g;
  double ext[3];
  for (int a = 0; a < 3; ++a) { g.min[a] = dec_f64(hb[a]); ext[a] = dec_f64(hb[3 + a]) - g.min[a]; }
  double cell = fr->far_dist / 4.0;
  double maxext = std::max(ext[0], std::max(ext[1], ext[2]));
  cell = std::max(cell, maxext * 1e-9);
  for (;;) {
    double total = 1;
    for (int a = 0; a < 3; ++a) total *= floor(ext[a] / cell) + 2.0;
    if (total <= (double)(1 << 24)) break;
    cell *= 1.26;
  }
  g.inv_cell = 1.0 / cell;
  int64_t ncells = 1;
  for (int a = 0; a < 3; ++a) { g.n[a] = (int)floor(ext[a] * g.inv_cell) + 1; ncells *= g.n[a]; }
  CK(cudaEventRecord(ctx->vev[0], st));
  if (!ctx->pts_binned || ctx->pts_cell != cell) {
    CK(ctx->cell_of.reserve((size_t)np * 4));
    CK(ctx->pts_cell_start.reserve((size_t)(ncells + 1) * 4));
    CK(ctx->cell_cursor.reserve((size_t)(ncells + 1) * 4));
    CK(ctx->sorted_idx.reserve((size_t)np * 4));
    CK(ctx->pts_sorted.reserve((size_t)np * 16));
    CK(cudaMemsetAsync(ctx->cell_cursor.p, 0, (size_t)(ncells + 1) * 4, st));
    k_pts_hist<<<ceil_div(np, 256), 256, 0, st>>>(ctx->pts.as<float>(), np, g, ctx->cell_of.as<int>(), ctx->cell_cursor.as<int>());
    LAUNCHED(ctx);
    { int rc = scan_exclusive<int>(ctx, ctx->cell_cursor.as<int>(), ncells, ctx->pts_cell_start.as<int>()); if (rc) return rc; }
    CK(cudaMemcpyAsync(ctx->cell_cursor.p, ctx->pts_cell_start.p, (size_t)ncells * 4, cudaMemcpyDeviceToDevice, st));
    k_pts_scatter<<<ceil_div(np, 256), 256, 0, st>>>(ctx->pts.as<float>(), np, ctx->cell_of.as<int>(), ctx->cell_cursor.as<int>(),
                                                    ctx->sorted_idx.as<int>(), ctx->pts_sorted.as<float4>());
    LAUNCHED(ctx);
    ctx->pts_binned = true;
    ctx->pts_cell = cell;
  }
  const float4* sp = ctx->pts_sorted.as<float4>();
  const int* pcs = ctx->pts_cell_start.as<int>();
  CK(cudaEventRecord(ctx->vev[1], st));

  // --- viewpoint records
  CK(ctx->vis_poses.reserve((size_t)nv * 56));
  CK(ctx->vis_views.reserve((size_t)nv * sizeof(ViewRec)));
  CK(cudaMemcpyAsync(ctx->vis_poses.p, poses, (size_t)nv * 56, cudaMemcpyHostToDevice, st));
  ViewRec* views = ctx->vis_views.as<ViewRec>();
  CK(ctx->vis_cnt.reserve((size_t)nv * 4 * 2));
  int* ray_count = ctx->vis_cnt.as<int>();
  int* vis_count = ray_count + nv;
  CK(ctx->vis_rayoff.reserve((size_t)(nv + 1) * 8));
  long long* ray_off = ctx->vis_rayoff.as<long long>();
  // ballot words per viewpoint (the count pass keeps its frustum verdicts for the fill pass)
  CK(ctx->vis_words.reserve((((size_t)nv * 4 + 7) & ~(size_t)7) + (size_t)(nv + 1) * 8));
  int* words = ctx->vis_words.as<int>();
  long long* word_off = (long long*)(ctx->vis_words.as<char>() + (((size_t)nv * 4 + 7) & ~(size_t)7));
  CK(ctx->vis_counters.reserve(8 * 8));
  long long* vis_ptr = ctx->vis_ptr.as<long long>();

  // The mask / ray / index buffers keep their size from call to call.  The pass is queued in one go with
  // the capacities it has; kernels refuse to run when a size outgrows its buffer, and the host sees all
  // three sizes at its single synchronisation at the end: a steady-state call waits for the device once.
  long long sizes[3] = {0, 0, 0};  // ballot words, rays, visible pairs
  unsigned long long steps = 0;
  for (int attempt = 0;; ++attempt) {
    const long long mask_cap = (long long)(ctx->vis_mask.cap / 4);
    const long long ray_cap = (long long)std::min(std::min(ctx->vis_rayview.cap / 4, ctx->vis_raypt.cap / 4), ctx->vis_flag.cap);
    const long long out_cap = (long long)(ctx->vis_idx.cap / 4);
    CK(cudaMemsetAsync(ctx->vis_counters.p, 0, 8 * 8, st));
    k_view_setup<<<ceil_div(nv, 128), 128, 0, st>>>(ctx->vis_poses.as<double>(), nv, vh, hw, np_dist, fp_dist, g, views);
    LAUNCHED(ctx);
    k_view_bound<<<ceil_div(nv * 32, 256), 256, 0, st>>>(views, (int)nv, g, pcs, words);
    LAUNCHED(ctx);
    { int rc = scan_exclusive<long long>(ctx, words, nv, word_off); if (rc) return rc; }
    uint32_t* vmask = ctx->vis_mask.as<uint32_t>();
    int* ray_view = ctx->vis_rayview.as<int>();
    int* ray_pt = ctx->vis_raypt.as<int>();
    uint8_t* vis = ctx->vis_flag.as<uint8_t>();
    CK(cudaMemsetAsync(ray_count, 0, (size_t)nv * 4 * 2, st));
    k_view_gather<false><<<ceil_div(nv * 32, 256), 256, 0, st>>>(views, 0, (int)nv, g, pcs, sp, ray_count, nullptr, nullptr,
                                                                nullptr, word_off, vmask, mask_cap, ray_cap);
    LAUNCHED(ctx);
    { int rc = scan_exclusive<long long>(ctx, ray_count, nv, ray_off); if (rc) return rc; }
    k_view_gather<true><<<ceil_div(nv * 32, 256), 256, 0, st>>>(views, 0, (int)nv, g, pcs, sp, nullptr, ray_off, ray_view, ray_pt,
                                                               word_off, vmask, mask_cap, ray_cap);
    LAUNCHED(ctx);
    CK(cudaEventRecord(ctx->vev[2], st));
    // --- occlusion rays
    {
      const int blocks = ctx->sm_count * VIEW_RAYS_RESIDENT * 2;  // two rounds of resident CTAs, handed out grid-stride
      if (count_steps)
        k_view_rays<true><<<blocks, 256, 0, st>>>(ctx->map, views, 0, sp, ray_view, ray_pt, ray_off + nv, ray_cap, vis, vis_count,
                                                  ctx->vis_counters.as<unsigned long long>());
      else
        k_view_rays<false><<<blocks, 256, 0, st>>>(ctx->map, views, 0, sp, ray_view, ray_pt, ray_off + nv, ray_cap, vis, vis_count,
                                                   ctx->vis_counters.as<unsigned long long>());
      LAUNCHED(ctx);
    }
    CK(cudaEventRecord(ctx->vev[3], st));
    { int rc = scan_exclusive<long long>(ctx, vis_count, nv, vis_ptr); if (rc) return rc; }
    k_view_compact<<<ceil_div(nv * 32, 256), 256, 0, st>>>((int)nv, ray_off, vis, ray_pt, sp, vis_ptr, ctx->vis_idx.as<int>(), ray_cap,
                                                          out_cap);
    LAUNCHED(ctx);
    k_segment_sort<<<(unsigned)std::min<int64_t>(nv, 148 * 8), 256, 8192 * 4, st>>>((int)nv, vis_ptr, ctx->vis_idx.as<int>(), out_cap);
    LAUNCHED(ctx);
    CK(cudaEventRecord(ctx->vev[4], st));
    CK(cudaMemcpyAsync(&sizes[0], word_off + nv, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&sizes[1], ray_off + nv, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&sizes[2], vis_ptr + nv, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&steps, ctx->vis_counters.p, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    bool again = false;
    if (sizes[0] > mask_cap) { CK(ctx->vis_mask.reserve((size_t)(sizes[0] + sizes[0] / 8 + 64) * 4)); again = true; }
    if (!again && sizes[1] > ray_cap) {
      const size_t want = (size_t)(sizes[1] + sizes[1] / 8 + 64);
      CK(ctx->vis_rayview.reserve(want * 4));
      CK(ctx->vis_raypt.reserve(want * 4));
      CK(ctx->vis_flag.reserve(want));
      again = true;
    }
    if (!again && sizes[2] > out_cap) { CK(ctx->vis_idx.reserve((size_t)(sizes[2] + sizes[2] / 8 + 64) * 4)); again = true; }
    if (!again) break;
    if (attempt > 4) return fail(ctx, SSPP_ERR_CUDA, "visibility: buffers still too small after growing them");
  }
  cudaEventElapsedTime(&ctx->vis_timings[0], ctx->vev[0], ctx->vev[4]);
  cudaEventElapsedTime(&ctx->vis_timings[1], ctx->vev[0], ctx->vev[1]);
  cudaEventElapsedTime(&ctx->vis_timings[2], ctx->vev[1], ctx->vev[2]);
  cudaEventElapsedTime(&ctx->vis_timings[3], ctx->vev[2], ctx->vev[3]);
  cudaEventElapsedTime(&ctx->vis_timings[4], ctx->vev[3], ctx->vev[4]);
  ctx->has_vis = true;
  ctx->vis_nv = nv;
  ctx->vis_total = sizes[2];
  if (stats) { stats->visible_pairs = sizes[2]; stats->rays = sizes[1]; stats->voxel_steps = count_steps ? (int64_t)steps : 0; }
  return SSPP_OK;
}

int sspp_gpu_visibility_timings(sspp_gpu_ctx* ctx, float ms[8]) {
  if (!ctx || !ms) return SSPP_ERR_INVALID;
  for (int i = 0; i < 8; ++i) ms[i] = ctx->vis_timings[i];
  return SSPP_OK;
}

int sspp_gpu_visibility_fetch(sspp_gpu_ctx* ctx, int64_t* view_ptr, int32_t* point_idx) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_vis) return fail(ctx, SSPP_ERR_STATE, "visibility_fetch: no visibility result");
  if (!view_ptr || (ctx->vis_total > 0 && !point_idx)) return fail(ctx, SSPP_ERR_INVALID, "visibility_fetch: null argument");
  CK(cudaSetDevice(ctx->device));
  CK(cudaMemcpyAsync(view_ptr, ctx->vis_ptr.p, (size_t)(ctx->vis_nv + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (ctx->vis_total)
    CK(cudaMemcpyAsync(point_idx, ctx->vis_idx.p, (size_t)ctx->vis_total * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return SSPP_OK;
}

int sspp_gpu_measure_l2_bandwidth(sspp_gpu_ctx* ctx, int64_t bytes, int iters, double* gbps) {
  if (!ctx || !gbps || bytes < 4096 || iters < 1) return fail(ctx, SSPP_ERR_INVALID, "measure_l2_bandwidth: bad argument");
  CK(cudaSetDevice(ctx->device));
  size_t n16 = (size_t)bytes / 16;
  CK(ctx->q0.reserve(n16 * 16 + 16));
  CK(cudaMemsetAsync(ctx->q0.p, 1, n16 * 16, ctx->stream));
  unsigned* sink = (unsigned*)((char*)ctx->q0.p + n16 * 16);
  k_l2_read<<<148 * 4, 512, 0, ctx->stream>>>(ctx->q0.as<uint4>(), n16, 2, sink);  // warm the cache
  LAUNCHED(ctx);
  CK(cudaEventRecord(ctx->ev[0], ctx->stream));
  k_l2_read<<<148 * 4, 512, 0, ctx->stream>>>(ctx->q0.as<uint4>(), n16, iters, sink);
  LAUNCHED(ctx);
  CK(cudaEventRecord(ctx->ev[1], ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]));
  *gbps = (double)n16 * 16.0 * iters / (ms * 1e-3) / 1e9;
  return SSPP_OK;
}

}  // extern "C"
