// context.cuh — the engine context, device buffers and the error / launch helpers every part uses.
// Part of the single translation unit sspp_gpu.cu (included there, in order); not a stand-alone header.
#pragma once

// ------------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------------
namespace {

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { want = bytes; e = cudaMalloc(&p, want); }
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return (T*)p; }
};

struct RowReach {
  double reach[3];    // radius + box/2 per axis: how far from a row's node any of its rays can get
};

}  // namespace

struct sspp_gpu_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  int64_t launches = 0;

  // map
  bool has_map = false;
  bool planes_dirty = true;
  DeviceMap map;
  int user_dims[3] = {0, 0, 0};
  size_t map_words = 0, cmap_words = 0;
  DevBuf occ, known, blocked, cblocked, map_stage;
  // incremental update (point-cloud insertion): float log-odds per voxel, per-cloud flag planes, first-point table
  DevBuf logodds, scan_free, scan_occ, pc_pts, pc_keys, pc_vals;
  bool has_logodds = false;
  float prob_hit_log = 0, prob_miss_log = 0, clamp_min_log = 0, clamp_max_log = 0, occ_thres_log = 0;

  // nodes
  int64_t n_nodes = 0;
  DevBuf xyz;
  bool bbox_pending = false;          // k_bbox of the current node set has been queued, result not read yet
  bool bbox_valid = false;
  unsigned long long bbox_host[7] = {0, 0, 0, 0, 0, 0, 0};
  int64_t shard_lo = -1, shard_hi = -1;  // row range whose bounding box is cached in shard_bbox_host
  unsigned long long shard_bbox_host[7] = {0, 0, 0, 0, 0, 0, 0};

  // connect scratch + result
  DevBuf bbox, cell_of, cell_start, cell_cursor, sorted_idx, sx, sy, sz, rows, row_flag, row_free, row_class;
  DevBuf cand_count, cand_off, candA, candB, acc, deg, row_ptr, col_idx, counters, scan_tmp, lattice, work_rows;
  RayLattice lattice_host;     // what ctx->lattice holds on the device (valid: lattice_valid)
  bool lattice_valid = false;
  int sm_count = 148;
  bool has_result = false;
  int64_t res_rows = 0, res_edges = 0;
  cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  float timings[8] = {0, 0, 0, 0, 0, 0, 0, 0};

  // chunked delivery of the CSR to host buffers (sspp_gpu_connect_nodes_host)
  cudaStream_t copy_stream = nullptr;
  cudaStream_t push_stream = nullptr;  // sspp_gpu_connect_to_peers: the chunk pushes, beside the next chunk's edge kernel
  cudaEvent_t chunk_ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  long long* pinned = nullptr;  // small pinned host block the chunk totals land in
  DevBuf chunk_base;

  // peer-addressable buffers (multi-GPU CSR assembly): created here / opened from a peer's IPC handle
  std::vector<void*> peer_created, peer_opened;

  // batched insertNode (pose de-duplication)
  DevBuf dd_poses, dd_xyz, dd_state, dd_pend, dd_dup, dd_keep, dd_misc;

  // generic query scratch
  DevBuf q0, q1, q2, q3, q4;

  // model points / visibility
  int64_t n_points = 0;
  DevBuf pts, vis_cnt, vis_ptr, vis_idx, vis_poses, vis_counters, vis_bbox, vis_words, vis_mask;
  bool has_vis = false;
  int64_t vis_nv = 0, vis_total = 0;
  // binning of the model points, kept while the points and the cell size do not change
  DevBuf pts_cell_start, pts_sorted, vis_views, vis_rayoff, vis_rayview, vis_raypt, vis_flag;
  bool pts_binned = false;
  double pts_cell = 0;
  unsigned long long pts_bbox_host[7] = {0, 0, 0, 0, 0, 0, 0};
  bool pts_bbox_valid = false;
  cudaEvent_t vev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  float vis_timings[8] = {0, 0, 0, 0, 0, 0, 0, 0};
};

namespace {

int fail(sspp_gpu_ctx* c, int code, const std::string& msg) {
  if (c) c->err = msg;
  return code;
}
int cuda_fail(sspp_gpu_ctx* c, cudaError_t e, const char* where) {
  return fail(c, SSPP_ERR_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}
#define CK(call)                                                   \
  do {                                                             \
    cudaError_t e__ = (call);                                      \
    if (e__ != cudaSuccess) return cuda_fail(ctx, e__, #call);     \
  } while (0)
#define LAUNCHED(ctx) do { ++(ctx)->launches; CK(cudaGetLastError()); } while (0)

inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

}  // namespace
