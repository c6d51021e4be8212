// sspp_gpu.cu — CUDA engine (sm_100a) and C ABI of include/sspp_gpu.h.
//
// Hot path (reference file:line in include/sspp_gpu.h):
//   map ingest      k_upload_boxes / k_set_box / k_dense_in  -> bit planes, k_blocked, k_pyramid
//   neighbour search k_bbox, k_cell_hist, scan, k_cell_scatter  (uniform grid, cell edge >= radius)
//                    k_count                                    (in-radius pairs per source row)
//   edge test        k_edges   one warp per source row: gather candidates, order them by list
//                              index, stage the blocked-plane tile in shared memory, run the
//                              swept-box ray lattice (lanes = rays) with ballot early exit
//   CSR              scan + k_compact
//   queries          k_check_segments, k_validity, k_cast_rays
// There is no CPU fallback: every entry point runs these kernels or fails with a status.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/sspp_gpu.h"
#include "engine.cuh"

using namespace sspp;

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

struct TileShape {
  int XW, SY, SZ;     // 0 = no staging
  double reach[3];    // radius + box/2 per axis
  double half[3];     // box/2 per axis (largest ray offset)
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
  DevBuf bbox, cell_of, cell_start, cell_cursor, sorted_idx, sx, sy, sz, rows, row_flag, row_free;
  DevBuf cand_count, cand_off, candA, candB, acc, deg, row_ptr, col_idx, counters, scan_tmp, lattice, work_rows;
  int sm_count = 148;
  bool has_result = false;
  int64_t res_rows = 0, res_edges = 0;
  cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  float timings[8] = {0, 0, 0, 0, 0, 0, 0, 0};

  // generic query scratch
  DevBuf q0, q1, q2, q3, q4;

  // model points / visibility
  int64_t n_points = 0;
  DevBuf pts, vis_cnt, vis_ptr, vis_idx, vis_poses, vis_counters;
  bool has_vis = false;
  int64_t vis_nv = 0, vis_total = 0;
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

// ------------------------------------------------------------------------------------------------
// exclusive scan (int32 in, T out, n+1 outputs): block sums -> one-block scan -> block scan
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__global__ void k_scan_block_sums(const int* __restrict__ in, int64_t n, long long* __restrict__ sums) {
  __shared__ long long sh[SCAN_THREADS / 32];
  int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
  long long v = 0;
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    int64_t i = base + (int64_t)k * SCAN_THREADS + threadIdx.x;
    if (i < n) v += in[i];
  }
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t = 0;
    for (int w = 0; w < SCAN_THREADS / 32; ++w) t += sh[w];
    sums[blockIdx.x] = t;
  }
}

__global__ void k_scan_sums(long long* sums, int nb) {
  // one block; sequential over chunks of blockDim.x with a running carry
  __shared__ long long sh[1024];
  __shared__ long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += blockDim.x) {
    int i = base + threadIdx.x;
    long long v = (i < nb) ? sums[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < (int)blockDim.x; o <<= 1) {
      long long t = (threadIdx.x >= (unsigned)o) ? sh[threadIdx.x - o] : 0;
      __syncthreads();
      sh[threadIdx.x] += t;
      __syncthreads();
    }
    long long incl = sh[threadIdx.x];
    if (i < nb) sums[i] = carry + incl - v;  // exclusive
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry += incl;
    __syncthreads();
  }
}

template <class T>
__global__ void k_scan_apply(const int* __restrict__ in, int64_t n, const long long* __restrict__ sums, T* __restrict__ out) {
  // items are laid out thread-major inside the tile so the local scan is a plain running sum
  __shared__ long long sh[SCAN_THREADS];
  int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
  int v[SCAN_ITEMS];
  long long t = 0;
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    int64_t i = base + k;
    v[k] = (i < n) ? in[i] : 0;
    t += v[k];
  }
  sh[threadIdx.x] = t;
  __syncthreads();
  for (int o = 1; o < SCAN_THREADS; o <<= 1) {
    long long u = (threadIdx.x >= (unsigned)o) ? sh[threadIdx.x - o] : 0;
    __syncthreads();
    sh[threadIdx.x] += u;
    __syncthreads();
  }
  long long run = sums[blockIdx.x] + sh[threadIdx.x] - t;
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    int64_t i = base + k;
    if (i < n) out[i] = (T)run;
    run += v[k];
    if (i == n - 1) out[n] = (T)run;
  }
}

template <class T>
__global__ void k_scan_empty(T* out) { out[0] = 0; }

// out must hold n+1 entries; out[n] = total
template <class T>
int scan_exclusive(sspp_gpu_ctx* ctx, const int* in, int64_t n, T* out) {
  if (n == 0) {
    k_scan_empty<T><<<1, 1, 0, ctx->stream>>>(out);
    LAUNCHED(ctx);
    return SSPP_OK;
  }
  int nb = ceil_div(n, SCAN_TILE);
  CK(ctx->scan_tmp.reserve((size_t)nb * sizeof(long long)));
  long long* sums = ctx->scan_tmp.as<long long>();
  k_scan_block_sums<<<nb, SCAN_THREADS, 0, ctx->stream>>>(in, n, sums);
  LAUNCHED(ctx);
  k_scan_sums<<<1, 1024, 0, ctx->stream>>>(sums, nb);
  LAUNCHED(ctx);
  k_scan_apply<T><<<nb, SCAN_THREADS, 0, ctx->stream>>>(in, n, sums, out);
  LAUNCHED(ctx);
  return SSPP_OK;
}

// ------------------------------------------------------------------------------------------------
// map kernels
// ------------------------------------------------------------------------------------------------

// host dense layout (x-fastest rows of `wpr` words, user dims) -> device layout (xw, z, y)
__global__ void k_dense_in(const uint32_t* __restrict__ src, int wpr, int uy, int uz, uint32_t* __restrict__ dst, int nxw,
                           int ny, int nz, int ux) {
  size_t total = (size_t)nxw * ny * nz;
  for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (size_t)gridDim.x * blockDim.x) {
    int y = (int)(w % ny);
    int z = (int)((w / ny) % nz);
    int xw = (int)(w / ((size_t)ny * nz));
    uint32_t v = 0;
    if (xw < wpr && y < uy && z < uz) {
      v = src[((size_t)z * uy + y) * wpr + xw];
      int valid = ux - xw * 32;  // bits beyond the user's x extent are not part of the map
      if (valid < 32) v &= (valid <= 0) ? 0u : ((1u << valid) - 1u);
    }
    dst[w] = v;
  }
}

__global__ void k_dense_out(const uint32_t* __restrict__ src, int nxw, int ny, int nz, uint32_t* __restrict__ dst, int wpr,
                            int uy, int uz, int ux) {
  size_t total = (size_t)wpr * uy * uz;
  for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (size_t)gridDim.x * blockDim.x) {
    int xw = (int)(w % wpr);
    int y = (int)((w / wpr) % uy);
    int z = (int)(w / ((size_t)wpr * uy));
    uint32_t v = src[((size_t)xw * nz + z) * ny + y];
    int valid = ux - xw * 32;
    if (valid < 32) v &= (valid <= 0) ? 0u : ((1u << valid) - 1u);
    dst[w] = v;
  }
}

// one warp per leaf box: rows (y,z) over lanes, x range as word masks
// L (may be null): the log-odds grid, kept in line with the planes — leaves written here sit at the
// clamping values, as setFree / setOccupied leave them
__global__ void k_upload_boxes(DeviceMap m, uint32_t* occ, uint32_t* known, const double* __restrict__ centres,
                               const double* __restrict__ sizes, const uint8_t* __restrict__ occupied, int64_t n, float* L,
                               float lo_v, float hi_v) {
  int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (b >= n) return;
  double s = sizes[b];
  int nv = (int)floor(s * m.ks.res_factor + 0.5);  // leaves are res * 2^j cubes
  if (nv < 1) nv = 1;
  int k0[3];
  for (int a = 0; a < 3; ++a) {
    // lower corner is a voxel boundary: round((c - s/2) / res) in key units
    double lo = (centres[3 * b + a] - 0.5 * s) * m.ks.res_factor;
    k0[a] = (int)floor(lo + 0.5) + 32768 - m.kmin[a];
  }
  bool is_occ = occupied[b] != 0;
  int x0 = k0[0], x1 = k0[0] + nv - 1;
  if (x1 < 0 || x0 >= m.dims[0]) return;
  x0 = max(x0, 0);
  x1 = min(x1, m.dims[0] - 1);
  int64_t nrows = (int64_t)nv * nv;
  for (int64_t r = lane; r < nrows; r += 32) {
    int y = k0[1] + (int)(r % nv), z = k0[2] + (int)(r / nv);
    if ((unsigned)y >= (unsigned)m.dims[1] || (unsigned)z >= (unsigned)m.dims[2]) continue;
    for (int xw = x0 >> 5; xw <= (x1 >> 5); ++xw) {
      int lo = max(x0, xw * 32) & 31, hi = min(x1, xw * 32 + 31) & 31;
      uint32_t mask = (hi == 31 ? 0xffffffffu : ((1u << (hi + 1)) - 1u)) & ~((1u << lo) - 1u);
      size_t w = map_word(m, xw, y, z);
      atomicOr(known + w, mask);
      if (is_occ) atomicOr(occ + w, mask); else atomicAnd(occ + w, ~mask);
      if (L)
        for (uint32_t rest = mask; rest; rest &= rest - 1) L[w * 32 + (__ffs(rest) - 1)] = is_occ ? hi_v : lo_v;
    }
  }
}

// setFree / setOccupied: explicit per-axis key lists (window coordinates) from the host lattice loop
__global__ void k_set_box(DeviceMap m, uint32_t* occ, uint32_t* known, const int* __restrict__ kx, int nx,
                          const int* __restrict__ ky, int ny, const int* __restrict__ kz, int nz, int is_occ, float* L,
                          float lo_v, float hi_v) {
  int64_t total = (int64_t)nx * ny * nz;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    int x = kx[t % nx], y = ky[(t / nx) % ny], z = kz[t / ((int64_t)nx * ny)];
    if ((unsigned)x >= (unsigned)m.dims[0] || (unsigned)y >= (unsigned)m.dims[1] || (unsigned)z >= (unsigned)m.dims[2]) continue;
    size_t w = map_word(m, x >> 5, y, z);
    uint32_t bit = 1u << (x & 31);
    atomicOr(known + w, bit);
    if (is_occ) atomicOr(occ + w, bit); else atomicAnd(occ + w, ~bit);
    if (L) L[w * 32 + (x & 31)] = is_occ ? hi_v : lo_v;
  }
}

__global__ void k_blocked(const uint32_t* __restrict__ occ, const uint32_t* __restrict__ known, uint32_t* __restrict__ blocked,
                          size_t words, int treat_unknown) {
  for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (size_t)gridDim.x * blockDim.x)
    blocked[w] = occ[w] | (treat_unknown ? ~known[w] : 0u);
}

// coarse word (cxw, cz, cy) <- OR over 4 fine words in x (128 voxels -> 32 coarse bits), 4 in y, 4 in z
__global__ void k_pyramid(DeviceMap m, const uint32_t* __restrict__ blocked, uint32_t* __restrict__ cblocked) {
  size_t total = (size_t)m.cnxw * m.cdims[1] * m.cdims[2];
  for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (size_t)gridDim.x * blockDim.x) {
    int cy = (int)(w % m.cdims[1]);
    int cz = (int)((w / m.cdims[1]) % m.cdims[2]);
    int cxw = (int)(w / ((size_t)m.cdims[1] * m.cdims[2]));
    uint32_t out = 0;
    for (int q = 0; q < 4; ++q) {
      uint32_t v = 0;
      for (int dz = 0; dz < 4; ++dz)
        for (int dy = 0; dy < 4; ++dy) v |= blocked[map_word(m, cxw * 4 + q, cy * 4 + dy, cz * 4 + dz)];
      // fold every nibble to one bit
      v |= v >> 1;
      v |= v >> 2;
      uint32_t c = 0;
      for (int nb = 0; nb < 8; ++nb) c |= ((v >> (4 * nb)) & 1u) << nb;
      out |= c << (8 * q);
    }
    cblocked[w] = out;
  }
}

// ---- incremental update: point-cloud insertion (OctomapWorld::insertPointcloudIntoMapImpl, upstream) ----
// log-odds grid: voxel (word w, bit b) -> logodds[w * 32 + b]

__global__ void k_logodds_init(const uint32_t* __restrict__ occ, const uint32_t* __restrict__ known, float* __restrict__ L,
                               size_t words, float lo, float hi) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < words * 32; i += (size_t)gridDim.x * blockDim.x) {
    const uint32_t bit = 1u << (i & 31);
    const size_t w = i >> 5;
    L[i] = (known[w] & bit) ? ((occ[w] & bit) ? hi : lo) : 0.0f;
  }
}

__device__ __forceinline__ unsigned long long pc_hash(unsigned long long k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return k;
}
// OcTreeBaseImpl::coordToKey (unchecked): the 16-bit truncation of floor(c / res) + 32768
__device__ __forceinline__ unsigned key_unchecked(const KeySpace& ks, float c) {
  return ((unsigned)__double2int_rd(d_mul(ks.res_factor, (double)c)) + 32768u) & 0xffffu;
}
__device__ __forceinline__ bool pc_in_range(float px, float py, float pz, float ox, float oy, float oz, double max_range) {
  return max_range < 0.0 || norm3f(f_sub(px, ox), f_sub(py, oy), f_sub(pz, oz)) <= max_range;
}

// pass 1: first in-range point (cloud order) of every end voxel
__global__ void k_pc_first(KeySpace ks, const float* __restrict__ pts, int64_t n, float ox, float oy, float oz, double max_range,
                           unsigned long long* keys, int* vals, unsigned long long mask) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float px = pts[3 * i], py = pts[3 * i + 1], pz = pts[3 * i + 2];
  if (!pc_in_range(px, py, pz, ox, oy, oz, max_range)) return;
  int kx, ky, kz;
  if (!coord_to_key(ks, (double)px, kx) || !coord_to_key(ks, (double)py, ky) || !coord_to_key(ks, (double)pz, kz)) return;
  const unsigned long long key = ((unsigned long long)kx << 32) | ((unsigned long long)ky << 16) | (unsigned long long)kz;
  unsigned long long slot = pc_hash(key) & mask;
  for (;;) {
    unsigned long long prev = atomicCAS(keys + slot, ~0ull, key);
    if (prev == ~0ull || prev == key) { atomicMin(vals + slot, (int)i); return; }
    slot = (slot + 1) & mask;
  }
}

struct MarkFree {
  const DeviceMap* m;
  uint32_t* plane;
  __device__ __forceinline__ int operator()(int kx, int ky, int kz) const {
    int x = kx - m->kmin[0], y = ky - m->kmin[1], z = kz - m->kmin[2];
    if ((unsigned)x < (unsigned)m->dims[0] && (unsigned)y < (unsigned)m->dims[1] && (unsigned)z < (unsigned)m->dims[2])
      atomicOr(plane + map_word(*m, x >> 5, y, z), 1u << (x & 31));
    return 0;
  }
};

// pass 2: one ray per point that is not shadowed by an earlier point of its voxel
__global__ void __launch_bounds__(128) k_pc_cast(DeviceMap m, const float* __restrict__ pts, int64_t n, float ox, float oy, float oz,
                                                 double max_range, const unsigned long long* __restrict__ keys,
                                                 const int* __restrict__ vals, unsigned long long mask, uint32_t* scan_free,
                                                 uint32_t* scan_occ) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float px = pts[3 * i], py = pts[3 * i + 1], pz = pts[3 * i + 2];
  {
    // "already in occupied_cells": some earlier in-range point ended in this voxel
    const unsigned long long key = ((unsigned long long)key_unchecked(m.ks, px) << 32) |
                                   ((unsigned long long)key_unchecked(m.ks, py) << 16) | (unsigned long long)key_unchecked(m.ks, pz);
    unsigned long long slot = pc_hash(key) & mask;
    for (;;) {
      const unsigned long long k = keys[slot];
      if (k == ~0ull) break;
      if (k == key) { if (vals[slot] < (int)i) return; break; }
      slot = (slot + 1) & mask;
    }
  }
  MarkFree mark;
  mark.m = &m;
  mark.plane = scan_free;
  if (pc_in_range(px, py, pz, ox, oy, oz, max_range)) {
    line_status(m.ks, ox, oy, oz, px, py, pz, mark);
    int kx, ky, kz;
    if (coord_to_key(m.ks, (double)px, kx) && coord_to_key(m.ks, (double)py, ky) && coord_to_key(m.ks, (double)pz, kz)) {
      mark.plane = scan_occ;
      mark(kx, ky, kz);
    }
  } else {
    // free space only, up to the range: origin + normalized(p - origin) * (float)range
    float dx = f_sub(px, ox), dy = f_sub(py, oy), dz = f_sub(pz, oz);
    const double len = norm3f(dx, dy, dz);
    if (len > 0) { const float fl = d2f(len); dx = f_div(dx, fl); dy = f_div(dy, fl); dz = f_div(dz, fl); }
    const float r = d2f(max_range);
    line_status(m.ks, ox, oy, oz, f_add(ox, f_mul(dx, r)), f_add(oy, f_mul(dy, r)), f_add(oz, f_mul(dz, r)), mark);
  }
}

// pass 3: every flagged voxel once — occupied wins over free; add, clamp, re-threshold; clear the flags
__global__ void k_pc_apply(uint32_t* occ, uint32_t* known, uint32_t* scan_free, uint32_t* scan_occ, float* __restrict__ L,
                           size_t words, float hit, float miss, float lo, float hi, float thres) {
  for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (size_t)gridDim.x * blockDim.x) {
    const uint32_t f = scan_free[w], o = scan_occ[w];
    uint32_t any = f | o;
    if (!any) continue;
    uint32_t ow = occ[w];
    for (uint32_t rest = any; rest; rest &= rest - 1) {
      const int b = __ffs(rest) - 1;
      float v = f_add(L[w * 32 + b], ((o >> b) & 1u) ? hit : miss);
      if (v < lo) v = lo; else if (v > hi) v = hi;
      L[w * 32 + b] = v;
      if (v >= thres) ow |= 1u << b; else ow &= ~(1u << b);
    }
    occ[w] = ow;
    known[w] |= any;
    scan_free[w] = 0;
    scan_occ[w] = 0;
  }
}

// device (xw,z,y) log-odds -> host x-fastest float grid, NaN where unknown
__global__ void k_logodds_out(const float* __restrict__ L, const uint32_t* __restrict__ known, int nxw, int ny, int nz,
                              float* __restrict__ dst, int ux, int uy, int uz) {
  size_t total = (size_t)ux * uy * uz;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int x = (int)(i % ux), y = (int)((i / ux) % uy), z = (int)(i / ((size_t)ux * uy));
    size_t w = ((size_t)(x >> 5) * nz + z) * ny + y;
    dst[i] = ((known[w] >> (x & 31)) & 1u) ? L[w * 32 + (x & 31)] : __int_as_float(0x7fc00000);
  }
}

// ------------------------------------------------------------------------------------------------
// neighbour search kernels
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long enc_f64(double v) {
  unsigned long long b = (unsigned long long)__double_as_longlong(v);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
inline double dec_f64(unsigned long long k) {
  unsigned long long b = (k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k;
  double d;
  memcpy(&d, &b, 8);
  return d;
}

// bbox[0..2] = min (encoded), bbox[3..5] = max (encoded), bbox[6] = count of non-finite coordinates
__global__ void k_bbox(const double* __restrict__ xyz, int64_t n, unsigned long long* bbox) {
  unsigned long long mn[3] = {~0ull, ~0ull, ~0ull}, mx[3] = {0, 0, 0};
  unsigned long long bad = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    for (int a = 0; a < 3; ++a) {
      double v = xyz[3 * i + a];
      if (!isfinite(v)) { ++bad; continue; }
      unsigned long long e = enc_f64(v);
      mn[a] = min(mn[a], e);
      mx[a] = max(mx[a], e);
    }
  }
  for (int a = 0; a < 3; ++a) {
    for (int o = 16; o; o >>= 1) {
      mn[a] = min(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = max(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  }
  for (int o = 16; o; o >>= 1) bad += __shfl_xor_sync(0xffffffffu, bad, o);
  if ((threadIdx.x & 31) == 0) {
    for (int a = 0; a < 3; ++a) { atomicMin(bbox + a, mn[a]); atomicMax(bbox + 3 + a, mx[a]); }
    if (bad) atomicAdd(bbox + 6, bad);
  }
}

// Nodes outside [g.min, g.max] (a shard only bins the nodes its rows can reach) get cell -1 and are
// left out of the sorted arrays.
__global__ void k_cell_hist(const double* __restrict__ xyz, int64_t n, NodeGrid g, int* __restrict__ cell_of, int* cell_count) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  if (x < g.min[0] || x > g.max[0] || y < g.min[1] || y > g.max[1] || z < g.min[2] || z > g.max[2]) {
    cell_of[i] = -1;
    return;
  }
  int cx = grid_cell_axis(g, x, 0), cy = grid_cell_axis(g, y, 1), cz = grid_cell_axis(g, z, 2);
  int c = (cz * g.n[1] + cy) * g.n[0] + cx;
  cell_of[i] = c;
  atomicAdd(cell_count + c, 1);
}

__global__ void k_cell_scatter(const double* __restrict__ xyz, int64_t n, const int* __restrict__ cell_of, int* cursor,
                               int* __restrict__ sorted_idx, double* __restrict__ sx, double* __restrict__ sy,
                               double* __restrict__ sz) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || cell_of[i] < 0) return;
  int pos = atomicAdd(cursor + cell_of[i], 1);
  sorted_idx[pos] = (int)i;
  sx[pos] = xyz[3 * i];
  sy[pos] = xyz[3 * i + 1];
  sz[pos] = xyz[3 * i + 2];
}

// Order the members of every cell by list index, so the whole pipeline is run-to-run deterministic
// (the scatter above lands them in atomic order).  Cells hold a handful of nodes: insertion sort.
__global__ void k_cell_sort(const int* __restrict__ cell_start, int ncells, int* sorted_idx, double* sx, double* sy, double* sz) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncells) return;
  int b = cell_start[c], e = cell_start[c + 1];
  for (int i = b + 1; i < e; ++i) {
    int k = sorted_idx[i];
    double x = sx[i], y = sy[i], z = sz[i];
    int j = i - 1;
    while (j >= b && sorted_idx[j] > k) {
      sorted_idx[j + 1] = sorted_idx[j]; sx[j + 1] = sx[j]; sy[j + 1] = sy[j]; sz[j + 1] = sz[j];
      --j;
    }
    sorted_idx[j + 1] = k; sx[j + 1] = x; sy[j + 1] = y; sz[j + 1] = z;
  }
}

// rows of a shard: flag sorted positions whose list index is in [lo, hi)
// n_sorted_ptr: how many nodes were binned (device side); positions beyond it hold nothing
__global__ void k_row_flags(const int* __restrict__ sorted_idx, int64_t n, const int* __restrict__ n_sorted_ptr, int64_t lo,
                            int64_t hi, int* __restrict__ flag) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n) flag[s] = (s < *n_sorted_ptr && sorted_idx[s] >= lo && sorted_idx[s] < hi) ? 1 : 0;
}
__global__ void k_row_scatter(const int* __restrict__ flag, const int* __restrict__ pos, int64_t n, int* __restrict__ rows) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n && flag[s]) rows[pos[s]] = (int)s;
}

// pair classification shared by count and gather: 0 = not connected, 1 = goes to the collision test,
// 2 = same position (connected without a test).  S = candidate, T = source row ("temp").
// The reference tests sqrt(d2) <= radius (Dist, include/sspp/utils.h:356-359).  sqrt is monotone and
// correctly rounded, so that is the same predicate as d2 <= r2max with r2max = the largest double
// whose square root does not exceed the radius — computed once on the host (radius_threshold()).
__device__ __forceinline__ int pair_class(double sx, double sy, double sz, double tx, double ty, double tz, double r2max,
                                          int mode) {
  double dx = d_sub(sx, tx), dy = d_sub(sy, ty), dz = d_sub(sz, tz);
  double d2 = d_add(d_add(d_mul(dx, dx), d_mul(dy, dy)), d_mul(dz, dz));
  if (!(d2 <= r2max)) return 0;
  bool differs;
  if (mode == 0) differs = (sx != tx || sy != ty || sz != tz);
  else differs = !(fabs(dx) < 1E-12 && fabs(dy) < 1E-12 && fabs(dz) < 1E-12);
  return differs ? 1 : 2;
}

// Is the whole reach of a row free at the coarse level?  Region = keys of (p -+ reach) grown by two
// voxels; every voxel a ray of this row can visit lies inside it (DESIGN.md "pyramid skip").
__device__ bool row_reach_is_free(const DeviceMap& m, double px, double py, double pz, const double reach[3]) {
  double p[3] = {px, py, pz};
  int lo[3], hi[3];
  for (int a = 0; a < 3; ++a) {
    int k0, k1;
    if (!coord_to_key(m.ks, p[a] - reach[a], k0) || !coord_to_key(m.ks, p[a] + reach[a], k1)) return false;
    lo[a] = k0 - 2 - m.kmin[a];
    hi[a] = k1 + 2 - m.kmin[a];
    if (lo[a] < 0 || hi[a] >= m.dims[a]) {
      if (m.treat_unknown) return false;  // outside the window = never observed = blocked
      lo[a] = max(lo[a], 0);
      hi[a] = min(hi[a], m.dims[a] - 1);
      if (lo[a] > hi[a]) return true;  // nothing of the reach is inside the window, and unknown is free
    }
    lo[a] >>= 2;
    hi[a] >>= 2;
  }
  for (int cz = lo[2]; cz <= hi[2]; ++cz)
    for (int cy = lo[1]; cy <= hi[1]; ++cy)
      for (int cxw = lo[0] >> 5; cxw <= (hi[0] >> 5); ++cxw) {
        int b0 = max(lo[0], cxw * 32) & 31, b1 = min(hi[0], cxw * 32 + 31) & 31;
        uint32_t mask = (b1 == 31 ? 0xffffffffu : ((1u << (b1 + 1)) - 1u)) & ~((1u << b0) - 1u);
        if (__ldg(m.cblocked + ((size_t)cxw * m.cdims[2] + cz) * m.cdims[1] + cy) & mask) return false;
      }
  return true;
}

struct ConnectArgs {
  NodeGrid g;
  const int* cell_start;
  const int* sorted_idx;
  const double* sx;
  const double* sy;
  const double* sz;
  const int* rows;  // sorted positions of the rows to process, or NULL = all
  int nrows;
  int64_t lo;  // first list index of this shard
  double r2max;  // largest d2 with sqrt(d2) <= radius
  int mode;
  int use_map;
  int use_pyramid;
  TileShape ts;
};

// counters: [0] in-radius pairs, [1] tested pairs, [2] pyramid rows, [3] voxel steps (instrumented),
//           [4] next work row (edge kernel's ticket), [5] pairs accepted by their key box,
//           [6] pairs that ran the ray lattice, [7] number of work rows
//
// One thread per source row walks the 27 neighbouring cells (9 contiguous ranges of the cell-sorted
// arrays).  FILL == false: count the in-radius pairs and test the row's reach against the pyramid.
// FILL == true (after the scan of the counts): write the pairs — (sorted position, list index with
// bit 31 = same position) — to cand[], in cell order; the edge kernel orders them by list index.
template <bool FILL>
__global__ void __launch_bounds__(256) k_count(ConnectArgs a, DeviceMap m, int* __restrict__ cand_count,
                                               uint8_t* __restrict__ row_free, unsigned long long* counters,
                                               const long long* __restrict__ cand_off, int2* __restrict__ cand,
                                               int* __restrict__ work_rows, long long pair_cap) {
  if (FILL && cand_off[a.nrows] > pair_cap) return;  // the pair buffers are too small: the host grows them and reruns
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long inr = 0, tested = 0, pyr = 0;
  bool is_work = false;
  if (t < a.nrows) {
    int s = a.rows ? a.rows[t] : t;
    double px = a.sx[s], py = a.sy[s], pz = a.sz[s];
    int r = (int)(a.sorted_idx[s] - a.lo);
    int cx = grid_cell_axis(a.g, px, 0), cy = grid_cell_axis(a.g, py, 1), cz = grid_cell_axis(a.g, pz, 2);
    int x0 = max(cx - 1, 0), x1 = min(cx + 1, a.g.n[0] - 1);
    int cnt = 0, tst = 0;
    long long w = FILL ? cand_off[r] : 0;
    for (int zz = max(cz - 1, 0); zz <= min(cz + 1, a.g.n[2] - 1); ++zz)
      for (int yy = max(cy - 1, 0); yy <= min(cy + 1, a.g.n[1] - 1); ++yy) {
        int base = (zz * a.g.n[1] + yy) * a.g.n[0];
        int b = a.cell_start[base + x0], e = a.cell_start[base + x1 + 1];
        for (int j = b; j < e; ++j) {
          if (j == s) continue;
          int pc = pair_class(a.sx[j], a.sy[j], a.sz[j], px, py, pz, a.r2max, a.mode);
          if (FILL) {
            if (pc) cand[w++] = make_int2(j, a.sorted_idx[j] | (pc == 2 ? (int)0x80000000 : 0));
          } else {
            cnt += (pc != 0);
            tst += (pc == 1);
          }
        }
      }
    if (!FILL) {
      cand_count[r] = cnt;
      uint8_t fr = 0;
      if (a.use_map && a.use_pyramid && tst > 0) fr = row_reach_is_free(m, px, py, pz, a.ts.reach) ? 1 : 0;
      row_free[r] = fr;
      inr = cnt; tested = tst; pyr = fr;
      is_work = a.use_map && !fr && cnt > 0;
    }
  }
  if (!FILL) {
    // rows that need the edge kernel, appended warp by warp (their order does not matter)
    const unsigned wb = __ballot_sync(0xffffffffu, is_work);
    if (wb) {
      const int lane = threadIdx.x & 31;
      unsigned long long base = 0;
      if (lane == 0) base = atomicAdd(counters + 7, (unsigned long long)__popc(wb));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (is_work) work_rows[base + __popc(wb & ((1u << lane) - 1u))] = t;
    }
    for (int o = 16; o; o >>= 1) {
      inr += __shfl_xor_sync(0xffffffffu, inr, o);
      tested += __shfl_xor_sync(0xffffffffu, tested, o);
      pyr += __shfl_xor_sync(0xffffffffu, pyr, o);
    }
    if ((threadIdx.x & 31) == 0) {
      if (inr) atomicAdd(counters + 0, inr);
      if (tested) atomicAdd(counters + 1, tested);
      if (pyr) atomicAdd(counters + 2, pyr);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// edge test: one warp per source row
// ------------------------------------------------------------------------------------------------
constexpr int EDGE_WARPS = 8;

// Does the key box [lo, hi] (absolute keys, inclusive) hold a blocked voxel?  "true" may also mean
// "cannot tell cheaply" (the box leaves the mapped window): callers then take the exact path.
__device__ __forceinline__ bool box_maybe_blocked(const DeviceMap& m, int lox, int hix, int loy, int hiy, int loz, int hiz) {
  const int x0 = lox - m.kmin[0], x1 = hix - m.kmin[0];
  const int y0 = loy - m.kmin[1], y1 = hiy - m.kmin[1];
  const int z0 = loz - m.kmin[2], z1 = hiz - m.kmin[2];
  if (x0 < 0 || y0 < 0 || z0 < 0 || x1 >= m.dims[0] || y1 >= m.dims[1] || z1 >= m.dims[2] || x0 > x1 || y0 > y1 || z0 > z1)
    return true;
  const int w0 = x0 >> 5, w1 = x1 >> 5;
  uint32_t any = 0;
  for (int w = w0; w <= w1; ++w) {
    const int b0 = (w == w0) ? (x0 & 31) : 0, b1 = (w == w1) ? (x1 & 31) : 31;
    const uint32_t mask = (b1 == 31 ? 0xffffffffu : ((1u << (b1 + 1)) - 1u)) & ~((1u << b0) - 1u);
    for (int z = z0; z <= z1; ++z) {
      const uint32_t* row = m.blocked + map_word(m, w, 0, z);
      for (int y = y0; y <= y1; ++y) any |= __ldg(row + y) & mask;
    }
  }
  return any != 0;
}

// the exact walk on the blocked plane, kept out of line: it is the rare path of line_status_fast and
// inlining it would cost the whole edge kernel its occupancy
__device__ __noinline__ int line_status_exact(const DeviceMap& m, float ox, float oy, float oz, float ex, float ey, float ez) {
  BlockedGlobal cell;
  cell.m = &m;
  return line_status(m.ks, ox, oy, oz, ex, ey, ez, cell);
}

__device__ __forceinline__ int line_status_fast(const DeviceMap& m, float ox, float oy, float oz, float ex, float ey, float ez) {
  int kx, ky, kz, ekx, eky, ekz;
  if (!coord_to_key(m.ks, (double)ox, kx) || !coord_to_key(m.ks, (double)oy, ky) || !coord_to_key(m.ks, (double)oz, kz) ||
      !coord_to_key(m.ks, (double)ex, ekx) || !coord_to_key(m.ks, (double)ey, eky) || !coord_to_key(m.ks, (double)ez, ekz))
    return CELL_FREE;
  if (kx == ekx && ky == eky && kz == ekz) return CELL_FREE;
  const int lox = min(kx, ekx) - (ex < ox ? 1 : 0), hix = max(kx, ekx) + (ex > ox ? 1 : 0);
  const int loy = min(ky, eky) - (ey < oy ? 1 : 0), hiy = max(ky, eky) + (ey > oy ? 1 : 0);
  const int loz = min(kz, ekz) - (ez < oz ? 1 : 0), hiz = max(kz, ekz) + (ez > oz ? 1 : 0);
  if (!box_maybe_blocked(m, lox, hix, loy, hiy, loz, hiz)) return CELL_FREE;
  return line_status_exact(m, ox, oy, oz, ex, ey, ez);
}

// Swept-box status of one segment over lanes: `G` lanes share a segment, chunks of G rays with a
// ballot after each chunk.  Returns, uniformly over the group, the status of the first non-free ray
// in lattice order (x outer, y, z inner) or CELL_FREE.  `active` = this lane's group has a segment.
// (o0x,o0y,o0z) = this lane's ray offset in the first chunk, hoisted by the caller.
// FAST: rays go through line_status_fast on the blocked plane (`cell` unused).
template <bool COUNT, bool FAST, class Cell>
__device__ __forceinline__ int swept_status(const DeviceMap& m, const RayLattice& lat, Cell& cell, bool active, int G,
                                            int sub, int lig, double o0x, double o0y, double o0z, double tx, double ty,
                                            double tz, double sx, double sy, double sz,
                                            unsigned long long* steps_acc = nullptr, bool fast = true) {
  const int nyz = lat.n[1] * lat.n[2];
  int result = CELL_FREE;
  bool done = !active;
  unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (sub * G));
  for (int q0 = 0; q0 < lat.rays; q0 += G) {
    int q = q0 + lig;
    int st = CELL_FREE;
    int nsteps = 0;
    if (!done && q < lat.rays) {
      double ox = o0x, oy = o0y, oz = o0z;
      if (q0) {
        int ix = q / nyz, rem = q - ix * nyz;
        int iy = rem / lat.n[2], iz = rem - iy * lat.n[2];
        ox = lat.off[0][ix]; oy = lat.off[1][iy]; oz = lat.off[2][iz];
      }
      const float fox = d2f(d_add(tx, ox)), foy = d2f(d_add(ty, oy)), foz = d2f(d_add(tz, oz));
      const float fex = d2f(d_add(sx, ox)), fey = d2f(d_add(sy, oy)), fez = d2f(d_add(sz, oz));
      if (FAST) st = fast ? line_status_fast(m, fox, foy, foz, fex, fey, fez) : line_status_exact(m, fox, foy, foz, fex, fey, fez);
      else st = line_status(m.ks, fox, foy, foz, fex, fey, fez, cell, COUNT ? &nsteps : nullptr);
    }
    unsigned b = __ballot_sync(0xffffffffu, st != CELL_FREE);
    unsigned mine = b & gmask;
    // lowest lane of my group that failed = first ray in lattice order (every lane shuffles)
    int first = mine ? (__ffs(mine) - 1) : (int)(threadIdx.x & 31);
    int got = __shfl_sync(0xffffffffu, st, first);
    if (COUNT && !done && (!mine || (int)(threadIdx.x & 31) <= first)) *steps_acc += (unsigned long long)nsteps;
    if (!done && mine) {
      result = got;
      done = true;
    }
    if (__all_sync(0xffffffffu, done)) break;
  }
  return result;
}

// lane's ray offset in the first chunk of the lattice
__device__ __forceinline__ void lane_offset(const RayLattice& lat, int lig, double& ox, double& oy, double& oz) {
  ox = oy = oz = 0.0;
  if (lig < lat.rays) {
    const int nyz = lat.n[1] * lat.n[2];
    int ix = lig / nyz, rem = lig - ix * nyz;
    int iy = rem / lat.n[2], iz = rem - iy * lat.n[2];
    ox = lat.off[0][ix]; oy = lat.off[1][iy]; oz = lat.off[2][iz];
  }
}

// One axis of the key box ALL rays of a pair stay in.  The lattice offsets run from off_first to
// off_last, and adding an offset and narrowing to float are both monotone, so the extreme rays
// bound the rest: start keys in [key(f(t+first)), key(f(t+last))], end keys likewise for s; plus the
// one-voxel overshoot beyond the end on the side the pair travels to (none when s == t: then every
// ray has a zero direction component on this axis).  False when a key leaves the 16-bit space.
__device__ __forceinline__ bool pair_key_box(const KeySpace& ks, double t, double s, double off_first, double off_last,
                                             int& lo, int& hi) {
  int o_lo, o_hi, e_lo, e_hi;
  if (!coord_to_key(ks, (double)d2f(d_add(t, off_first)), o_lo) || !coord_to_key(ks, (double)d2f(d_add(t, off_last)), o_hi) ||
      !coord_to_key(ks, (double)d2f(d_add(s, off_first)), e_lo) || !coord_to_key(ks, (double)d2f(d_add(s, off_last)), e_hi))
    return false;
  if (s > t) { lo = o_lo; hi = e_hi + 1; }
  else if (s < t) { lo = e_lo - 1; hi = o_hi; }
  else { lo = o_lo; hi = o_hi; }
  return true;
}

// Cheap verdict on one ray before any DDA arithmetic: RAY_FREE / RAY_BLOCKED are final, RAY_WALK
// means the exact walk has to decide.  `fast` enables the key-box shortcut of line_status_fast.
enum { RAY_FREE = 0, RAY_BLOCKED = 1, RAY_WALK = 2 };
__device__ __forceinline__ int ray_precheck(const DeviceMap& m, bool fast, float ox, float oy, float oz, float ex, float ey,
                                            float ez) {
  int kx, ky, kz, ekx, eky, ekz;
  if (!coord_to_key(m.ks, (double)ox, kx) || !coord_to_key(m.ks, (double)oy, ky) || !coord_to_key(m.ks, (double)oz, kz) ||
      !coord_to_key(m.ks, (double)ex, ekx) || !coord_to_key(m.ks, (double)ey, eky) || !coord_to_key(m.ks, (double)ez, ekz))
    return RAY_FREE;  // out of key range: upstream leaves the ray empty
  if (kx == ekx && ky == eky && kz == ekz) return RAY_FREE;
  if (fast) {
    const int lox = min(kx, ekx) - (ex < ox ? 1 : 0), hix = max(kx, ekx) + (ex > ox ? 1 : 0);
    const int loy = min(ky, eky) - (ey < oy ? 1 : 0), hiy = max(ky, eky) + (ey > oy ? 1 : 0);
    const int loz = min(kz, ekz) - (ez < oz ? 1 : 0), hiz = max(kz, ekz) + (ez > oz ? 1 : 0);
    if (!box_maybe_blocked(m, lox, hix, loy, hiy, loz, hiz)) return RAY_FREE;
  }
  if (map_blocked(m, kx, ky, kz)) return RAY_BLOCKED;  // the start voxel is the first one the walk tests
  return RAY_WALK;
}

constexpr int RAY_TABLE_MAX = 256;
struct EdgeSmemSort {
  int2 buf[EDGE_WARPS][32];
};
struct EdgeSmem {
  RayLattice lat;
  double ray_off[RAY_TABLE_MAX][3];  // offset of ray q, when the lattice has <= RAY_TABLE_MAX rays
  EdgeSmemSort sort;
  double far_end[EDGE_WARPS][32][3];  // far-end coordinates of the batch's pairs
  int slots[EDGE_WARPS][32];          // batch slots that still need rays, compacted
  unsigned queue[EDGE_WARPS][64];     // (slot << 20 | ray) waiting for the exact walk
  int okey[EDGE_WARPS][32][3];        // origin keys of the first 32 rays of the row (they do not depend on the pair)
};

__device__ __forceinline__ void ray_offset(const EdgeSmem& sm, int q, double& ox, double& oy, double& oz) {
  if (sm.lat.rays <= RAY_TABLE_MAX) {
    ox = sm.ray_off[q][0]; oy = sm.ray_off[q][1]; oz = sm.ray_off[q][2];
  } else {
    const int nyz = sm.lat.n[1] * sm.lat.n[2];
    int ix = q / nyz, rem = q - ix * nyz;
    int iy = rem / sm.lat.n[2], iz = rem - iy * sm.lat.n[2];
    ox = sm.lat.off[0][ix]; oy = sm.lat.off[1][iy]; oz = sm.lat.off[2][iz];
  }
}

// Exact walks for the first `count` (<= 32) queue entries of this warp; returns the slots found blocked.
__device__ __forceinline__ unsigned run_walks(const DeviceMap& m, const EdgeSmem& sm, int warp, int lane, int count,
                                              unsigned blockedmask, double px, double py, double pz) {
  unsigned bit = 0;
  if (lane < count) {
    const unsigned e = sm.queue[warp][lane];
    const int k = (int)(e >> 20), q = (int)(e & 0xfffffu);
    if (!((blockedmask >> k) & 1u)) {
      double ox, oy, oz;
      ray_offset(sm, q, ox, oy, oz);
      const double* f = sm.far_end[warp][k];
      int st = line_status_exact(m, d2f(d_add(px, ox)), d2f(d_add(py, oy)), d2f(d_add(pz, oz)), d2f(d_add(f[0], ox)),
                                 d2f(d_add(f[1], oy)), d2f(d_add(f[2], oz)));
      if (st != CELL_FREE) bit = 1u << k;
    }
  }
  return __reduce_or_sync(0xffffffffu, bit);
}

// Edge test, one warp per source row.  Work is shed in stages, none of which can change a result
// (DESIGN.md §5): (1) rows whose reach is free in the coarse pyramid were flagged by k_count;
// (2) per pair, one lane each: the key box all its rays stay in holds no blocked voxel -> accepted;
// (3) per ray: the ray's own key box is free -> ray free; (4) the exact DDA walk.
// The instrumented variant (COUNT) runs every ray through the exact walk.
#ifndef EDGE_MIN_BLOCKS
#define EDGE_MIN_BLOCKS 4
#endif
// order one row's pairs by list index (reference adjacency is in list order): rank sort.
// n <= 32: returns this lane's pair after sorting (in registers); else sorts candA -> candB in global memory.
__device__ __forceinline__ int2 sort_row_pairs(EdgeSmemSort& sb, int warp, int lane, const int2* __restrict__ candA,
                                               int2* candB, long long off, int n) {
  int2 c0 = make_int2(0, 0);
  if (n <= 32) {
    int2 mine = (lane < n) ? candA[off + lane] : make_int2(0, 0x7fffffff);
    int key = mine.y & 0x7fffffff;
    int rank = 0;
    for (int k = 0; k < n; ++k) rank += ((__shfl_sync(0xffffffffu, mine.y, k) & 0x7fffffff) < key);
    if (lane < n) sb.buf[warp][rank] = mine;
    __syncwarp();
    if (lane < n) c0 = sb.buf[warp][lane];
    __syncwarp();
  } else {
    for (int e = lane; e < n; e += 32) {
      int2 mine = candA[off + e];
      int key = mine.y & 0x7fffffff;
      int rank = 0;
      for (int k = 0; k < n; ++k) rank += ((candA[off + k].y & 0x7fffffff) < key);
      candB[off + rank] = mine;
    }
    __syncwarp();
  }
  return c0;
}

// Rows the edge kernel does not have to look at — no map, reach free in the pyramid, or no pair at
// all: every pair is an edge.  One warp per row: sort by list index, write.
__global__ void __launch_bounds__(256) k_rows_accept(ConnectArgs a, const long long* __restrict__ cand_off,
                                                     const uint8_t* __restrict__ row_free, const int2* __restrict__ candA,
                                                     int2* candB, int* __restrict__ acc, int* __restrict__ deg, long long pair_cap) {
  __shared__ EdgeSmemSort sb;
  if (cand_off[a.nrows] > pair_cap) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x * EDGE_WARPS + warp;
  if (t >= a.nrows) return;
  const int s = a.rows ? a.rows[t] : t;
  const int r = (int)(a.sorted_idx[s] - a.lo);
  const long long off = cand_off[r];
  const int n = (int)(cand_off[r + 1] - off);
  if (a.use_map && !row_free[r] && n > 0) return;  // a work row: k_edges writes it
  if (n == 0) {
    if (lane == 0) deg[r] = 0;
    return;
  }
  int2 c0 = sort_row_pairs(sb, warp, lane, candA, candB, off, n);
  if (n <= 32) {
    if (lane < n) acc[off + lane] = c0.y & 0x7fffffff;
  } else {
    for (int e = lane; e < n; e += 32) acc[off + e] = candB[off + e].y & 0x7fffffff;
  }
  if (lane == 0) deg[r] = n;
}

template <bool COUNT>
__device__ __forceinline__ void edge_row(const ConnectArgs& a, const DeviceMap& m, EdgeSmem& sm, const int warp, const int lane,
                                         const int t, const long long* __restrict__ cand_off,
                                         const int2* __restrict__ candA, int2* candB, int* __restrict__ acc,
                                         int* __restrict__ deg, unsigned long long* counters) {
  const RayLattice& lat = sm.lat;
  const int s = a.rows ? a.rows[t] : t;
  const int r = (int)(a.sorted_idx[s] - a.lo);
  const long long off = cand_off[r];
  const int n = (int)(cand_off[r + 1] - off);
  const int2 c0 = sort_row_pairs(sm.sort, warp, lane, candA, candB, off, n);

  const double px = a.sx[s], py = a.sy[s], pz = a.sz[s];
  const int R = lat.rays;
  const bool fast = a.use_pyramid != 0;
  // Row-level verdicts: every ray starts at p + offset whatever the pair, and the start voxel is the
  // first one a walk tests.  A ray whose start voxel is blocked therefore rejects every pair of this
  // row for which that ray is not empty (its end key differs from its start key).
  unsigned startmask = 0;
  if (!COUNT && fast && R <= 32) {
    bool blk = false;
    if (lane < R) {
      double ox, oy, oz;
      ray_offset(sm, lane, ox, oy, oz);
      int kx, ky, kz;
      if (coord_to_key(m.ks, (double)d2f(d_add(px, ox)), kx) && coord_to_key(m.ks, (double)d2f(d_add(py, oy)), ky) &&
          coord_to_key(m.ks, (double)d2f(d_add(pz, oz)), kz)) {
        sm.okey[warp][lane][0] = kx; sm.okey[warp][lane][1] = ky; sm.okey[warp][lane][2] = kz;
        blk = map_blocked(m, kx, ky, kz) != 0;
      }
    }
    startmask = __ballot_sync(0xffffffffu, blk);
    __syncwarp();
  }
  int nacc = 0;
  unsigned n_box = 0, n_lat = 0;
  unsigned long long nsteps = 0;
  for (int base = 0; base < n; base += 32) {
    const int nb = min(32, n - base);
    int2 c = (n <= 32) ? c0 : ((lane < nb) ? candB[off + base + lane] : make_int2(0, 0));
    const bool mine_valid = lane < nb;
    const bool same = mine_valid && (c.y & 0x80000000);
    // this lane's pair: coordinates of the far end
    double qx = 0, qy = 0, qz = 0;
    if (mine_valid) { qx = a.sx[c.x]; qy = a.sy[c.x]; qz = a.sz[c.x]; }
    bool rejected = false;
    bool needs = mine_valid && !same;
    // stage 2a: a ray of this pair that starts in a blocked voxel and is not empty rejects the pair
    if (needs && startmask) {
      for (unsigned mb = startmask; mb; mb &= mb - 1) {
        const int q = __ffs(mb) - 1;
        double ox, oy, oz;
        ray_offset(sm, q, ox, oy, oz);
        int ekx, eky, ekz;
        if (coord_to_key(m.ks, (double)d2f(d_add(qx, ox)), ekx) && coord_to_key(m.ks, (double)d2f(d_add(qy, oy)), eky) &&
            coord_to_key(m.ks, (double)d2f(d_add(qz, oz)), ekz) &&
            (ekx != sm.okey[warp][q][0] || eky != sm.okey[warp][q][1] || ekz != sm.okey[warp][q][2])) {
          needs = false;
          rejected = true;
          break;
        }
      }
    }
    // stage 2b: the pair's key box (one lane per pair)
    if (!COUNT && fast && needs) {
      int lox, hix, loy, hiy, loz, hiz;
      if (pair_key_box(m.ks, px, qx, lat.off[0][0], lat.off[0][lat.n[0] - 1], lox, hix) &&
          pair_key_box(m.ks, py, qy, lat.off[1][0], lat.off[1][lat.n[1] - 1], loy, hiy) &&
          pair_key_box(m.ks, pz, qz, lat.off[2][0], lat.off[2][lat.n[2] - 1], loz, hiz) &&
          !box_maybe_blocked(m, lox, hix, loy, hiy, loz, hiz))
        needs = false;
    }
    unsigned need = __ballot_sync(0xffffffffu, needs);
    unsigned accmask = __ballot_sync(0xffffffffu, mine_valid && !needs && !rejected);  // same-position or box-free pairs
    n_box += __popc(accmask & ~__ballot_sync(0xffffffffu, same));
    n_lat += __popc(need);
    if (COUNT) {
      // instrumented run: every ray through the exact walk, counted with the reference's
      // sequential semantics (rays in lattice order, stop after the first non-free ray)
      BlockedGlobal glob_fn;
      glob_fn.m = &m;
      const int G = R >= 32 ? 32 : R, EPI = 32 / G;
      const int sub = lane / G, lig = lane - sub * G;
      double o0x, o0y, o0z;
      lane_offset(lat, lig, o0x, o0y, o0z);
      while (need) {
        const int kk = __fns(need, 0, sub + 1);
        const bool have = (sub < EPI) && kk >= 0 && kk < 32;
        const int src = have ? kk : 0;
        const double ex = __shfl_sync(0xffffffffu, qx, src), ey = __shfl_sync(0xffffffffu, qy, src),
                     ez = __shfl_sync(0xffffffffu, qz, src);
        int st = swept_status<true, false>(m, lat, glob_fn, have, G, sub, lig, o0x, o0y, o0z, px, py, pz, ex, ey, ez, &nsteps);
        unsigned took = 0, okm = 0;
        for (int g = 0; g < EPI; ++g) {
          int kg = __shfl_sync(0xffffffffu, have ? kk : -1, g * G);
          int sg = __shfl_sync(0xffffffffu, st, g * G);
          if (kg >= 0) {
            took |= 1u << kg;
            if (sg == CELL_FREE) okm |= 1u << kg;
          }
        }
        accmask |= okm;
        need &= ~took;
      }
    } else if (need) {
      // stages 3-4: all (pair, ray) items of the batch flattened over the lanes.  A cheap verdict per
      // ray first; rays that need the exact walk are queued and walked 32 at a time, so the warp
      // stays full whatever the number of rays per pair, and a pair found blocked drops its other rays.
      if (needs) sm.slots[warp][__popc(need & ((1u << lane) - 1u))] = lane;
      if (mine_valid) { sm.far_end[warp][lane][0] = qx; sm.far_end[warp][lane][1] = qy; sm.far_end[warp][lane][2] = qz; }
      __syncwarp();
      const int W = __popc(need) * R;
      unsigned blockedmask = 0;
      int qn = 0;
      for (int w0 = 0; w0 < W; w0 += 32) {
        const int w = w0 + lane;
        int verdict = RAY_FREE, k = 0, q = 0;
        if (w < W) {
          const int p = w / R;
          q = w - p * R;
          k = sm.slots[warp][p];
          if (!((blockedmask >> k) & 1u)) {
            double ox, oy, oz;
            ray_offset(sm, q, ox, oy, oz);
            const double* f = sm.far_end[warp][k];
            verdict = ray_precheck(m, fast, d2f(d_add(px, ox)), d2f(d_add(py, oy)), d2f(d_add(pz, oz)), d2f(d_add(f[0], ox)),
                                   d2f(d_add(f[1], oy)), d2f(d_add(f[2], oz)));
          }
        }
        blockedmask |= __reduce_or_sync(0xffffffffu, verdict == RAY_BLOCKED ? (1u << k) : 0u);
        const bool push = verdict == RAY_WALK && !((blockedmask >> k) & 1u);
        const unsigned pb = __ballot_sync(0xffffffffu, push);
        if (push) sm.queue[warp][qn + __popc(pb & ((1u << lane) - 1u))] = ((unsigned)k << 20) | (unsigned)q;
        qn += __popc(pb);
        __syncwarp();
        if (qn >= 32) {
          blockedmask |= run_walks(m, sm, warp, lane, 32, blockedmask, px, py, pz);
          const unsigned moved = (lane + 32 < qn) ? sm.queue[warp][lane + 32] : 0u;
          __syncwarp();
          sm.queue[warp][lane] = moved;
          qn -= 32;
          __syncwarp();
        }
      }
      if (qn > 0) blockedmask |= run_walks(m, sm, warp, lane, qn, blockedmask, px, py, pz);
      __syncwarp();
      accmask |= need & ~blockedmask;
    }
    if ((accmask >> lane) & 1u) acc[off + nacc + __popc(accmask & ((1u << lane) - 1u))] = c.y & 0x7fffffff;
    nacc += __popc(accmask);
  }
  if (lane == 0) {
    deg[r] = nacc;
    if (n_box) atomicAdd(counters + 5, (unsigned long long)n_box);
    if (n_lat) atomicAdd(counters + 6, (unsigned long long)n_lat);
  }
  if (COUNT) {
    for (int o = 16; o; o >>= 1) nsteps += __shfl_xor_sync(0xffffffffu, nsteps, o);
    if (lane == 0 && nsteps) atomicAdd(counters + 3, nsteps);
  }
  __syncwarp();
}

// Persistent edge kernel: a fixed grid (EDGE_MIN_BLOCKS CTAs per SM); every warp draws the next work
// row with a ticket, so a row that needs many walks does not hold seven finished warps hostage.
template <bool COUNT>
__global__ void __launch_bounds__(EDGE_WARPS * 32, EDGE_MIN_BLOCKS)
k_edges(ConnectArgs a, DeviceMap m, const RayLattice* __restrict__ lat_g, const long long* __restrict__ cand_off,
        const int* __restrict__ work_rows, const int2* __restrict__ candA, int2* candB, int* __restrict__ acc,
        int* __restrict__ deg, unsigned long long* counters, long long pair_cap) {
  __shared__ EdgeSmem sm;
  if (cand_off[a.nrows] > pair_cap) return;
  RayLattice& lat = sm.lat;
  for (int i = threadIdx.x; i < (int)(sizeof(RayLattice) / 4); i += blockDim.x) ((uint32_t*)&lat)[i] = ((const uint32_t*)lat_g)[i];
  __syncthreads();
  if (lat.rays <= RAY_TABLE_MAX) {
    const int nyz = lat.n[1] * lat.n[2];
    for (int q = threadIdx.x; q < lat.rays; q += blockDim.x) {
      int ix = q / nyz, rem = q - ix * nyz;
      int iy = rem / lat.n[2], iz = rem - iy * lat.n[2];
      sm.ray_off[q][0] = lat.off[0][ix]; sm.ray_off[q][1] = lat.off[1][iy]; sm.ray_off[q][2] = lat.off[2][iz];
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned long long n_work = counters[7];
  for (;;) {
    unsigned long long i = 0;
    if (lane == 0) i = atomicAdd(counters + 4, 1ull);
    i = __shfl_sync(0xffffffffu, i, 0);
    if (i >= n_work) break;
    edge_row<COUNT>(a, m, sm, warp, lane, work_rows[i], cand_off, candA, candB, acc, deg, counters);
  }
}

// CSR compaction: one warp per row, coalesced copy of the accepted prefix
__global__ void k_compact(const long long* __restrict__ cand_off, const int* __restrict__ deg,
                          const long long* __restrict__ row_ptr, const int* __restrict__ acc, int* __restrict__ col,
                          int nrows, long long pair_cap) {
  if (cand_off[nrows] > pair_cap) return;
  int r = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  int lane = threadIdx.x & 31;
  if (r >= nrows) return;
  long long src = cand_off[r], dst = row_ptr[r];
  int d = deg[r];
  for (int e = lane; e < d; e += 32) col[dst + e] = acc[src + e];
}

// Multi-GPU CSR assembly: every rank holds the all-gathered, padded blocks of all ranks
// (row_ptr blocks of stride_r entries whose entry [rows_g] is the block's edge count; column blocks
// of stride_e entries).  One CTA per block rebases and copies: the whole-graph CSR without a host
// round trip.  bases[g] = first row of block g; out_row_ptr[total_rows] is written by the last block.
__global__ void k_assemble_blocks(const long long* __restrict__ rp_all, const int* __restrict__ col_all, int world,
                                  long long stride_r, long long stride_e, const long long* __restrict__ row_base,
                                  long long* __restrict__ out_rp, int* __restrict__ out_col, long long* __restrict__ overflow) {
  const int g = blockIdx.y;
  const long long rows = row_base[g + 1] - row_base[g];
  long long ebase = 0;
  for (int k = 0; k < g; ++k) ebase += rp_all[k * stride_r + (row_base[k + 1] - row_base[k])];
  const long long* rp = rp_all + g * stride_r;
  const long long edges = rp[rows];
  if (edges > stride_e) {  // a block did not fit its slot: tell the host, which regathers with larger slots
    if (threadIdx.x == 0 && blockIdx.x == 0) atomicMax((unsigned long long*)overflow, (unsigned long long)edges);
    return;
  }
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  for (long long i = tid; i < rows; i += nth) out_rp[row_base[g] + i] = rp[i] + ebase;
  if (g == world - 1 && tid == 0) out_rp[row_base[g] + rows] = ebase + edges;
  const int* col = col_all + g * stride_e;
  for (long long i = tid; i < edges; i += nth) out_col[ebase + i] = col[i];
}

// ------------------------------------------------------------------------------------------------
// batched queries
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_check_segments(DeviceMap m, const RayLattice* __restrict__ lat_g,
                                                        const double* __restrict__ starts, const double* __restrict__ ends,
                                                        int64_t n, uint8_t* __restrict__ status) {
  __shared__ RayLattice lat;
  for (int i = threadIdx.x; i < (int)(sizeof(RayLattice) / 4); i += blockDim.x) ((uint32_t*)&lat)[i] = ((const uint32_t*)lat_g)[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int R = lat.rays;
  const int G = R >= 32 ? 32 : R;
  const int EPI = 32 / G;
  const int sub = lane / G, lig = lane - sub * G;
  int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int64_t seg = warp * EPI + sub;
  bool active = (sub < EPI) && seg < n;
  double tx = 0, ty = 0, tz = 0, sx = 0, sy = 0, sz = 0;
  if (active) {
    tx = starts[3 * seg]; ty = starts[3 * seg + 1]; tz = starts[3 * seg + 2];
    sx = ends[3 * seg]; sy = ends[3 * seg + 1]; sz = ends[3 * seg + 2];
  }
  CellFromPlanes cell;
  cell.m = &m;
  double o0x, o0y, o0z;
  lane_offset(lat, lig, o0x, o0y, o0z);
  int st = swept_status<false, false>(m, lat, cell, active, G, sub, lig, o0x, o0y, o0z, tx, ty, tz, sx, sy, sz);
  if (active && lig == 0) status[seg] = (uint8_t)st;
}

// validity spec: centre voxel, then the key box of the float-narrowed corners
__global__ void k_validity(DeviceMap m, const double* __restrict__ pts, int64_t n, double bx, double by, double bz,
                           uint8_t* __restrict__ status) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double p[3] = {pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]};
  double box[3] = {bx, by, bz};
  int kc[3];
  if (!coord_to_key(m.ks, p[0], kc[0]) || !coord_to_key(m.ks, p[1], kc[1]) || !coord_to_key(m.ks, p[2], kc[2])) {
    status[i] = CELL_UNKNOWN;
    return;
  }
  int st = map_cell_status(m, kc[0], kc[1], kc[2]);
  if (st != CELL_FREE) { status[i] = (uint8_t)st; return; }
  int lo[3], hi[3];
  for (int a = 0; a < 3; ++a) {
    float fl = d2f(d_sub(p[a], d_div(box[a], 2.0)));
    float fh = d2f(d_add(p[a], d_div(box[a], 2.0)));
    if (!coord_to_key(m.ks, (double)fl, lo[a]) || !coord_to_key(m.ks, (double)fh, hi[a])) { status[i] = CELL_UNKNOWN; return; }
  }
  bool unknown = false;
  for (int z = lo[2]; z <= hi[2]; ++z)
    for (int y = lo[1]; y <= hi[1]; ++y)
      for (int x = lo[0]; x <= hi[0]; ++x) {
        int c = map_cell_status(m, x, y, z);
        if (c == CELL_OCCUPIED) { status[i] = CELL_OCCUPIED; return; }
        if (c == CELL_UNKNOWN) unknown = true;
      }
  status[i] = unknown ? CELL_UNKNOWN : CELL_FREE;
}

__global__ void k_cast_rays(DeviceMap m, const float* __restrict__ org, const float* __restrict__ dir, int64_t n,
                            int ignore_unknown, const double* __restrict__ max_range, uint8_t* __restrict__ hit,
                            int* __restrict__ hit_keys) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  CellStatusFn cell;
  cell.m = &m;
  int hx, hy, hz;
  bool h = cast_ray(m.ks, org[3 * i], org[3 * i + 1], org[3 * i + 2], dir[3 * i], dir[3 * i + 1], dir[3 * i + 2],
                    ignore_unknown != 0, max_range ? max_range[i] : -1.0, cell, hx, hy, hz);
  hit[i] = h ? 1 : 0;
  if (hit_keys) {
    hit_keys[3 * i] = h ? hx : -1;
    hit_keys[3 * i + 1] = h ? hy : -1;
    hit_keys[3 * i + 2] = h ? hz : -1;
  }
}


// ------------------------------------------------------------------------------------------------
// coverage visibility (spec v1): frustum planes -> in-frustum pairs -> one castRay each
// ------------------------------------------------------------------------------------------------
struct PointGrid {
  double min[3];
  double inv_cell;
  int n[3];
};
__device__ __forceinline__ int pgrid_axis(const PointGrid& g, double v, int a) {
  int c = (int)floor((v - g.min[a]) * g.inv_cell);
  return c < 0 ? 0 : (c >= g.n[a] ? g.n[a] - 1 : c);
}

struct ViewRec {       // one per viewpoint, built on the device from the pose
  float pl[6][4];      // left, right, top, bottom, far, near
  float T[3];
  int c0[3], c1[3];    // inclusive cell range of the frustum's bounding box in the point grid
};

__device__ __forceinline__ void cross3f(const float a[3], const float b[3], float o[3]) {
  o[0] = f_sub(f_mul(a[1], b[2]), f_mul(a[2], b[1]));
  o[1] = f_sub(f_mul(a[2], b[0]), f_mul(a[0], b[2]));
  o[2] = f_sub(f_mul(a[0], b[1]), f_mul(a[1], b[0]));
}
__device__ __forceinline__ float dot3f(const float a[3], const float b[3]) {
  return f_add(f_add(f_mul(a[0], b[0]), f_mul(a[1], b[1])), f_mul(a[2], b[2]));
}

// fr = (vh, hw, near, far): 2*tan(vfov/2), 2*tan(hfov/2) evaluated on the host
__global__ void k_view_setup(const double* __restrict__ poses, int64_t nv, float vh, float hw, float np_dist, float fp_dist,
                             PointGrid g, ViewRec* __restrict__ out) {
  int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nv) return;
  const double* P = poses + 7 * v;
  double x = P[3], y = P[4], z = P[5], w = P[6];
  // tf::Matrix3x3::setRotation in double
  double d = d_add(d_add(d_add(d_mul(x, x), d_mul(y, y)), d_mul(z, z)), d_mul(w, w));
  double s = d_div(2.0, d);
  double xs = d_mul(x, s), ys = d_mul(y, s), zs = d_mul(z, s);
  double wx = d_mul(w, xs), wy = d_mul(w, ys), wz = d_mul(w, zs);
  double xx = d_mul(x, xs), xy = d_mul(x, ys), xz = d_mul(x, zs);
  double yy = d_mul(y, ys), yz = d_mul(y, zs), zz = d_mul(z, zs);
  float view[3] = {d2f(d_sub(1.0, d_add(yy, zz))), d2f(d_add(xy, wz)), d2f(d_sub(xz, wy))};
  float up[3] = {d2f(d_sub(xy, wz)), d2f(d_sub(1.0, d_add(xx, zz))), d2f(d_add(yz, wx))};
  float right[3] = {d2f(d_add(xz, wy)), d2f(d_sub(yz, wx)), d2f(d_sub(1.0, d_add(xx, yy)))};
  float T[3] = {d2f(P[0]), d2f(P[1]), d2f(P[2])};
  float np_h = f_mul(vh, np_dist), np_w = f_mul(hw, np_dist), fp_h = f_mul(vh, fp_dist), fp_w = f_mul(hw, fp_dist);
  float fp_c[3], np_c[3], fp_tl[3], fp_tr[3], fp_bl[3], fp_br[3], np_tl[3], np_tr[3], np_bl[3], np_br[3];
  for (int i = 0; i < 3; ++i) {
    fp_c[i] = f_add(T[i], f_mul(view[i], fp_dist));
    np_c[i] = f_add(T[i], f_mul(view[i], np_dist));
    float fu = f_div(f_mul(up[i], fp_h), 2.0f), fr = f_div(f_mul(right[i], fp_w), 2.0f);
    float nu = f_div(f_mul(up[i], np_h), 2.0f), nr = f_div(f_mul(right[i], np_w), 2.0f);
    fp_tl[i] = f_sub(f_add(fp_c[i], fu), fr);
    fp_tr[i] = f_add(f_add(fp_c[i], fu), fr);
    fp_bl[i] = f_sub(f_sub(fp_c[i], fu), fr);
    fp_br[i] = f_add(f_sub(fp_c[i], fu), fr);
    np_tl[i] = f_sub(f_add(np_c[i], nu), nr);
    np_tr[i] = f_add(f_add(np_c[i], nu), nr);
    np_bl[i] = f_sub(f_sub(np_c[i], nu), nr);
    np_br[i] = f_add(f_sub(np_c[i], nu), nr);
  }
  ViewRec R;
  float e1[3], e2[3], a[3], b[3], c[3], dd[3];
  for (int i = 0; i < 3; ++i) { e1[i] = f_sub(fp_bl[i], fp_br[i]); e2[i] = f_sub(fp_tr[i], fp_br[i]); }
  cross3f(e1, e2, R.pl[4]);
  R.pl[4][3] = -dot3f(fp_c, R.pl[4]);
  for (int i = 0; i < 3; ++i) { e1[i] = f_sub(np_tr[i], np_br[i]); e2[i] = f_sub(np_bl[i], np_br[i]); }
  cross3f(e1, e2, R.pl[5]);
  R.pl[5][3] = -dot3f(np_c, R.pl[5]);
  for (int i = 0; i < 3; ++i) {
    a[i] = f_sub(fp_bl[i], T[i]); b[i] = f_sub(fp_br[i], T[i]); c[i] = f_sub(fp_tr[i], T[i]); dd[i] = f_sub(fp_tl[i], T[i]);
  }
  cross3f(dd, a, R.pl[0]);
  cross3f(b, c, R.pl[1]);
  cross3f(c, dd, R.pl[2]);
  cross3f(a, b, R.pl[3]);
  for (int k = 0; k < 4; ++k) R.pl[k][3] = -dot3f(T, R.pl[k]);
  for (int i = 0; i < 3; ++i) R.T[i] = T[i];
  // bounding box of the eight corners, grown by a rounding allowance, in grid cells
  for (int i = 0; i < 3; ++i) {
    float lo = fminf(fminf(fminf(fp_tl[i], fp_tr[i]), fminf(fp_bl[i], fp_br[i])), fminf(fminf(np_tl[i], np_tr[i]), fminf(np_bl[i], np_br[i])));
    float hi = fmaxf(fmaxf(fmaxf(fp_tl[i], fp_tr[i]), fmaxf(fp_bl[i], fp_br[i])), fmaxf(fmaxf(np_tl[i], np_tr[i]), fmaxf(np_bl[i], np_br[i])));
    double eps = 1e-3 + 1e-4 * ((double)fabsf(T[i]) + (double)fp_dist);
    R.c0[i] = pgrid_axis(g, (double)lo - eps, i);
    R.c1[i] = pgrid_axis(g, (double)hi + eps, i);
    if (!(hi >= lo)) { R.c0[i] = 0; R.c1[i] = g.n[i] - 1; }  // NaN pose: scan everything, planes reject all
  }
  out[v] = R;
}

__device__ __forceinline__ bool in_frustum_dev(const ViewRec& R, float px, float py, float pz) {
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float val = f_add(f_add(f_add(f_mul(px, R.pl[k][0]), f_mul(py, R.pl[k][1])), f_mul(pz, R.pl[k][2])), R.pl[k][3]);
    if (!(val <= 0.0f)) return false;
  }
  return true;
}

__global__ void k_pts_bbox(const float* __restrict__ xyz, int64_t n, unsigned long long* bbox) {
  unsigned long long mn[3] = {~0ull, ~0ull, ~0ull}, mx[3] = {0, 0, 0};
  unsigned long long bad = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    for (int a = 0; a < 3; ++a) {
      double v = (double)xyz[3 * i + a];
      if (!isfinite(v)) { ++bad; continue; }
      unsigned long long e = enc_f64(v);
      mn[a] = min(mn[a], e);
      mx[a] = max(mx[a], e);
    }
  for (int a = 0; a < 3; ++a)
    for (int o = 16; o; o >>= 1) {
      mn[a] = min(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
      mx[a] = max(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
    }
  for (int o = 16; o; o >>= 1) bad += __shfl_xor_sync(0xffffffffu, bad, o);
  if ((threadIdx.x & 31) == 0) {
    for (int a = 0; a < 3; ++a) { atomicMin(bbox + a, mn[a]); atomicMax(bbox + 3 + a, mx[a]); }
    if (bad) atomicAdd(bbox + 6, bad);
  }
}

__global__ void k_pts_hist(const float* __restrict__ xyz, int64_t n, PointGrid g, int* __restrict__ cell_of, int* cell_count) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int cx = pgrid_axis(g, (double)xyz[3 * i], 0), cy = pgrid_axis(g, (double)xyz[3 * i + 1], 1), cz = pgrid_axis(g, (double)xyz[3 * i + 2], 2);
  int c = (cz * g.n[1] + cy) * g.n[0] + cx;
  cell_of[i] = c;
  atomicAdd(cell_count + c, 1);
}

__global__ void k_pts_scatter(const float* __restrict__ xyz, int64_t n, const int* __restrict__ cell_of, int* cursor,
                              int* __restrict__ sorted_idx, float4* __restrict__ sp) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || cell_of[i] < 0) return;
  int pos = atomicAdd(cursor + cell_of[i], 1);
  sorted_idx[pos] = (int)i;
  sp[pos] = make_float4(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], __int_as_float((int)i));
}

// one warp per viewpoint: in-frustum points of the cells its bounding box touches.
// FILL == false: count only.  FILL == true: write (viewpoint, binned position) per ray.
template <bool FILL>
__global__ void __launch_bounds__(256) k_view_gather(const ViewRec* __restrict__ views, int64_t v0, int nv, PointGrid g,
                                                     const int* __restrict__ cell_start, const float4* __restrict__ sp,
                                                     int* __restrict__ ray_count, const long long* __restrict__ ray_off,
                                                     int* __restrict__ ray_view, int* __restrict__ ray_pt) {
  int w = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  int lane = threadIdx.x & 31;
  if (w >= nv) return;
  const ViewRec R = views[v0 + w];
  long long base = FILL ? ray_off[w] : 0;
  int filled = 0;
  for (int cz = R.c0[2]; cz <= R.c1[2]; ++cz)
    for (int cy = R.c0[1]; cy <= R.c1[1]; ++cy) {
      int cb = (cz * g.n[1] + cy) * g.n[0];
      int b = cell_start[cb + R.c0[0]], e = cell_start[cb + R.c1[0] + 1];
      for (int j0 = b; j0 < e; j0 += 32) {
        int j = j0 + lane;
        bool in = false;
        if (j < e) {
          float4 p = sp[j];
          in = in_frustum_dev(R, p.x, p.y, p.z);
        }
        unsigned bal = __ballot_sync(0xffffffffu, in);
        if (FILL && in) {
          long long o = base + filled + __popc(bal & ((1u << lane) - 1u));
          ray_view[o] = w;
          ray_pt[o] = j;
        }
        filled += __popc(bal);
      }
    }
  if (!FILL && lane == 0) ray_count[w] = filled;
}

// one thread per occlusion ray
__global__ void __launch_bounds__(256) k_view_rays(DeviceMap m, const ViewRec* __restrict__ views, int64_t v0,
                                                   const float4* __restrict__ sp, const int* __restrict__ ray_view,
                                                   const int* __restrict__ ray_pt, long long nrays, uint8_t* __restrict__ vis,
                                                   int* vis_count) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  bool visible = false;
  int w = -1;
  if (i < nrays) {
    w = ray_view[i];
    const ViewRec& R = views[v0 + w];
    float4 p = sp[ray_pt[i]];
    float ox = R.T[0], oy = R.T[1], oz = R.T[2];
    float dx = f_sub(p.x, ox), dy = f_sub(p.y, oy), dz = f_sub(p.z, oz);
    CellStatusFn cell;
    cell.m = &m;
    int hx, hy, hz;
    bool hit = cast_ray(m.ks, ox, oy, oz, dx, dy, dz, true, norm3f(dx, dy, dz), cell, hx, hy, hz);
    if (!hit) visible = true;
    else {
      int kx, ky, kz;
      visible = coord_to_key(m.ks, (double)p.x, kx) && coord_to_key(m.ks, (double)p.y, ky) &&
                coord_to_key(m.ks, (double)p.z, kz) && kx == hx && ky == hy && kz == hz;
    }
    vis[i] = visible ? 1 : 0;
  }
  // rays of one viewpoint are contiguous: aggregate the count per warp where the warp is uniform
  unsigned bal = __ballot_sync(0xffffffffu, visible);
  int w0 = __shfl_sync(0xffffffffu, w, 0);
  bool uniform = __all_sync(0xffffffffu, w == w0 || w < 0);
  if (uniform) {
    if ((threadIdx.x & 31) == 0 && bal && w0 >= 0) atomicAdd(vis_count + w0, __popc(bal));
  } else if (visible) {
    atomicAdd(vis_count + w, 1);
  }
}

// one warp per viewpoint: compact the visible rays' ORIGINAL point indices
__global__ void k_view_compact(int nv, const long long* __restrict__ ray_off, const uint8_t* __restrict__ vis,
                               const int* __restrict__ ray_pt, const float4* __restrict__ sp,
                               const long long* __restrict__ vis_ptr, int* __restrict__ out) {
  int w = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  int lane = threadIdx.x & 31;
  if (w >= nv) return;
  long long b = ray_off[w], e = ray_off[w + 1], o = vis_ptr[w];
  for (long long j0 = b; j0 < e; j0 += 32) {
    long long j = j0 + lane;
    bool v = (j < e) && vis[j];
    unsigned bal = __ballot_sync(0xffffffffu, v);
    if (v) out[o + __popc(bal & ((1u << lane) - 1u))] = __float_as_int(sp[ray_pt[j]].w);
    o += __popc(bal);
  }
}

// ascending sort of every viewpoint's index list: one CTA per viewpoint, all-ascending bitonic
// network (flip + half-cleaners), virtual +inf padding beyond the segment
__device__ __forceinline__ void cmpswap(int* a, long long i, long long j, long long len) {
  if (j < len) {
    int x = a[i], y = a[j];
    if (x > y) { a[i] = y; a[j] = x; }
  }
}
__global__ void __launch_bounds__(256) k_segment_sort(int nv, const long long* __restrict__ ptr, int* data) {
  extern __shared__ int sh[];
  const int SM_CAP = 8192;
  for (int w = blockIdx.x; w < nv; w += gridDim.x) {
    long long b = ptr[w], len = ptr[w + 1] - b;
    if (len < 2) continue;
    int* a = data + b;
    bool in_smem = len <= SM_CAP;
    if (in_smem) {
      for (int i = threadIdx.x; i < len; i += blockDim.x) sh[i] = a[i];
      a = sh;
    }
    __syncthreads();
    long long P = 1;
    while (P < len) P <<= 1;
    for (long long k = 2; k <= P; k <<= 1) {
      // flip: i in the lower half of its k-block is compared with its mirror
      for (long long t = threadIdx.x; t < P / 2; t += blockDim.x) {
        long long blk = t / (k / 2), o = t % (k / 2);
        long long i = blk * k + o, j = blk * k + (k - 1 - o);
        if (i < len) cmpswap(a, i, j, len);
      }
      __syncthreads();
      for (long long h = k / 4; h >= 1; h >>= 1) {
        for (long long t = threadIdx.x; t < P / 2; t += blockDim.x) {
          long long blk = t / h, o = t % h;
          long long i = blk * 2 * h + o, j = i + h;
          if (i < len) cmpswap(a, i, j, len);
        }
        __syncthreads();
      }
    }
    if (in_smem) {
      int* g = data + b;
      for (int i = threadIdx.x; i < len; i += blockDim.x) g[i] = sh[i];
    }
    __syncthreads();
  }
}

// L2-resident read bandwidth probe
__global__ void k_l2_read(const uint4* __restrict__ buf, size_t n16, int iters, unsigned* sink) {
  unsigned acc = 0;
  for (int it = 0; it < iters; ++it)
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
      uint4 v;
      asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));
      acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
  if (acc == 0x12345678u) *sink = acc;
}

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
  CK(ctx->lattice.reserve(sizeof(RayLattice)));
  // `lat` is pageable host memory: the call returns once it has been copied to the driver's staging
  // buffer, so the caller's stack object may go away afterwards
  CK(cudaMemcpyAsync(ctx->lattice.p, &lat, sizeof(RayLattice), cudaMemcpyHostToDevice, ctx->stream));
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
                    &ctx->scan_occ, &ctx->pc_pts, &ctx->pc_keys, &ctx->pc_vals, &ctx->xyz, &ctx->bbox,
                    &ctx->cell_of, &ctx->cell_start, &ctx->cell_cursor, &ctx->sorted_idx, &ctx->sx, &ctx->sy, &ctx->sz,
                    &ctx->rows, &ctx->row_flag, &ctx->row_free, &ctx->cand_count, &ctx->cand_off, &ctx->candA, &ctx->candB,
                    &ctx->acc, &ctx->deg, &ctx->row_ptr, &ctx->col_idx, &ctx->counters, &ctx->scan_tmp, &ctx->lattice, &ctx->work_rows,
                    &ctx->q0, &ctx->q1, &ctx->q2, &ctx->q3, &ctx->q4, &ctx->pts, &ctx->vis_cnt, &ctx->vis_ptr,
                    &ctx->vis_idx, &ctx->vis_poses, &ctx->vis_counters};
  for (DevBuf* b : bufs) b->release();
  for (int i = 0; i < 8; ++i) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
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
int sspp_gpu_connect(sspp_gpu_ctx* ctx, const sspp_connect_params* prm, sspp_connect_stats* stats) {
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
  TileShape ts;
  ts.XW = ts.SY = ts.SZ = 0;
  ts.reach[0] = ts.reach[1] = ts.reach[2] = 0;
  ts.half[0] = ts.half[1] = ts.half[2] = 0;
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
      ts.half[a] = prm->box[a] * 0.5;
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
  CK(ctx->sorted_idx.reserve((size_t)n * 4));
  CK(ctx->sx.reserve((size_t)n * 8));
  CK(ctx->sy.reserve((size_t)n * 8));
  CK(ctx->sz.reserve((size_t)n * 8));
  CK(cudaMemsetAsync(ctx->cell_cursor.p, 0, (size_t)(ncells + 1) * 4, st));
  k_cell_hist<<<ceil_div(n, 256), 256, 0, st>>>(ctx->xyz.as<double>(), n, g, ctx->cell_of.as<int>(), ctx->cell_cursor.as<int>());
  LAUNCHED(ctx);
  {
    int rc = scan_exclusive<int>(ctx, ctx->cell_cursor.as<int>(), ncells, ctx->cell_start.as<int>());
    if (rc) return rc;
  }
  CK(cudaMemcpyAsync(ctx->cell_cursor.p, ctx->cell_start.p, (size_t)ncells * 4, cudaMemcpyDeviceToDevice, st));
  k_cell_scatter<<<ceil_div(n, 256), 256, 0, st>>>(ctx->xyz.as<double>(), n, ctx->cell_of.as<int>(), ctx->cell_cursor.as<int>(),
                                                   ctx->sorted_idx.as<int>(), ctx->sx.as<double>(), ctx->sy.as<double>(),
                                                   ctx->sz.as<double>());
  LAUNCHED(ctx);
  k_cell_sort<<<ceil_div(ncells, 256), 256, 0, st>>>(ctx->cell_start.as<int>(), (int)ncells, ctx->sorted_idx.as<int>(),
                                                     ctx->sx.as<double>(), ctx->sy.as<double>(), ctx->sz.as<double>());
  LAUNCHED(ctx);

  // rows of this shard, in cell order
  const int* rows = nullptr;
  if (sharded) {
    CK(ctx->row_flag.reserve((size_t)n * 4));
    CK(ctx->rows.reserve((size_t)(n + 1) * 4 + (size_t)nrows * 4));
    int* pos = ctx->rows.as<int>();
    int* rows_w = pos + (n + 1);
    k_row_flags<<<ceil_div(n, 256), 256, 0, st>>>(ctx->sorted_idx.as<int>(), n, ctx->cell_start.as<int>() + ncells, lo, hi,
                                                 ctx->row_flag.as<int>());
    LAUNCHED(ctx);
    int rc = scan_exclusive<int>(ctx, ctx->row_flag.as<int>(), n, pos);
    if (rc) return rc;
    k_row_scatter<<<ceil_div(n, 256), 256, 0, st>>>(ctx->row_flag.as<int>(), pos, n, rows_w);
    LAUNCHED(ctx);
    rows = rows_w;
  }
  CK(cudaEventRecord(ctx->ev[1], st));

  // --- 3. count in-radius pairs per row
  ConnectArgs a;
  a.g = g;
  a.cell_start = ctx->cell_start.as<int>();
  a.sorted_idx = ctx->sorted_idx.as<int>();
  a.sx = ctx->sx.as<double>();
  a.sy = ctx->sy.as<double>();
  a.sz = ctx->sz.as<double>();
  a.rows = rows;
  a.nrows = (int)nrows;
  a.lo = lo;
  a.r2max = radius_threshold(prm->radius);
  a.mode = prm->position_mode;
  a.use_map = use_map;
  const bool count_steps = use_map && prm->count_steps;
  a.use_pyramid = (prm->use_pyramid && !count_steps) ? 1 : 0;
  a.ts = ts;
  CK(ctx->cand_count.reserve((size_t)nrows * 4));
  CK(ctx->row_free.reserve((size_t)nrows));
  CK(ctx->cand_off.reserve((size_t)(nrows + 1) * 8));
  CK(ctx->counters.reserve(8 * 8));
  CK(cudaMemsetAsync(ctx->counters.p, 0, 8 * 8, st));
  CK(ctx->work_rows.reserve((size_t)nrows * 4));
  CK(ctx->deg.reserve((size_t)nrows * 4));
  k_count<false><<<ceil_div(nrows, 256), 256, 0, st>>>(a, ctx->map, ctx->cand_count.as<int>(), ctx->row_free.as<uint8_t>(),
                                                      ctx->counters.as<unsigned long long>(), nullptr, nullptr,
                                                      ctx->work_rows.as<int>(), 0);
  LAUNCHED(ctx);
  {
    int rc = scan_exclusive<long long>(ctx, ctx->cand_count.as<int>(), nrows, ctx->cand_off.as<long long>());
    if (rc) return rc;
  }
  CK(cudaEventRecord(ctx->ev[2], st));

  // --- 4/5. edge test and CSR.  The pair buffers keep their size from call to call; the kernels
  // refuse to run when the pair count exceeds it, which the host sees at the single synchronisation
  // at the end (it then grows the buffers and repeats this phase).  Only the first call, with no
  // buffers yet, has to wait for the count here.
  auto pair_capacity = [&]() -> long long {
    return (long long)std::min(std::min(ctx->candA.cap / 8, ctx->candB.cap / 8), std::min(ctx->acc.cap / 4, ctx->col_idx.cap / 4));
  };
  auto reserve_pairs = [&](long long c) -> cudaError_t {
    const size_t want = (size_t)std::max<long long>(c + c / 16 + 1024, 1);
    cudaError_t e;
    if ((e = ctx->candA.reserve(want * 8)) != cudaSuccess) return e;
    if ((e = ctx->candB.reserve(want * 8)) != cudaSuccess) return e;
    if ((e = ctx->acc.reserve(want * 4)) != cudaSuccess) return e;
    return ctx->col_idx.reserve(want * 4);
  };
  long long C = -1;
  if (pair_capacity() == 0) {
    CK(cudaMemcpyAsync(&C, ctx->cand_off.as<long long>() + nrows, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(reserve_pairs(C));
  }
  unsigned long long hc[8];
  long long E = 0;
  for (int attempt = 0;; ++attempt) {
    const long long cap = pair_capacity();
    k_count<true><<<ceil_div(nrows, 256), 256, 0, st>>>(a, ctx->map, nullptr, nullptr, nullptr, ctx->cand_off.as<long long>(),
                                                       ctx->candA.as<int2>(), nullptr, cap);
    LAUNCHED(ctx);
    k_rows_accept<<<ceil_div(nrows, EDGE_WARPS), EDGE_WARPS * 32, 0, st>>>(a, ctx->cand_off.as<long long>(),
                                                                          ctx->row_free.as<uint8_t>(), ctx->candA.as<int2>(),
                                                                          ctx->candB.as<int2>(), ctx->acc.as<int>(),
                                                                          ctx->deg.as<int>(), cap);
    LAUNCHED(ctx);
    CK(cudaEventRecord(ctx->ev[5], st));
    if (use_map) {
      const int blocks = ctx->sm_count * EDGE_MIN_BLOCKS;
      if (count_steps)
        k_edges<true><<<blocks, EDGE_WARPS * 32, 0, st>>>(a, ctx->map, ctx->lattice.as<RayLattice>(),
                                                          ctx->cand_off.as<long long>(), ctx->work_rows.as<int>(),
                                                          ctx->candA.as<int2>(), ctx->candB.as<int2>(), ctx->acc.as<int>(),
                                                          ctx->deg.as<int>(), ctx->counters.as<unsigned long long>(), cap);
      else
        k_edges<false><<<blocks, EDGE_WARPS * 32, 0, st>>>(a, ctx->map, ctx->lattice.as<RayLattice>(),
                                                           ctx->cand_off.as<long long>(), ctx->work_rows.as<int>(),
                                                           ctx->candA.as<int2>(), ctx->candB.as<int2>(), ctx->acc.as<int>(),
                                                           ctx->deg.as<int>(), ctx->counters.as<unsigned long long>(), cap);
      LAUNCHED(ctx);
    }
    CK(cudaEventRecord(ctx->ev[3], st));
    {
      int rc = scan_exclusive<long long>(ctx, ctx->deg.as<int>(), nrows, ctx->row_ptr.as<long long>());
      if (rc) return rc;
    }
    k_compact<<<ceil_div(nrows * 32, 256), 256, 0, st>>>(ctx->cand_off.as<long long>(), ctx->deg.as<int>(),
                                                        ctx->row_ptr.as<long long>(), ctx->acc.as<int>(), ctx->col_idx.as<int>(),
                                                        (int)nrows, cap);
    LAUNCHED(ctx);
    CK(cudaEventRecord(ctx->ev[4], st));
    CK(cudaMemcpyAsync(&E, ctx->row_ptr.as<long long>() + nrows, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&C, ctx->cand_off.as<long long>() + nrows, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(hc, ctx->counters.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (C <= cap) break;
    if (attempt > 0) return fail(ctx, SSPP_ERR_CUDA, "connect: pair buffers still too small after growing them");
    CK(reserve_pairs(C));
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
  ctx->has_result = true;
  ctx->res_rows = nrows;
  ctx->res_edges = E;
  return SSPP_OK;
}

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

int sspp_gpu_connect_result_device(sspp_gpu_ctx* ctx, const int64_t** row_ptr_device, const int32_t** col_idx_device) {
  if (!ctx) return SSPP_ERR_INVALID;
  if (!ctx->has_result) return fail(ctx, SSPP_ERR_STATE, "no connect result");
  if (row_ptr_device) *row_ptr_device = ctx->row_ptr.as<int64_t>();
  if (col_idx_device) *col_idx_device = ctx->col_idx.as<int32_t>();
  return SSPP_OK;
}

int sspp_gpu_connect_nodes_host(sspp_gpu_ctx* ctx, const double* xyz, int64_t n, const sspp_connect_params* params,
                                int64_t* row_ptr, int32_t* col_idx, int64_t col_capacity, sspp_connect_stats* stats) {
  sspp_connect_stats local;
  if (!stats) stats = &local;
  int rc = set_nodes_impl(ctx, xyz, n, cudaMemcpyHostToDevice);
  if (rc) return rc;
  rc = sspp_gpu_connect(ctx, params, stats);
  if (rc) return rc;
  if (stats->edges > col_capacity) return fail(ctx, SSPP_ERR_RANGE, "connect_nodes_host: col_idx capacity too small");
  return sspp_gpu_connect_fetch(ctx, row_ptr, col_idx);
}

int sspp_gpu_connect_timings(sspp_gpu_ctx* ctx, float ms[8]) {
  if (!ctx || !ms) return SSPP_ERR_INVALID;
  for (int i = 0; i < 8; ++i) ms[i] = ctx->timings[i];
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
  if (!g_noalloc_addr[ctx->device]) {
    unsigned char* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) return cuda_fail(ctx, e, "cudaGetSymbolAddress(g_noalloc_scratch)");
    g_noalloc_addr[ctx->device] = ptr;
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
  CK(ctx->pts.reserve((size_t)std::max<int64_t>(n, 1) * 12));
  if (n) CK(cudaMemcpyAsync(ctx->pts.p, xyz, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  ctx->n_points = n;
  ctx->has_vis = false;
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
  const int64_t np = ctx->n_points;
  CK(ctx->vis_ptr.reserve((size_t)(nv + 1) * 8));
  if (nv == 0 || np == 0) {
    CK(cudaMemsetAsync(ctx->vis_ptr.p, 0, (size_t)(nv + 1) * 8, st));
    CK(ctx->vis_idx.reserve(4));
    CK(cudaStreamSynchronize(st));
    ctx->has_vis = true; ctx->vis_nv = nv; ctx->vis_total = 0;
    return SSPP_OK;
  }
  // host half of the spec: tan() of the half angles, narrowed exactly as the oracle states
  float vfov_rad = float(fr->vfov_deg * M_PI / 180);
  float hfov_rad = float(fr->hfov_deg * M_PI / 180);
  float vh = float(2 * tan(vfov_rad / 2));
  float hw = float(2 * tan(hfov_rad / 2));
  float np_dist = (float)fr->near_dist, fp_dist = (float)fr->far_dist;

  // --- bin the model points (cell edge = far/4)
  CK(ctx->bbox.reserve(7 * 8));
  unsigned long long init[7] = {~0ull, ~0ull, ~0ull, 0, 0, 0, 0};
  CK(cudaMemcpyAsync(ctx->bbox.p, init, sizeof(init), cudaMemcpyHostToDevice, st));
  k_pts_bbox<<<std::min(ceil_div(np, 256), 148 * 8), 256, 0, st>>>(ctx->pts.as<float>(), np, ctx->bbox.as<unsigned long long>());
  LAUNCHED(ctx);
  unsigned long long hb[7];
  CK(cudaMemcpyAsync(hb, ctx->bbox.p, sizeof(hb), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
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
  CK(ctx->cell_of.reserve((size_t)np * 4));
  CK(ctx->cell_start.reserve((size_t)(ncells + 1) * 4));
  CK(ctx->cell_cursor.reserve((size_t)(ncells + 1) * 4));
  CK(ctx->sorted_idx.reserve((size_t)np * 4));
  CK(ctx->q0.reserve((size_t)np * 16));
  float4* sp = ctx->q0.as<float4>();
  CK(cudaMemsetAsync(ctx->cell_cursor.p, 0, (size_t)(ncells + 1) * 4, st));
  k_pts_hist<<<ceil_div(np, 256), 256, 0, st>>>(ctx->pts.as<float>(), np, g, ctx->cell_of.as<int>(), ctx->cell_cursor.as<int>());
  LAUNCHED(ctx);
  { int rc = scan_exclusive<int>(ctx, ctx->cell_cursor.as<int>(), ncells, ctx->cell_start.as<int>()); if (rc) return rc; }
  CK(cudaMemcpyAsync(ctx->cell_cursor.p, ctx->cell_start.p, (size_t)ncells * 4, cudaMemcpyDeviceToDevice, st));
  k_pts_scatter<<<ceil_div(np, 256), 256, 0, st>>>(ctx->pts.as<float>(), np, ctx->cell_of.as<int>(), ctx->cell_cursor.as<int>(),
                                                  ctx->sorted_idx.as<int>(), sp);
  LAUNCHED(ctx);

  // --- viewpoint records
  CK(ctx->vis_poses.reserve((size_t)nv * 56));
  CK(ctx->q1.reserve((size_t)nv * sizeof(ViewRec)));
  CK(cudaMemcpyAsync(ctx->vis_poses.p, poses, (size_t)nv * 56, cudaMemcpyHostToDevice, st));
  ViewRec* views = ctx->q1.as<ViewRec>();
  k_view_setup<<<ceil_div(nv, 128), 128, 0, st>>>(ctx->vis_poses.as<double>(), nv, vh, hw, np_dist, fp_dist, g, views);
  LAUNCHED(ctx);

  // --- in-frustum pairs per viewpoint
  CK(ctx->vis_cnt.reserve((size_t)nv * 4 * 2));
  int* ray_count = ctx->vis_cnt.as<int>();
  int* vis_count = ray_count + nv;
  CK(ctx->q2.reserve((size_t)(nv + 1) * 8));
  long long* ray_off = ctx->q2.as<long long>();
  k_view_gather<false><<<ceil_div(nv * 32, 256), 256, 0, st>>>(views, 0, (int)nv, g, ctx->cell_start.as<int>(), sp, ray_count,
                                                              nullptr, nullptr, nullptr);
  LAUNCHED(ctx);
  { int rc = scan_exclusive<long long>(ctx, ray_count, nv, ray_off); if (rc) return rc; }
  long long nrays = 0;
  CK(cudaMemcpyAsync(&nrays, ray_off + nv, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  CK(ctx->q3.reserve((size_t)std::max<long long>(nrays, 1) * 4));
  CK(ctx->q4.reserve((size_t)std::max<long long>(nrays, 1) * 4));
  CK(ctx->row_free.reserve((size_t)std::max<long long>(nrays, 1)));
  int* ray_view = ctx->q3.as<int>();
  int* ray_pt = ctx->q4.as<int>();
  uint8_t* vis = ctx->row_free.as<uint8_t>();
  k_view_gather<true><<<ceil_div(nv * 32, 256), 256, 0, st>>>(views, 0, (int)nv, g, ctx->cell_start.as<int>(), sp, nullptr,
                                                             ray_off, ray_view, ray_pt);
  LAUNCHED(ctx);

  // --- occlusion rays
  CK(cudaMemsetAsync(vis_count, 0, (size_t)nv * 4, st));
  if (nrays) {
    k_view_rays<<<(unsigned)((nrays + 255) / 256), 256, 0, st>>>(ctx->map, views, 0, sp, ray_view, ray_pt, nrays, vis, vis_count);
    LAUNCHED(ctx);
  }
  long long* vis_ptr = ctx->vis_ptr.as<long long>();
  { int rc = scan_exclusive<long long>(ctx, vis_count, nv, vis_ptr); if (rc) return rc; }
  long long total = 0;
  CK(cudaMemcpyAsync(&total, vis_ptr + nv, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  CK(ctx->vis_idx.reserve((size_t)std::max<long long>(total, 1) * 4));
  k_view_compact<<<ceil_div(nv * 32, 256), 256, 0, st>>>((int)nv, ray_off, vis, ray_pt, sp, vis_ptr, ctx->vis_idx.as<int>());
  LAUNCHED(ctx);
  k_segment_sort<<<(unsigned)std::min<int64_t>(nv, 148 * 8), 256, 8192 * 4, st>>>((int)nv, vis_ptr, ctx->vis_idx.as<int>());
  LAUNCHED(ctx);
  CK(cudaStreamSynchronize(st));
  ctx->has_vis = true;
  ctx->vis_nv = nv;
  ctx->vis_total = total;
  if (stats) { stats->visible_pairs = total; stats->rays = nrays; }
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
