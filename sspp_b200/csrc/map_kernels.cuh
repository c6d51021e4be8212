// map_kernels.cuh — map ingest: layout conversion, leaf boxes, setFree/setOccupied, blocked plane + pyramid, point-cloud insertion.
// Part of the single translation unit sspp_gpu.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

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


}  // namespace
