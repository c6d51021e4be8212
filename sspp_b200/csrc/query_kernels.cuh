// query_kernels.cuh — batched segment / validity / castRay queries and the L2 bandwidth probe.
// Part of the single translation unit sspp_gpu.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

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


}  // namespace
