// engine.cuh — device-side data layout shared by the kernels of sspp_gpu.cu.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "dda_core.cuh"

namespace sspp {

// Dense occupancy map in HBM.
//
// Voxel (x,y,z) of the window (x = key_x - kmin[0] ...) is bit (x & 31) of 32-bit word
//     word(xw = x >> 5, y, z) = ((xw * NZ) + z) * NY + y                       ["x-packed, y-fastest"]
// so that the key boxes the edge kernel scans — one or two words in x, a few rows in y and z — are
// short runs of consecutive words along y, and a DDA step along y or z stays in the same or the
// next 32-byte sector.  Three planes of that shape:
//   occ      1 = occupied leaf            (octomap: isNodeOccupied)
//   known    1 = a leaf exists here       (octomap: search() != NULL)
//   blocked  occ | (treat_unknown ? ~known : 0)   — what a swept-robot ray stops at
// and one coarse level over 4x4x4 voxel blocks with the same word layout:
//   cblocked 1 = some voxel of the block is blocked.
struct DeviceMap {
  KeySpace ks;
  int kmin[3];   // key of window voxel (0,0,0)
  int dims[3];   // padded window size in voxels: x multiple of 128, y and z multiples of 4
  int nxw;       // dims[0] / 32
  int cdims[3];  // dims / 4
  int cnxw;      // cdims[0] / 32
  int treat_unknown;
  const uint32_t* occ;
  const uint32_t* known;
  const uint32_t* blocked;
  const uint32_t* cblocked;
};

__device__ __forceinline__ size_t map_word(const DeviceMap& m, int xw, int y, int z) {
  return ((size_t)xw * m.dims[2] + z) * m.dims[1] + y;
}

// CellStatus of one key straight from HBM/L2 (no staging).  Keys outside the window were never observed.
__device__ __forceinline__ int map_cell_status(const DeviceMap& m, int kx, int ky, int kz) {
  int x = kx - m.kmin[0], y = ky - m.kmin[1], z = kz - m.kmin[2];
  if ((unsigned)x >= (unsigned)m.dims[0] || (unsigned)y >= (unsigned)m.dims[1] || (unsigned)z >= (unsigned)m.dims[2])
    return CELL_UNKNOWN;
  size_t w = map_word(m, x >> 5, y, z);
  uint32_t bit = 1u << (x & 31);
  if (!(__ldg(m.known + w) & bit)) return CELL_UNKNOWN;
  return (__ldg(m.occ + w) & bit) ? CELL_OCCUPIED : CELL_FREE;
}

// 1 when a swept-robot ray must stop at this key.
__device__ __forceinline__ int map_blocked(const DeviceMap& m, int kx, int ky, int kz) {
  int x = kx - m.kmin[0], y = ky - m.kmin[1], z = kz - m.kmin[2];
  if ((unsigned)x >= (unsigned)m.dims[0] || (unsigned)y >= (unsigned)m.dims[1] || (unsigned)z >= (unsigned)m.dims[2])
    return m.treat_unknown;
  return (__ldg(m.blocked + map_word(m, x >> 5, y, z)) >> (x & 31)) & 1u;
}

// getLineStatus cell functor on the raw planes: unknown only stops the walk when it counts as occupied.
struct CellFromPlanes {
  const DeviceMap* m;
  __device__ __forceinline__ int operator()(int kx, int ky, int kz) const {
    int st = map_cell_status(*m, kx, ky, kz);
    if (st == CELL_UNKNOWN && !m->treat_unknown) return CELL_FREE;
    return st;
  }
};

// castRay cell functor: the plain status.
struct CellStatusFn {
  const DeviceMap* m;
  __device__ __forceinline__ int operator()(int kx, int ky, int kz) const { return map_cell_status(*m, kx, ky, kz); }
};

// castRay cell functor when unknown cells are ignored: only the occupied plane matters (a key
// outside the window is unknown, i.e. passed through)
struct CellOccupiedOnly {
  const DeviceMap* m;
  __device__ __forceinline__ int operator()(int kx, int ky, int kz) const {
    int x = kx - m->kmin[0], y = ky - m->kmin[1], z = kz - m->kmin[2];
    if ((unsigned)x >= (unsigned)m->dims[0] || (unsigned)y >= (unsigned)m->dims[1] || (unsigned)z >= (unsigned)m->dims[2])
      return CELL_FREE;
    return ((__ldg(m->occ + map_word(*m, x >> 5, y, z)) >> (x & 31)) & 1u) ? CELL_OCCUPIED : CELL_FREE;
  }
};

// The same for a walk that provably stays inside the mapped window: no bounds test, 32-bit word index
// (the planes of any map that fits the 16-bit key space hold fewer than 2^31 words only up to 4096^3 / 32:
// the caller checks map_words).
struct CellOccupiedInside {
  const uint32_t* occ;
  int kx0, ky0, kz0;
  unsigned ny, nz;
  __device__ __forceinline__ int operator()(int kx, int ky, int kz) const {
    const unsigned x = (unsigned)(kx - kx0), y = (unsigned)(ky - ky0), z = (unsigned)(kz - kz0);
    const unsigned w = ((x >> 5) * nz + z) * ny + y;
    return ((__ldg(occ + w) >> (x & 31)) & 1u) ? CELL_OCCUPIED : CELL_FREE;
  }
};

struct BlockedGlobal {
  const DeviceMap* m;
  __device__ __forceinline__ int operator()(int kx, int ky, int kz) const { return map_blocked(*m, kx, ky, kz); }
};

// Uniform grid over the node set for the radius search (cell edge >= radius, so 27 cells suffice).
struct NodeGrid {
  double min[3];
  double max[3];  // nodes outside [min, max] are not binned (a shard's window; the whole set otherwise)
  double inv_cell;
  int n[3];  // cells per axis; cell id = (cz * n[1] + cy) * n[0] + cx
};

__device__ __forceinline__ int grid_cell_axis(const NodeGrid& g, double v, int a) {
  int c = (int)floor((v - g.min[a]) * g.inv_cell);
  return c < 0 ? 0 : (c >= g.n[a] ? g.n[a] - 1 : c);
}

#define SSPP_MAX_LATTICE 64  // ray offsets per axis the kernels keep in constant memory

struct RayLattice {
  double off[3][SSPP_MAX_LATTICE];
  int n[3];
  int rays;  // n[0]*n[1]*n[2]
};

}  // namespace sspp
