// dda_core.cuh — the voxel-walk arithmetic of the engine, written once for device and host.
//
// These routines reproduce, operation for operation, the float/double operand mix of the
// un-vendored libraries the reference calls on its hot path (upstream = OctoMap,
// volumetric_mapping octomap_world; reference call site src/distance_heuristic.cpp:109):
//   line_status()  = OctomapWorld::getLineStatus  over  OcTreeBaseImpl::computeRayKeys
//   cast_ray()     = OccupancyOcTreeBase::castRay
// Every rounding step is spelled with a round-to-nearest intrinsic on the device so that no
// multiply-add is ever fused (the reference builds for baseline x86-64: no FMA, SSE2 doubles;
// CMakeLists.txt:6-8).  On the host the same functions compile as plain C++ (build with
// -ffp-contract=off); tests/ use that build to check this file against the oracle without a GPU.
#pragma once

#include <float.h>
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define SSPP_HD __host__ __device__ __forceinline__
#else
#define SSPP_HD inline
#endif

namespace sspp {

enum { CELL_FREE = 0, CELL_OCCUPIED = 1, CELL_UNKNOWN = 2 };

#if defined(__CUDA_ARCH__)
SSPP_HD float f_add(float a, float b) { return __fadd_rn(a, b); }
SSPP_HD float f_sub(float a, float b) { return __fsub_rn(a, b); }
SSPP_HD float f_mul(float a, float b) { return __fmul_rn(a, b); }
SSPP_HD float f_div(float a, float b) { return __fdiv_rn(a, b); }
SSPP_HD double d_add(double a, double b) { return __dadd_rn(a, b); }
SSPP_HD double d_sub(double a, double b) { return __dsub_rn(a, b); }
SSPP_HD double d_mul(double a, double b) { return __dmul_rn(a, b); }
SSPP_HD double d_div(double a, double b) { return __ddiv_rn(a, b); }
SSPP_HD double d_sqrt(double a) { return __dsqrt_rn(a); }
SSPP_HD float d2f(double a) { return __double2float_rn(a); }
SSPP_HD float f_sqrt(float a) { return __fsqrt_rn(a); }
#else
SSPP_HD float f_add(float a, float b) { return a + b; }
SSPP_HD float f_sub(float a, float b) { return a - b; }
SSPP_HD float f_mul(float a, float b) { return a * b; }
SSPP_HD float f_div(float a, float b) { return a / b; }
SSPP_HD double d_add(double a, double b) { return a + b; }
SSPP_HD double d_sub(double a, double b) { return a - b; }
SSPP_HD double d_mul(double a, double b) { return a * b; }
SSPP_HD double d_div(double a, double b) { return a / b; }
SSPP_HD double d_sqrt(double a) { return sqrt(a); }
SSPP_HD float d2f(double a) { return (float)a; }
SSPP_HD float f_sqrt(float a) { return sqrtf(a); }
#endif

// Constants of one map, computed once on the host in double.
struct KeySpace {
  double res;         // OcTree::resolution
  double res_factor;  // 1.0 / resolution
  float half_res_f;   // (float)(resolution * 0.5): the border shift computeRayKeys narrows to float
  double half_res_d;  // resolution * 0.5: the border shift castRay keeps in double
};

// upstream OcTreeBaseImpl::coordToKeyChecked(double): floor(res_factor * c) + 32768, valid in [0, 65536)
SSPP_HD bool coord_to_key(const KeySpace& ks, double c, int& key) {
#if defined(__CUDA_ARCH__)
  // floor + conversion in one saturating instruction; out-of-range values saturate far outside
  // [0, 65536) and NaN (which converts to 0) is refused explicitly
  const int k = __double2int_rd(d_mul(ks.res_factor, c));
  const unsigned u = (unsigned)k + 32768u;
  if (u >= 65536u || c != c) return false;
  key = (int)u;
  return true;
#else
  double s = floor(d_mul(ks.res_factor, c));
  // out-of-range values (and NaN) must fail without an undefined double->int conversion
  if (!(s >= -32768.0 && s < 32768.0)) return false;
  key = (int)s + 32768;
  return true;
#endif
}
// upstream OcTreeBaseImpl::keyToCoord(key_type): (double(key - 32768) + 0.5) * resolution
SSPP_HD double key_to_coord(const KeySpace& ks, int key) { return d_mul(d_add((double)(key - 32768), 0.5), ks.res); }

// Euclidean distance exactly as Dist() evaluates it (include/sspp/utils.h:356-359): double, left to right.
SSPP_HD double dist3(double ax, double ay, double az, double bx, double by, double bz) {
  double dx = d_sub(ax, bx), dy = d_sub(ay, by), dz = d_sub(az, bz);
  return d_sqrt(d_add(d_add(d_mul(dx, dx), d_mul(dy, dy)), d_mul(dz, dz)));
}

// Vector3::norm(): float products and sums, double sqrt.
SSPP_HD double norm3f(float x, float y, float z) {
  float s = f_add(f_add(f_mul(x, x), f_mul(y, y)), f_mul(z, z));
  return d_sqrt((double)s);
}

// One axis of the DDA set-up shared by computeRayKeys and castRay.
SSPP_HD void dda_axis(const KeySpace& ks, float dir, float origin, int key, bool border_in_float, int& step, double& tmax,
                      double& tdelta) {
  if (dir > 0.0f) step = 1;
  else if (dir < 0.0f) step = -1;
  else step = 0;
  if (step != 0) {
    double border = key_to_coord(ks, key);
    if (border_in_float) border = d_add(border, (double)(step > 0 ? ks.half_res_f : -ks.half_res_f));
    else border = d_add(border, step > 0 ? ks.half_res_d : -ks.half_res_d);
    tmax = d_div(d_sub(border, (double)origin), (double)dir);
    tdelta = d_div(ks.res, (double)fabsf(dir));
  } else {
    tmax = DBL_MAX;
    tdelta = DBL_MAX;
  }
}

// getLineStatus over computeRayKeys.  `cell(kx,ky,kz)` returns a CellStatus-like int: 0 lets the walk
// continue, anything else ends it and is returned (the caller folds "unknown counts only when
// treat_unknown_as_occupied" into the functor).  Start voxel tested, end voxel not; a segment whose
// ends share a voxel, or with an end outside the 16-bit key range, tests nothing.  `steps`
// (optional) counts cell() calls.
//
// Two rewrites that leave every result unchanged:
//  * length = (float)sqrt((double)s) upstream; a correctly rounded float sqrt gives the same float
//    (double rounding is innocuous for sqrt when the wide format has >= 2p+2 bits: 53 >= 50).
//  * upstream finds argmin(tMax) to advance, then min(tMax) to compare with the length, then
//    argmin again on the unchanged values: one argmin per iteration serves both (the axis the tie
//    rule picks always holds the minimum value).
//
// line_status_from: the same walk for a caller that already holds the origin key (kx, ky, kz valid) and, with
// start_known_free, already knows that the start voxel does not stop the walk — the edge kernel's row stage
// has both for every ray of a row.  line_status derives the origin key and calls it.
template <class Cell>
SSPP_HD int line_status_from(const KeySpace& ks, float ox, float oy, float oz, float ex, float ey, float ez, int kx, int ky, int kz,
                             bool start_known_free, Cell& cell, int* steps = nullptr) {
  int ekx, eky, ekz;
  if (!coord_to_key(ks, (double)ex, ekx) || !coord_to_key(ks, (double)ey, eky) || !coord_to_key(ks, (double)ez, ekz))
    return CELL_FREE;
  if (kx == ekx && ky == eky && kz == ekz) return CELL_FREE;
  int st;
  if (!start_known_free) {
    if (steps) ++*steps;
    st = cell(kx, ky, kz);
    if (st) return st;
  }

  float dx = f_sub(ex, ox), dy = f_sub(ey, oy), dz = f_sub(ez, oz);
  float length = f_sqrt(f_add(f_add(f_mul(dx, dx), f_mul(dy, dy)), f_mul(dz, dz)));
  dx = f_div(dx, length);
  dy = f_div(dy, length);
  dz = f_div(dz, length);
  int sx, sy, sz;
  double tx, ty, tz, dtx, dty, dtz;
  dda_axis(ks, dx, ox, kx, true, sx, tx, dtx);
  dda_axis(ks, dy, oy, ky, true, sy, ty, dty);
  dda_axis(ks, dz, oz, kz, true, sz, tz, dtz);
  const double dlen = (double)length;
  // upstream's tie rule: x only if strictly below both, y only if strictly below z, else z
  bool d0 = (tx < ty) && (tx < tz);
  bool d1 = !(tx < ty) && (ty < tz);
  for (int it = 0; it < 262144; ++it) {
    const bool d2 = !d0 && !d1;
    const double tsel = d0 ? tx : (d1 ? ty : tz);
    const double dsel = d0 ? dtx : (d1 ? dty : dtz);
    const double tn = d_add(tsel, dsel);
    tx = d0 ? tn : tx;
    ty = d1 ? tn : ty;
    tz = d2 ? tn : tz;
    kx = (kx + (d0 ? sx : 0)) & 0xFFFF;
    ky = (ky + (d1 ? sy : 0)) & 0xFFFF;
    kz = (kz + (d2 ? sz : 0)) & 0xFFFF;
    if (kx == ekx && ky == eky && kz == ekz) break;
    d0 = (tx < ty) && (tx < tz);
    d1 = !(tx < ty) && (ty < tz);
    const double m = d0 ? tx : (d1 ? ty : tz);  // == std::min(std::min(t0,t1),t2)
    if (m > dlen) break;
    if (steps) ++*steps;
    st = cell(kx, ky, kz);
    if (st) return st;
  }
  return CELL_FREE;
}

template <class Cell>
SSPP_HD int line_status(const KeySpace& ks, float ox, float oy, float oz, float ex, float ey, float ez, Cell& cell,
                        int* steps = nullptr) {
  int kx, ky, kz;
  if (!coord_to_key(ks, (double)ox, kx) || !coord_to_key(ks, (double)oy, ky) || !coord_to_key(ks, (double)oz, kz)) return CELL_FREE;
  return line_status_from(ks, ox, oy, oz, ex, ey, ez, kx, ky, kz, false, cell, steps);
}

// OccupancyOcTreeBase::castRay.  `cell` returns CELL_FREE / CELL_OCCUPIED / CELL_UNKNOWN for a key.
// Returns true when an occupied voxel was reached (hit_k* = its key).
// EDGE = false: the caller has shown that the walk cannot come near the border of the 16-bit key space (origin key
// further than max_range / res + 3 voxels from 0 and 65535 on every axis), so upstream's "leaving the key space"
// test, which then can never fire, is left out of the loop.
template <bool EDGE = true, class Cell>
SSPP_HD bool cast_ray(const KeySpace& ks, float ox, float oy, float oz, float dirx, float diry, float dirz,
                      bool ignore_unknown, double max_range, Cell& cell, int& hx, int& hy, int& hz, int* steps = nullptr) {
  int kx, ky, kz;
  hx = hy = hz = -1;
  if (!coord_to_key(ks, (double)ox, kx) || !coord_to_key(ks, (double)oy, ky) || !coord_to_key(ks, (double)oz, kz))
    return false;
  if (steps) ++*steps;
  int st = cell(kx, ky, kz);
  if (st == CELL_OCCUPIED) { hx = kx; hy = ky; hz = kz; return true; }
  if (st == CELL_UNKNOWN && !ignore_unknown) return false;
  // directionP.normalized(): divide by (float)norm() when the norm is positive
  double len = norm3f(dirx, diry, dirz);
  if (len > 0) {
    float fl = d2f(len);
    dirx = f_div(dirx, fl); diry = f_div(diry, fl); dirz = f_div(dirz, fl);
  }
  int sx, sy, sz;
  double tx, ty, tz, dtx, dty, dtz;
  dda_axis(ks, dirx, ox, kx, false, sx, tx, dtx);
  dda_axis(ks, diry, oy, ky, false, sy, ty, dty);
  dda_axis(ks, dirz, oz, kz, false, sz, tz, dtz);
  if (sx == 0 && sy == 0 && sz == 0) return false;
  const bool max_range_set = (max_range > 0.0);
  const double maxrange_sq = d_mul(max_range, max_range);
  // squared distance of the current voxel centre from the origin, one term per axis: upstream rebuilds
  // all three every step; only the axis that moved changes, so the other two terms are kept
  // (each term is the float product (c - o)^2 widened to double; the sum keeps upstream's order).
  // The test cannot fire while the ray is still well short of its range: a step crosses a voxel face at ray
  // parameter t (the tMax it picks; the direction is normalised, so t is a distance from the origin), the voxel
  // it enters holds that crossing point, and a voxel's centre is never more than 0.87 res from a point of the
  // voxel - so the centre is within t + 0.87 res of the origin.  While t <= max_range - res the bookkeeping is
  // skipped (the 0.13 res left over dwarfs every rounding error involved), and the three terms are rebuilt from
  // the current keys when it starts (each term depends on the current key of its axis only, so nothing is lost;
  // t never decreases, so once started it stays on).
  const double range_guard = d_sub(max_range, ks.res);
  bool ranged = false;  // qx, qy, qz hold the terms of the current voxel
  double qx = 0.0, qy = 0.0, qz = 0.0;
  for (int it = 0; it < 262144; ++it) {
    // upstream's tie rule: x only if strictly below both, y only if strictly below z, else z.  The
    // three per-axis branches are written as selects: the lanes of a warp step along different axes,
    // and three divergent branches would each run at a third of the warp.
    const bool d0 = (tx < ty) && (tx < tz);
    const bool d1 = !(tx < ty) && (ty < tz);
    const bool d2 = !d0 && !d1;
    const double tcur = d0 ? tx : (d1 ? ty : tz);  // == min(tx, ty, tz): where the ray leaves the current voxel
    const bool check = max_range_set && (ranged || tcur > range_guard);
    if (check && !ranged) {
      float ddx = f_sub(d2f(key_to_coord(ks, kx)), ox), ddy = f_sub(d2f(key_to_coord(ks, ky)), oy),
            ddz = f_sub(d2f(key_to_coord(ks, kz)), oz);
      qx = (double)f_mul(ddx, ddx);
      qy = (double)f_mul(ddy, ddy);
      qz = (double)f_mul(ddz, ddz);
      ranged = true;
    }
    const int ksel = d0 ? kx : (d1 ? ky : kz);
    const int ssel = d0 ? sx : (d1 ? sy : sz);
    if (EDGE && ((ssel < 0 && ksel == 0) || (ssel > 0 && ksel == 65535))) return false;  // leaving the key space
    const int kn = ksel + ssel;
    const double tn = d_add(tcur, d0 ? dtx : (d1 ? dty : dtz));
    kx = d0 ? kn : kx;
    ky = d1 ? kn : ky;
    kz = d2 ? kn : kz;
    tx = d0 ? tn : tx;
    ty = d1 ? tn : ty;
    tz = d2 ? tn : tz;
    if (check) {
      const float dd = f_sub(d2f(key_to_coord(ks, kn)), d0 ? ox : (d1 ? oy : oz));
      const double qn = (double)f_mul(dd, dd);
      qx = d0 ? qn : qx;
      qy = d1 ? qn : qy;
      qz = d2 ? qn : qz;
      double dsq = 0.0;
      dsq = d_add(dsq, qx);
      dsq = d_add(dsq, qy);
      dsq = d_add(dsq, qz);
      if (dsq > maxrange_sq) return false;
    }
    if (steps) ++*steps;
    st = cell(kx, ky, kz);
    if (st == CELL_OCCUPIED) { hx = kx; hy = ky; hz = kz; return true; }
    if (st == CELL_UNKNOWN && !ignore_unknown) return false;
  }
  return false;
}

// The parallel-ray lattice OctomapWorld::getLineStatusBoundingBox sweeps over the robot box
// (upstream): per axis disc = size / ceil((size + 0.001) / res), offsets from -size/2 while
// <= +size/2 by `double +=`.  The accumulated rounding decides how many offsets there are (e.g.
// four, not five, along a 0.3 m axis at res 0.15), so the table is always generated by this loop.
// Host only.  Returns the count, or -1 when it exceeds cap.
inline int ray_lattice_axis(double size, double res, double* out, int cap) {
  const double epsilon = 0.001;
  double disc = size / ceil((size + epsilon) / res);
  if (disc <= 0.0) disc = 1.0;
  double half = size * 0.5;
  int n = 0;
  for (double v = -half; v <= half; v += disc) {
    if (n >= cap) return -1;
    out[n++] = v;
  }
  return n;
}

}  // namespace sspp
