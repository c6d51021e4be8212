// visibility_kernels.cuh — coverage visibility: frustum planes, point binning, in-frustum gather, occlusion rays, per-viewpoint sort.
// Part of the single translation unit sspp_gpu.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

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

// The points a viewpoint has to test: for every (z, y) row of cells inside the frustum's bounding box, the run of x
// cells the frustum can reach.  Each of the six planes bounds x on that row: a point (x, y, z) of the row passes
// plane k only if a x + min over the row's y/z extent of (b y + c z) + d <= 0, which is a half-line in x.  The frustum
// test itself runs in float with its own rounding, so the bound is relaxed by `tol`, forty times the worst rounding
// error of that float expression, and the row's y/z extent is widened by a thousandth of a cell: a cell is only
// left out when no point binned into it can pass the float test.  Returns the binned range [b, e) of the row.
__device__ __noinline__ void view_row_range(const ViewRec* __restrict__ Rg, const PointGrid& g, const int* __restrict__ cell_start,
                                            int cz, int cy, int& b, int& e) {
  const ViewRec& R = *Rg;  // in global memory: the callers keep their register copy of the record to themselves
  const double cell = 1.0 / g.inv_cell, widen = cell * 1e-3, big = 1e30;
  const double y0 = cy == 0 ? -big : g.min[1] + cy * cell - widen, y1 = cy == g.n[1] - 1 ? big : g.min[1] + (cy + 1) * cell + widen;
  const double z0 = cz == 0 ? -big : g.min[2] + cz * cell - widen, z1 = cz == g.n[2] - 1 ? big : g.min[2] + (cz + 1) * cell + widen;
  const double xm = fmax(fabs(g.min[0]), fabs(g.min[0] + g.n[0] * cell)), ym = fmax(fabs(g.min[1]), fabs(g.min[1] + g.n[1] * cell)),
               zm = fmax(fabs(g.min[2]), fabs(g.min[2] + g.n[2] * cell));
  double xlo = -big, xhi = big;
  bool empty = false;
#pragma unroll 1
  for (int k = 0; k < 6; ++k) {
    const double pa = R.pl[k][0], pb = R.pl[k][1], pc = R.pl[k][2], pd = R.pl[k][3];
    const double tol = 1e-5 * (fabs(pa) * xm + fabs(pb) * ym + fabs(pc) * zm + fabs(pd)) + 1e-9;
    const double myz = fmin(pb * y0, pb * y1) + fmin(pc * z0, pc * z1);
    const double rhs = tol - myz - pd;  // a x <= rhs is necessary for the point to pass plane k
    if (pa > 0.0) xhi = fmin(xhi, rhs / pa);
    else if (pa < 0.0) xlo = fmax(xlo, rhs / pa);
    else if (rhs < 0.0) empty = true;
  }
  int ca = R.c0[0], cb = R.c1[0];
  if (xlo > -big) ca = max(ca, pgrid_axis(g, xlo - widen, 0));
  if (xhi < big) cb = min(cb, pgrid_axis(g, xhi + widen, 0));
  if (empty || ca > cb || !(xlo <= xhi)) { b = e = 0; return; }
  const int base = (cz * g.n[1] + cy) * g.n[0];
  b = cell_start[base + ca];
  e = cell_start[base + cb + 1];
}

// Ballot words a viewpoint's gather loop produces: one per 32 candidate points of every row of its cell box.  The
// count pass stores its ballots there and the fill pass reads them back instead of loading the points and running
// the six-plane test a second time.  One warp per viewpoint, one lane per row.
__global__ void __launch_bounds__(256) k_view_bound(const ViewRec* __restrict__ views, int nv, PointGrid g,
                                                    const int* __restrict__ cell_start, int* __restrict__ words) {
  const int w = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (w >= nv) return;
  const ViewRec& R = views[w];
  const int nyr = R.c1[1] - R.c0[1] + 1, nrows = nyr * (R.c1[2] - R.c0[2] + 1);
  int n = 0;
  for (int r = lane; r < nrows; r += 32) {
    int b, e;
    view_row_range(views + w, g, cell_start, R.c0[2] + r / nyr, R.c0[1] + r % nyr, b, e);
    n += (e - b + 31) >> 5;
  }
  for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if (lane == 0) words[w] = n;
}

// one warp per viewpoint: in-frustum points of the cells its frustum can reach.
// FILL == false: run the frustum test, keep the ballots, count.  FILL == true (after the scan of the
// counts): write (viewpoint, binned position) per ray from the stored ballots.
#ifndef GATHER_MIN_BLOCKS
#define GATHER_MIN_BLOCKS 6
#endif
template <bool FILL>
__global__ void __launch_bounds__(256, GATHER_MIN_BLOCKS) k_view_gather(const ViewRec* __restrict__ views, int64_t v0, int nv, PointGrid g,
                                                     const int* __restrict__ cell_start, const float4* __restrict__ sp,
                                                     int* __restrict__ ray_count, const long long* __restrict__ ray_off,
                                                     int* __restrict__ ray_view, int* __restrict__ ray_pt,
                                                     const long long* __restrict__ word_off, uint32_t* __restrict__ mask,
                                                     long long mask_cap, long long ray_cap) {
  const int w = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  int lane = threadIdx.x & 31;
  if (w >= nv) return;
  // the buffers keep their size from call to call; when this call's sizes outgrow them every kernel of
  // the pass returns at once and the host, at its one synchronisation, grows them and repeats the pass
  if (word_off[nv] > mask_cap || (FILL && ray_off[nv] > ray_cap)) return;
  const ViewRec R = views[v0 + w];
  long long base = FILL ? ray_off[w] : 0;
  uint32_t* mw = mask + word_off[w];
  int filled = 0, it = 0;
  const int nyr = R.c1[1] - R.c0[1] + 1, nrows = nyr * (R.c1[2] - R.c0[2] + 1);
  for (int r0 = 0; r0 < nrows; r0 += 32) {
    // the lanes work out the runs of 32 rows side by side, then the warp walks them one after the other
    int rb = 0, re = 0;
    if (r0 + lane < nrows) view_row_range(views + v0 + w, g, cell_start, R.c0[2] + (r0 + lane) / nyr, R.c0[1] + (r0 + lane) % nyr, rb, re);
    const int nr = min(32, nrows - r0);
    for (int rr = 0; rr < nr; ++rr) {
      const int b = __shfl_sync(0xffffffffu, rb, rr), e = __shfl_sync(0xffffffffu, re, rr);
      if (FILL) {
        for (int j0 = b; j0 < e; j0 += 32, ++it) {
          const int j = j0 + lane;
          const unsigned bal = mw[it];
          if ((bal >> lane) & 1u) {
            long long o = base + filled + __popc(bal & ((1u << lane) - 1u));
            ray_view[o] = w;
            ray_pt[o] = j;
          }
          filled += __popc(bal);
        }
      } else {
        // four chunks of 32 points per turn: the four loads of a lane are in flight together
        for (int j0 = b; j0 < e; j0 += 128) {
          float4 p[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int j = j0 + 32 * u + lane;
            p[u] = (j < e) ? sp[j] : make_float4(0.f, 0.f, 0.f, 0.f);
          }
          unsigned mine = 0;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int j = j0 + 32 * u + lane;
            const bool in = (j < e) && in_frustum_dev(R, p[u].x, p[u].y, p[u].z);
            const unsigned bal = __ballot_sync(0xffffffffu, in);
            if (lane == u) mine = bal;
            filled += __popc(bal);
          }
          const int nchunks = min(4, (e - j0 + 31) >> 5);
          if (lane < nchunks) mw[it + lane] = mine;
          it += nchunks;
        }
      }
    }
  }
  if (!FILL && lane == 0) ray_count[w] = filled;
}

// one thread per occlusion ray, blocks of 256 rays handed out grid-stride (the ray count lives on the device:
// ray_off[nv]).  COUNT: the instrumented variant also counts the cells castRay looks up.
// Tuning switch VIEW_RAYS_MIN_BLOCKS: CTAs per SM the ray kernel is compiled for.  Left undefined the compiler settles on
// 64 registers = four CTAs per SM by itself, and that build is the one all numbers are from (asking for four explicitly
// changes the instruction schedule: 4.6 ms against 4.35 on the bench scene; five and six are slower still).
#ifdef VIEW_RAYS_MIN_BLOCKS
#define VIEW_RAYS_BOUNDS __launch_bounds__(256, VIEW_RAYS_MIN_BLOCKS)
#define VIEW_RAYS_RESIDENT VIEW_RAYS_MIN_BLOCKS
#else
#define VIEW_RAYS_BOUNDS __launch_bounds__(256)
#define VIEW_RAYS_RESIDENT 4
#endif
template <bool COUNT>
__global__ void VIEW_RAYS_BOUNDS k_view_rays(DeviceMap m, const ViewRec* __restrict__ views, int64_t v0,
                                                   const float4* __restrict__ sp, const int* __restrict__ ray_view,
                                                   const int* __restrict__ ray_pt, const long long* __restrict__ ray_total,
                                                   long long ray_cap, uint8_t* __restrict__ vis, int* vis_count,
                                                   unsigned long long* steps_total) {
  const long long nrays = *ray_total;
  if (nrays > ray_cap) return;
  const bool small_map = (long long)(m.dims[0] >> 5) * m.dims[1] * m.dims[2] < (1ll << 31);
  unsigned long long my_steps = 0;
  for (long long i0 = (long long)blockIdx.x * blockDim.x; i0 < nrays; i0 += (long long)gridDim.x * blockDim.x) {
    const long long i = i0 + threadIdx.x;
    bool visible = false;
    int w = -1;
    const bool have = i < nrays;
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    float ox = 0.f, oy = 0.f, oz = 0.f;
    if (have) {
      w = ray_view[i];
      const ViewRec& R = views[v0 + w];
      p = sp[ray_pt[i]];
      ox = R.T[0]; oy = R.T[1]; oz = R.T[2];
    }
    const float dx = f_sub(p.x, ox), dy = f_sub(p.y, oy), dz = f_sub(p.z, oz);
    const double range = norm3f(dx, dy, dz);
    // A ray whose whole reach — origin key +- (range / res + 3) voxels — lies inside the mapped window (and so well
    // inside the key space) needs neither the window test per voxel nor the key-space test per step.  The warp takes
    // that path only when all its rays qualify (they share a viewpoint, so they almost always agree).
    bool inside = !have;
    if (have && small_map) {
      int kx, ky, kz;
      if (coord_to_key(m.ks, (double)ox, kx) && coord_to_key(m.ks, (double)oy, ky) && coord_to_key(m.ks, (double)oz, kz)) {
        const int reach = (int)fmin(ceil(range * m.ks.res_factor), 100000.0) + 3;
        inside = kx - reach >= m.kmin[0] && kx + reach < m.kmin[0] + m.dims[0] && ky - reach >= m.kmin[1] &&
                 ky + reach < m.kmin[1] + m.dims[1] && kz - reach >= m.kmin[2] && kz + reach < m.kmin[2] + m.dims[2] &&
                 m.kmin[0] > 0 && m.kmin[1] > 0 && m.kmin[2] > 0;
      }
    }
    inside = __all_sync(0xffffffffu, inside);
    if (have) {
      int hx, hy, hz, nsteps = 0;
      bool hit;
      if (inside) {
        CellOccupiedInside cell;
        cell.occ = m.occ; cell.kx0 = m.kmin[0]; cell.ky0 = m.kmin[1]; cell.kz0 = m.kmin[2];
        cell.ny = (unsigned)m.dims[1]; cell.nz = (unsigned)m.dims[2];
        hit = cast_ray<false>(m.ks, ox, oy, oz, dx, dy, dz, true, range, cell, hx, hy, hz, COUNT ? &nsteps : nullptr);
      } else {
        CellOccupiedOnly cell;  // ignoreUnknown = true: unknown cells are walked through
        cell.m = &m;
        hit = cast_ray(m.ks, ox, oy, oz, dx, dy, dz, true, range, cell, hx, hy, hz, COUNT ? &nsteps : nullptr);
      }
      if (COUNT) my_steps += (unsigned long long)nsteps;
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
  if (COUNT) {
    for (int o = 16; o; o >>= 1) my_steps += __shfl_xor_sync(0xffffffffu, my_steps, o);
    if ((threadIdx.x & 31) == 0 && my_steps) atomicAdd(steps_total, my_steps);
  }
}

// one warp per viewpoint: compact the visible rays' ORIGINAL point indices
__global__ void k_view_compact(int nv, const long long* __restrict__ ray_off, const uint8_t* __restrict__ vis,
                               const int* __restrict__ ray_pt, const float4* __restrict__ sp,
                               const long long* __restrict__ vis_ptr, int* __restrict__ out, long long ray_cap, long long out_cap) {
  int w = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  int lane = threadIdx.x & 31;
  if (w >= nv) return;
  if (ray_off[nv] > ray_cap || vis_ptr[nv] > out_cap) return;
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
__global__ void __launch_bounds__(256) k_segment_sort(int nv, const long long* __restrict__ ptr, int* data, long long data_cap) {
  extern __shared__ int sh[];
  const int SM_CAP = 8192;
  if (ptr[nv] > data_cap) return;
  for (int w = blockIdx.x; w < nv; w += gridDim.x) {
    const long long b = ptr[w], len = ptr[w + 1] - b;
    if (len < 2) continue;
    if (len <= 2 * 256) {
      // short segments (the usual case: a viewpoint sees a hundred points or so): rank sort, two barriers.
      // The indices of one viewpoint are distinct, so the rank of a value is the number of smaller ones;
      // every thread reads the same shared word at the same time (a broadcast).
      const int n = (int)len;
      int* g = data + b;
      for (int i = threadIdx.x; i < n; i += blockDim.x) sh[i] = g[i];
      __syncthreads();
      for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int v = sh[i];
        int rank = 0;
        for (int k = 0; k < n; ++k) rank += (sh[k] < v);
        g[rank] = v;
      }
      __syncthreads();
      continue;
    }
    int* a = data + b;
    const bool in_smem = len <= SM_CAP;
    if (in_smem) {
      for (int i = threadIdx.x; i < len; i += blockDim.x) sh[i] = a[i];
      a = sh;
    }
    __syncthreads();
    // bitonic network on the next power of two; blocks are powers of two, so the index arithmetic is shifts
    int lp = 1;
    while ((1ll << lp) < len) ++lp;
    const long long half = 1ll << (lp - 1);
    for (int lk = 1; lk <= lp; ++lk) {
      // flip: i in the lower half of its 2^lk block is compared with its mirror
      for (long long t = threadIdx.x; t < half; t += blockDim.x) {
        const long long blk = t >> (lk - 1), o = t & ((1ll << (lk - 1)) - 1);
        const long long i = (blk << lk) + o, j = (blk << lk) + ((1ll << lk) - 1 - o);
        if (i < len) cmpswap(a, i, j, len);
      }
      __syncthreads();
      for (int lh = lk - 2; lh >= 0; --lh) {
        for (long long t = threadIdx.x; t < half; t += blockDim.x) {
          const long long blk = t >> lh, o = t & ((1ll << lh) - 1);
          const long long i = (blk << (lh + 1)) + o, j = i + (1ll << lh);
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


}  // namespace
