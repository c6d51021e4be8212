// ORACLE — TEST INFRASTRUCTURE ONLY (see octomap_restated.hpp for the parity statement).
// C entry points over the restated octomap / octomap_world plus CPU restatements of the
// reference's own loops on the hot path, for tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py.  Nothing in sspp_b200/ may load this.
//
//   orc_connect            <- PathPlanner::connectNodes          src/pathplanner.cpp:527-567
//                             PathPlanner::connectClustersInternalNodes  :612-656 (mode 1)
//                             DistanceHeuristic::isConnectionConditionSatisfied
//                                                                 src/distance_heuristic.cpp:99-127
//                             Dist                                include/sspp/utils.h:356-359
//   orc_dedup_poses        <- SearchSpace::insertNode / nodeExists   src/searchspace.cpp:69-105,164-175
//                             samePosition / isEqual                 include/sspp/utils.h:386-390,231-239
//   orc_visibility         <- AUTHORED SPEC (CoverageHeuristic has no source in the reference,
//                             SURVEY.md F1): PCL-FrustumCulling-style six-plane test in float,
//                             then octomap castRay towards the point.  Stated in full below.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <thread>
#include <vector>

#include "octomap_restated.hpp"

using orc::OcTreeKey;
using orc::OctomapWorld;
using orc::point3d;

// ------------------------------------------------------------------------------------------
// Visibility spec v1 (authored; frozen here, the CUDA kernel follows it to the bit).
//   Inputs: sensor pose (t in double^3, q = (x,y,z,w) in double^4), frustum (hfov_deg, vfov_deg,
//   near, far) as Sensors holds them (include/sspp/sensors.h:61), model points in float^3.
//   1. R = tf::Matrix3x3::setRotation(q) in double; view = col 0, up = col 1, right = col 2
//      (PCL camera convention, which is why Sensors::robot2sensorTransformation appends Rx(+90deg),
//      src/sensors.cpp:74-80); all narrowed to float; T = (float)t.
//   2. vh = (float)(2*tan(vfov_rad/2)), hw = (float)(2*tan(hfov_rad/2)) with
//      vfov_rad = (float)(vfov*M_PI/180) — evaluated on the host; np_h = vh*near etc. in float.
//   3. Six planes (left,right,top,bottom,far,near) from the far/near corner points exactly as
//      written in frustum_planes() below, float arithmetic, no contraction.
//   4. p is in the frustum iff for each plane ((p.x*a + p.y*b) + p.z*c) + d <= 0 in float.
//   5. An in-frustum point casts ONE ray: castRay(origin = T, direction = p - T (float),
//      ignoreUnknown = true, maxRange = (p - T).norm()).  p is VISIBLE iff castRay returns false
//      (nothing occupied met within range) or the occupied voxel it stops in has the key of p.
// ------------------------------------------------------------------------------------------
namespace {

struct Frustum {
  float pl[6][4];
  float T[3];
};

inline void cross3(const float a[3], const float b[3], float o[3]) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}
inline float dot3(const float a[3], const float b[3]) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }

void frustum_planes(const double pose[7], const float fr[4] /*vh, hw, near, far*/, Frustum& F) {
  double x = pose[3], y = pose[4], z = pose[5], w = pose[6];
  double d = x * x + y * y + z * z + w * w;
  double s = 2.0 / d;
  double xs = x * s, ys = y * s, zs = z * s;
  double wx = w * xs, wy = w * ys, wz = w * zs;
  double xx = x * xs, xy = x * ys, xz = x * zs;
  double yy = y * ys, yz = y * zs, zz = z * zs;
  float view[3] = {(float)(1.0 - (yy + zz)), (float)(xy + wz), (float)(xz - wy)};
  float up[3] = {(float)(xy - wz), (float)(1.0 - (xx + zz)), (float)(yz + wx)};
  float right[3] = {(float)(xz + wy), (float)(yz - wx), (float)(1.0 - (xx + yy))};
  float T[3] = {(float)pose[0], (float)pose[1], (float)pose[2]};
  float vh = fr[0], hw = fr[1], np_dist = fr[2], fp_dist = fr[3];
  float np_h = vh * np_dist, np_w = hw * np_dist, fp_h = vh * fp_dist, fp_w = hw * fp_dist;
  float fp_c[3], np_c[3], fp_tl[3], fp_tr[3], fp_bl[3], fp_br[3], np_tr[3], np_bl[3], np_br[3];
  for (int i = 0; i < 3; ++i) {
    fp_c[i] = T[i] + view[i] * fp_dist;
    np_c[i] = T[i] + view[i] * np_dist;
    float fu = up[i] * fp_h / 2, fr_ = right[i] * fp_w / 2;
    float nu = up[i] * np_h / 2, nr = right[i] * np_w / 2;
    fp_tl[i] = (fp_c[i] + fu) - fr_;
    fp_tr[i] = (fp_c[i] + fu) + fr_;
    fp_bl[i] = (fp_c[i] - fu) - fr_;
    fp_br[i] = (fp_c[i] - fu) + fr_;
    np_tr[i] = (np_c[i] + nu) + nr;
    np_bl[i] = (np_c[i] - nu) - nr;
    np_br[i] = (np_c[i] - nu) + nr;
  }
  float e1[3], e2[3], a[3], b[3], c[3], dd[3];
  for (int i = 0; i < 3; ++i) { e1[i] = fp_bl[i] - fp_br[i]; e2[i] = fp_tr[i] - fp_br[i]; }
  cross3(e1, e2, F.pl[4]);
  F.pl[4][3] = -dot3(fp_c, F.pl[4]);
  for (int i = 0; i < 3; ++i) { e1[i] = np_tr[i] - np_br[i]; e2[i] = np_bl[i] - np_br[i]; }
  cross3(e1, e2, F.pl[5]);
  F.pl[5][3] = -dot3(np_c, F.pl[5]);
  for (int i = 0; i < 3; ++i) {
    a[i] = fp_bl[i] - T[i]; b[i] = fp_br[i] - T[i]; c[i] = fp_tr[i] - T[i]; dd[i] = fp_tl[i] - T[i];
  }
  cross3(dd, a, F.pl[0]);  // left
  cross3(b, c, F.pl[1]);   // right
  cross3(c, dd, F.pl[2]);  // top
  cross3(a, b, F.pl[3]);   // bottom
  for (int k = 0; k < 4; ++k) F.pl[k][3] = -dot3(T, F.pl[k]);
  for (int i = 0; i < 3; ++i) F.T[i] = T[i];
}

inline bool in_frustum(const Frustum& F, const float p[3]) {
  for (int k = 0; k < 6; ++k) {
    float v = ((p[0] * F.pl[k][0] + p[1] * F.pl[k][1]) + p[2] * F.pl[k][2]) + F.pl[k][3];
    if (!(v <= 0)) return false;
  }
  return true;
}

inline double dist3(const double* a, const double* b) {
  // Dist(): sqrt((ax-bx)^2 + (ay-by)^2 + (az-bz)^2), left to right, double (utils.h:356-359)
  return std::sqrt((a[0] - b[0]) * (a[0] - b[0]) + (a[1] - b[1]) * (a[1] - b[1]) + (a[2] - b[2]) * (a[2] - b[2]));
}

}  // namespace

extern "C" {

void* orc_world_create(double resolution, int treat_unknown_as_occupied) {
  return new OctomapWorld(resolution, treat_unknown_as_occupied != 0);
}
void orc_world_destroy(void* w) { delete (OctomapWorld*)w; }

void orc_world_set_box(void* w, const double pos[3], const double box[3], int occupied) {
  OctomapWorld* W = (OctomapWorld*)w;
  if (occupied) W->setOccupied(pos, box); else W->setFree(pos, box);
}

// direct leaf insertion by key (lazy), e.g. a voxelised point cloud
void orc_world_set_keys(void* w, const int32_t* keys, long n, int occupied) {
  OctomapWorld* W = (OctomapWorld*)w;
  float v = occupied ? W->tree.clamping_thres_max : W->tree.clamping_thres_min;
  for (long i = 0; i < n; ++i)
    W->tree.setNodeValue(OcTreeKey((orc::key_type)keys[3 * i], (orc::key_type)keys[3 * i + 1], (orc::key_type)keys[3 * i + 2]),
                         v, true);
  W->tree.updateInnerOccupancy();
}

// insertPointcloud for world-frame float points (see OctomapWorld::insertPointcloud)
void orc_world_insert_pointcloud(void* w, const float* pts, long n, const double origin[3], double max_range) {
  ((OctomapWorld*)w)->insertPointcloud(pts, n, origin, max_range);
}

// log-odds of every voxel of a key window (NaN where nothing was observed), x fastest
void orc_world_export_logodds(void* w, const int32_t kmin[3], const int32_t dims[3], float* out) {
  OctomapWorld* W = (OctomapWorld*)w;
  size_t i = 0;
  for (int z = 0; z < dims[2]; ++z)
    for (int y = 0; y < dims[1]; ++y)
      for (int x = 0; x < dims[0]; ++x, ++i) {
        const orc::OcTreeNode* n = W->tree.search(OcTreeKey((orc::key_type)(kmin[0] + x), (orc::key_type)(kmin[1] + y), (orc::key_type)(kmin[2] + z)));
        out[i] = n ? n->value : std::nanf("");
      }
}

void orc_world_prune(void* w) { ((OctomapWorld*)w)->tree.prune(); }
long orc_world_num_tree_nodes(void* w) { return (long)((OctomapWorld*)w)->tree.size(); }

void orc_world_build_dense(void* w, const int32_t kmin[3], const int32_t dims[3], const uint32_t* occ_bits,
                           const uint32_t* known_bits, int prune) {
  OctomapWorld* W = (OctomapWorld*)w;
  W->tree.buildFromDense(kmin, dims, occ_bits, known_bits, W->tree.clamping_thres_max, W->tree.clamping_thres_min,
                         prune != 0);
}

// getAllOccupiedBoxes / getAllFreeBoxes; call with centres == NULL to count
long orc_world_export_boxes(void* w, int occupied, double* centres, double* sizes, long cap) {
  std::vector<double> c, s;
  ((OctomapWorld*)w)->getAllBoxes(occupied != 0, c, s);
  long n = (long)s.size();
  if (centres && sizes) {
    long m = std::min(n, cap);
    std::copy(c.begin(), c.begin() + 3 * m, centres);
    std::copy(s.begin(), s.begin() + m, sizes);
  }
  return n;
}

// per-voxel search() over a key window -> x-fastest dense bit planes (32-bit words per row)
void orc_world_export_dense(void* w, const int32_t kmin[3], const int32_t dims[3], uint32_t* occ_bits,
                            uint32_t* known_bits) {
  OctomapWorld* W = (OctomapWorld*)w;
  int wpr = (dims[0] + 31) / 32;
  size_t words = (size_t)wpr * dims[1] * dims[2];
  std::fill(occ_bits, occ_bits + words, 0u);
  std::fill(known_bits, known_bits + words, 0u);
  for (int z = 0; z < dims[2]; ++z)
    for (int y = 0; y < dims[1]; ++y)
      for (int x = 0; x < dims[0]; ++x) {
        int kx = kmin[0] + x, ky = kmin[1] + y, kz = kmin[2] + z;
        if (kx < 0 || ky < 0 || kz < 0 || kx > 65535 || ky > 65535 || kz > 65535) continue;
        const orc::OcTreeNode* n = W->tree.search(OcTreeKey((orc::key_type)kx, (orc::key_type)ky, (orc::key_type)kz));
        if (!n) continue;
        size_t wi = ((size_t)z * dims[1] + y) * wpr + (x >> 5);
        known_bits[wi] |= 1u << (x & 31);
        if (W->tree.isNodeOccupied(n)) occ_bits[wi] |= 1u << (x & 31);
      }
}

// the ray lattice of getLineStatusBoundingBox; each table must hold >= 4096 entries
void orc_ray_lattice(const double box[3], double resolution, double* offx, double* offy, double* offz, int32_t n[3]) {
  std::vector<double> off[3];
  OctomapWorld::rayLattice(box, resolution, off);
  double* dst[3] = {offx, offy, offz};
  for (int a = 0; a < 3; ++a) {
    n[a] = (int32_t)off[a].size();
    std::copy(off[a].begin(), off[a].end(), dst[a]);
  }
}

// computeRayKeys for one float segment; returns the number of keys (written up to cap), -1 if out of range
long orc_ray_keys(void* w, const float origin[3], const float end[3], int32_t* keys, long cap) {
  OctomapWorld* W = (OctomapWorld*)w;
  std::vector<OcTreeKey> ray;
  bool ok = W->tree.computeRayKeys(point3d(origin[0], origin[1], origin[2]), point3d(end[0], end[1], end[2]), ray);
  if (!ok) return -1;
  long m = std::min((long)ray.size(), cap);
  for (long i = 0; i < m; ++i) { keys[3 * i] = ray[i][0]; keys[3 * i + 1] = ray[i][1]; keys[3 * i + 2] = ray[i][2]; }
  return (long)ray.size();
}

void orc_line_status(void* w, const double* starts, const double* ends, long n, uint8_t* status) {
  OctomapWorld* W = (OctomapWorld*)w;
  for (long i = 0; i < n; ++i) status[i] = (uint8_t)W->getLineStatus(starts + 3 * i, ends + 3 * i);
}

// batched getLineStatusBoundingBox; steps_out (optional) = voxel lookups made in total
void orc_line_status_bbox(void* w, const double* starts, const double* ends, long n, const double box[3],
                          uint8_t* status, long* steps_out, long* rays_out) {
  OctomapWorld* W = (OctomapWorld*)w;
  long s0 = OctomapWorld::tl_steps(), r0 = OctomapWorld::tl_rays();
  for (long i = 0; i < n; ++i) status[i] = (uint8_t)W->getLineStatusBoundingBox(starts + 3 * i, ends + 3 * i, box);
  if (steps_out) *steps_out = OctomapWorld::tl_steps() - s0;
  if (rays_out) *rays_out = OctomapWorld::tl_rays() - r0;
}

// same, split over `threads` host threads, each with its own query scratch (courtesy multi-core leg)
void orc_line_status_bbox_mt(void* w, const double* starts, const double* ends, long n, const double box[3],
                             uint8_t* status, int threads) {
  OctomapWorld* W = (OctomapWorld*)w;
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) {
    pool.push_back(std::thread([=]() {
      long lo = n * t / threads, hi = n * (t + 1) / threads;
      for (long i = lo; i < hi; ++i)
        status[i] = (uint8_t)W->getLineStatusBoundingBox(starts + 3 * i, ends + 3 * i, box);
    }));
  }
  for (size_t t = 0; t < pool.size(); ++t) pool[t].join();
}

void orc_cell_status_point(void* w, const double* pts, long n, uint8_t* status) {
  OctomapWorld* W = (OctomapWorld*)w;
  for (long i = 0; i < n; ++i) status[i] = (uint8_t)W->getCellStatusPoint(pts + 3 * i);
}

void orc_cell_status_bbox(void* w, const double* pts, long n, const double box[3], uint8_t* status) {
  OctomapWorld* W = (OctomapWorld*)w;
  for (long i = 0; i < n; ++i) status[i] = (uint8_t)W->getCellStatusBoundingBox(pts + 3 * i, box);
}

// PathPlanner::connectNodes restated: all ordered pairs in LIST ORDER (the caller passes nodes in
// list order), `<=` radius test on the double Euclidean distance, same-position pairs connected
// without a collision test, otherwise getLineStatusBoundingBox(temp -> S) == kFree.  w == NULL is
// the manager==NULL case (every in-radius pair accepted, distance_heuristic.cpp:101,126).
//   mode 0: connectNodes (exact != on the three coordinates, pathplanner.cpp:545)
//   mode 1: connectClustersInternalNodes (isPositionEqual 1e-12, pathplanner.cpp:630)
// row_ptr has n+1 entries; col_idx is filled up to cap; returns E.  counters[0] = candidate pairs
// handed to the collision test, [1] = same-position pairs, [2] = voxel lookups, [3] = rays.
long orc_connect(void* w, const double* xyz, long n, double radius, const double box[3], int mode, int64_t* row_ptr,
                 int32_t* col_idx, long cap, long* counters) {
  OctomapWorld* W = (OctomapWorld*)w;
  long E = 0, cand = 0, same = 0;
  long s0 = OctomapWorld::tl_steps(), r0 = OctomapWorld::tl_rays();
  for (long t = 0; t < n; ++t) {
    row_ptr[t] = E;
    const double* temp = xyz + 3 * t;
    for (long s = 0; s < n; ++s) {
      const double* S = xyz + 3 * s;
      double distance = dist3(S, temp);
      if (distance <= radius && s != t) {
        bool differs = (mode == 0)
                           ? (S[0] != temp[0] || S[1] != temp[1] || S[2] != temp[2])
                           : !(std::fabs(S[0] - temp[0]) < 1E-12 && std::fabs(S[1] - temp[1]) < 1E-12 &&
                               std::fabs(S[2] - temp[2]) < 1E-12);
        bool ok;
        if (differs) {
          ++cand;
          ok = W ? (W->getLineStatusBoundingBox(temp, S, box) == OctomapWorld::kFree) : true;
        } else {
          ++same;
          ok = true;
        }
        if (ok) {
          if (E < cap) col_idx[E] = (int32_t)s;
          ++E;
        }
      }
    }
  }
  row_ptr[n] = E;
  if (counters) {
    counters[0] = cand; counters[1] = same;
    counters[2] = OctomapWorld::tl_steps() - s0; counters[3] = OctomapWorld::tl_rays() - r0;
  }
  return E;
}

// castRay batch.  hit[i] = 1 when castRay returned true; hit_keys = key it stopped in (or -1s);
// ends = the `end` out-parameter (voxel centre, float).
void orc_cast_rays(void* w, const float* origins, const float* dirs, long n, int ignore_unknown, const double* max_range,
                   uint8_t* hit, int32_t* hit_keys, float* ends) {
  OctomapWorld* W = (OctomapWorld*)w;
  for (long i = 0; i < n; ++i) {
    point3d o(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]);
    point3d d(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
    point3d e;
    OcTreeKey hk;
    bool h = W->tree.castRay(o, d, e, ignore_unknown != 0, max_range ? max_range[i] : -1.0, &hk);
    hit[i] = h ? 1 : 0;
    if (hit_keys) for (int a = 0; a < 3; ++a) hit_keys[3 * i + a] = h ? (int32_t)hk[a] : -1;
    if (ends) for (int a = 0; a < 3; ++a) ends[3 * i + a] = e(a);
  }
}

// host half of the frustum spec: (hfov_deg, vfov_deg, near, far) -> (vh, hw, near, far) floats
void orc_frustum_consts(double hfov_deg, double vfov_deg, double near_d, double far_d, float out[4]) {
  float vfov_rad = float(vfov_deg * M_PI / 180);
  float hfov_rad = float(hfov_deg * M_PI / 180);
  out[0] = float(2 * std::tan(vfov_rad / 2));
  out[1] = float(2 * std::tan(hfov_rad / 2));
  out[2] = (float)near_d;
  out[3] = (float)far_d;
}

void orc_frustum_planes(const double pose[7], const float fr[4], float planes[24]) {
  Frustum F;
  frustum_planes(pose, fr, F);
  for (int k = 0; k < 6; ++k) for (int j = 0; j < 4; ++j) planes[4 * k + j] = F.pl[k][j];
}

// Visibility spec v1 over all (viewpoint, point) pairs.  view_ptr has nv+1 entries; idx (ascending
// point index per viewpoint) filled up to cap; returns the total number of visible pairs.
// counters[0] = occlusion rays cast (in-frustum pairs), [1] = voxel lookups.
long orc_visibility(void* w, const double* poses, long nv, const float fr[4], const float* pts, long np,
                    int64_t* view_ptr, int32_t* idx, long cap, long* counters) {
  OctomapWorld* W = (OctomapWorld*)w;
  long total = 0, rays = 0, steps = 0;
  for (long v = 0; v < nv; ++v) {
    view_ptr[v] = total;
    Frustum F;
    frustum_planes(poses + 7 * v, fr, F);
    point3d o(F.T[0], F.T[1], F.T[2]);
    for (long j = 0; j < np; ++j) {
      const float* p = pts + 3 * j;
      if (!in_frustum(F, p)) continue;
      ++rays;
      point3d pp(p[0], p[1], p[2]);
      point3d dir = pp - o;
      point3d end;
      OcTreeKey hk, pk;
      bool hit = W->tree.castRay(o, dir, end, true, dir.norm(), &hk, &steps);
      bool visible;
      if (!hit) visible = true;
      else visible = W->tree.coordToKeyChecked(pp, pk) && pk == hk;
      if (visible) {
        if (total < cap) idx[total] = (int32_t)j;
        ++total;
      }
    }
  }
  view_ptr[nv] = total;
  if (counters) { counters[0] = rays; counters[1] = steps; }
  return total;
}

// SearchSpace::insertNode over a batch, in insertion order (src/searchspace.cpp:69-105): pose i is
// stored unless nodeExists (:164-175) finds a stored pose equal to it — samePosition, i.e. all seven
// components with fabs(a-b) < 1E-12 (include/sspp/utils.h:231-239,386-390).  nodeExists walks the list
// from the head, and the head is the LATEST stored node, so the node it returns is the matching
// stored pose with the largest insertion index.  keep[i] = 1 when pose i was stored; dup_of[i] = -1
// for stored poses, else the insertion index of the node insertNode returned.  Returns the number stored.
long orc_dedup_poses(const double* poses, long n, uint8_t* keep, int64_t* dup_of) {
  std::vector<long> stored;  // insertion indices, oldest first
  for (long i = 0; i < n; ++i) {
    const double* a = poses + 7 * i;
    long found = -1;
    for (long s = (long)stored.size() - 1; s >= 0 && found < 0; --s) {
      const double* b = poses + 7 * stored[s];
      bool same = true;
      for (int c = 0; c < 7 && same; ++c) same = std::fabs(a[c] - b[c]) < 1E-12;
      if (same) found = stored[s];
    }
    keep[i] = found < 0 ? 1 : 0;
    if (dup_of) dup_of[i] = found;
    if (found < 0) stored.push_back(i);
  }
  return (long)stored.size();
}

}  // extern "C"
