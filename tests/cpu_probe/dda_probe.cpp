// CPU build of the engine's own voxel-walk header (sspp_b200/csrc/dda_core.cuh), for tests only:
// lets `-m "not gpu"` tests check the arithmetic the kernels run against the oracle without a GPU.
// It is never linked into the product library and is not a fallback path.
#include <cstdint>
#include <cstring>

#include "../../sspp_b200/csrc/dda_core.cuh"

namespace {
struct Dense {
  const uint32_t* occ;
  const uint32_t* known;
  int kmin[3], dims[3], wpr, treat;
  int status(int kx, int ky, int kz) const {
    int x = kx - kmin[0], y = ky - kmin[1], z = kz - kmin[2];
    if (x < 0 || y < 0 || z < 0 || x >= dims[0] || y >= dims[1] || z >= dims[2]) return sspp::CELL_UNKNOWN;
    size_t w = ((size_t)z * dims[1] + y) * wpr + (x >> 5);
    uint32_t bit = 1u << (x & 31);
    if (!(known[w] & bit)) return sspp::CELL_UNKNOWN;
    return (occ[w] & bit) ? sspp::CELL_OCCUPIED : sspp::CELL_FREE;
  }
};
struct LineCell {
  const Dense* d;
  int operator()(int kx, int ky, int kz) const {
    int st = d->status(kx, ky, kz);
    if (st == sspp::CELL_UNKNOWN && !d->treat) return sspp::CELL_FREE;
    return st;
  }
};
struct RayCell {
  const Dense* d;
  int operator()(int kx, int ky, int kz) const { return d->status(kx, ky, kz); }
};
sspp::KeySpace keyspace(double res) {
  sspp::KeySpace ks;
  ks.res = res;
  ks.res_factor = 1.0 / res;
  ks.half_res_f = (float)(res * 0.5);
  ks.half_res_d = res * 0.5;
  return ks;
}
Dense dense(const int32_t kmin[3], const int32_t dims[3], const uint32_t* occ, const uint32_t* known, int treat) {
  Dense d;
  d.occ = occ; d.known = known; d.treat = treat;
  for (int a = 0; a < 3; ++a) { d.kmin[a] = kmin[a]; d.dims[a] = dims[a]; }
  d.wpr = (dims[0] + 31) / 32;
  return d;
}
}  // namespace

extern "C" {

int probe_ray_lattice(double size, double res, double* out, int cap) { return sspp::ray_lattice_axis(size, res, out, cap); }

void probe_line_status_bbox(double res, const int32_t kmin[3], const int32_t dims[3], const uint32_t* occ,
                            const uint32_t* known, int treat, const double* starts, const double* ends, long n,
                            const double box[3], uint8_t* status, long* steps_out) {
  sspp::KeySpace ks = keyspace(res);
  Dense d = dense(kmin, dims, occ, known, treat);
  LineCell cell;
  cell.d = &d;
  double off[3][64];
  int cnt[3];
  for (int a = 0; a < 3; ++a) cnt[a] = sspp::ray_lattice_axis(box[a], res, off[a], 64);
  long steps = 0;
  for (long i = 0; i < n; ++i) {
    int st = sspp::CELL_FREE;
    for (int ix = 0; ix < cnt[0] && !st; ++ix)
      for (int iy = 0; iy < cnt[1] && !st; ++iy)
        for (int iz = 0; iz < cnt[2] && !st; ++iz) {
          int s = 0;
          st = sspp::line_status(ks, (float)(starts[3 * i] + off[0][ix]), (float)(starts[3 * i + 1] + off[1][iy]),
                                 (float)(starts[3 * i + 2] + off[2][iz]), (float)(ends[3 * i] + off[0][ix]),
                                 (float)(ends[3 * i + 1] + off[1][iy]), (float)(ends[3 * i + 2] + off[2][iz]), cell, &s);
          steps += s;
        }
    status[i] = (uint8_t)st;
  }
  if (steps_out) *steps_out = steps;
}

void probe_cast_rays(double res, const int32_t kmin[3], const int32_t dims[3], const uint32_t* occ, const uint32_t* known,
                     const float* org, const float* dir, long n, int ignore_unknown, const double* max_range, uint8_t* hit,
                     int32_t* hit_keys) {
  sspp::KeySpace ks = keyspace(res);
  Dense d = dense(kmin, dims, occ, known, 0);
  RayCell cell;
  cell.d = &d;
  for (long i = 0; i < n; ++i) {
    int hx, hy, hz;
    bool h = sspp::cast_ray(ks, org[3 * i], org[3 * i + 1], org[3 * i + 2], dir[3 * i], dir[3 * i + 1], dir[3 * i + 2],
                            ignore_unknown != 0, max_range ? max_range[i] : -1.0, cell, hx, hy, hz);
    hit[i] = h ? 1 : 0;
    hit_keys[3 * i] = h ? hx : -1;
    hit_keys[3 * i + 1] = h ? hy : -1;
    hit_keys[3 * i + 2] = h ? hz : -1;
  }
}

}  // extern "C"
