// connect_kernels.cuh — the connectNodes pass: node binning, in-radius pair count / fill, the edge kernel, CSR compaction, multi-GPU block assembly.
// Part of the single translation unit sspp_gpu.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

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

// One record per binned node, in cell order: position and list index together, 32 bytes, so that a
// candidate costs one aligned 32-byte read (two 128-bit loads) instead of four scattered ones.
struct __align__(32) NodeRec {
  double x, y, z;
  long long idx;  // list index of the node
};
__device__ __forceinline__ NodeRec load_node(const NodeRec* p) {  // read-only path: the loads may be hoisted over stores
  const double2 a = __ldg(reinterpret_cast<const double2*>(p));
  const double2 b = __ldg(reinterpret_cast<const double2*>(p) + 1);
  NodeRec r;
  r.x = a.x; r.y = a.y; r.z = b.x;
  r.idx = __double_as_longlong(b.y);
  return r;
}

__global__ void k_cell_scatter(const double* __restrict__ xyz, int64_t n, const int* __restrict__ cell_of, int* cursor,
                               NodeRec* __restrict__ sn) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || cell_of[i] < 0) return;
  int pos = atomicAdd(cursor + cell_of[i], 1);
  NodeRec r;
  r.x = xyz[3 * i]; r.y = xyz[3 * i + 1]; r.z = xyz[3 * i + 2];
  r.idx = i;
  sn[pos] = r;
}

// Order the members of every cell by list index, so the whole pipeline is run-to-run deterministic
// (the scatter above lands them in atomic order).  Cells hold a handful of nodes: insertion sort.
__global__ void k_cell_sort(const int* __restrict__ cell_start, int ncells, NodeRec* sn) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncells) return;
  int b = cell_start[c], e = cell_start[c + 1];
  for (int i = b + 1; i < e; ++i) {
    NodeRec k = sn[i];
    int j = i - 1;
    while (j >= b && sn[j].idx > k.idx) {
      sn[j + 1] = sn[j];
      --j;
    }
    sn[j + 1] = k;
  }
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

constexpr int SSPP_MAX_CHUNKS = 8;
struct ConnectArgs {
  NodeGrid g;
  const int* cell_start;
  const NodeRec* sn;  // binned nodes, cell order
  // Rows to process: list indices [lo, hi).  The row kernels run over the sorted positions [0, nscan) and skip
  // what is not a row of this shard: positions beyond the binned count (*n_sorted: a shard bins only the nodes
  // its rows can reach) and nodes outside [lo, hi).  nrows = hi - lo sizes every per-row array.
  const int* n_sorted;  // device: number of binned nodes, or NULL = nscan
  int nscan;
  int nrows;
  int64_t lo, hi;
  double r2max;  // largest d2 with sqrt(d2) <= radius
  int mode;
  int use_map;
  int use_pyramid;
  RowReach ts;
  // Pair storage.  slot > 0: row r owns the fixed slot [r*slot, r*slot + slot) of the pair buffers and
  // cand_count[r] says how many entries it holds — the counting pass writes the pairs itself, no scan
  // and no second neighbour pass; a row with more than `slot` pairs raises counters[8] and the host
  // repeats the pass with scanned offsets (slot == 0: row r owns [cand_off[r], cand_off[r+1])).
  int slot;
  const int* cand_count;
  // Row chunks (list-index ranges [chunk_lo[c], chunk_lo[c+1]) relative to `lo`): the edge test and the
  // CSR compaction can run chunk by chunk so that a finished chunk travels to the host while the next
  // one is computed.  nchunks == 1: one pass.  Chunk c keeps its work list at work_rows + chunk_lo[c]
  // (a chunk never has more work rows than rows), its length in counters[16 + c], its ticket in [32 + c].
  int nchunks;
  int chunk_lo[SSPP_MAX_CHUNKS + 1];
  // Exchange fused into the pass (sspp_gpu_connect_fused; slot mode only): every row of the shard goes through the edge
  // kernel, which stores the row's accepted columns, sorted, and its degree straight into the slot-format buffers of ALL
  // ranks (fslot[k] / fdeg[k], indexed by GLOBAL row: frow0 + local row) through NVLink peer pointers, row by row, while
  // the rest of the pass is still running.  Rows whose reach is free in the pyramid take the accept-all path there.
  unsigned ray_magic;  // 2^32 / rays + 1 when ray_stride > 1, else 0
  int ray_stride;  // order in which k_edges takes the rays of a pair: ray (t * ray_stride) % rays at turn position t; 1 = lattice order
  int fused, fworld;
  long long frow0;
  int* fslot[SSPP_MAX_PEERS];
  int* fdeg[SSPP_MAX_PEERS];
  const uint8_t* row_free_dev;
  // Diagnostics (sspp_gpu_connect_row_classes): how every row was answered — 0 no pair, 1 reach free in the
  // coarse pyramid (k_count), 3 pair by pair in k_edges (2 is reserved)
  uint8_t* row_class;
};

// where row r keeps its pairs
__device__ __forceinline__ void row_span(const ConnectArgs& a, const long long* __restrict__ cand_off, int r, long long& off,
                                         int& n) {
  if (a.slot) {
    off = (long long)r * a.slot;
    n = a.cand_count[r];
  } else {
    off = cand_off[r];
    n = (int)(cand_off[r + 1] - off);
  }
}
// true when the pair buffers do not hold this pass's pairs (too small / a slot overflowed): every
// consumer kernel then returns at once and the host repeats the phase
__device__ __forceinline__ bool pairs_unusable(const ConnectArgs& a, const long long* __restrict__ cand_off,
                                               const unsigned long long* counters, long long pair_cap) {
  return a.slot ? (counters[8] != 0) : (cand_off[a.nrows] > pair_cap);
}

// counters: [0] in-radius pairs, [1] tested pairs, [2] pyramid rows, [3] voxel steps (instrumented),
//           [4] next work row (edge kernel's ticket), [5] pairs accepted by their key box,
//           [6] pairs that ran the ray lattice, [8] rows that overflowed their slot,
//           [16 + c] work rows of chunk c, [32 + c] next work row of chunk c (edge kernel's ticket)
//
// One thread per source row walks the 27 neighbouring cells (9 contiguous ranges of the cell-sorted
// arrays).  FILL == false: count the in-radius pairs and test the row's reach against the pyramid;
// in slot mode also write the pairs into the row's slot (then no FILL pass runs).
// FILL == true (after the scan of the counts): write the pairs — (sorted position, list index with
// bit 31 = same position) — to cand[], in cell order; the edge kernel orders them by list index.
#ifndef COUNT_MIN_BLOCKS
#define COUNT_MIN_BLOCKS 5  // 48 registers; six (40) and eight (32) measured slower: 0.329 / 0.374 ms against 0.315
#endif
template <bool FILL>
__global__ void __launch_bounds__(256, COUNT_MIN_BLOCKS) k_count(ConnectArgs a, DeviceMap m, int* __restrict__ cand_count,
                                               uint8_t* __restrict__ row_free, unsigned long long* counters,
                                               const long long* __restrict__ cand_off, int2* __restrict__ cand,
                                               int* __restrict__ work_rows, int* __restrict__ deg, long long pair_cap) {
  if (FILL && cand_off[a.nrows] > pair_cap) return;  // the pair buffers are too small: the host grows them and reruns
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long inr = 0, tested = 0, pyr = 0;
  bool is_work = false;
  int chunk = 0;
  NodeRec me;
  me.idx = -1;
  if (t < a.nscan && (!a.n_sorted || t < *a.n_sorted)) me = load_node(a.sn + t);
  if (me.idx >= a.lo && me.idx < a.hi) {
    const int s = t;
    const double px = me.x, py = me.y, pz = me.z;
    int r = (int)(me.idx - a.lo);
    int cx = grid_cell_axis(a.g, px, 0), cy = grid_cell_axis(a.g, py, 1), cz = grid_cell_axis(a.g, pz, 2);
    int x0 = max(cx - 1, 0), x1 = min(cx + 1, a.g.n[0] - 1);
    int cnt = 0, tst = 0;
    long long w = FILL ? cand_off[r] : (long long)r * a.slot;
    const bool slots = !FILL && a.slot > 0;
    for (int zz = max(cz - 1, 0); zz <= min(cz + 1, a.g.n[2] - 1); ++zz)
      for (int yy = max(cy - 1, 0); yy <= min(cy + 1, a.g.n[1] - 1); ++yy) {
        int base = (zz * a.g.n[1] + yy) * a.g.n[0];
        int b = a.cell_start[base + x0], e = a.cell_start[base + x1 + 1];
        // two candidates per turn: both records are requested before either is looked at (the pass is bound by
        // the latency of these loads)
        for (int j = b; j < e; j += 2) {
          const bool two = j + 1 < e;
          const NodeRec o0 = load_node(a.sn + j);
          const NodeRec o1 = load_node(a.sn + (two ? j + 1 : j));
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            if (u == 1 && !two) break;
            const int jj = j + u;
            if (jj == s) continue;
            const NodeRec& o = u ? o1 : o0;
            int pc = pair_class(o.x, o.y, o.z, px, py, pz, a.r2max, a.mode);
            if (FILL) {
              if (pc) cand[w++] = make_int2(jj, (int)o.idx | (pc == 2 ? (int)0x80000000 : 0));
            } else {
              if (slots && pc && cnt < a.slot) cand[w + cnt] = make_int2(jj, (int)o.idx | (pc == 2 ? (int)0x80000000 : 0));
              cnt += (pc != 0);
              tst += (pc == 1);
            }
          }
        }
      }
    if (!FILL) {
      cand_count[r] = cnt;
      if (slots && cnt > a.slot) atomicAdd(counters + 8, 1ull);
      uint8_t fr = 0;
      if (a.use_map && a.use_pyramid && tst > 0) fr = row_reach_is_free(m, px, py, pz, a.ts.reach) ? 1 : 0;
      row_free[r] = fr;
      inr = cnt; tested = tst; pyr = fr;
      is_work = a.fused ? true : (a.use_map && !fr && cnt > 0);
      a.row_class[r] = cnt == 0 ? 0 : (fr || !a.use_map ? 1 : 3);
      while (chunk + 1 < a.nchunks && r >= a.chunk_lo[chunk + 1]) ++chunk;
      // slot mode: a row that needs no ray keeps all its pairs; k_compact orders them on the way out
      if (slots && !is_work) deg[r] = cnt;
    }
  }
  if (!FILL) {
    // rows that need the edge kernel, appended to their chunk's list, one atomic per warp and chunk
    // (the order inside a list does not matter)
    const unsigned grp = __match_any_sync(0xffffffffu, is_work ? chunk : -1);
    if (is_work) {
      const int lane = threadIdx.x & 31;
      const int leader = __ffs(grp) - 1;
      unsigned long long base = 0;
      if (lane == leader) base = atomicAdd(counters + 16 + chunk, (unsigned long long)__popc(grp));
      base = __shfl_sync(grp, base, leader);
      work_rows[a.chunk_lo[chunk] + base + __popc(grp & ((1u << lane) - 1u))] = t;
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
#ifndef SSPP_EDGE_WARPS
#define SSPP_EDGE_WARPS 8  // tuning switch: warps per CTA of the edge kernel
#endif
constexpr int EDGE_WARPS = SSPP_EDGE_WARPS;
#ifndef SSPP_PAIR_COARSE
#define SSPP_PAIR_COARSE 1  // tuning switch: 0 = pair boxes go to the fine plane without asking the 4x4x4 pyramid first
#endif

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
    uint32_t acc = 0;
    const uint32_t* col = m.blocked + map_word(m, w, y0, z0);
    for (int z = z0; z <= z1; ++z, col += m.dims[1]) {
      // the y run is a few consecutive words: eight predicated loads at a time, all in flight together
      for (int yb = 0; yb <= y1 - y0; yb += 8) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (yb + k <= y1 - y0) acc |= __ldg(col + yb + k);
      }
    }
    any |= acc & mask;
  }
  return any != 0;
}

// The same question asked of the 4x4x4 pyramid first: a box whose coarse cells are all clear holds no
// blocked voxel (a coarse bit is the OR of its 64 voxels), and that costs a few words instead of a few
// dozen.  "true" = some coarse cell is set or the box leaves the window: the fine scan decides.
__device__ __forceinline__ bool box_maybe_blocked_coarse(const DeviceMap& m, int lox, int hix, int loy, int hiy, int loz, int hiz) {
  const int x0 = lox - m.kmin[0], x1 = hix - m.kmin[0];
  const int y0 = loy - m.kmin[1], y1 = hiy - m.kmin[1];
  const int z0 = loz - m.kmin[2], z1 = hiz - m.kmin[2];
  if (x0 < 0 || y0 < 0 || z0 < 0 || x1 >= m.dims[0] || y1 >= m.dims[1] || z1 >= m.dims[2] || x0 > x1 || y0 > y1 || z0 > z1)
    return true;
  const int cx0 = x0 >> 2, cx1 = x1 >> 2, cy0 = y0 >> 2, cy1 = y1 >> 2, cz0 = z0 >> 2, cz1 = z1 >> 2;
  uint32_t any = 0;
  for (int w = cx0 >> 5; w <= (cx1 >> 5); ++w) {
    const int b0 = max(cx0, w * 32) & 31, b1 = min(cx1, w * 32 + 31) & 31;
    const uint32_t mask = (b1 == 31 ? 0xffffffffu : ((1u << (b1 + 1)) - 1u)) & ~((1u << b0) - 1u);
    uint32_t acc = 0;
    for (int z = cz0; z <= cz1; ++z) {
      const uint32_t* col = m.cblocked + ((size_t)w * m.cdims[2] + z) * m.cdims[1];
      for (int y = cy0; y <= cy1; ++y) acc |= __ldg(col + y);
    }
    any |= acc & mask;
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

// Turn table of the walk stage: with P pairs open a turn takes rays[P] = 32 / P rays of each; magic[P] = 2^16 / P + 1
// turns a lane number into (ray of the turn, pair): lane / P == (lane * magic[P]) >> 16 for lane < 32.
struct TurnTable {
  unsigned rays[33];
  unsigned magic[33];
};
constexpr TurnTable make_turn_table() {
  TurnTable t{};
  for (unsigned p = 1; p <= 32; ++p) {
    t.rays[p] = 32u / p;
    t.magic[p] = 65536u / p + 1u;
  }
  return t;
}
__constant__ TurnTable c_turn = make_turn_table();

// One ray of the lattice stage, out of line (the walk is the bulk of the edge kernel and inlining it would cost the
// kernel its occupancy).  okey: the ray's origin key, whose voxel the row stage has already found free (kx < 0: origin
// outside the key space — upstream leaves such a ray empty), or NULL = derive the key and test the start voxel here.
__device__ __noinline__ int ray_status(const DeviceMap& m, float ox, float oy, float oz, float ex, float ey, float ez, const int* okey) {
  BlockedGlobal cell;
  cell.m = &m;
  if (!okey) return line_status(m.ks, ox, oy, oz, ex, ey, ez, cell);
  if (okey[0] < 0) return CELL_FREE;
  return line_status_from(m.ks, ox, oy, oz, ex, ey, ez, okey[0], okey[1], okey[2], true, cell);
}

constexpr int RAY_TABLE_MAX = 256;
constexpr int START_RAYS_MAX = 128;  // lattices up to this many rays get the row-level start-voxel stage (27, 48, 100 ...)
inline size_t edge_okey_bytes(int rays) { return (size_t)EDGE_WARPS * 3 * sizeof(int) * (size_t)(rays < START_RAYS_MAX ? rays : START_RAYS_MAX); }
struct EdgeSmemSort {
  int2 buf[EDGE_WARPS][32];
};
struct EdgeSmem {
  RayLattice lat;
  double ray_off[RAY_TABLE_MAX][3];  // offset of ray q, when the lattice has <= RAY_TABLE_MAX rays
  EdgeSmemSort sort;
  double far_end[EDGE_WARPS][32][3];  // far-end coordinates of the batch's pairs
  int slots[EDGE_WARPS][32];          // batch slots that still need rays, compacted
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
    const int key = mine.y & 0x7fffffff;
    int rank = 0;
#pragma unroll 4
    for (int k = 0; k < n; ++k) rank += (__shfl_sync(0xffffffffu, key, k) < key);
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
                                                     int2* candB, int* __restrict__ acc, int* __restrict__ deg,
                                                     const unsigned long long* counters, long long pair_cap) {
  __shared__ EdgeSmemSort sb;
  if (pairs_unusable(a, cand_off, counters, pair_cap)) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x * EDGE_WARPS + warp;
  if (t >= a.nscan || (a.n_sorted && t >= *a.n_sorted)) return;
  const int s = t;
  const long long idx = a.sn[s].idx;
  if (idx < a.lo || idx >= a.hi) return;
  const int r = (int)(idx - a.lo);
  long long off;
  int n;
  row_span(a, cand_off, r, off, n);
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

// NW: 32-ray words of the start-voxel stage the kernel is compiled for (1: lattices up to 32 rays, the common
// case, with the mask in one register; 4: up to START_RAYS_MAX)
template <bool COUNT, int NW, bool FUSED>
__device__ __forceinline__ void edge_row(const ConnectArgs& a, const DeviceMap& m, EdgeSmem& sm, int* okey, const int warp, const int lane,
                                         const int t, const long long* __restrict__ cand_off,
                                         const int2* __restrict__ candA, int2* candB, int* __restrict__ acc,
                                         int* __restrict__ deg, unsigned long long* counters, unsigned& n_box, unsigned& n_lat, unsigned long long& n_edges) {
  const RayLattice& lat = sm.lat;
  const int s = t;  // work lists hold sorted positions
  const NodeRec me = load_node(a.sn + s);
  const int r = (int)(me.idx - a.lo);
  long long off;
  int n;
  row_span(a, cand_off, r, off, n);
  const int2 c0 = sort_row_pairs(sm.sort, warp, lane, candA, candB, off, n);

  const double px = me.x, py = me.y, pz = me.z;
  const int R = lat.rays;
  const bool accept_all = FUSED && (!a.use_map || a.row_free_dev[r] != 0);  // nothing to test: every pair is an edge
  const bool fast = a.use_pyramid != 0 && !accept_all;
  // Row-level verdicts: every ray starts at p + offset whatever the pair, and the start voxel is the
  // first one a walk tests.  A ray whose start voxel is blocked therefore rejects every pair of this
  // row for which that ray is not empty (its end key differs from its start key).
  unsigned startmask[NW];
#pragma unroll
  for (int c = 0; c < NW; ++c) startmask[c] = 0;
  bool any_start = false;
  const bool start_stage = !COUNT && fast && R <= NW * 32 && !accept_all;
  if (start_stage) {
#pragma unroll
    for (int c = 0; c < NW; ++c) {
      if (c * 32 >= R) break;
      const int q = c * 32 + lane;
      bool blk = false;
      if (q < R) {
        double ox, oy, oz;
        ray_offset(sm, q, ox, oy, oz);
        int kx = -1, ky = -1, kz = -1;
        if (coord_to_key(m.ks, (double)d2f(d_add(px, ox)), kx) && coord_to_key(m.ks, (double)d2f(d_add(py, oy)), ky) &&
            coord_to_key(m.ks, (double)d2f(d_add(pz, oz)), kz))
          blk = map_blocked(m, kx, ky, kz) != 0;
        else
          kx = -1;
        okey[3 * q] = kx; okey[3 * q + 1] = ky; okey[3 * q + 2] = kz;
      }
      startmask[c] = __ballot_sync(0xffffffffu, blk);
      any_start |= startmask[c] != 0;
    }
    __syncwarp();
  }
  int nacc = 0;
  unsigned long long nsteps = 0;
  for (int base = 0; base < n; base += 32) {
    const int nb = min(32, n - base);
    int2 c = (n <= 32) ? c0 : ((lane < nb) ? candB[off + base + lane] : make_int2(0, 0));
    const bool mine_valid = lane < nb;
    const bool same = mine_valid && (c.y & 0x80000000);
    // this lane's pair: coordinates of the far end
    double qx = 0, qy = 0, qz = 0;
    if (mine_valid) { const NodeRec o = load_node(a.sn + c.x); qx = o.x; qy = o.y; qz = o.z; }
    bool rejected = false;
    bool needs = mine_valid && !same && !accept_all;
    // stage 2a: a ray of this pair that starts in a blocked voxel and is not empty rejects the pair
    if (needs && any_start) {
#pragma unroll
      for (int c = 0; c < NW; ++c) {
        for (unsigned mb = startmask[c]; mb && needs; mb &= mb - 1) {
          const int q = c * 32 + __ffs(mb) - 1;
          double ox, oy, oz;
          ray_offset(sm, q, ox, oy, oz);
          int ekx, eky, ekz;
          if (coord_to_key(m.ks, (double)d2f(d_add(qx, ox)), ekx) && coord_to_key(m.ks, (double)d2f(d_add(qy, oy)), eky) &&
              coord_to_key(m.ks, (double)d2f(d_add(qz, oz)), ekz) &&
              (ekx != okey[3 * q] || eky != okey[3 * q + 1] || ekz != okey[3 * q + 2])) {
            needs = false;
            rejected = true;
          }
        }
      }
    }
    // stage 2b: the pair's key box (one lane per pair)
    if (!COUNT && fast && needs) {
      int lox, hix, loy, hiy, loz, hiz;
      if (pair_key_box(m.ks, px, qx, lat.off[0][0], lat.off[0][lat.n[0] - 1], lox, hix) &&
          pair_key_box(m.ks, py, qy, lat.off[1][0], lat.off[1][lat.n[1] - 1], loy, hiy) &&
          pair_key_box(m.ks, pz, qz, lat.off[2][0], lat.off[2][lat.n[2] - 1], loz, hiz) &&
          ((SSPP_PAIR_COARSE && !box_maybe_blocked_coarse(m, lox, hix, loy, hiy, loz, hiz)) || !box_maybe_blocked(m, lox, hix, loy, hiy, loz, hiz)))
        needs = false;
    }
    unsigned need = __ballot_sync(0xffffffffu, needs);
    unsigned accmask = __ballot_sync(0xffffffffu, mine_valid && !needs && !rejected);  // same-position or box-free pairs
    if (!accept_all) n_box += __popc(accmask & ~__ballot_sync(0xffffffffu, same));
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
      // stage 3: the exact walks.  The pairs still open advance through the ray lattice in lock step: every turn
      // takes the next k = 32 / (open pairs) rays of each of them (one (pair, ray) item per lane, walked by the exact
      // DDA), then the pairs a ray found blocked leave and the rest are packed again - the status of a pair is "any ray
      // not free" (src/distance_heuristic.cpp:104-111 stops at the first such ray), so the order of the rays is free,
      // and a blocked pair costs one turn instead of its whole lattice.  The rays are taken in a strided order
      // (a.ray_stride, coprime with the lattice size) so that the first turn samples the whole box rather than one
      // face of it.  (A key-box test per ray before the walk was measured to cost more than the walks it saves - S4a
      // 1.57 -> 1.36 ms without it - and is gone.)
      if (mine_valid) { sm.far_end[warp][lane][0] = qx; sm.far_end[warp][lane][1] = qy; sm.far_end[warp][lane][2] = qz; }
      const bool have_okey = start_stage;  // the start stage above ran for this row
      const unsigned smagic = a.ray_magic;  // 2^32 / R + 1 for (x * stride) % R (x * stride < 2^32 / R), 0 = lattice order
      unsigned open = need, blockedmask = 0;
      int q0 = 0;
      while (open) {
        const int P = __popc(open);
        __syncwarp();  // the previous turn has read the list
        if ((open >> lane) & 1u) sm.slots[warp][__popc(open & ((1u << lane) - 1u))] = lane;
        __syncwarp();
        const int k = min(R - q0, (int)c_turn.rays[P]);
        const unsigned pmagic = c_turn.magic[P];  // lane / P for lane < 32
        unsigned bit = 0;
        if (lane < P * k) {
          const int t = (int)(((unsigned)lane * pmagic) >> 16), kk = sm.slots[warp][lane - t * P];
          int q = q0 + t;
          if (smagic) {
            const unsigned x = (unsigned)q * (unsigned)a.ray_stride;
            q = (int)(x - (unsigned)R * __umulhi(x, smagic));
          }
          double ox, oy, oz;
          ray_offset(sm, q, ox, oy, oz);
          const double* f = sm.far_end[warp][kk];
          const int st = ray_status(m, d2f(d_add(px, ox)), d2f(d_add(py, oy)), d2f(d_add(pz, oz)), d2f(d_add(f[0], ox)),
                                    d2f(d_add(f[1], oy)), d2f(d_add(f[2], oz)), have_okey ? okey + 3 * q : nullptr);
          if (st != CELL_FREE) bit = 1u << kk;
        }
        const unsigned b = __reduce_or_sync(0xffffffffu, bit);
        blockedmask |= b;
        open &= ~b;
        q0 += k;
        if (q0 >= R) break;
      }
      __syncwarp();
      accmask |= need & ~blockedmask;
    }
    if ((accmask >> lane) & 1u) {
      const int pos = nacc + __popc(accmask & ((1u << lane) - 1u)), v = c.y & 0x7fffffff;
      acc[off + pos] = v;
      if (FUSED) {
        const long long g = (a.frow0 + r) * (long long)a.slot + pos;
        for (int k = 0; k < a.fworld; ++k) a.fslot[k][g] = v;
      }
    }
    nacc += __popc(accmask);
  }
  if (FUSED) {
    if (lane < a.fworld) a.fdeg[lane][a.frow0 + r] = nacc;
    n_edges += (unsigned)nacc;
  }
  if (lane == 0) {
    deg[r] = nacc;
  }
  if (COUNT) {
    for (int o = 16; o; o >>= 1) nsteps += __shfl_xor_sync(0xffffffffu, nsteps, o);
    if (lane == 0 && nsteps) atomicAdd(counters + 3, nsteps);
  }
  __syncwarp();
}

// Persistent edge kernel: a fixed grid (EDGE_MIN_BLOCKS CTAs per SM); every warp draws the next work
// row with a ticket, so a row that needs many walks does not hold seven finished warps hostage.
#ifndef EDGE_MIN_BLOCKS_SMALL
#define EDGE_MIN_BLOCKS_SMALL 6
#endif
// CTAs per SM: the kernel for lattices of up to 32 rays is compiled for six (40 registers; S4a 1.03 ms at four, 0.98 at
// five, 0.95 at six, 1.04 / 1.00 at seven / eight: the extra resident warps cover more of the load latency than the
// tighter register budget costs, until the spills take over); the kernels with the four-word start mask (S5: 9.52 ms
// at four, 9.60 / 9.57 at five / six; BASELINE configs[1] 0.655 / 0.673 / 0.694), the fused and the instrumented one
// keep four.
constexpr int edge_min_blocks(bool count, int nw, bool fused) { return (!count && nw == 1 && !fused) ? EDGE_MIN_BLOCKS_SMALL : EDGE_MIN_BLOCKS; }
template <bool COUNT, int NW, bool FUSED>
__global__ void __launch_bounds__(EDGE_WARPS * 32, edge_min_blocks(COUNT, NW, FUSED))
k_edges(ConnectArgs a, DeviceMap m, const RayLattice* __restrict__ lat_g, const long long* __restrict__ cand_off,
        const int* __restrict__ work_rows, const int2* __restrict__ candA, int2* candB, int* __restrict__ acc,
        int* __restrict__ deg, unsigned long long* counters, long long pair_cap, int chunk) {
  __shared__ EdgeSmem sm;
  // origin keys of the row's rays (they do not depend on the pair), 3 ints per ray and warp: dynamic, sized by the
  // lattice (edge_okey_bytes), so that the common 27-ray lattice leaves the space to L1
  extern __shared__ int okey_all[];
  if (pairs_unusable(a, cand_off, counters, pair_cap)) return;
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
  const unsigned long long n_work = counters[16 + chunk];
  unsigned n_box = 0, n_lat = 0;
  unsigned long long n_edges = 0;
  work_rows += a.chunk_lo[chunk];
  for (;;) {
    unsigned long long i = 0;
    if (lane == 0) i = atomicAdd(counters + 32 + chunk, 1ull);
    i = __shfl_sync(0xffffffffu, i, 0);
    if (i >= n_work) break;
    edge_row<COUNT, NW, FUSED>(a, m, sm, okey_all + warp * 3 * min(lat.rays, START_RAYS_MAX), warp, lane, work_rows[i], cand_off, candA,
                               candB, acc, deg, counters, n_box, n_lat, n_edges);
  }
  // statistics: one atomic per warp and kernel, not per row
  if (lane == 0) {
    if (n_box) atomicAdd(counters + 5, (unsigned long long)n_box);
    if (n_lat) atomicAdd(counters + 6, (unsigned long long)n_lat);
    if (FUSED && n_edges) atomicAdd(counters + 9, n_edges);  // [9] edges of this shard (no local CSR in fused mode)
  }
}

// CSR compaction: eight lanes per row (rows hold a dozen or two edges).  Work rows: coalesced copy of
// the accepted prefix k_edges left in acc[].  Slot mode, rows that needed no ray: all pairs of the slot
// are edges — rank-sorted by list index straight from the pair slot (the row never went through acc[]).
constexpr int COMPACT_LANES = 8;
// Rank sort of one row inside its eight lanes: every lane keeps its NI keys (d <= 8 NI; slot <= 64, so NI <= 8) in
// registers and the group passes all keys round by shuffle, so the slot is read once.  List indices are distinct:
// the rank of a key is the number of smaller keys.
template <int NI>
__device__ __forceinline__ void compact_rank_sort(const int2* __restrict__ pairs, int d, int sub, int* __restrict__ out) {
  const unsigned gmask = 0xffu << ((threadIdx.x & 31) & ~7);
  int keys[NI], rank[NI];
#pragma unroll
  for (int i = 0; i < NI; ++i) {
    const int e = sub + COMPACT_LANES * i;
    keys[i] = e < d ? (pairs[e].y & 0x7fffffff) : 0x7fffffff;
    rank[i] = 0;
  }
#pragma unroll
  for (int i = 0; i < NI; ++i) {
#pragma unroll
    for (int l = 0; l < COMPACT_LANES; ++l) {
      const int v = __shfl_sync(gmask, keys[i], l, COMPACT_LANES);
#pragma unroll
      for (int j = 0; j < NI; ++j) rank[j] += (v < keys[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < NI; ++j)
    if (sub + COMPACT_LANES * j < d) out[rank[j]] = keys[j];
}
__global__ void k_compact(ConnectArgs a, const long long* __restrict__ cand_off, const int* __restrict__ deg,
                          const long long* __restrict__ row_ptr, const int* __restrict__ acc, const int2* __restrict__ candA,
                          const uint8_t* __restrict__ row_free, int* __restrict__ col, const unsigned long long* counters,
                          long long pair_cap, int r_begin, int r_end) {
  if (pairs_unusable(a, cand_off, counters, pair_cap)) return;
  const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int r = r_begin + (int)(gt / COMPACT_LANES);
  const int sub = (int)(gt % COMPACT_LANES);
  if (r >= r_end) return;
  const long long src = a.slot ? (long long)r * a.slot : cand_off[r], dst = row_ptr[r];
  const int d = deg[r];
  if (a.slot && !(a.use_map && !row_free[r])) {
    switch ((d + COMPACT_LANES - 1) / COMPACT_LANES) {
      case 0: break;
      case 1: compact_rank_sort<1>(candA + src, d, sub, col + dst); break;
      case 2: compact_rank_sort<2>(candA + src, d, sub, col + dst); break;
      case 3: compact_rank_sort<3>(candA + src, d, sub, col + dst); break;
      case 4: compact_rank_sort<4>(candA + src, d, sub, col + dst); break;
      case 5: compact_rank_sort<5>(candA + src, d, sub, col + dst); break;
      case 6: compact_rank_sort<6>(candA + src, d, sub, col + dst); break;
      case 7: compact_rank_sort<7>(candA + src, d, sub, col + dst); break;
      default: compact_rank_sort<8>(candA + src, d, sub, col + dst); break;
    }
    return;
  }
  for (int e = sub; e < d; e += COMPACT_LANES) col[dst + e] = acc[src + e];
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

// Multi-GPU CSR assembly fused with its all-gather: this rank's block goes straight into the
// whole-graph CSR of EVERY rank through NVLink peer memory (plain stores to peer pointers), rebased
// on the way, so no gathered staging copy and no separate packing pass exist.  Every rank computes
// the same edge prefix from the same all-gathered counts, so all ranks take the same overflow
// decision without talking.  Column stores are 16 bytes wide, aligned on the destination (the
// destination offset ebase is the same for all peers, so one alignment fits all of them); peers are
// visited starting with the rank's right-hand neighbour so that the ranks do not all write to the
// same GPU at once.
struct PushPeers {
  long long* rp[SSPP_MAX_PEERS];
  int* col[SSPP_MAX_PEERS];
  int world, rank;
};

__global__ void __launch_bounds__(256)
k_push_csr(PushPeers p, const long long* __restrict__ edge_counts, const long long* __restrict__ row_base,
           const long long* __restrict__ rp, const int* __restrict__ col, long long col_capacity,
           long long* __restrict__ overflow) {
  long long ebase = 0, total = 0;
  for (int k = 0; k < p.world; ++k) {
    long long c = edge_counts[k];
    if (k < p.rank) ebase += c;
    total += c;
  }
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  if (total > col_capacity) {
    if (tid == 0) atomicMax((unsigned long long*)overflow, (unsigned long long)total);
    return;
  }
  const long long edges = edge_counts[p.rank];
  const long long r0 = row_base[p.rank], rows = row_base[p.rank + 1] - r0;
  for (long long i = tid; i < rows; i += nth) {
    long long v = rp[i] + ebase;
    for (int j = 0; j < p.world; ++j) {
      int k = p.rank + 1 + j;
      if (k >= p.world) k -= p.world;
      p.rp[k][r0 + i] = v;
    }
  }
  if (p.rank == p.world - 1 && tid == 0)
    for (int k = 0; k < p.world; ++k) p.rp[k][r0 + rows] = total;
  long long head = (4 - (ebase & 3)) & 3;
  if (head > edges) head = edges;
  const long long nvec = (edges - head) >> 2, tail0 = head + (nvec << 2);
  for (long long i = tid; i < head + (edges - tail0); i += nth) {
    long long e = i < head ? i : tail0 + (i - head);
    int v = col[e];
    for (int k = 0; k < p.world; ++k) p.col[k][ebase + e] = v;
  }
  for (long long q = tid; q < nvec; q += nth) {
    const int* s = col + head + (q << 2);
    int4 v = make_int4(s[0], s[1], s[2], s[3]);
    for (int j = 0; j < p.world; ++j) {
      int k = p.rank + 1 + j;
      if (k >= p.world) k -= p.world;
      *reinterpret_cast<int4*>(p.col[k] + ebase + head + (q << 2)) = v;
    }
  }
}


// Chunk-wise variant for the exchange that overlaps the edge test (sspp_gpu_connect_to_peers): one finished row
// chunk of this rank's block — rows [r0, r1) and the edges [chunk_base[c], chunk_base[c + 1]) the chunk scan left
// on the device — goes into slot [rank] of EVERY rank's block buffers (block-relative row pointers, columns at the
// same offset as in the local block), by plain stores through NVLink peer pointers.  Everything the kernel needs
// is in device memory, so the host queues it behind the chunk's compaction without waiting for anything; a few
// dozen CTAs are enough to keep the links busy, which lets it run beside the next chunk's edge kernel.  A chunk
// that would end beyond the slot (col_stride) is not sent: the host learns the block's size at the end of the
// pass and reports it.  `last`: the chunk also carries row_ptr[rows] = the block's edge count.
__global__ void __launch_bounds__(256)
k_push_chunk(PushPeers p, const long long* __restrict__ chunk_base, int c, int last, const long long* __restrict__ rp,
             const int* __restrict__ col, int r0, int r1, long long col_stride) {
  const long long e0 = chunk_base[c], e1 = chunk_base[c + 1];
  if (e1 > col_stride) return;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  const long long nrp = (long long)(r1 - r0) + (last ? 1 : 0);
  for (long long i = tid; i < nrp; i += nth) {
    const long long v = rp[r0 + i];
    for (int j = 0; j < p.world; ++j) {
      int k = p.rank + 1 + j;
      if (k >= p.world) k -= p.world;
      p.rp[k][r0 + i] = v;
    }
  }
  const long long edges = e1 - e0;
  long long head = (4 - (e0 & 3)) & 3;  // slots start 16-byte aligned and the offset e0 is the same in all of them
  if (head > edges) head = edges;
  const long long nvec = (edges - head) >> 2, tail0 = head + (nvec << 2);
  for (long long i = tid; i < head + (edges - tail0); i += nth) {
    const long long e = e0 + (i < head ? i : tail0 + (i - head));
    const int v = col[e];
    for (int k = 0; k < p.world; ++k) p.col[k][e] = v;
  }
  for (long long q = tid; q < nvec; q += nth) {
    const long long e = e0 + head + (q << 2);
    const int4 v = *reinterpret_cast<const int4*>(col + e);
    for (int j = 0; j < p.world; ++j) {
      int k = p.rank + 1 + j;
      if (k >= p.world) k -= p.world;
      *reinterpret_cast<int4*>(p.col[k] + e) = v;
    }
  }
}

// Row weights for the balanced partition: in-radius pairs, times `heavy` for rows that need the edge test.
__global__ void k_row_weights(const int* __restrict__ cand_count, const uint8_t* __restrict__ row_free, int use_map, int heavy,
                              int64_t n, int* __restrict__ w) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const int c = cand_count[r];
  w[r] = (use_map && !row_free[r]) ? c * heavy : c;
}
// cuts[g] = first row whose weight prefix reaches g/world of the total (prefix: n + 1 exclusive sums)
__global__ void k_find_cuts(const long long* __restrict__ prefix, int64_t n, int world, long long* __restrict__ cuts) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g > world) return;
  if (g == 0) { cuts[0] = 0; return; }
  if (g == world) { cuts[world] = n; return; }
  const long long total = prefix[n];
  const long long target = (long long)(((__int128)total * g) / world);
  int64_t lo = 0, hi = n;  // smallest r with prefix[r] >= target
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (prefix[mid] >= target) hi = mid; else lo = mid + 1;
  }
  cuts[g] = lo;
}

// After the fused exchange: the slot-format rows of ALL ranks (n rows of `slot` ints, deg[r] of them valid and sorted)
// become the whole-graph CSR — row_ptr from a scan of the degrees, eight lanes per row copy the columns.
__global__ void k_pack_slots(const int* __restrict__ slots, const int* __restrict__ deg, const long long* __restrict__ row_ptr,
                             long long n, int slot, int* __restrict__ col, long long col_capacity, long long* __restrict__ overflow) {
  if (row_ptr[n] > col_capacity) {
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicMax((unsigned long long*)overflow, (unsigned long long)row_ptr[n]);
    return;
  }
  const long long gt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long r = gt / COMPACT_LANES;
  const int sub = (int)(gt % COMPACT_LANES);
  if (r >= n) return;
  const int d = deg[r];
  const int* src = slots + r * (long long)slot;
  int* dst = col + row_ptr[r];
  for (int e = sub; e < d; e += COMPACT_LANES) dst[e] = src[e];
}

}  // namespace
