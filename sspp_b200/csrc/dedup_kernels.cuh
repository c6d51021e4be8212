// dedup_kernels.cuh — SearchSpace::insertNode over a batch (reference src/searchspace.cpp:69-105 with
// nodeExists :164-175 and samePosition / isEqual, include/sspp/utils.h:386-390,231-239).
// Part of the single translation unit sspp_gpu.cu.
//
// The reference stores pose i unless an already STORED pose equals it in all seven components
// (fabs(a-b) < 1E-12 each); it finds that pose by walking the whole list, O(N) per insert.  The
// verdict of pose i therefore depends on the verdicts of the earlier poses it matches (the test is
// not transitive: a ~ b and b ~ c do not give a ~ c), which is the only sequential part.  Here:
//   1. the positions are binned into a uniform grid whose cell edge is far above the tolerance, so all
//      matches of a pose sit in its own or an adjacent cell (the node binning of the connection pass);
//   2. k_dedup_first: a pose with no earlier match is stored — final at once (almost every pose);
//      the others go on a pending list;
//   3. k_dedup_round, repeated until the list is resolved: a pending pose with a STORED earlier match is
//      refused; one whose earlier matches are all refused is stored; otherwise it waits.  The pending
//      pose with the smallest index always resolves, so the loop ends, and every verdict equals the
//      sequential one by induction over the insertion index;
//   4. k_dedup_final: for a refused pose, the node insertNode returns — nodeExists walks from the head
//      (the latest stored node), so it is the stored match with the LARGEST insertion index.
#pragma once

namespace {

enum { DD_UNKNOWN = 0, DD_STORED = 1, DD_REFUSED = 2 };

__global__ void k_pose_xyz(const double* __restrict__ poses, int64_t n, double* __restrict__ xyz) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  xyz[3 * i] = poses[7 * i];
  xyz[3 * i + 1] = poses[7 * i + 1];
  xyz[3 * i + 2] = poses[7 * i + 2];
}

// isEqual on the four orientation components (the position has been compared by the caller)
__device__ __forceinline__ bool same_orientation(const double* __restrict__ poses, long long i, long long j, double tol) {
  const double* a = poses + 7 * i + 3;
  const double* b = poses + 7 * j + 3;
  return fabs(d_sub(a[0], b[0])) < tol && fabs(d_sub(a[1], b[1])) < tol && fabs(d_sub(a[2], b[2])) < tol &&
         fabs(d_sub(a[3], b[3])) < tol;
}

// Visits every EARLIER pose equal to the pose at sorted position s; f(j) returns false to stop.
template <class F>
__device__ __forceinline__ void for_earlier_matches(const NodeGrid& g, const int* __restrict__ cell_start,
                                                    const NodeRec* __restrict__ sn, const double* __restrict__ poses, double tol,
                                                    const NodeRec& T, F f) {
  const int cx = grid_cell_axis(g, T.x, 0), cy = grid_cell_axis(g, T.y, 1), cz = grid_cell_axis(g, T.z, 2);
  for (int z = max(cz - 1, 0); z <= min(cz + 1, g.n[2] - 1); ++z)
    for (int y = max(cy - 1, 0); y <= min(cy + 1, g.n[1] - 1); ++y) {
      const int x0 = max(cx - 1, 0), x1 = min(cx + 1, g.n[0] - 1);
      const int row = (z * g.n[1] + y) * g.n[0];
      const int b = cell_start[row + x0], e = cell_start[row + x1 + 1];  // the x-run of cells is contiguous
      for (int q = b; q < e; ++q) {
        const NodeRec S = load_node(sn + q);
        if (S.idx >= T.idx) continue;
        if (!(fabs(d_sub(T.x, S.x)) < tol && fabs(d_sub(T.y, S.y)) < tol && fabs(d_sub(T.z, S.z)) < tol)) continue;
        if (!same_orientation(poses, T.idx, S.idx, tol)) continue;
        if (!f(S.idx)) return;
      }
    }
}

// counters: [0] pending poses appended, [1] poses still unknown after the current round
__global__ void k_dedup_first(NodeGrid g, const int* __restrict__ cell_start, const NodeRec* __restrict__ sn, int64_t n,
                              const double* __restrict__ poses, double tol, uint8_t* __restrict__ state, int* __restrict__ pend,
                              unsigned long long* counters) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const NodeRec T = load_node(sn + s);
  bool earlier = false;
  for_earlier_matches(g, cell_start, sn, poses, tol, T, [&](long long) { earlier = true; return false; });
  if (!earlier) {
    state[T.idx] = DD_STORED;
  } else {
    state[T.idx] = DD_UNKNOWN;
    pend[atomicAdd(counters, 1ull)] = (int)s;
  }
}

__global__ void k_dedup_round(NodeGrid g, const int* __restrict__ cell_start, const NodeRec* __restrict__ sn,
                              const double* __restrict__ poses, double tol, volatile uint8_t* state, const int* __restrict__ pend,
                              int64_t npend, unsigned long long* counters) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= npend) return;
  const NodeRec T = load_node(sn + pend[t]);
  if (state[T.idx] != DD_UNKNOWN) return;
  bool stored_match = false, waiting = false;
  for_earlier_matches(g, cell_start, sn, poses, tol, T, [&](long long j) {
    const int sj = state[j];  // a verdict written during this round is final, so reading it early is harmless
    if (sj == DD_STORED) { stored_match = true; return false; }
    if (sj == DD_UNKNOWN) waiting = true;
    return true;
  });
  if (stored_match) state[T.idx] = DD_REFUSED;
  else if (!waiting) state[T.idx] = DD_STORED;
  else atomicAdd(counters + 1, 1ull);
}

__global__ void k_dedup_init_dup(int64_t n, long long* __restrict__ dup_of) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dup_of[i] = -1;
}

__global__ void k_dedup_final(NodeGrid g, const int* __restrict__ cell_start, const NodeRec* __restrict__ sn,
                              const double* __restrict__ poses, double tol, const uint8_t* __restrict__ state,
                              const int* __restrict__ pend, int64_t npend, long long* __restrict__ dup_of) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= npend) return;
  const NodeRec T = load_node(sn + pend[t]);
  if (state[T.idx] != DD_REFUSED) return;
  long long latest = -1;
  for_earlier_matches(g, cell_start, sn, poses, tol, T, [&](long long j) {
    if (state[j] == DD_STORED && j > latest) latest = j;
    return true;
  });
  dup_of[T.idx] = latest;
}

__global__ void k_dedup_keep(const uint8_t* __restrict__ state, int64_t n, uint8_t* __restrict__ keep, unsigned long long* counters) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool k = i < n && state[i] == DD_STORED;
  if (i < n) keep[i] = k ? 1 : 0;
  const unsigned b = __ballot_sync(0xffffffffu, k);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(counters + 2, (unsigned long long)__popc(b));
}

}  // namespace
