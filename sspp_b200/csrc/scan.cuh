// scan.cuh — exclusive scan used for cell starts, pair offsets and CSR row pointers.
// Part of the single translation unit sspp_gpu.cu (included there, in order); not a stand-alone header.
#pragma once

namespace {

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

// base (may be NULL): a device scalar added to every output; total_out (may be NULL) receives out[n]
template <class T>
__global__ void k_scan_apply(const int* __restrict__ in, int64_t n, const long long* __restrict__ sums, T* __restrict__ out,
                             const long long* __restrict__ base, long long* total_out) {
  // items are laid out thread-major inside the tile so the local scan is a plain running sum
  __shared__ long long sh[SCAN_THREADS];
  int64_t first = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
  int v[SCAN_ITEMS];
  long long t = 0;
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    int64_t i = first + k;
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
  long long run = sums[blockIdx.x] + sh[threadIdx.x] - t + (base ? *base : 0);
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    int64_t i = first + k;
    if (i < n) out[i] = (T)run;
    run += v[k];
    if (i == n - 1) {
      out[n] = (T)run;
      if (total_out) *total_out = run;
    }
  }
}

template <class T>
__global__ void k_scan_empty(T* out, const long long* base, long long* total_out) {
  const long long b = base ? *base : 0;
  out[0] = (T)b;
  if (total_out) *total_out = b;
}

// out must hold n+1 entries; out[n] = total.  With `base` (a device scalar) the scan starts there
// instead of at 0, and total_out receives out[n]: a chain of scans over consecutive pieces of one array.
template <class T>
int scan_exclusive(sspp_gpu_ctx* ctx, const int* in, int64_t n, T* out, const long long* base = nullptr,
                   long long* total_out = nullptr) {
  if (n == 0) {
    k_scan_empty<T><<<1, 1, 0, ctx->stream>>>(out, base, total_out);
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
  k_scan_apply<T><<<nb, SCAN_THREADS, 0, ctx->stream>>>(in, n, sums, out, base, total_out);
  LAUNCHED(ctx);
  return SSPP_OK;
}


}  // namespace
