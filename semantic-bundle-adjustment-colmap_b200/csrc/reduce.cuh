// Deterministic (fixed-tree) block / grid reductions shared by the kernels.
#pragma once
#include "common.cuh"

namespace sba {

__device__ inline double WarpSum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

// block sum of two values -> out[2*blockIdx.x + {0,1}]
template <int THREADS>
__device__ inline void BlockSum2(double a, double b, double* out) {
  __shared__ double sa[THREADS / 32], sb[THREADS / 32];
  a = WarpSum(a); b = WarpSum(b);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) { sa[w] = a; sb[w] = b; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ta = 0.0, tb = 0.0;
#pragma unroll
    for (int i = 0; i < THREADS / 32; ++i) { ta += sa[i]; tb += sb[i]; }
    out[2 * blockIdx.x] = ta; out[2 * blockIdx.x + 1] = tb;
  }
}

// Grid-wide sum of N values per thread without a second launch: every block writes its partial sums, takes a
// ticket, and the block that draws the last ticket adds all partials in a FIXED order (strided per thread,
// then the same shuffle tree) -- the result does not depend on which block happened to finish last.
// out[i] may be null.  `ticket` must be zero on entry and is left at zero.
template <int THREADS, int N>
__device__ inline void GridSum(double (&v)[N], double* __restrict__ partial, unsigned* ticket, double* const (&out)[N]) {
  __shared__ double sm[THREADS / 32][N];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < N; ++i) { const double t = WarpSum(v[i]); if (lane == 0) sm[w][i] = t; }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < N; ++i) {
      double t = 0.0;
#pragma unroll
      for (int k = 0; k < THREADS / 32; ++k) t += sm[k][i];
      partial[(size_t)N * blockIdx.x + i] = t;
    }
    __threadfence();
    is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  double acc[N];
#pragma unroll
  for (int i = 0; i < N; ++i) acc[i] = 0.0;
  for (unsigned b = threadIdx.x; b < gridDim.x; b += THREADS) {
#pragma unroll
    for (int i = 0; i < N; ++i) acc[i] += __ldcg(partial + (size_t)N * b + i);
  }
#pragma unroll
  for (int i = 0; i < N; ++i) { const double t = WarpSum(acc[i]); if (lane == 0) sm[w][i] = t; }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < N; ++i) {
      double t = 0.0;
#pragma unroll
      for (int k = 0; k < THREADS / 32; ++k) t += sm[k][i];
      if (out[i]) *out[i] = t;
    }
    *ticket = 0u;
  }
}

// fixed-order final reduction of per-block pairs -> out0 (sum of [2i]), out1 (sum of [2i+1])
static __global__ void __launch_bounds__(1024) finalize_pair_kernel(const double* __restrict__ partial, int n,
                                                                    double* out0, double* out1, double sign0) {
  __shared__ double sa[32], sb[32];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) { a += partial[2 * i]; b += partial[2 * i + 1]; }
  a = WarpSum(a); b = WarpSum(b);
  if ((threadIdx.x & 31) == 0) { sa[threadIdx.x >> 5] = a; sb[threadIdx.x >> 5] = b; }
  __syncthreads();
  if (threadIdx.x < 32) {
    a = sa[threadIdx.x]; b = sb[threadIdx.x];
    a = WarpSum(a); b = WarpSum(b);
    if (threadIdx.x == 0) { if (out0) *out0 = sign0 * a; if (out1) *out1 = b; }
  }
}

}  // namespace sba
