// Warp-tile helpers for the point-sorted observation array.
//
// A tile is a run of <=32 observations made of whole points (one lane per
// observation, segmented shuffle reductions keyed by point), or a single point
// with more than 32 observations (the warp strides over it).  All reductions
// use a fixed tree, so results are bit-reproducible run to run.
// This is the device counterpart of the per-point loops Ceres' SchurEliminator
// runs over e-block chunks for the problem built at
//   /root/reference/src/optim/bundle_adjustment.cc:326-530.
#pragma once
#include "common.cuh"

namespace sba {

struct Seg {
  int lane;
  int head;   // first lane of my segment
  int last;   // last lane of my segment
  bool is_head;
};

// key: point id of this lane's observation (unique negative value for idle lanes)
__device__ inline Seg MakeSeg(int key) {
  Seg s;
  s.lane = threadIdx.x & 31;
  const int prev = __shfl_up_sync(0xffffffffu, key, 1);
  s.is_head = (s.lane == 0) || (prev != key);
  const unsigned heads = __ballot_sync(0xffffffffu, s.is_head);
  const unsigned le = heads & (0xffffffffu >> (31 - s.lane));
  s.head = 31 - __clz(le);
  const unsigned above = (s.lane == 31) ? 0u : (heads >> (s.lane + 1));
  s.last = above ? (s.lane + __ffs(above) - 1) : 31;
  return s;
}

// sum over the segment, delivered to every lane of the segment
template <int N>
__device__ inline void SegSum(const Seg& s, double (&v)[N]) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const bool take = (s.lane + o) <= s.last;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const double t = __shfl_down_sync(0xffffffffu, v[i], o);
      if (take) v[i] += t;
    }
  }
#pragma unroll
  for (int i = 0; i < N; ++i) v[i] = __shfl_sync(0xffffffffu, v[i], s.head);
}

template <int N>
__device__ inline void WarpAllSum(double (&v)[N]) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
  }
}

// inverse of a symmetric positive definite 3x3 given as (c00,c01,c02,c11,c12,c22)
// through its Cholesky factor (what Eigen's LLT-based InvertPSDMatrix does on the
// Ceres side); returns false when not positive definite.
// `fac` (optional) receives M = L^{-1} (lower: m00, m10, m11, m20, m21, m22) with inv = M^T M -- the explicit Schur
// build multiplies by M^T once instead of by inv (schur.cu).
__device__ inline bool InvertSym3(const double c[6], double inv[6], double* fac = nullptr) {
  const double l00s = c[0];
  if (!(l00s > 0.0)) return false;
  const double l00 = sqrt(l00s), i00 = 1.0 / l00;
  const double l10 = c[1] * i00, l20 = c[2] * i00;
  const double l11s = c[3] - l10 * l10;
  if (!(l11s > 0.0)) return false;
  const double l11 = sqrt(l11s), i11 = 1.0 / l11;
  const double l21 = (c[4] - l20 * l10) * i11;
  const double l22s = c[5] - l20 * l20 - l21 * l21;
  if (!(l22s > 0.0)) return false;
  const double l22 = sqrt(l22s), i22 = 1.0 / l22;
  // M = L^{-1} (lower)
  const double m10 = -l10 * i00 * i11;
  const double m21 = -l21 * i11 * i22;
  const double m20 = -(l20 * i00 + l21 * m10) * i22;
  // inv = M^T M
  inv[0] = i00 * i00 + m10 * m10 + m20 * m20;
  inv[1] = m10 * i11 + m20 * m21;
  inv[2] = m20 * i22;
  inv[3] = i11 * i11 + m21 * m21;
  inv[4] = m21 * i22;
  inv[5] = i22 * i22;
  if (fac) { fac[0] = i00; fac[1] = m10; fac[2] = i11; fac[3] = m20; fac[4] = m21; fac[5] = i22; }
  return true;
}

__device__ inline void Sym3Mul(const double* __restrict__ s, const double u[3], double out[3]) {
  out[0] = s[0] * u[0] + s[1] * u[1] + s[2] * u[2];
  out[1] = s[1] * u[0] + s[3] * u[1] + s[4] * u[2];
  out[2] = s[2] * u[0] + s[4] * u[1] + s[5] * u[2];
}

}  // namespace sba
