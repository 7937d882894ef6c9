// Dense-Jacobian linearisation of reprojection blocks over (pose a, pose b, camera, point) -- the block kind of the rig
// solve (rig_solve.cu), also used for the landmark blocks of the cylinder solve (cylinder_solve.cu).  Kernels are
// `static` / templates: every translation unit that includes this header gets its own copy (no relocatable device code).
#pragma once
#include "dense_lm.h"
#include "rig_device.cuh"

namespace sba {

constexpr int64_t kRigMaxJacDoubles = (int64_t)1 << 28;  // 2 GiB of dense Jacobian

struct RigView {
  int n_pose, n_cam, n_pt, n;
  int64_t n_obs;
  const double *pq, *pt, *cam, *X;  // poses q [n_pose*4], t [n_pose*3]; camera params [n_cam*12]; points [n_pt*3]
  const int* model;
  const int *qcol, *tcol, *pcol, *xcol;  // tangent columns: q start [n_pose], t [n_pose*3], params [n_cam*12], X start [n_pt]; -1 constant
  const int *b_a, *b_b, *b_c, *b_p;      // per block: rig pose, relative pose (both into the pose arrays), camera, point
  const double* b_xy;
  double weight;  // ScaledLoss(loss, weight): rho, rho', rho'' scaled (1 for the rig solve)
};

static __device__ inline void RigLoss(int type, double a, double s, double rho[3]) { LossEval(type, a, s, rho); }

// residuals (and the loss-corrected tangent Jacobian rows, written into the dense J) of every block; per-CTA cost partials
template <bool JAC>
__global__ void __launch_bounds__(128) rig_lin_kernel(RigView V, LossView L, double* __restrict__ Jf, double* __restrict__ rf,
                                                      double* __restrict__ cost_partial) {
  const int64_t k = (int64_t)blockIdx.x * 128 + threadIdx.x;
  double cost = 0.0;
  if (k < V.n_obs) {
    const int ia = V.b_a[k], ib = V.b_b[k], ic = V.b_c[k], ip = V.b_p[k];
    double w[4], z[4], tg[3], tl[3], X[3], prm[12], r[2], J[2 * kRigJac];
#pragma unroll
    for (int j = 0; j < 4; ++j) { w[j] = V.pq[4 * (size_t)ia + j]; z[j] = V.pq[4 * (size_t)ib + j]; }
#pragma unroll
    for (int j = 0; j < 3; ++j) { tg[j] = V.pt[3 * (size_t)ia + j]; tl[j] = V.pt[3 * (size_t)ib + j]; X[j] = V.X[3 * (size_t)ip + j]; }
#pragma unroll
    for (int j = 0; j < 12; ++j) prm[j] = V.cam[12 * (size_t)ic + j];
    const double xy[2] = {V.b_xy[2 * k], V.b_xy[2 * k + 1]};
    RigBlock(w, tg, z, tl, X, prm, V.model[ic], xy, r, JAC ? J : nullptr);
    const double s = r[0] * r[0] + r[1] * r[1];
    double rho[3];
    RigLoss(L.type, L.scale, s, rho);
    rho[0] *= V.weight; rho[1] *= V.weight; rho[2] *= V.weight;
    cost = 0.5 * rho[0];
    if (JAC) {
      // corrector (ceres corrector.cc), as csrc/eval.cu
      double sqrt_rho1 = 1.0, residual_scaling = 1.0, alpha_sq_norm = 0.0;
      if (L.apply) {
        sqrt_rho1 = sqrt(rho[1]);
        if (s == 0.0 || rho[2] <= 0.0) {
          residual_scaling = sqrt_rho1;
        } else {
          const double Dd = 1.0 + 2.0 * s * rho[2] / rho[1];
          const double alpha = 1.0 - sqrt(Dd);
          residual_scaling = sqrt_rho1 / (1.0 - alpha);
          alpha_sq_norm = alpha / s;
        }
      }
      double* row0 = Jf + (size_t)(2 * k) * V.n;
      double* row1 = row0 + V.n;
      auto put = [&](int col, double j0, double j1) {
        if (col < 0) return;
        if (alpha_sq_norm != 0.0) {
          const double rtj = r[0] * j0 + r[1] * j1;
          j0 -= alpha_sq_norm * r[0] * rtj;
          j1 -= alpha_sq_norm * r[1] * rtj;
        }
        row0[col] = sqrt_rho1 * j0; row1[col] = sqrt_rho1 * j1;
      };
      const double* A = J;            // row 0
      const double* B = J + kRigJac;  // row 1
      double l0[3], l1[3];
      if (V.qcol[ia] >= 0) {
        QuatTangent(w, A, l0); QuatTangent(w, B, l1);
        for (int j = 0; j < 3; ++j) put(V.qcol[ia] + j, l0[j], l1[j]);
      }
      for (int j = 0; j < 3; ++j) put(V.tcol[3 * ia + j], A[4 + j], B[4 + j]);
      if (V.qcol[ib] >= 0) {
        QuatTangent(z, A + 7, l0); QuatTangent(z, B + 7, l1);
        for (int j = 0; j < 3; ++j) put(V.qcol[ib] + j, l0[j], l1[j]);
      }
      for (int j = 0; j < 3; ++j) put(V.tcol[3 * ib + j], A[11 + j], B[11 + j]);
      if (V.xcol[ip] >= 0)
        for (int j = 0; j < 3; ++j) put(V.xcol[ip] + j, A[14 + j], B[14 + j]);
      for (int j = 0; j < 12; ++j) put(V.pcol[12 * ic + j], A[17 + j], B[17 + j]);
      rf[2 * k] = r[0] * residual_scaling;
      rf[2 * k + 1] = r[1] * residual_scaling;
    }
  }
  __shared__ double sm[128];
  sm[threadIdx.x] = cost;
  __syncthreads();
  for (int o = 64; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) cost_partial[blockIdx.x] = sm[0];
}

// out[c] = sum_b partial[b * stride + c], c < count: one CTA, fixed order
static __global__ void __launch_bounds__(256) rig_sum_kernel(const double* __restrict__ partial, int n_part, int stride, int count,
                                                      double* __restrict__ out) {
  __shared__ double sm[256];
  for (int c = 0; c < count; ++c) {
    double a = 0.0;
    for (int b = threadIdx.x; b < n_part; b += 256) a += partial[(size_t)b * stride + c];
    sm[threadIdx.x] = a;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) out[c] = sm[0];
    __syncthreads();
  }
}

// lower triangle of H = J^T J, one 32 x 32 tile per CTA, rows consumed in order
static __global__ void __launch_bounds__(1024) rig_jtj_kernel(int n, int64_t m, const double* __restrict__ J, double* __restrict__ H,
                                                              int accumulate) {
  if (blockIdx.x > blockIdx.y) return;  // tile (row block y, column block x), x <= y
  __shared__ double As[32][33], Bs[32][33];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int i = blockIdx.y * 32 + ty, j = blockIdx.x * 32 + tx;
  const int ai = blockIdx.y * 32 + tx;  // column of J this thread stages for the i side
  double acc = 0.0;
  for (int64_t r0 = 0; r0 < m; r0 += 32) {
    const int64_t r = r0 + ty;
    As[ty][tx] = (r < m && ai < n) ? J[(size_t)r * n + ai] : 0.0;
    Bs[ty][tx] = (r < m && j < n) ? J[(size_t)r * n + j] : 0.0;
    __syncthreads();
#pragma unroll 8
    for (int q = 0; q < 32; ++q) acc += As[q][ty] * Bs[q][tx];
    __syncthreads();
  }
  if (i < n && j < n && j <= i) H[(size_t)i * n + j] = accumulate ? H[(size_t)i * n + j] + acc : acc;
}

// g = J^T r
static __global__ void rig_jtr_kernel(int n, int64_t m, const double* __restrict__ J, const double* __restrict__ r, double* __restrict__ g,
                                      int accumulate) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  double a = 0.0;
  for (int64_t q = 0; q < m; ++q) a += J[(size_t)q * n + j] * r[q];
  g[j] = accumulate ? g[j] + a : a;
}


}  // namespace sba
