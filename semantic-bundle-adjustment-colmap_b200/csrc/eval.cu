// K-EVAL: per-observation reprojection residual + analytic Jacobian kernel (sm_100a, fp64).
//
// One thread per observation (point-sorted order).  What is computed is the
// reference functor
//   BundleAdjustmentCostFunction<CM>::operator()           /root/reference/src/base/cost_functions.h:57-81
//   BundleAdjustmentConstantPoseCostFunction<CM>           /root/reference/src/base/cost_functions.h:115-140
// followed by what Ceres does to the block before handing it to the linear
// solver: J_local = J_ambient * PlusJacobian
// (QuaternionManifold / SubsetManifold, cost_functions.h:270-287), loss corrector
// (bundle_adjustment.cc:53-68) and Jacobi column scaling.
//
// HBM traffic per observation (algorithmic): 8 B indices + 16 B xy in,
// 64 B ER record + 16*DC B F record out (tile-blocked 32 B units, common.cuh); poses / intrinsics /
// points are gathered through L2 (n_img*64 B + n_cam*96 B + n_pt*24 B per evaluation).
#include <algorithm>
#include <cstdlib>
#include <mutex>
#include <string>

#include "device_math.cuh"
#include "reduce.cuh"
#include "solver.h"

namespace sba {

constexpr int kEvalThreads = 128;

// Extension X1 (no reference implementation): the third residual row
//   r_s = weight * bilinear(map_of_image, x, y)
// sampled at the projection (x, y) with the convention of Bitmap::InterpolateBilinear
// (/root/reference/src/util/bitmap.cc:238-269: FreeImage scanlines are bottom-up, x0 = floor(x), the sample is
// rejected when a neighbour falls outside the image).  Its derivative w.r.t. (x, y) is that of the bilinear
// interpolant inside the cell -- differences of the four texels -- and the row then shares the chain rule of the
// reprojection rows: d r_s / d s = (g_x, g_y) . d (x, y) / d s.  A rejected sample contributes r_s = 0, J = 0.
__device__ __forceinline__ bool BilinearWithGrad(const float* __restrict__ map, int H, int W, double x, double y,
                                                 double* val, double* gx, double* gy) {
  const double inv_y = H - 1 - y;
  const double fx = floor(x), fy = floor(inv_y);
  // (NaN fails every comparison: rejected)
  if (!(fx >= 0.0 && fx + 1.0 < (double)W && fy >= 0.0 && fy + 1.0 < (double)H)) return false;
  const int x0 = (int)fx, y0 = (int)fy;
  const double dx = x - x0, dy = inv_y - y0, dx_1 = 1 - dx, dy_1 = 1 - dy;
  const int row0 = H - 1 - y0, row1 = row0 - 1;  // scanline y0 + 1 is one row UP in the top-down map
  const int tpr = (W + kMapTileW - 1) / kMapTileW;
  const double a0 = __ldg(map + MapTexel(row0, x0, tpr)), a1 = __ldg(map + MapTexel(row0, x0 + 1, tpr)),
               b0 = __ldg(map + MapTexel(row1, x0, tpr)), b1 = __ldg(map + MapTexel(row1, x0 + 1, tpr));
  const double v0 = dx_1 * a0 + dx * a1;
  const double v1 = dx_1 * b0 + dx * b1;
  *val = dy_1 * v0 + dy * v1;
  *gx = dy_1 * (a1 - a0) + dy * (b1 - b0);
  *gy = -(v1 - v0);  // d / d y = - d / d inv_y
  return true;
}

struct ObsGeom {
  double p[3];      // R(q) X + t
  double R[9];      // d p / d X  (the un-normalised rotation polynomial)
  double G[9];      // d p / d (quaternion tangent), row-major 3x3
};

// ceres::UnitQuaternionRotatePoint polynomial (no normalisation) and its
// derivatives, with the QuaternionManifold PlusJacobian folded in.
__device__ inline void RotateWithJac(const double q[4], const double t[3], const double X[3], ObsGeom* g,
                                     bool need_jac) {
  const double q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
  const double t2 = q0 * q1, t3 = q0 * q2, t4 = q0 * q3, t5 = -q1 * q1, t6 = q1 * q2, t7 = q1 * q3,
               t8 = -q2 * q2, t9 = q2 * q3, t1 = -q3 * q3;
  g->p[0] = 2.0 * ((t8 + t1) * X[0] + (t6 - t4) * X[1] + (t3 + t7) * X[2]) + X[0] + t[0];
  g->p[1] = 2.0 * ((t4 + t6) * X[0] + (t5 + t1) * X[1] + (t9 - t2) * X[2]) + X[1] + t[1];
  g->p[2] = 2.0 * ((t7 - t3) * X[0] + (t2 + t9) * X[1] + (t5 + t8) * X[2]) + X[2] + t[2];
  if (!need_jac) return;
  g->R[0] = 1.0 + 2.0 * (t8 + t1); g->R[1] = 2.0 * (t6 - t4);       g->R[2] = 2.0 * (t3 + t7);
  g->R[3] = 2.0 * (t4 + t6);       g->R[4] = 1.0 + 2.0 * (t5 + t1); g->R[5] = 2.0 * (t9 - t2);
  g->R[6] = 2.0 * (t7 - t3);       g->R[7] = 2.0 * (t2 + t9);       g->R[8] = 1.0 + 2.0 * (t5 + t8);
  // d p / d q (3x4), columns q0..q3
  const double X0 = X[0], X1 = X[1], X2 = X[2];
  const double D[12] = {
      2.0 * (-q3 * X1 + q2 * X2), 2.0 * (q2 * X1 + q3 * X2),             2.0 * (-2.0 * q2 * X0 + q1 * X1 + q0 * X2), 2.0 * (-2.0 * q3 * X0 - q0 * X1 + q1 * X2),
      2.0 * (q3 * X0 - q1 * X2),  2.0 * (q2 * X0 - 2.0 * q1 * X1 - q0 * X2), 2.0 * (q1 * X0 + q3 * X2),              2.0 * (q0 * X0 - 2.0 * q3 * X1 + q2 * X2),
      2.0 * (-q2 * X0 + q1 * X1), 2.0 * (q3 * X0 + q0 * X1 - 2.0 * q1 * X2), 2.0 * (-q0 * X0 + q3 * X1 - 2.0 * q2 * X2), 2.0 * (q1 * X0 + q2 * X1)};
  // PlusJacobian rows (w,x,y,z): [-q1,-q2,-q3],[q0,q3,-q2],[-q3,q0,q1],[q2,-q1,q0]
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const double a = D[4 * r], b = D[4 * r + 1], c = D[4 * r + 2], d = D[4 * r + 3];
    g->G[3 * r + 0] = -a * q1 + b * q0 - c * q3 + d * q2;
    g->G[3 * r + 1] = -a * q2 + b * q3 + c * q0 - d * q1;
    g->G[3 * r + 2] = -a * q3 - b * q2 + c * q1 + d * q0;
  }
}

// value of Jk[amb] (row 0) / Jk[12 + amb] (row 1) without a dynamic register index
template <int NP>
__device__ __forceinline__ void PickParam(const double (&Jk)[24], int amb, double* j0, double* j1) {
  double a = 0.0, b = 0.0;
#pragma unroll
  for (int t = 0; t < NP; ++t) { a = (t == amb) ? Jk[t] : a; b = (t == amb) ? Jk[12 + t] : b; }
  *j0 = a; *j1 = b;
}

// scale_i [n_img*6] / scale_k [n_cam*12, by intrinsics slot]: Jacobi scale of the slot, 0 where the slot is
// constant or unused (the plain evaluation passes 1 / 0 masks).  Folding "is this slot free" into the scale
// replaces the per-slot column-index lookups: an observation gathers its pose (two 32 B loads), its pose
// scales (three 16 B loads), its camera record and one packed slot map -- a third of the scattered loads the
// index-table version needed, which is what bounded it (L1 wavefronts, profiles/r01_ncu_full_c4_eval_v1.txt).
// The records leave as 32 B units of the tile-blocked layout: every store instruction of a warp writes 1 KB
// contiguous, straight from registers.
#ifndef SBA_EVAL_MINB
#define SBA_EVAL_MINB 4
#endif
// three-row instantiations: the three small models fit 128 registers without spills (4 CTAs / SM); the others keep 3
constexpr int EvalResident(int model, int nr) { return nr == 3 ? (model <= 2 ? 4 : 3) : SBA_EVAL_MINB; }  // CTAs per SM (launch bounds)
template <int MODEL, int NR>
__global__ void __launch_bounds__(kEvalThreads, EvalResident(MODEL, NR))
eval_kernel(ParamsView P, StructView S, LossView L, const double* __restrict__ scale_i,
            const double* __restrict__ scale_k, const unsigned long long* __restrict__ k_pack,
            const double* __restrict__ scale_p, D4* __restrict__ ER, D4* __restrict__ F, Row3View R3,
            double* __restrict__ cost_partial, double* __restrict__ Jq_ambient, unsigned* ticket,
            DeviceScalars* sc) {
  constexpr int NP = MODEL >= 0 ? ModelNumParams(MODEL >= 0 ? MODEL : 0) : 12;
  constexpr int NS = NR == 3 ? (6 + NP + 3) / 4 * 4 : 4;  // third-row slots held in registers, padded to whole units
  double cost = 0.0, fixed = 0.0;
  // grid-stride: the grid is sized to the machine (a few CTAs per SM), not to the problem, so the
  // per-block partial sums that the last block folds stay in the hundreds however large the scene is
  // The streamed per-observation inputs of the NEXT trip are requested before this trip's arithmetic, so
  // the DRAM latency of the index loads (which the pose / point gathers depend on) is off the chain.
  const int stride = gridDim.x * kEvalThreads;
  int k = blockIdx.x * kEvalThreads + threadIdx.x;
  int i_n = 0, pt_n = 0; double2 xy_n = make_double2(0.0, 0.0); bool act_n = false;
  if (k < S.n_obs) {
    i_n = S.obs_img[k]; pt_n = S.obs_pt[k];
    xy_n = __ldcs(reinterpret_cast<const double2*>(S.obs_xy) + k); act_n = S.obs_active[k] != 0;
  }
  for (; k < S.n_obs; k += stride) {
    const int i = i_n, pt = pt_n;
    const double2 xy = xy_n;
    const bool active = act_n;
    if (k + stride < S.n_obs) {
      i_n = S.obs_img[k + stride]; pt_n = S.obs_pt[k + stride];
      xy_n = __ldcs(reinterpret_cast<const double2*>(S.obs_xy) + k + stride); act_n = S.obs_active[k + stride] != 0;
    }
    const int c = S.img_cam[i];
    const int model = MODEL >= 0 ? MODEL : S.model[c];
    const D4* pp = reinterpret_cast<const D4*>(P.pose + 8 * (size_t)i);
    const D4 a0 = Ld256(pp), a1 = Ld256(pp + 1);
    const double q[4] = {a0.x, a0.y, a0.z, a0.w};
    const double t[3] = {a1.x, a1.y, a1.z};
    const double X[3] = {P.X[3 * (size_t)pt], P.X[3 * (size_t)pt + 1], P.X[3 * (size_t)pt + 2]};
    double prm[12];
    {
      const D4* cp = reinterpret_cast<const D4*>(P.cam + 12 * (size_t)c);
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        if (4 * j < NP) { const D4 v = Ld256(cp + j); prm[4 * j] = v.x; prm[4 * j + 1] = v.y; prm[4 * j + 2] = v.z; prm[4 * j + 3] = v.w; }
        else { prm[4 * j] = 0.0; prm[4 * j + 1] = 0.0; prm[4 * j + 2] = 0.0; prm[4 * j + 3] = 0.0; }
      }
    }
    ObsGeom g;
    RotateWithJac(q, t, X, &g, true);
    const double iz = 1.0 / g.p[2];
    const double u = g.p[0] * iz, v = g.p[1] * iz;
    double x, y, A[4], Jk[24];
    if (MODEL >= 0) WorldToImageJac<(MODEL >= 0 ? MODEL : 0)>(prm, u, v, &x, &y, A, Jk);
    else WorldToImageJacDyn(model, prm, u, v, &x, &y, A, Jk);
    double r0 = x - xy.x, r1 = y - xy.y;
    // AB = A * d(u,v)/dp   (2x3)
    const double AB[6] = {A[0] * iz, A[1] * iz, -(A[0] * u + A[1] * v) * iz,
                          A[2] * iz, A[3] * iz, -(A[2] * u + A[3] * v) * iz};
    if (Jq_ambient) {  // raw 2x4 block w.r.t. (w,x,y,z): parity surface only
      const double q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3], X0 = X[0], X1 = X[1], X2 = X[2];
      const double D[12] = {
          2.0 * (-q3 * X1 + q2 * X2), 2.0 * (q2 * X1 + q3 * X2),             2.0 * (-2.0 * q2 * X0 + q1 * X1 + q0 * X2), 2.0 * (-2.0 * q3 * X0 - q0 * X1 + q1 * X2),
          2.0 * (q3 * X0 - q1 * X2),  2.0 * (q2 * X0 - 2.0 * q1 * X1 - q0 * X2), 2.0 * (q1 * X0 + q3 * X2),              2.0 * (q0 * X0 - 2.0 * q3 * X1 + q2 * X2),
          2.0 * (-q2 * X0 + q1 * X1), 2.0 * (q3 * X0 + q0 * X1 - 2.0 * q1 * X2), 2.0 * (-q0 * X0 + q3 * X1 - 2.0 * q2 * X2), 2.0 * (q1 * X0 + q2 * X1)};
      for (int row = 0; row < 2; ++row)
        for (int j = 0; j < 4; ++j)
          Jq_ambient[(size_t)k * 8 + row * 4 + j] =
              AB[3 * row] * D[j] + AB[3 * row + 1] * D[4 + j] + AB[3 * row + 2] * D[8 + j];
    }
    // third row: value and (g_x, g_y) = weight * d map / d (x, y)
    double rs = 0.0, gx = 0.0, gy = 0.0;
    if (NR == 3) {
      double val;
      if (BilinearWithGrad(R3.maps + (size_t)i * MapTiledSize(R3.H, R3.W), R3.H, R3.W, x, y, &val, &gx, &gy)) {
        rs = R3.weight * val; gx *= R3.weight; gy *= R3.weight;
      } else { gx = 0.0; gy = 0.0; }
    }
    const double s = r0 * r0 + r1 * r1 + (NR == 3 ? rs * rs : 0.0);
    double rho[3];
    LossEval(L.type, L.scale, s, rho);
    const int dch = S.dc >> 1;
    const int nq = FSUnits(S.dc);
    const D4 zero = {0.0, 0.0, 0.0, 0.0};
    if (!active) {
      fixed += 0.5 * rho[0];
      for (int jj = 0; jj < dch; ++jj) St256(F + FUnit(k, jj, dch), zero);
      // the raw residual is still exported for the parity surface; E = 0 keeps it out of every product
      St256(ER + EUnit(k, 0), zero);
      St256(ER + EUnit(k, 1), D4{0.0, 0.0, r0, r1});
      if (NR == 3) {
        for (int q = 0; q < nq; ++q) St256(R3.FS + FSUnit(k, q, nq), zero);
        St256(R3.ES + k, D4{0.0, 0.0, 0.0, rs});
      }
    } else {
      cost += 0.5 * rho[0];
      // corrector (ceres corrector.cc)
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
      // (j2: the third-row entry of the same column; untouched storage when NR == 2)
      auto correct = [&](double& j0, double& j1, double& j2, double scale) {
        if (alpha_sq_norm != 0.0) {
          const double rtj = r0 * j0 + r1 * j1 + (NR == 3 ? rs * j2 : 0.0);
          j0 -= alpha_sq_norm * r0 * rtj;
          j1 -= alpha_sq_norm * r1 * rtj;
          if (NR == 3) j2 -= alpha_sq_norm * rs * rtj;
        }
        j0 = sqrt_rho1 * j0 * scale; j1 = sqrt_rho1 * j1 * scale;
        if (NR == 3) j2 = sqrt_rho1 * j2 * scale;
      };
      // third row of A * d(u,v)/dp:  (g_x, g_y) . AB
      const double ABs[3] = {gx * AB[0] + gy * AB[3], gx * AB[1] + gy * AB[4], gx * AB[2] + gy * AB[5]};
      double f2[NS];
#pragma unroll
      for (int j = 0; j < NS; ++j) f2[j] = 0.0;
      // pose slots: 0..2 quaternion tangent, 3..5 tvec
      {
        const double2* sp2 = reinterpret_cast<const double2*>(scale_i + 6 * (size_t)i);
        const double2 s01 = __ldg(sp2), s23 = __ldg(sp2 + 1), s45 = __ldg(sp2 + 2);
        const double si[6] = {s01.x, s01.y, s23.x, s23.y, s45.x, s45.y};
        double f0[6], f1[6];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          f0[j] = AB[0] * g.G[j] + AB[1] * g.G[3 + j] + AB[2] * g.G[6 + j];
          f1[j] = AB[3] * g.G[j] + AB[4] * g.G[3 + j] + AB[5] * g.G[6 + j];
          f0[3 + j] = AB[j]; f1[3 + j] = AB[3 + j];
          if (NR == 3) {
            f2[j] = ABs[0] * g.G[j] + ABs[1] * g.G[3 + j] + ABs[2] * g.G[6 + j];
            f2[3 + j] = ABs[j];
          }
        }
#pragma unroll
        for (int j = 0; j < 6; ++j) {
          if (si[j] == 0.0) { f0[j] = 0.0; f1[j] = 0.0; f2[NR == 3 ? j : 0] = 0.0; }  // constant slot: exactly zero, whatever the value
          else correct(f0[j], f1[j], f2[NR == 3 ? j : 0], si[j]);
        }
#pragma unroll
        for (int jj = 0; jj < 3; ++jj) St256(F + FUnit(k, jj, dch), D4{f0[2 * jj], f1[2 * jj], f0[2 * jj + 1], f1[2 * jj + 1]});
      }
      // intrinsics slots
      if (dch > 3) {
        const unsigned long long pack = k_pack[c];
        const double* sk = scale_k + 12 * (size_t)c;
#pragma unroll
        for (int jj = 3; jj < 3 + (NP + 1) / 2; ++jj) {
          if (jj < dch) {
            double f0[2], f1[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int slot = 2 * jj - 6 + h;
              const double scl = sk[slot];
              PickParam<NP>(Jk, (int)((pack >> (4 * slot)) & 15ull), &f0[h], &f1[h]);
              double& fs = f2[NR == 3 ? 6 + slot : 0];
              if (NR == 3) fs = gx * f0[h] + gy * f1[h];
              if (scl == 0.0) { f0[h] = 0.0; f1[h] = 0.0; fs = 0.0; }
              else correct(f0[h], f1[h], fs, scl);
            }
            St256(F + FUnit(k, jj, dch), D4{f0[0], f1[0], f0[1], f1[1]});
          }
        }
        // units of the dc tier beyond this model's parameters (e.g. RADIAL, 5 free: dc = 14, slots 12-13) hold no
        // column; they are multiplied by zeros downstream and must not be stale memory
        for (int jj = 3 + (NP + 1) / 2; jj < dch; ++jj) St256(F + FUnit(k, jj, dch), zero);
      }
      // point block
      double E[6], Es[3] = {0.0, 0.0, 0.0};
      const int po = S.p_off[pt];
      if (po >= 0) {
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          double j0 = AB[0] * g.R[j] + AB[1] * g.R[3 + j] + AB[2] * g.R[6 + j];
          double j1 = AB[3] * g.R[j] + AB[4] * g.R[3 + j] + AB[5] * g.R[6 + j];
          if (NR == 3) Es[j] = ABs[0] * g.R[j] + ABs[1] * g.R[3 + j] + ABs[2] * g.R[6 + j];
          correct(j0, j1, Es[j], scale_p ? scale_p[3 * (size_t)po + j] : 1.0);
          E[j] = j0; E[3 + j] = j1;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 6; ++j) E[j] = 0.0;
      }
      r0 *= residual_scaling; r1 *= residual_scaling;
      St256(ER + EUnit(k, 0), D4{E[0], E[1], E[2], E[3]});
      St256(ER + EUnit(k, 1), D4{E[4], E[5], r0, r1});
      if (NR == 3) {
        rs *= residual_scaling;
#pragma unroll
        for (int q = 0; q < NS / 4; ++q)
          if (q < nq) St256(R3.FS + FSUnit(k, q, nq), D4{f2[4 * q], f2[4 * q + 1], f2[4 * q + 2], f2[4 * q + 3]});
        for (int q = NS / 4; q < nq; ++q) St256(R3.FS + FSUnit(k, q, nq), zero);
        St256(R3.ES + k, D4{Es[0], Es[1], Es[2], rs});
      }
    }
  }
  double v[2] = {cost, fixed};
  double* const out[2] = {&sc->cost, &sc->fixed_cost};
  GridSum<kEvalThreads, 2>(v, cost_partial, ticket, out);
}

// residual-only evaluation of the candidate point (Evaluator::Evaluate with no jacobian)
template <int MODEL, int NR>
__global__ void __launch_bounds__(kEvalThreads)
cost_kernel(ParamsView P, StructView S, LossView L, Row3View R3, double* __restrict__ cost_partial, unsigned* ticket,
            double* out_cost) {
  double cost = 0.0, fixed = 0.0;
  // the streamed inputs of the next trip are requested before this trip's gathers (as in eval_kernel)
  const int stride = gridDim.x * kEvalThreads;
  int k = blockIdx.x * kEvalThreads + threadIdx.x;
  int i_n = 0, pt_n = 0; double2 xy_n = make_double2(0.0, 0.0); bool act_n = false;
  if (k < S.n_obs) {
    i_n = S.obs_img[k]; pt_n = S.obs_pt[k];
    xy_n = __ldcs(reinterpret_cast<const double2*>(S.obs_xy) + k); act_n = S.obs_active[k] != 0;
  }
  for (; k < S.n_obs; k += stride) {
    const int i = i_n, pt = pt_n;
    const double2 xy = xy_n;
    const bool active = act_n;
    if (k + stride < S.n_obs) {
      i_n = S.obs_img[k + stride]; pt_n = S.obs_pt[k + stride];
      xy_n = __ldcs(reinterpret_cast<const double2*>(S.obs_xy) + k + stride); act_n = S.obs_active[k + stride] != 0;
    }
    const int c = S.img_cam[i];
    const int model = MODEL >= 0 ? MODEL : S.model[c];
    const double2* pp = reinterpret_cast<const double2*>(P.pose + 8 * (size_t)i);
    const double2 a0 = __ldg(pp), a1 = __ldg(pp + 1), a2 = __ldg(pp + 2), a3 = __ldg(pp + 3);
    const double q[4] = {a0.x, a0.y, a1.x, a1.y};
    const double t[3] = {a2.x, a2.y, a3.x};
    const double X[3] = {P.X[3 * (size_t)pt], P.X[3 * (size_t)pt + 1], P.X[3 * (size_t)pt + 2]};
    double prm[12];
    {
      const double2* cp = reinterpret_cast<const double2*>(P.cam + 12 * (size_t)c);
      const int np = MODEL >= 0 ? ModelNumParams(MODEL) : 12;
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        if (2 * j < np) { const double2 v = __ldg(cp + j); prm[2 * j] = v.x; prm[2 * j + 1] = v.y; }
        else { prm[2 * j] = 0.0; prm[2 * j + 1] = 0.0; }
      }
    }
    ObsGeom g;
    RotateWithJac(q, t, X, &g, false);
    const double iz = 1.0 / g.p[2];
    double x, y;
    if (MODEL >= 0) WorldToImageVal<(MODEL >= 0 ? MODEL : 0)>(prm, g.p[0] * iz, g.p[1] * iz, &x, &y);
    else WorldToImageValDyn(model, prm, g.p[0] * iz, g.p[1] * iz, &x, &y);
    const double r0 = x - xy.x, r1 = y - xy.y;
    double rs = 0.0;
    if (NR == 3) {
      double val, gx, gy;
      if (BilinearWithGrad(R3.maps + (size_t)i * MapTiledSize(R3.H, R3.W), R3.H, R3.W, x, y, &val, &gx, &gy)) rs = R3.weight * val;
    }
    double rho[3];
    LossEval(L.type, L.scale, r0 * r0 + r1 * r1 + rs * rs, rho);
    if (active) cost += 0.5 * rho[0]; else fixed += 0.5 * rho[0];
  }
  double v[2] = {cost, fixed};
  double* const out[2] = {out_cost, nullptr};
  GridSum<kEvalThreads, 2>(v, cost_partial, ticket, out);
}


template <int MODEL>
static void LaunchEvalT(Problem* p, const LossView& loss, bool use_scale, bool candidate, int blocks) {
#define SBA_EVAL_ARGS                                                                                                  \
  p->Params(candidate), p->Struct(), loss, use_scale ? p->d_scale_i.p : p->d_mask_i.p,                                \
      use_scale ? p->d_scale_k.p : p->d_mask_k.p, p->d_k_pack.p, use_scale ? p->d_scale_p.p : nullptr, p->d_ER.p,      \
      p->d_F.p, p->Row3(), p->d_cost_partial.p, p->jq_out, p->d_tickets.p + 0, p->d_scalars.p
  // one resident wave: the kernel is grid-stride with uniform work per observation, and more CTAs than fit at once only
  // add a tail (C4: 0.298 ms at 16 CTAs per SM, 0.288 ms at the 4 that are resident)
  if (p->nr == 3) eval_kernel<MODEL, 3><<<std::min(blocks, p->n_sm * EvalResident(MODEL, 3)), kEvalThreads, 0, p->stream>>>(SBA_EVAL_ARGS);
  else eval_kernel<MODEL, 2><<<std::min(blocks, p->n_sm * EvalResident(MODEL, 2)), kEvalThreads, 0, p->stream>>>(SBA_EVAL_ARGS);
#undef SBA_EVAL_ARGS
}
template <int MODEL>
static void LaunchCostT(Problem* p, const LossView& loss, bool candidate, int blocks, double* d_out_cost) {
  // (two rows, measured at C4: 0.095 ms at 8 CTAs per SM against 0.117 ms at 16; three rows want the parallelism for
  // their texel loads: C5 0.735 ms at 16 against 0.797 ms at 8)
  if (p->nr != 3) blocks = std::min(blocks, p->n_sm * 8);
  if (p->nr == 3)
    cost_kernel<MODEL, 3><<<blocks, kEvalThreads, 0, p->stream>>>(p->Params(candidate), p->Struct(), loss, p->Row3(),
                                                                  p->d_cost_partial.p, p->d_tickets.p + 1, d_out_cost);
  else
    cost_kernel<MODEL, 2><<<blocks, kEvalThreads, 0, p->stream>>>(p->Params(candidate), p->Struct(), loss, p->Row3(),
                                                                  p->d_cost_partial.p, p->d_tickets.p + 1, d_out_cost);
}

#define SBA_MODEL_SWITCH(m, CALL)                                            \
  switch (m) {                                                               \
    case 0: CALL(0); break;  case 1: CALL(1); break;  case 2: CALL(2); break; \
    case 3: CALL(3); break;  case 4: CALL(4); break;  case 5: CALL(5); break; \
    case 6: CALL(6); break;  case 7: CALL(7); break;  case 8: CALL(8); break; \
    case 9: CALL(9); break;  case 10: CALL(10); break; default: CALL(-1); break; \
  }

// CTAs per launch: enough to fill every SM several times over, never more than the problem needs
static int EvalGrid(const Problem* p) { return std::min(DivUp(p->n_obs, kEvalThreads), p->n_sm * 16); }

int LaunchEvaluate(Problem* p, const LossView& loss, bool use_scale, bool candidate, bool defer_allreduce) {
  p->colnorm_current = false;  // a new Jacobian: d_sq_c is stale until ColumnNorms / Gradient(with_norms)
  p->ete_valid = false;        // ... and the cached E_p^T E_p until the next pass over ER (PointBlocks)
  const int blocks = EvalGrid(p);
  if (blocks > 0) {
#define CALL(M) LaunchEvalT<M>(p, loss, use_scale, candidate, blocks)
    SBA_MODEL_SWITCH(p->uniform_model, CALL)
#undef CALL
    p->launches++;
  } else {  // empty shard: the sums are zero
    SBA_CUDA(cudaMemsetAsync(&p->d_scalars.p->cost, 0, 2 * sizeof(double), p->stream));
  }
  SBA_CUDA(cudaGetLastError());
  p->linearized = true;
  p->cost_unreduced = defer_allreduce && p->comm != nullptr;
  if (p->cost_unreduced) return SBA_OK;
  return AllReduceSum(p, &p->d_scalars.p->cost, 2);  // cost, fixed_cost are adjacent
}

int LaunchCostOnly(Problem* p, const LossView& loss, bool candidate, double* d_out_cost, bool with_step_scalars) {
  const int blocks = EvalGrid(p);
  if (blocks > 0) {
#define CALL(M) LaunchCostT<M>(p, loss, candidate, blocks, d_out_cost)
    SBA_MODEL_SWITCH(p->uniform_model, CALL)
#undef CALL
    p->launches++;
  } else {
    SBA_CUDA(cudaMemsetAsync(d_out_cost, 0, sizeof(double), p->stream));
  }
  SBA_CUDA(cudaGetLastError());
  // with_step_scalars: d_out_cost is &scalars->cand_cost and model_cost_change, step2_p, x2_p behind it still hold this
  // rank's partial sums (BackSubstitute / ApplyStep deferred theirs): one collective for the four
  return AllReduceSum(p, d_out_cost, with_step_scalars ? 4 : 1);
}


// ---- cost evaluation that allocates nothing (sba_cost_no_alloc, include/sba_b200.h) ----------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is cost_kernel with its scratch --
// the per-CTA partial sums, the arrival ticket and the two totals of GridSum -- in a file-scope device array instead of
// the handle's cudaMalloc'ed d_cost_partial / d_tickets / d_scalars.  Every GPU that loads the library holds the array
// (kNoAllocScratchBytes = 65 792 B) until the process exits.  ALL calls on a device share it: two calls that overlap on
// the GPU (different streams, different handles, different host threads) race on the partial sums and the ticket; calls
// are serialised per process by a mutex and each call ends with a stream synchronisation, which is the supported use.
constexpr int kNoAllocMaxCtas = SBA_NO_ALLOC_MAX_CTAS;       // launch grids beyond this are rejected
constexpr int kNoAllocHead = 32;                             // doubles: [0] cost, [1] fixed cost, [2] ticket (as u32), pad to 256 B
constexpr size_t kNoAllocScratchBytes = (size_t)(kNoAllocHead + 2 * kNoAllocMaxCtas) * sizeof(double);
static_assert(kNoAllocScratchBytes == SBA_NO_ALLOC_SCRATCH_BYTES, "header constant");
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocHead + 2 * kNoAllocMaxCtas];

template <int MODEL, int NR>
__global__ void __launch_bounds__(kEvalThreads)
cost_noalloc_kernel(ParamsView P, StructView S, LossView L, Row3View R3, double* __restrict__ scratch) {
  double cost = 0.0, fixed = 0.0;
  const int stride = gridDim.x * kEvalThreads;
  for (int k = blockIdx.x * kEvalThreads + threadIdx.x; k < S.n_obs; k += stride) {
    const int i = S.obs_img[k], pt = S.obs_pt[k];
    const double2 xy = __ldcs(reinterpret_cast<const double2*>(S.obs_xy) + k);
    const bool active = S.obs_active[k] != 0;
    const int c = S.img_cam[i];
    const int model = MODEL >= 0 ? MODEL : S.model[c];
    const double2* pp = reinterpret_cast<const double2*>(P.pose + 8 * (size_t)i);
    const double2 a0 = __ldg(pp), a1 = __ldg(pp + 1), a2 = __ldg(pp + 2), a3 = __ldg(pp + 3);
    const double q[4] = {a0.x, a0.y, a1.x, a1.y};
    const double t[3] = {a2.x, a2.y, a3.x};
    const double X[3] = {P.X[3 * (size_t)pt], P.X[3 * (size_t)pt + 1], P.X[3 * (size_t)pt + 2]};
    double prm[12];
    {
      const double2* cp = reinterpret_cast<const double2*>(P.cam + 12 * (size_t)c);
      const int np = MODEL >= 0 ? ModelNumParams(MODEL) : 12;
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        if (2 * j < np) { const double2 v = __ldg(cp + j); prm[2 * j] = v.x; prm[2 * j + 1] = v.y; }
        else { prm[2 * j] = 0.0; prm[2 * j + 1] = 0.0; }
      }
    }
    ObsGeom g;
    RotateWithJac(q, t, X, &g, false);
    const double iz = 1.0 / g.p[2];
    double x, y;
    if (MODEL >= 0) WorldToImageVal<(MODEL >= 0 ? MODEL : 0)>(prm, g.p[0] * iz, g.p[1] * iz, &x, &y);
    else WorldToImageValDyn(model, prm, g.p[0] * iz, g.p[1] * iz, &x, &y);
    const double r0 = x - xy.x, r1 = y - xy.y;
    double rs = 0.0;
    if (NR == 3) {
      double val, gx, gy;
      if (BilinearWithGrad(R3.maps + (size_t)i * MapTiledSize(R3.H, R3.W), R3.H, R3.W, x, y, &val, &gx, &gy)) rs = R3.weight * val;
    }
    double rho[3];
    LossEval(L.type, L.scale, r0 * r0 + r1 * r1 + rs * rs, rho);
    if (active) cost += 0.5 * rho[0]; else fixed += 0.5 * rho[0];
  }
  double v[2] = {cost, fixed};
  double* const out[2] = {scratch, scratch + 1};
  GridSum<kEvalThreads, 2>(v, scratch + kNoAllocHead, reinterpret_cast<unsigned*>(scratch + 2), out);
}

__global__ void noalloc_total_kernel(const double* __restrict__ scratch, double* __restrict__ out) { *out = scratch[0] + scratch[1]; }

// Device address of g_noalloc_scratch on the CURRENT device.  In host code the array's name is only a stand-in: the
// address a kernel can use comes from cudaGetSymbolAddress, asked once per device.
static int NoAllocScratchAddress(int device, double** out) {
  constexpr int kMaxDevices = 64;
  static double* addr[kMaxDevices] = {};
  if (device < 0 || device >= kMaxDevices) { SetError(SBA_ERR_INVALID_ARGUMENT, "device ordinal out of range"); return SBA_ERR_INVALID_ARGUMENT; }
  if (!addr[device]) {
    double* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) {
      SetError(SBA_ERR_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
      return SBA_ERR_CUDA;
    }
    addr[device] = ptr;
  }
  *out = addr[device];
  return SBA_OK;
}

// 0.5 sum rho over the residual blocks of the handle's CURRENT parameters (active + constant-only blocks, as the cost
// sba_evaluate returns), with no cudaMalloc / cudaFree / allocating library call.  cost_host / cost_device: either may be
// null.  Returns after the stream has been synchronised.
int CostNoAlloc(Problem* p, const LossView& loss, double* cost_host, double* cost_device) {
  static std::mutex mu;  // one call at a time per process: the scratch array is shared
  std::lock_guard<std::mutex> lk(mu);
  if (p->comm || p->shard_count > 1) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "sba_cost_no_alloc: sharded handles are not supported (the sum would be this shard's part)");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  const int blocks = EvalGrid(p);
  if (blocks > kNoAllocMaxCtas) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "sba_cost_no_alloc: the launch needs " + std::to_string(blocks) + " CTAs, the static scratch holds " +
                                           std::to_string(kNoAllocMaxCtas) + " (SBA_NO_ALLOC_MAX_CTAS)");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  int prev = -1;
  SBA_CUDA(cudaGetDevice(&prev));
  if (prev != p->device) SBA_CUDA(cudaSetDevice(p->device));
  double* scratch = nullptr;
  int rc = NoAllocScratchAddress(p->device, &scratch);
  cudaStream_t st = p->stream;
  cudaError_t e = cudaSuccess;
  if (rc == SBA_OK) {
    // the array keeps the last call's contents: totals and ticket start at zero (a call that died mid-way must not
    // leave a ticket behind); the partial sums are written before they are read
    e = cudaMemsetAsync(scratch, 0, kNoAllocHead * sizeof(double), st);
    if (e == cudaSuccess && blocks > 0) {
#define CALL(M)                                                                                                          \
  do {                                                                                                                   \
    if (p->nr == 3) cost_noalloc_kernel<M, 3><<<blocks, kEvalThreads, 0, st>>>(p->Params(false), p->Struct(), loss, p->Row3(), scratch); \
    else cost_noalloc_kernel<M, 2><<<blocks, kEvalThreads, 0, st>>>(p->Params(false), p->Struct(), loss, p->Row3(), scratch);           \
  } while (0)
      SBA_MODEL_SWITCH(p->uniform_model, CALL)
#undef CALL
      p->launches++;
      e = cudaGetLastError();
    }
    double h[2] = {0.0, 0.0};
    if (e == cudaSuccess && cost_device) { noalloc_total_kernel<<<1, 1, 0, st>>>(scratch, cost_device); e = cudaGetLastError(); }
    if (e == cudaSuccess && cost_host) e = cudaMemcpyAsync(h, scratch, sizeof(h), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e == cudaSuccess && cost_host) *cost_host = h[0] + h[1];
  }
  if (prev != p->device && prev >= 0) cudaSetDevice(prev);
  if (rc) return rc;
  if (e != cudaSuccess) { SetError(SBA_ERR_CUDA, std::string("sba_cost_no_alloc: ") + cudaGetErrorString(e)); return SBA_ERR_CUDA; }
  return SBA_OK;
}

}  // namespace sba
