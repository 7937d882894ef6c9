// Point-block Schur elimination pieces, implicit reduced-camera-matrix product,
// back-substitution and the camera-side segmented reductions.
//
// These kernels re-create on the device what ceres::Solve does for the problem the
// reference builds (/root/reference/src/optim/bundle_adjustment.cc:306, :326-530):
// SchurEliminator / ImplicitSchurComplement arithmetic per SURVEY.md App. B.5,
// Jacobi scaling and the LM diagonal per App. B.3/B.4.
//
// Conventions: per observation k (point-sorted): ER[k] = {E 2x3 row-major, r(2)},
// F[k] = double2[DC].  Camera-space vectors exist in two forms: compact
// (n_cdim) and padded (n_img*6 | n_cam*12, zeros at constant slots) -- the padded
// form lets an observation fetch its pose's six entries with three 16 B loads.
#pragma once
#include "reduce.cuh"
#include "solver.h"
#include "tiles.cuh"

namespace sba {

#define SBA_DC_SWITCH(dc, CALL)                                                       \
  switch (dc) {                                                                       \
    case 6: CALL(6); break; case 8: CALL(8); break; case 10: CALL(10); break;         \
    case 14: CALL(14); break; default: CALL(18); break;                               \
  }

// ---- compact <-> padded camera vectors ----------------------------------------------
__global__ void expand_kernel(StructView S, const double* __restrict__ x, double* __restrict__ xi,
                              double* __restrict__ xc) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int ni = S.n_img * 6;
  if (t < ni) {
    const int col = S.col_img[t];
    xi[t] = col >= 0 ? x[col] : 0.0;
  } else if (t < ni + S.n_cam * 12) {
    const int col = S.col_cam[t - ni];
    xc[t - ni] = col >= 0 ? x[col] : 0.0;
  }
}

// DC x NR dispatch: CALL(DC, NR)
#define SBA_DCNR_CASE(dcv, nr, CALL) case dcv: if ((nr) == 3) { CALL(dcv, 3); } else { CALL(dcv, 2); } break;
#define SBA_DCNR_SWITCH(dc, nr, CALL)                                                              \
  switch (dc) {                                                                                    \
    SBA_DCNR_CASE(6, nr, CALL) SBA_DCNR_CASE(8, nr, CALL) SBA_DCNR_CASE(10, nr, CALL)              \
    SBA_DCNR_CASE(14, nr, CALL)                                                                    \
    default: if ((nr) == 3) { CALL(18, 3); } else { CALL(18, 2); } break;                          \
  }

// One observation's camera-side Jacobian block: a[j] = (d r0 / d s_j, d r1 / d s_j), s[j] = d r_s / d s_j (NR == 3)
template <int DC, int NR> struct FRow { double2 a[DC]; double s[NR == 3 ? DC : 1]; };

template <int DC, int NR>
__device__ inline void LoadRow(const D4* __restrict__ F, const D4* __restrict__ FS, int k, FRow<DC, NR>& row) {
#pragma unroll
  for (int jj = 0; jj < DC / 2; ++jj) {
    const D4 v = Ld256(F + FUnit(k, jj, DC / 2));
    row.a[2 * jj] = make_double2(v.x, v.y); row.a[2 * jj + 1] = make_double2(v.z, v.w);
  }
  if constexpr (NR == 3) {
    constexpr int NQ = FSUnits(DC);
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      const D4 v = Ld256(FS + FSUnit(k, q, NQ));
      row.s[4 * q] = v.x;
      if (4 * q + 1 < DC) row.s[4 * q + 1] = v.y;
      if (4 * q + 2 < DC) row.s[4 * q + 2] = v.z;
      if (4 * q + 3 < DC) row.s[4 * q + 3] = v.w;
    }
  }
}

// f = F_k * x  (NR-vector), x in padded form
template <int DC, int NR>
__device__ inline void RowTimesX(const FRow<DC, NR>& row, const double* __restrict__ xi, const double* __restrict__ xc,
                                 int i, int c, double (&f)[NR]) {
  const double2* pi = reinterpret_cast<const double2*>(xi + 6 * (size_t)i);
#pragma unroll
  for (int r = 0; r < NR; ++r) f[r] = 0.0;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const double2 v = __ldg(pi + j);
    f[0] += row.a[2 * j].x * v.x + row.a[2 * j + 1].x * v.y;
    f[1] += row.a[2 * j].y * v.x + row.a[2 * j + 1].y * v.y;
    if constexpr (NR == 3) f[2] += row.s[2 * j] * v.x + row.s[2 * j + 1] * v.y;
  }
  if (DC > 6) {
    const double2* pc = reinterpret_cast<const double2*>(xc + 12 * (size_t)c);
#pragma unroll
    for (int j = 0; j < (DC - 6) / 2; ++j) {
      const double2 v = __ldg(pc + j);
      f[0] += row.a[6 + 2 * j].x * v.x + row.a[7 + 2 * j].x * v.y;
      f[1] += row.a[6 + 2 * j].y * v.x + row.a[7 + 2 * j].y * v.y;
      if constexpr (NR == 3) f[2] += row.s[6 + 2 * j] * v.x + row.s[7 + 2 * j] * v.y;
    }
  }
}

// E (NR x 3, row-major) and r (NR) of one observation
template <int NR> struct ERec { double E[3 * NR]; double r[NR]; };
template <int NR>
__device__ inline ERec<NR> LoadER(const D4* __restrict__ ER, const D4* __restrict__ ES, int k) {
  const D4 a = Ld256(ER + EUnit(k, 0)), b = Ld256(ER + EUnit(k, 1));
  ERec<NR> e;
  e.E[0] = a.x; e.E[1] = a.y; e.E[2] = a.z; e.E[3] = a.w; e.E[4] = b.x; e.E[5] = b.y;
  e.r[0] = b.z; e.r[1] = b.w;
  if constexpr (NR == 3) {
    const D4 c = Ld256(ES + k);
    e.E[6] = c.x; e.E[7] = c.y; e.E[8] = c.z; e.r[2] = c.w;
  }
  return e;
}
// u += E^T f,  f -= E w
template <int NR>
__device__ __forceinline__ void EtAdd(const ERec<NR>& e, const double (&f)[NR], double (&u)[3]) {
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int r = 0; r < NR; ++r) u[a] += e.E[3 * r + a] * f[r];
  }
}
template <int NR>
__device__ __forceinline__ void ESub(const ERec<NR>& e, const double (&w)[3], double (&f)[NR]) {
#pragma unroll
  for (int r = 0; r < NR; ++r) f[r] -= e.E[3 * r] * w[0] + e.E[3 * r + 1] * w[1] + e.E[3 * r + 2] * w[2];
}
// the per-observation NR-vector scratch: (f0, f1) in v2[k], f2 in v1[k]
template <int NR>
__device__ __forceinline__ void StoreVec(double2* __restrict__ v2, double* __restrict__ v1, int k, const double (&f)[NR]) {
  v2[k] = make_double2(f[0], f[1]);
  if constexpr (NR == 3) v1[k] = f[2];
}
template <int NR>
__device__ __forceinline__ void LoadVec(const double2* __restrict__ v2, const double* __restrict__ v1, int k, double (&f)[NR]) {
  const double2 t = v2[k];
  f[0] = t.x; f[1] = t.y;
  if constexpr (NR == 3) f[2] = v1[k];
}

// ---- point-side squared column norms (Jacobi scaling / LM diagonal) -----------------
template <int NR>
__global__ void __launch_bounds__(128) point_sqnorm_kernel(StructView S, TileView T, const D4* __restrict__ ER,
                                                           const D4* __restrict__ ES, double* __restrict__ np) {
  const int tile = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (tile >= T.n_tiles) return;
  const int lane = threadIdx.x & 31;
  const int s0 = T.tile_start[tile], n = T.tile_start[tile + 1] - s0;
  auto sq = [](const ERec<NR>& e, int a) {
    double v = e.E[a] * e.E[a] + e.E[3 + a] * e.E[3 + a];
    if constexpr (NR == 3) v += e.E[6 + a] * e.E[6 + a];
    return v;
  };
  if (n <= 32) {
    const bool valid = lane < n;
    const int k = s0 + lane;
    const int pt = valid ? S.obs_pt[k] : -1 - lane;
    double v[3] = {0.0, 0.0, 0.0};
    if (valid) {
      const ERec<NR> e = LoadER<NR>(ER, ES, k);
#pragma unroll
      for (int a = 0; a < 3; ++a) v[a] = sq(e, a);
    }
    const Seg sg = MakeSeg(pt);
    SegSum<3>(sg, v);
    if (valid && sg.is_head) {
      const int po = S.p_off[pt];
      if (po >= 0) { np[3 * (size_t)po] = v[0]; np[3 * (size_t)po + 1] = v[1]; np[3 * (size_t)po + 2] = v[2]; }
    }
  } else {
    const int pt = S.obs_pt[s0];
    double v[3] = {0.0, 0.0, 0.0};
    for (int k = s0 + lane; k < s0 + n; k += 32) {
      const ERec<NR> e = LoadER<NR>(ER, ES, k);
#pragma unroll
      for (int a = 0; a < 3; ++a) v[a] += sq(e, a);
    }
    WarpAllSum<3>(v);
    const int po = S.p_off[pt];
    if (lane == 0 && po >= 0) { np[3 * (size_t)po] = v[0]; np[3 * (size_t)po + 1] = v[1]; np[3 * (size_t)po + 2] = v[2]; }
  }
}

// ---- point blocks: C_p = E^T E + D_p^2, its inverse, and g_p = E^T r -----------------
template <int NR>
__global__ void __launch_bounds__(128) point_blocks_kernel(StructView S, TileView T, const D4* __restrict__ ER,
                                                           const D4* __restrict__ ES,
                                                           const double* __restrict__ Dp, double* __restrict__ Cinv,
                                                           double* __restrict__ Cfac,
                                                           double* __restrict__ gp, double* __restrict__ EtE, DeviceScalars* sc) {
  const int tile = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (tile >= T.n_tiles) return;
  const int lane = threadIdx.x & 31;
  const int s0 = T.tile_start[tile], n = T.tile_start[tile + 1] - s0;
  double v[9];
#pragma unroll
  for (int a = 0; a < 9; ++a) v[a] = 0.0;
  auto accum = [&](const ERec<NR>& e) {
    v[0] += e.E[0] * e.E[0] + e.E[3] * e.E[3];
    v[1] += e.E[0] * e.E[1] + e.E[3] * e.E[4];
    v[2] += e.E[0] * e.E[2] + e.E[3] * e.E[5];
    v[3] += e.E[1] * e.E[1] + e.E[4] * e.E[4];
    v[4] += e.E[1] * e.E[2] + e.E[4] * e.E[5];
    v[5] += e.E[2] * e.E[2] + e.E[5] * e.E[5];
    v[6] += e.E[0] * e.r[0] + e.E[3] * e.r[1];
    v[7] += e.E[1] * e.r[0] + e.E[4] * e.r[1];
    v[8] += e.E[2] * e.r[0] + e.E[5] * e.r[1];
    if constexpr (NR == 3) {
      v[0] += e.E[6] * e.E[6]; v[1] += e.E[6] * e.E[7]; v[2] += e.E[6] * e.E[8];
      v[3] += e.E[7] * e.E[7]; v[4] += e.E[7] * e.E[8]; v[5] += e.E[8] * e.E[8];
      v[6] += e.E[6] * e.r[2]; v[7] += e.E[7] * e.r[2]; v[8] += e.E[8] * e.r[2];
    }
  };
  int pt; bool writer;
  if (n <= 32) {
    const bool valid = lane < n;
    pt = valid ? S.obs_pt[s0 + lane] : -1 - lane;
    if (valid) accum(LoadER<NR>(ER, ES, s0 + lane));
    const Seg sg = MakeSeg(pt);
    SegSum<9>(sg, v);
    writer = valid && sg.is_head;
  } else {
    pt = S.obs_pt[s0];
    for (int k = s0 + lane; k < s0 + n; k += 32) accum(LoadER<NR>(ER, ES, k));
    WarpAllSum<9>(v);
    writer = lane == 0;
  }
  if (writer) {
    const int po = S.p_off[pt];
    if (po >= 0) {
      if (EtE) {
#pragma unroll
        for (int a = 0; a < 6; ++a) EtE[6 * (size_t)po + a] = v[a];
      }
      if (Dp) {
        const double d0 = Dp[3 * (size_t)po], d1 = Dp[3 * (size_t)po + 1], d2 = Dp[3 * (size_t)po + 2];
        const double c[6] = {v[0] + d0 * d0, v[1], v[2], v[3] + d1 * d1, v[4], v[5] + d2 * d2};
        double inv[6], fac[6];
        if (!InvertSym3(c, inv, fac)) {
          sc->solver_failed = 1;
#pragma unroll
          for (int a = 0; a < 6; ++a) { inv[a] = 0.0; fac[a] = 0.0; }
        }
#pragma unroll
        for (int a = 0; a < 6; ++a) Cinv[6 * (size_t)po + a] = inv[a];
        if (Cfac) {
#pragma unroll
          for (int a = 0; a < 6; ++a) Cfac[6 * (size_t)po + a] = fac[a];
        }
      }
      gp[3 * (size_t)po] = v[6]; gp[3 * (size_t)po + 1] = v[7]; gp[3 * (size_t)po + 2] = v[8];
    }
  }
}

// The same blocks from the cached E_p^T E_p when only the LM diagonal changed (every iteration changes D_p; E changes
// with an accepted step only, and the pass that follows it -- the gradient's -- leaves E^T E behind): one thread per
// variable point, 72 B in and 96 B out instead of a pass over all ER records (C4: 0.144 -> ~0.04 ms).
__global__ void __launch_bounds__(256) point_blocks_refresh_kernel(int n_pvar, const double* __restrict__ EtE,
                                                                   const double* __restrict__ Dp, double* __restrict__ Cinv,
                                                                   double* __restrict__ Cfac, DeviceScalars* sc) {
  const int po = blockIdx.x * 256 + threadIdx.x;
  if (po >= n_pvar) return;
  const double* v = EtE + 6 * (size_t)po;
  const double d0 = Dp[3 * (size_t)po], d1 = Dp[3 * (size_t)po + 1], d2 = Dp[3 * (size_t)po + 2];
  const double c[6] = {v[0] + d0 * d0, v[1], v[2], v[3] + d1 * d1, v[4], v[5] + d2 * d2};
  double inv[6], fac[6];
  if (!InvertSym3(c, inv, fac)) {
    sc->solver_failed = 1;
#pragma unroll
    for (int a = 0; a < 6; ++a) { inv[a] = 0.0; fac[a] = 0.0; }
  }
#pragma unroll
  for (int a = 0; a < 6; ++a) { Cinv[6 * (size_t)po + a] = inv[a]; Cfac[6 * (size_t)po + a] = fac[a]; }
}

// ---- e_k = r_k - E_k C_p^{-1} g_p  (right-hand side of the reduced system, before F^T) ----
template <int NR>
__global__ void __launch_bounds__(256) rhs_e_kernel(StructView S, const D4* __restrict__ ER, const D4* __restrict__ ES,
                                                    const double* __restrict__ Cinv, const double* __restrict__ gp,
                                                    double2* __restrict__ out, double* __restrict__ out3) {
  const int k = blockIdx.x * 256 + threadIdx.x;
  if (k >= S.n_obs) return;
  const ERec<NR> e = LoadER<NR>(ER, ES, k);
  double f[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) f[r] = e.r[r];
  const int po = S.p_off[S.obs_pt[k]];
  if (po >= 0) {
    const double g[3] = {gp[3 * (size_t)po], gp[3 * (size_t)po + 1], gp[3 * (size_t)po + 2]};
    double w[3];
    Sym3Mul(Cinv + 6 * (size_t)po, g, w);
    ESub<NR>(e, w, f);
  }
  StoreVec<NR>(out, out3, k, f);
}

// ---- implicit Schur product, point pass:  g_k = F_k x - E_k C_p^{-1} sum_{j in p} E_j^T F_j x ----
template <int DC, int NR>
__global__ void __launch_bounds__(128) matvec_point_kernel(StructView S, TileView T, const D4* __restrict__ ER,
                                                           const D4* __restrict__ F, const D4* __restrict__ ES,
                                                           const D4* __restrict__ FS, const double* __restrict__ Cinv,
                                                           const double* __restrict__ xi, const double* __restrict__ xc,
                                                           double2* __restrict__ out, double* __restrict__ out3,
                                                           const int* __restrict__ skip) {
  if (skip && *skip != 0) return;  // CG already stopped (the host polls its state only every few iterations)
  const int tile = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (tile >= T.n_tiles) return;
  const int lane = threadIdx.x & 31;
  const int s0 = T.tile_start[tile], n = T.tile_start[tile + 1] - s0;
  if (n <= 32) {
    const bool valid = lane < n;
    const int k = s0 + lane;
    int pt = -1 - lane;
    double f[NR];
#pragma unroll
    for (int r = 0; r < NR; ++r) f[r] = 0.0;
    ERec<NR> e;
    double u[3] = {0.0, 0.0, 0.0};
    if (valid) {
      pt = S.obs_pt[k];
      const int i = S.obs_img[k];
      FRow<DC, NR> row;
      LoadRow<DC, NR>(F, FS, k, row);
      RowTimesX<DC, NR>(row, xi, xc, i, DC > 6 ? S.img_cam[i] : 0, f);
      e = LoadER<NR>(ER, ES, k);
      EtAdd<NR>(e, f, u);
    }
    const Seg sg = MakeSeg(pt);
    SegSum<3>(sg, u);
    if (valid) {
      const int po = S.p_off[pt];
      if (po >= 0) {
        double w[3];
        Sym3Mul(Cinv + 6 * (size_t)po, u, w);
        ESub<NR>(e, w, f);
      }
      StoreVec<NR>(out, out3, k, f);
    }
  } else {  // one long point: two strided passes
    const int pt = S.obs_pt[s0];
    const int po = S.p_off[pt];
    double u[3] = {0.0, 0.0, 0.0};
    for (int k = s0 + lane; k < s0 + n; k += 32) {
      const int i = S.obs_img[k];
      FRow<DC, NR> row;
      LoadRow<DC, NR>(F, FS, k, row);
      double f[NR];
      RowTimesX<DC, NR>(row, xi, xc, i, DC > 6 ? S.img_cam[i] : 0, f);
      const ERec<NR> e = LoadER<NR>(ER, ES, k);
      EtAdd<NR>(e, f, u);
      StoreVec<NR>(out, out3, k, f);
    }
    WarpAllSum<3>(u);
    if (po >= 0) {
      double w[3];
      Sym3Mul(Cinv + 6 * (size_t)po, u, w);
      for (int k = s0 + lane; k < s0 + n; k += 32) {
        const ERec<NR> e = LoadER<NR>(ER, ES, k);
        double f[NR];
        LoadVec<NR>(out, out3, k, f);
        ESub<NR>(e, w, f);
        StoreVec<NR>(out, out3, k, f);
      }
    }
  }
}

// ---- back-substitution y_p = C_p^{-1}(g_p - sum E^T F y_c), fused with the model cost change
//      -(J step)^T (r + J step / 2), step = -y   ----
template <int DC, int NR>
__global__ void __launch_bounds__(128) backsub_kernel(StructView S, TileView T, const D4* __restrict__ ER,
                                                      const D4* __restrict__ F, const D4* __restrict__ ES,
                                                      const D4* __restrict__ FS, const double* __restrict__ Cinv,
                                                      const double* __restrict__ gp, const double* __restrict__ yi,
                                                      const double* __restrict__ yc, double* __restrict__ yp,
                                                      double* __restrict__ partial, unsigned* ticket,
                                                      double* model_cost_change) {
  const int lane = threadIdx.x & 31;
  double acc = 0.0;
  // m = -(F y_c) - E w;  acc += m . (r + m / 2)
  auto model = [&](const ERec<NR>& e, const double (&f)[NR], const double (&w)[3], bool has_w) {
    double m[NR];
#pragma unroll
    for (int r = 0; r < NR; ++r) m[r] = -f[r];
    if (has_w) ESub<NR>(e, w, m);
#pragma unroll
    for (int r = 0; r < NR; ++r) acc += m[r] * (e.r[r] + 0.5 * m[r]);
  };
  for (int tile = blockIdx.x * 4 + (threadIdx.x >> 5); tile < T.n_tiles; tile += gridDim.x * 4) {
    const int s0 = T.tile_start[tile], n = T.tile_start[tile + 1] - s0;
    if (n <= 32) {
      const bool valid = lane < n;
      const int k = s0 + lane;
      int pt = -1 - lane;
      double f[NR];
#pragma unroll
      for (int r = 0; r < NR; ++r) f[r] = 0.0;
      ERec<NR> e;
      double u[3] = {0.0, 0.0, 0.0};
      if (valid) {
        pt = S.obs_pt[k];
        const int i = S.obs_img[k];
        FRow<DC, NR> row;
        LoadRow<DC, NR>(F, FS, k, row);
        RowTimesX<DC, NR>(row, yi, yc, i, DC > 6 ? S.img_cam[i] : 0, f);
        e = LoadER<NR>(ER, ES, k);
        EtAdd<NR>(e, f, u);
      }
      const Seg sg = MakeSeg(pt);
      SegSum<3>(sg, u);
      if (valid) {
        const int po = S.p_off[pt];
        double w[3] = {0.0, 0.0, 0.0};
        if (po >= 0) {
          const double d[3] = {gp[3 * (size_t)po] - u[0], gp[3 * (size_t)po + 1] - u[1], gp[3 * (size_t)po + 2] - u[2]};
          Sym3Mul(Cinv + 6 * (size_t)po, d, w);
          if (sg.is_head) { yp[3 * (size_t)po] = w[0]; yp[3 * (size_t)po + 1] = w[1]; yp[3 * (size_t)po + 2] = w[2]; }
        }
        model(e, f, w, po >= 0);
      }
    } else {
      const int pt = S.obs_pt[s0];
      const int po = S.p_off[pt];
      double u[3] = {0.0, 0.0, 0.0};
      for (int k = s0 + lane; k < s0 + n; k += 32) {
        const int i = S.obs_img[k];
        FRow<DC, NR> row;
        LoadRow<DC, NR>(F, FS, k, row);
        double f[NR];
        RowTimesX<DC, NR>(row, yi, yc, i, DC > 6 ? S.img_cam[i] : 0, f);
        const ERec<NR> e = LoadER<NR>(ER, ES, k);
        EtAdd<NR>(e, f, u);
      }
      WarpAllSum<3>(u);
      double w[3] = {0.0, 0.0, 0.0};
      if (po >= 0) {
        const double d[3] = {gp[3 * (size_t)po] - u[0], gp[3 * (size_t)po + 1] - u[1], gp[3 * (size_t)po + 2] - u[2]};
        Sym3Mul(Cinv + 6 * (size_t)po, d, w);
        if (lane == 0) { yp[3 * (size_t)po] = w[0]; yp[3 * (size_t)po + 1] = w[1]; yp[3 * (size_t)po + 2] = w[2]; }
      }
      for (int k = s0 + lane; k < s0 + n; k += 32) {
        const int i = S.obs_img[k];
        FRow<DC, NR> row;
        LoadRow<DC, NR>(F, FS, k, row);
        double f[NR];
        RowTimesX<DC, NR>(row, yi, yc, i, DC > 6 ? S.img_cam[i] : 0, f);
        const ERec<NR> e = LoadER<NR>(ER, ES, k);
        model(e, f, w, true);  // (w = 0 for a constant point: E is zero there anyway)
      }
    }
  }
  // model cost change = -sum; grid-wide sum finished by the last block (reduce.cuh)
  double v[1] = {-acc};
  double* const out[1] = {model_cost_change};
  GridSum<128, 1>(v, partial, ticket, out);
}

// ---- camera-side segmented reduction over image-major chunks ------------------------
// MODE 0: F^T v (v per observation)   MODE 1: F^T r (r from ER)   MODE 2: column square norms
// MODE 3: F^T r into partial AND the column square norms into partial2, one pass over F (after a new linearisation both
//         are wanted: the gradient now, the norms by the next LM diagonal)
template <int DC, int MODE, int NR>
__global__ void __launch_bounds__(256) cam_reduce_kernel(ChunkView C, const D4* __restrict__ F,
                                                         const D4* __restrict__ ER, const D4* __restrict__ FS,
                                                         const D4* __restrict__ ES, const double2* __restrict__ v,
                                                         const double* __restrict__ v3,
                                                         double* __restrict__ partial, const int* __restrict__ skip,
                                                         double* __restrict__ partial2 = nullptr) {
  if (skip && *skip != 0) return;
  const int ch = blockIdx.x;
  const int s0 = C.chunk_start[ch], n = C.chunk_start[ch + 1] - s0;
  double acc[DC], acc2[MODE == 3 ? DC : 1];
#pragma unroll
  for (int j = 0; j < DC; ++j) acc[j] = 0.0;
#pragma unroll
  for (int j = 0; j < (MODE == 3 ? DC : 1); ++j) acc2[j] = 0.0;
  if (threadIdx.x < n) {
    const int k = C.cam_order[s0 + threadIdx.x];
    FRow<DC, NR> row;
    LoadRow<DC, NR>(F, FS, k, row);
    double w[NR];
    if (MODE == 0) {
      LoadVec<NR>(v, v3, k, w);
    } else if (MODE == 1 || MODE == 3) {
      const double2 t = reinterpret_cast<const double2*>(ER + EUnit(k, 1))[1];  // r = unit 1, z w
      w[0] = t.x; w[1] = t.y;
      if constexpr (NR == 3) w[2] = reinterpret_cast<const double*>(ES + k)[3];
    }
    if (MODE != 2) {
#pragma unroll
      for (int j = 0; j < DC; ++j) {
        acc[j] = row.a[j].x * w[0] + row.a[j].y * w[1];
        if constexpr (NR == 3) acc[j] += row.s[j] * w[2];
      }
    }
    if (MODE == 2 || MODE == 3) {
#pragma unroll
      for (int j = 0; j < DC; ++j) {
        double q = row.a[j].x * row.a[j].x + row.a[j].y * row.a[j].y;
        if constexpr (NR == 3) q += row.s[j] * row.s[j];
        if (MODE == 2) acc[j] = q; else acc2[MODE == 3 ? j : 0] = q;
      }
    }
  }
  __shared__ double sm[8][DC];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < DC; ++j) {
    double a = acc[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
    if (lane == 0) sm[w][j] = a;
  }
  __syncthreads();
  if (threadIdx.x < DC) {
    double a = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) a += sm[i][threadIdx.x];
    partial[(size_t)ch * DC + threadIdx.x] = a;
  }
  if (MODE == 3) {  // same fixed-order reduction for the norms
    __syncthreads();
#pragma unroll
    for (int j = 0; j < DC; ++j) {
      double a = acc2[MODE == 3 ? j : 0];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
      if (lane == 0) sm[w][j] = a;
    }
    __syncthreads();
    if (threadIdx.x < DC) {
      double a = 0.0;
#pragma unroll
      for (int i = 0; i < 8; ++i) a += sm[i][threadIdx.x];
      partial2[(size_t)ch * DC + threadIdx.x] = a;
    }
  }
}

// sums the chunk partials of each image (pose slots) / each camera (intrinsics slots)
// in a fixed order into the compact camera-space vector; optionally adds D_c^2 x.
__global__ void __launch_bounds__(128) cam_finalize_kernel(StructView S, ChunkView C, const double* __restrict__ partial,
                                                           double* __restrict__ out, const double* __restrict__ Dc,
                                                           const double* __restrict__ x) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int ni = S.n_img * 6;
  int col; double a = 0.0;
  if (t < ni) {
    col = S.col_img[t];
    if (col < 0) return;
    const int i = t / 6, j = t - 6 * i;
    for (int ch = C.img_chunk_ptr[i]; ch < C.img_chunk_ptr[i + 1]; ++ch) a += partial[(size_t)ch * S.dc + j];
  } else if (t < ni + S.n_cam * 12) {
    const int u = t - ni;
    col = S.col_cam[u];
    if (col < 0) return;
    const int c = u / 12, j = u - 12 * c;
    for (int q = C.cam_img_ptr[c]; q < C.cam_img_ptr[c + 1]; ++q) {
      const int i = C.cam_img_list[q];
      for (int ch = C.img_chunk_ptr[i]; ch < C.img_chunk_ptr[i + 1]; ++ch) a += partial[(size_t)ch * S.dc + 6 + j];
    }
  } else {
    return;
  }
  if (Dc) a += Dc[col] * Dc[col] * x[col];
  out[col] = a;
}

__global__ void add_diag_kernel(int n, double* __restrict__ y, const double* __restrict__ Dc, const double* __restrict__ x) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) y[j] += Dc[j] * Dc[j] * x[j];
}

// ---- Jacobi scaling ---------------------------------------------
__global__ void jacobi_scale_kernel(int n, const double* __restrict__ sqnorm, double* __restrict__ scale) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) scale[j] = 1.0 / (1.0 + sqrt(sqnorm[j]));
}
template <int DC, int NR>
__global__ void __launch_bounds__(256) scale_jacobian_kernel(StructView S, const double* __restrict__ si,
                                                             const double* __restrict__ sc, const double* __restrict__ sp,
                                                             D4* __restrict__ F, D4* __restrict__ ER,
                                                             D4* __restrict__ FS, D4* __restrict__ ES) {
  const int k = blockIdx.x * 256 + threadIdx.x;
  if (k >= S.n_obs) return;
  const int i = S.obs_img[k];
  const int c = DC > 6 ? S.img_cam[i] : 0;
  double sl[DC];  // scale of every slot
#pragma unroll
  for (int j = 0; j < 6; ++j) sl[j] = si[6 * (size_t)i + j];
#pragma unroll
  for (int j = 6; j < DC; ++j) sl[j] = sc[12 * (size_t)c + j - 6];
#pragma unroll
  for (int jj = 0; jj < DC / 2; ++jj) {
    D4 v = Ld256(F + FUnit(k, jj, DC / 2));
    v.x *= sl[2 * jj]; v.y *= sl[2 * jj]; v.z *= sl[2 * jj + 1]; v.w *= sl[2 * jj + 1];
    St256(F + FUnit(k, jj, DC / 2), v);
  }
  if constexpr (NR == 3) {
    constexpr int NQ = FSUnits(DC);
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      D4 v = Ld256(FS + FSUnit(k, q, NQ));
      v.x *= sl[4 * q];
      v.y *= 4 * q + 1 < DC ? sl[4 * q + 1 < DC ? 4 * q + 1 : 0] : 0.0;
      v.z *= 4 * q + 2 < DC ? sl[4 * q + 2 < DC ? 4 * q + 2 : 0] : 0.0;
      v.w *= 4 * q + 3 < DC ? sl[4 * q + 3 < DC ? 4 * q + 3 : 0] : 0.0;
      St256(FS + FSUnit(k, q, NQ), v);
    }
  }
  const int po = S.p_off[S.obs_pt[k]];
  if (po >= 0) {
    const double s0 = sp[3 * (size_t)po], s1 = sp[3 * (size_t)po + 1], s2 = sp[3 * (size_t)po + 2];
    D4 a = Ld256(ER + EUnit(k, 0)), b = Ld256(ER + EUnit(k, 1));
    a.x *= s0; a.y *= s1; a.z *= s2; a.w *= s0; b.x *= s1; b.y *= s2;  // E row-major 2x3: (E0 E1 E2 | E3 E4 E5)
    St256(ER + EUnit(k, 0), a); St256(ER + EUnit(k, 1), b);
    if constexpr (NR == 3) {
      D4 e = Ld256(ES + k);
      e.x *= s0; e.y *= s1; e.z *= s2;
      St256(ES + k, e);
    }
  }
}

// LM diagonal: diag = clamp(colnorm^2), D = sqrt(diag / radius)
// camera columns [0, nc) and point columns [nc, nc + np) in one launch
__global__ void lm_diag_kernel(int nc, int np, const double* __restrict__ sq_c, double* __restrict__ diag_c,
                               double* __restrict__ D_c, const double* __restrict__ sq_p, double* __restrict__ diag_p,
                               double* __restrict__ D_p, int refresh, double lo, double hi, const DeviceScalars* sc) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nc + np) return;
  const double* sq = sq_c; double* diag = diag_c; double* D = D_c;
  if (j >= nc) { j -= nc; sq = sq_p; diag = diag_p; D = D_p; }
  double d = diag[j];
  if (refresh) { d = fmin(fmax(sq[j], lo), hi); diag[j] = d; }
  D[j] = sqrt(d / sc->radius);
}

// max |g_j / scale_j| over the camera part and the point part in one launch.  max is order-independent, and
// non-negative doubles order like their bit patterns, so an integer atomicMax gives a deterministic result.
// *out must be zero on entry.
__global__ void __launch_bounds__(256) grad_maxnorm_kernel(int nc, const double* __restrict__ gc, const double* __restrict__ sc_,
                                                           int np, const double* __restrict__ gp, const double* __restrict__ sp,
                                                           double* out) {
  double m = 0.0;
  for (int j = blockIdx.x * 256 + threadIdx.x; j < nc + np; j += gridDim.x * 256) {
    const double v = j < nc ? gc[j] / sc_[j] : gp[j - nc] / sp[j - nc];
    m = fmax(m, fabs(v));  // fmax drops NaN, as std::max(m, |v|) over a NaN-free gradient does
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_down_sync(0xffffffffu, m, o));
  __shared__ double sm[8];
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < 8; ++i) m = fmax(m, sm[i]);
    atomicMax(reinterpret_cast<unsigned long long*>(out), (unsigned long long)__double_as_longlong(m));
  }
}

// max |g_j / scale_j| over a vector -> partial maxima
__global__ void __launch_bounds__(256) absmax_kernel(int n, const double* __restrict__ g, const double* __restrict__ scale,
                                                     double* __restrict__ partial) {
  double m = 0.0;
  for (int j = blockIdx.x * 256 + threadIdx.x; j < n; j += gridDim.x * 256) m = fmax(m, fabs(scale ? g[j] / scale[j] : g[j]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_down_sync(0xffffffffu, m, o));
  __shared__ double sm[8];
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < 8; ++i) m = fmax(m, sm[i]);
    partial[blockIdx.x] = m;
  }
}
__global__ void absmax_final_kernel(int n, const double* __restrict__ partial, double* out, int accumulate) {
  double m = accumulate ? *out : 0.0;
  for (int i = 0; i < n; ++i) m = fmax(m, partial[i]);
  *out = m;
}
__global__ void sum_partial_neg_kernel(int n, const double* __restrict__ partial, double* out) {
  // fixed-order sum of partial[2*i]; writes the negated total (model cost change)
  __shared__ double sm[32];
  double a = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) a += partial[2 * i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x < 32) {
    a = sm[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
    if (threadIdx.x == 0) *out = -a;
  }
}
}  // namespace sba
