// Host launchers for the elimination / product / update kernels (linalg_kernels.cuh)
// and the device-resident trust-region bookkeeping.
// Reference: the inside of ceres::Solve as called at
//   /root/reference/src/optim/bundle_adjustment.cc:306   (SURVEY.md App. B.3-B.5).
#include "device_math.cuh"
#include "linalg_kernels.cuh"
#include "reduce.cuh"

namespace sba {

static int Expand(Problem* p, const double* x, double* xi, double* xc) {
  const int n = p->n_img * 6 + p->n_cam * 12;
  expand_kernel<<<DivUp(n, 256), 256, 0, p->stream>>>(p->Struct(), x, xi, xc);
  p->launches++;
  return SBA_OK;
}

// mode 0: F^T v, 1: F^T r, 2: column square norms.  out (compact) = sum over LOCAL
// observations, all-reduced over ranks, plus Dc^2 x when Dc != nullptr.
static int CamReduce(Problem* p, int mode, const double2* v, double* out, const double* Dc, const double* x,
                     const int* skip = nullptr, bool defer_allreduce = false) {
  if (p->n_cdim == 0) return SBA_OK;
  if (p->n_chunks > 0) {
#define CALL(DC, NR)                                                                                                    \
  do {                                                                                                                  \
    if (mode == 0) cam_reduce_kernel<DC, 0, NR><<<p->n_chunks, 256, 0, p->stream>>>(p->Chunks(), p->d_F.p, p->d_ER.p, p->d_FS.p, p->d_ES.p, v, p->d_g1.p, p->d_chunk_partial.p, skip); \
    else if (mode == 1) cam_reduce_kernel<DC, 1, NR><<<p->n_chunks, 256, 0, p->stream>>>(p->Chunks(), p->d_F.p, p->d_ER.p, p->d_FS.p, p->d_ES.p, v, p->d_g1.p, p->d_chunk_partial.p, skip); \
    else cam_reduce_kernel<DC, 2, NR><<<p->n_chunks, 256, 0, p->stream>>>(p->Chunks(), p->d_F.p, p->d_ER.p, p->d_FS.p, p->d_ES.p, v, p->d_g1.p, p->d_chunk_partial.p, skip); \
  } while (0)
    SBA_DCNR_SWITCH(p->dc, p->nr, CALL)
#undef CALL
    p->launches++;
  }
  const bool multi = p->comm != nullptr;
  const int n = p->n_img * 6 + p->n_cam * 12;
  cam_finalize_kernel<<<DivUp(n, 128), 128, 0, p->stream>>>(p->Struct(), p->Chunks(), p->d_chunk_partial.p, out,
                                                           multi ? nullptr : Dc, x);
  p->launches++;
  if (multi && !defer_allreduce) {
    int rc = AllReduceSum(p, out, p->n_cdim);
    if (rc) return rc;
    if (Dc) { add_diag_kernel<<<DivUp(p->n_cdim, 256), 256, 0, p->stream>>>(p->n_cdim, out, Dc, x); p->launches++; }
  }
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

// chunk partials [n_chunks][dc] (written by a kernel of another file) -> compact camera-space vector `out`
int FinalizeCameraVector(Problem* p, const double* d_partial, double* d_out) {
  if (p->n_cdim == 0) return SBA_OK;
  const int n = p->n_img * 6 + p->n_cam * 12;
  cam_finalize_kernel<<<DivUp(n, 128), 128, 0, p->stream>>>(p->Struct(), p->Chunks(), d_partial, d_out, nullptr, nullptr);
  p->launches++;
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

int ColumnNorms(Problem* p, double* d_nc, double* d_np, bool camera_part) {
  int rc = camera_part ? CamReduce(p, 2, nullptr, d_nc, nullptr, nullptr) : SBA_OK;
  if (rc) return rc;
  if (p->n_pvar > 0 && p->n_tiles > 0) {
    if (p->nr == 3) point_sqnorm_kernel<3><<<DivUp(p->n_tiles, 4), 128, 0, p->stream>>>(p->Struct(), p->Tiles(), p->d_ER.p, p->d_ES.p, d_np);
    else point_sqnorm_kernel<2><<<DivUp(p->n_tiles, 4), 128, 0, p->stream>>>(p->Struct(), p->Tiles(), p->d_ER.p, p->d_ES.p, d_np);
    p->launches++;
  }
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

// Jacobi scaling at iteration 0: scale = 1/(1+sqrt(colnorm^2)); the stored Jacobian is
// rescaled in place, later evaluations apply the scale when storing.
int InitJacobiScale(Problem* p) {
  int rc = ColumnNorms(p, p->d_sq_c.p, p->d_sq_p.p);
  if (rc) return rc;
  if (p->n_cdim > 0) { jacobi_scale_kernel<<<DivUp(p->n_cdim, 256), 256, 0, p->stream>>>(p->n_cdim, p->d_sq_c.p, p->d_scale_c.p); p->launches++; }
  if (p->n_pvar > 0) { jacobi_scale_kernel<<<DivUp(3 * p->n_pvar, 256), 256, 0, p->stream>>>(3 * p->n_pvar, p->d_sq_p.p, p->d_scale_p.p); p->launches++; }
  Expand(p, p->d_scale_c.p, p->d_scale_i.p, p->d_scale_k.p);
  if (p->n_obs > 0) {
#define CALL(DC, NR) scale_jacobian_kernel<DC, NR><<<DivUp(p->n_obs, 256), 256, 0, p->stream>>>(p->Struct(), p->d_scale_i.p, p->d_scale_k.p, p->d_scale_p.p, p->d_F.p, p->d_ER.p, p->d_FS.p, p->d_ES.p)
    SBA_DCNR_SWITCH(p->dc, p->nr, CALL)
#undef CALL
    p->launches++;
  }
  p->colnorm_current = false;  // the stored Jacobian was rescaled
  p->ete_valid = false;
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

// gradient max-norm of the UNSCALED problem: max |(J^T r)_j| = max |g_scaled_j / scale_j|.
// The point part g_p = E^T r must be current (PointBlocks).
// `with_norms`: the same pass over F also leaves the camera column square norms in d_sq_c (cam_reduce MODE 3), which
// the next LM diagonal would otherwise recompute with a pass of its own.
int Gradient(Problem* p, bool with_norms) {
  int rc = SBA_OK;
  if (with_norms && p->n_cdim > 0) {  // (every rank takes the same branch: the collectives below must pair up)
    double* part2 = p->d_chunk_partial.p + (size_t)p->n_chunks * kMaxDC;
#define CALL(DC, NR) cam_reduce_kernel<DC, 3, NR><<<p->n_chunks, 256, 0, p->stream>>>(p->Chunks(), p->d_F.p, p->d_ER.p, p->d_FS.p, p->d_ES.p, nullptr, nullptr, p->d_chunk_partial.p, nullptr, part2)
    if (p->n_chunks > 0) { SBA_DCNR_SWITCH(p->dc, p->nr, CALL) }
#undef CALL
    const int n = p->n_img * 6 + p->n_cam * 12;
    cam_finalize_kernel<<<DivUp(n, 128), 128, 0, p->stream>>>(p->Struct(), p->Chunks(), p->d_chunk_partial.p, p->d_grad_c.p, nullptr, nullptr);
    cam_finalize_kernel<<<DivUp(n, 128), 128, 0, p->stream>>>(p->Struct(), p->Chunks(), part2, p->d_sq_c.p, nullptr, nullptr);
    p->launches += 3;
    SBA_CUDA(cudaGetLastError());
    // d_sq_c lies right behind d_grad_c, and two more doubles behind it take the cost sums along when they are pending
    const size_t n2 = 2 * (size_t)p->n_cdim;
    double* cost = &p->d_scalars.p->cost;
    if (p->cost_unreduced) SBA_CUDA(cudaMemcpyAsync(p->d_grad_c.p + n2, cost, 2 * sizeof(double), cudaMemcpyDeviceToDevice, p->stream));
    if ((rc = AllReduceSum(p, p->d_grad_c.p, n2 + (p->cost_unreduced ? 2 : 0)))) return rc;
    if (p->cost_unreduced) SBA_CUDA(cudaMemcpyAsync(cost, p->d_grad_c.p + n2, 2 * sizeof(double), cudaMemcpyDeviceToDevice, p->stream));
    p->cost_unreduced = false;
    p->colnorm_current = true;
  } else {
    if (p->cost_unreduced) { if ((rc = AllReduceSum(p, &p->d_scalars.p->cost, 2))) return rc; p->cost_unreduced = false; }
    rc = CamReduce(p, 1, nullptr, p->d_grad_c.p, nullptr, nullptr);
  }
  if (rc) return rc;
  double* out = &p->d_scalars.p->gradient_max_norm;
  SBA_CUDA(cudaMemsetAsync(out, 0, sizeof(double), p->stream));
  const int nb = std::min(1184, std::max(1, DivUp((int64_t)p->n_cdim + 3 * (int64_t)p->n_pvar, 256)));
  grad_maxnorm_kernel<<<nb, 256, 0, p->stream>>>(p->n_cdim, p->d_grad_c.p, p->d_scale_c.p, 3 * p->n_pvar, p->d_gp.p,
                                                 p->d_scale_p.p, out);
  p->launches += 1;
  SBA_CUDA(cudaGetLastError());
  return AllReduceMax(p, out, 1);
}

__global__ void fill_kernel(int n, double* __restrict__ x, double v) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) x[j] = v;
}
int FillOnes(Problem* p) {
  auto fill = [&](double* x, size_t n) {
    if (n == 0) return;
    fill_kernel<<<DivUp(n, 256), 256, 0, p->stream>>>((int)n, x, 1.0);
    p->launches++;
  };
  fill(p->d_scale_c.p, p->n_cdim); fill(p->d_scale_p.p, 3 * (size_t)p->n_pvar);
  fill(p->d_scale_i.p, 6 * (size_t)p->n_img); fill(p->d_scale_k.p, 12 * (size_t)p->n_cam);
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

int PrepareLMDiagonal(Problem* p, const sba_solver_options& o) {
  const int refresh = p->h_scalars->reuse_diagonal ? 0 : 1;
  if (refresh) {  // the camera norms may already be there: Gradient(with_norms) of the linearisation that made this J
    int rc = ColumnNorms(p, p->d_sq_c.p, p->d_sq_p.p, !p->colnorm_current);
    if (rc) return rc;
    p->colnorm_current = false;  // (host bookkeeping only: a replayed graph keeps the launch sequence it was captured with)
  }
  if (p->n_cdim + p->n_pvar > 0) {
    lm_diag_kernel<<<DivUp((int64_t)p->n_cdim + 3 * (int64_t)p->n_pvar, 256), 256, 0, p->stream>>>(
        p->n_cdim, 3 * p->n_pvar, p->d_sq_c.p, p->d_diag_c.p, p->d_Dc.p, p->d_sq_p.p, p->d_diag_p.p, p->d_Dp.p, refresh,
        o.min_lm_diagonal, o.max_lm_diagonal, p->d_scalars.p);
    p->launches++;
  }
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

// grad_only: only g_p = E^T r (no LM diagonal needed yet)
int PointBlocks(Problem* p, bool grad_only) {
  if (p->n_pvar > 0 && p->n_tiles > 0) {
    if (!grad_only && p->ete_valid) {  // E unchanged since the last pass over ER: only D_p^2 is new
      point_blocks_refresh_kernel<<<DivUp(p->n_pvar, 256), 256, 0, p->stream>>>(p->n_pvar, p->d_EtE.p, p->d_Dp.p, p->d_Cinv.p,
                                                                                 p->d_Cfac.p, p->d_scalars.p);
    } else if (p->nr == 3) {
      point_blocks_kernel<3><<<DivUp(p->n_tiles, 4), 128, 0, p->stream>>>(p->Struct(), p->Tiles(), p->d_ER.p, p->d_ES.p,
                                                                        grad_only ? nullptr : p->d_Dp.p, p->d_Cinv.p,
                                                                        p->d_Cfac.p, p->d_gp.p, p->d_EtE.p, p->d_scalars.p);
    } else {
      point_blocks_kernel<2><<<DivUp(p->n_tiles, 4), 128, 0, p->stream>>>(p->Struct(), p->Tiles(), p->d_ER.p, p->d_ES.p,
                                                                        grad_only ? nullptr : p->d_Dp.p, p->d_Cinv.p,
                                                                        p->d_Cfac.p, p->d_gp.p, p->d_EtE.p, p->d_scalars.p);
    }
    p->ete_valid = true;
    p->launches++;
  }
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

// rhs = F^T (r - E C^{-1} g)   (reduced system right-hand side; Ceres solves S y = rhs, step = -y)
int ReducedRhs(Problem* p, bool defer_allreduce) {
  if (p->n_obs > 0) {
    if (p->nr == 3)
      rhs_e_kernel<3><<<DivUp(p->n_obs, 256), 256, 0, p->stream>>>(p->Struct(), p->d_ER.p, p->d_ES.p, p->d_Cinv.p, p->d_gp.p,
                                                                 reinterpret_cast<double2*>(p->d_g2.p), p->d_g1.p);
    else
      rhs_e_kernel<2><<<DivUp(p->n_obs, 256), 256, 0, p->stream>>>(p->Struct(), p->d_ER.p, p->d_ES.p, p->d_Cinv.p, p->d_gp.p,
                                                                 reinterpret_cast<double2*>(p->d_g2.p), p->d_g1.p);
    p->launches++;
  }
  defer_allreduce = defer_allreduce && p->comm != nullptr;
  const int rc = CamReduce(p, 0, reinterpret_cast<const double2*>(p->d_g2.p), p->d_rhs.p, nullptr, nullptr, nullptr, defer_allreduce);
  p->rhs_unreduced = defer_allreduce;
  return rc;
}

// y = S x with S = F^T F + Dc^2 - sum_p (F^T E)_p C_p^{-1} (E^T F)_p, never formed
// (ImplicitSchurComplement::RightMultiply).
int SchurMatVec(Problem* p, const double* d_x, double* d_y, const int* skip) {
  if (p->bsr.enabled) return BlockSchurMatVec(p, d_x, d_y, skip);  // explicit block-sparse S (bsr.cu)
  Expand(p, d_x, p->d_xi.p, p->d_xc.p);
  if (p->n_tiles > 0) {
#define CALL(DC, NR) matvec_point_kernel<DC, NR><<<DivUp(p->n_tiles, 4), 128, 0, p->stream>>>(p->Struct(), p->Tiles(), p->d_ER.p, p->d_F.p, p->d_ES.p, p->d_FS.p, p->d_Cinv.p, p->d_xi.p, p->d_xc.p, reinterpret_cast<double2*>(p->d_g2.p), p->d_g1.p, skip)
    SBA_DCNR_SWITCH(p->dc, p->nr, CALL)
#undef CALL
    p->launches++;
  }
  return CamReduce(p, 0, reinterpret_cast<const double2*>(p->d_g2.p), d_y, p->d_Dc.p, d_x, skip);
}

// y_p and model_cost_change from y_c (d_yc)
int BackSubstitute(Problem* p, bool defer_allreduce) {
  Expand(p, p->d_yc.p, p->d_xi.p, p->d_xc.p);
  const int nb = std::min(DivUp(p->n_tiles, 4), p->n_sm * 16);  // grid-stride over warp tiles
  if (nb > 0) {
#define CALL(DC, NR) backsub_kernel<DC, NR><<<nb, 128, 0, p->stream>>>(p->Struct(), p->Tiles(), p->d_ER.p, p->d_F.p, p->d_ES.p, p->d_FS.p, p->d_Cinv.p, p->d_gp.p, p->d_xi.p, p->d_xc.p, p->d_yp.p, p->d_cost_partial.p, p->d_tickets.p + 2, &p->d_scalars.p->model_cost_change)
    SBA_DCNR_SWITCH(p->dc, p->nr, CALL)
#undef CALL
    p->launches++;
  } else {
    SBA_CUDA(cudaMemsetAsync(&p->d_scalars.p->model_cost_change, 0, sizeof(double), p->stream));
  }
  SBA_CUDA(cudaGetLastError());
  return defer_allreduce ? SBA_OK : AllReduceSum(p, &p->d_scalars.p->model_cost_change, 1);
}

// ---- x_plus = Plus(x, -y * scale) on every variable block (manifolds) ----
__global__ void __launch_bounds__(128) apply_step_kernel(StructView S, ParamsView P, const double* __restrict__ yc,
                                                         const double* __restrict__ yp, const double* __restrict__ scale_c,
                                                         const double* __restrict__ scale_p, double* __restrict__ pose2,
                                                         double* __restrict__ cam2, double* __restrict__ X2,
                                                         int count, double* __restrict__ partial, unsigned* ticket,
                                                         DeviceScalars* sc) {
  double step2_c = 0.0, x2_c = 0.0, step2_p = 0.0, x2_p = 0.0;
  for (int t = blockIdx.x * 128 + threadIdx.x; t < count; t += gridDim.x * 128) {
    double step2 = 0.0, x2 = 0.0;
    if (t < S.n_img) {
      const int i = t;
      double rec[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) rec[j] = P.pose[8 * (size_t)i + j];
      const int* ci = S.col_img + 6 * (size_t)i;
      if (ci[0] >= 0) {
        double d[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) d[j] = -yc[ci[j]] * scale_c[ci[j]];
        const double n = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        double q[4] = {rec[0], rec[1], rec[2], rec[3]};
        if (n != 0.0) {  // QuaternionManifold::Plus: [cos|d|, sin|d|/|d| d] * x
          const double s = sin(n) / n;
          const double z0 = cos(n), z1 = s * d[0], z2 = s * d[1], z3 = s * d[2];
          q[0] = z0 * rec[0] - z1 * rec[1] - z2 * rec[2] - z3 * rec[3];
          q[1] = z0 * rec[1] + z1 * rec[0] + z2 * rec[3] - z3 * rec[2];
          q[2] = z0 * rec[2] - z1 * rec[3] + z2 * rec[0] + z3 * rec[1];
          q[3] = z0 * rec[3] + z1 * rec[2] - z2 * rec[1] + z3 * rec[0];
        }
        double tt[3] = {rec[4], rec[5], rec[6]};
#pragma unroll
        for (int j = 0; j < 3; ++j) if (ci[3 + j] >= 0) tt[j] += -yc[ci[3 + j]] * scale_c[ci[3 + j]];
#pragma unroll
        for (int j = 0; j < 4; ++j) { step2 += (q[j] - rec[j]) * (q[j] - rec[j]); x2 += rec[j] * rec[j]; }
        if (ci[3] >= 0 || ci[4] >= 0 || ci[5] >= 0) {
#pragma unroll
          for (int j = 0; j < 3; ++j) { step2 += (tt[j] - rec[4 + j]) * (tt[j] - rec[4 + j]); x2 += rec[4 + j] * rec[4 + j]; }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) rec[j] = q[j];
#pragma unroll
        for (int j = 0; j < 3; ++j) rec[4 + j] = tt[j];
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) pose2[8 * (size_t)i + j] = rec[j];
    } else if (t < S.n_img + S.n_cam) {
      const int c = t - S.n_img;
      double prm[12];
#pragma unroll
      for (int j = 0; j < 12; ++j) prm[j] = P.cam[12 * (size_t)c + j];
      const int kd = S.k_dim[c];
      if (kd > 0 && S.col_cam[12 * (size_t)c] >= 0) {
        const int np = ModelNumParams(S.model[c]);
        for (int j = 0; j < np; ++j) x2 += prm[j] * prm[j];
        for (int j = 0; j < kd; ++j) {
          const int col = S.col_cam[12 * (size_t)c + j];
          const double d = -yc[col] * scale_c[col];
          cam2[12 * (size_t)c + S.k_cols[12 * (size_t)c + j]] = P.cam[12 * (size_t)c + S.k_cols[12 * (size_t)c + j]] + d;
          step2 += d * d;
        }
        // constant entries copied below
        for (int j = 0; j < 12; ++j) if (S.k_slot[12 * (size_t)c + j] < 0) cam2[12 * (size_t)c + j] = prm[j];
      } else {
#pragma unroll
        for (int j = 0; j < 12; ++j) cam2[12 * (size_t)c + j] = prm[j];
      }
    } else {
      const int q = t - S.n_img - S.n_cam;
      const int po = S.p_off[q];
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const double x = P.X[3 * (size_t)q + a];
        double xn = x;
        if (po >= 0) {
          const double d = -yp[3 * (size_t)po + a] * scale_p[3 * (size_t)po + a];
          xn = x + d;
          step2 += (xn - x) * (xn - x); x2 += x * x;
        }
        X2[3 * (size_t)q + a] = xn;
      }
    }
    // camera-side blocks are replicated on every rank, point blocks are sharded: keep their norms apart
    if (t < S.n_img + S.n_cam) { step2_c += step2; x2_c += x2; } else { step2_p += step2; x2_p += x2; }
  }
  double v[4] = {step2_c, x2_c, step2_p, x2_p};
  double* const out[4] = {&sc->step2_c, &sc->x2_c, &sc->step2_p, &sc->x2_p};
  GridSum<128, 4>(v, partial, ticket, out);
}

int ApplyStep(Problem* p, bool defer_allreduce) {
  DeviceScalars* sc = p->d_scalars.p;
  const int count = p->n_img + p->n_cam + p->n_pt;
  const int nb = std::min(DivUp(count, 128), p->n_sm * 16);
  if (nb > 0) {
    apply_step_kernel<<<nb, 128, 0, p->stream>>>(p->Struct(), p->Params(false), p->d_yc.p, p->d_yp.p, p->d_scale_c.p,
                                                 p->d_scale_p.p, p->d_pose2.p, p->d_cam2.p, p->d_X2.p, count,
                                                 p->d_cost_partial.p, p->d_tickets.p + 3, sc);
    p->launches++;
  }
  SBA_CUDA(cudaGetLastError());
  return defer_allreduce ? SBA_OK : AllReduceSum(p, &sc->step2_p, 2);
}

// ---- trust-region bookkeeping on the device (LevenbergMarquardtStrategy + TrustRegionMinimizer,
//      SURVEY.md App. B.4).  phase 0: after the linear solve; phase 1: after the candidate cost. ----
__global__ void tr_update_kernel(DeviceScalars* s, int phase, double min_relative_decrease, double max_radius,
                                 double parameter_tolerance, double function_tolerance) {
  if (phase == 0) {
    const bool valid = !s->solver_failed && (s->model_cost_change > 0.0);
    s->step_valid = valid ? 1 : 0;
    s->step_accepted = 0;
    s->converged = 0;
    if (!valid) {
      s->radius = s->radius / s->decrease_factor;
      s->decrease_factor *= 2.0;
      s->reuse_diagonal = 1;
      s->cost_change = 0.0; s->relative_decrease = 0.0; s->step_norm = 0.0;
    }
    s->solver_failed = 0;
  } else {
    if (!s->step_valid) return;  // phase 0 rejected the step: the candidate that was enqueued anyway is ignored
    double cand = s->cand_cost;
    if (!isfinite(cand)) cand = 1.7976931348623157e308;
    s->step_norm = sqrt(s->step2_c + s->step2_p);
    s->x_norm = sqrt(s->x2_c + s->x2_p);
    s->cost_change = s->cost - cand;
    if (s->step_norm <= parameter_tolerance * (s->x_norm + parameter_tolerance)) { s->converged = 1; return; }
    if (fabs(s->cost_change) <= function_tolerance * s->cost) { s->converged = 2; return; }
    const double rd = s->cost_change / s->model_cost_change;
    s->relative_decrease = rd;
    if (rd > min_relative_decrease) {
      s->step_accepted = 1;
      const double t = 2.0 * rd - 1.0;
      s->radius = fmin(max_radius, s->radius / fmax(1.0 / 3.0, 1.0 - t * t * t));
      s->decrease_factor = 2.0;
      s->reuse_diagonal = 0;
    } else {
      s->step_accepted = 0;
      s->radius = s->radius / s->decrease_factor;
      s->decrease_factor *= 2.0;
      s->reuse_diagonal = 1;
    }
  }
}

int TrustRegionUpdate(Problem* p, const sba_solver_options& o, int phase, bool sync) {
  tr_update_kernel<<<1, 1, 0, p->stream>>>(p->d_scalars.p, phase, o.min_relative_decrease, o.max_trust_region_radius,
                                           o.parameter_tolerance, o.function_tolerance);
  p->launches++;
  if (!sync) return SBA_OK;
  SBA_CUDA(cudaMemcpyAsync(p->h_scalars, p->d_scalars.p, sizeof(DeviceScalars), cudaMemcpyDeviceToHost, p->stream));
  SBA_CUDA(cudaStreamSynchronize(p->stream));
  return SBA_OK;
}

}  // namespace sba
