// ORACLE (test infrastructure only -- never linked into the product library).
// See ba_oracle.h for the reference locations this file restates.
#include "ba_oracle.h"

#include <omp.h>

#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>

#include "jet.h"
#include "projection.h"
#include "rotation.h"

namespace oracle {

// ---------------------------------------------------------------------------
// Problem structure: what Ceres' reduced program looks like for the residual
// blocks BundleAdjuster::SetUp adds (bundle_adjustment.cc:326-530).
// ---------------------------------------------------------------------------
void BAProblem::Build(const sba_problem_desc& d) {
  n_img = d.num_images; n_cam = d.num_cameras; n_pt = d.num_points; n_obs = d.num_observations;
  if (d.sem_maps && d.sem_height > 0 && d.sem_width > 0 && d.num_images > 0) {
    nr = 3; sem_maps = d.sem_maps; sem_h = d.sem_height; sem_w = d.sem_width; sem_weight = d.sem_weight;
  } else {
    nr = 2; sem_maps = nullptr;
  }
  q.assign(d.image_qvec, d.image_qvec + 4 * (size_t)n_img);
  t.assign(d.image_tvec, d.image_tvec + 3 * (size_t)n_img);
  cam.assign(d.camera_params, d.camera_params + SBA_MAX_CAMERA_PARAMS * (size_t)n_cam);
  X.assign(d.points, d.points + 3 * (size_t)n_pt);
  img_cam.assign(d.image_camera, d.image_camera + n_img);
  model.assign(d.camera_model, d.camera_model + n_cam);
  pose_const.assign(n_img, 0); tvec_mask.assign(n_img, 0); pt_const.assign(n_pt, 0); cam_mask.assign(n_cam, 0);
  if (d.image_pose_constant) pose_const.assign(d.image_pose_constant, d.image_pose_constant + n_img);
  if (d.image_tvec_constant_mask) tvec_mask.assign(d.image_tvec_constant_mask, d.image_tvec_constant_mask + n_img);
  if (d.point_constant) pt_const.assign(d.point_constant, d.point_constant + n_pt);
  if (d.camera_param_constant_mask) cam_mask.assign(d.camera_param_constant_mask, d.camera_param_constant_mask + n_cam);
  obs_img.assign(d.obs_image, d.obs_image + n_obs);
  obs_pt.assign(d.obs_point, d.obs_point + n_obs);
  obs_xy.assign(d.obs_xy, d.obs_xy + 2 * n_obs);

  std::vector<uint8_t> img_used(n_img, 0), cam_used(n_cam, 0), pt_used(n_pt, 0);
  for (int64_t k = 0; k < n_obs; ++k) {
    img_used[obs_img[k]] = 1; cam_used[img_cam[obs_img[k]]] = 1; pt_used[obs_pt[k]] = 1;
  }
  q_off.assign(n_img, -1); t_off.assign(n_img, -1); t_dim.assign(n_img, 0);
  k_off.assign(n_cam, -1); k_dim.assign(n_cam, 0);
  t_cols.assign(3 * (size_t)n_img, -1); k_cols.assign(SBA_MAX_CAMERA_PARAMS * (size_t)n_cam, -1);
  blk_off.clear(); blk_dim.clear();
  num_parameter_blocks_reduced = num_parameters_reduced = num_effective_parameters_reduced = 0;
  int off = 0;
  for (int i = 0; i < n_img; ++i) {
    if (!img_used[i] || pose_const[i]) continue;
    q_off[i] = off; blk_off.push_back(off); blk_dim.push_back(3); off += 3;
    num_parameter_blocks_reduced++; num_parameters_reduced += 4; num_effective_parameters_reduced += 3;
    int td = 0;
    for (int c = 0; c < 3; ++c) if (!((tvec_mask[i] >> c) & 1)) t_cols[3 * i + td++] = c;
    t_dim[i] = td;
    if (td > 0) {
      t_off[i] = off; blk_off.push_back(off); blk_dim.push_back(td); off += td;
      num_parameter_blocks_reduced++; num_parameters_reduced += 3; num_effective_parameters_reduced += td;
    }
  }
  int kmax = 0;
  for (int c = 0; c < n_cam; ++c) {
    if (!cam_used[c]) continue;
    const int P = NumParams(model[c]);
    int kd = 0;
    for (int j = 0; j < P; ++j) if (!((cam_mask[c] >> j) & 1)) k_cols[SBA_MAX_CAMERA_PARAMS * c + kd++] = j;
    k_dim[c] = kd;
    if (kd > 0) {
      k_off[c] = off; blk_off.push_back(off); blk_dim.push_back(kd); off += kd;
      num_parameter_blocks_reduced++; num_parameters_reduced += P; num_effective_parameters_reduced += kd;
      kmax = std::max(kmax, kd);
    }
  }
  n_cdim = off;
  dc_max = 6 + kmax;
  p_off.assign(n_pt, -1); n_pvar = 0;
  for (int p = 0; p < n_pt; ++p) {
    if (!pt_used[p] || pt_const[p]) continue;
    p_off[p] = n_pvar++;
    num_parameter_blocks_reduced++; num_parameters_reduced += 3; num_effective_parameters_reduced += 3;
  }
  obs_active.assign(n_obs, 0);
  num_residual_blocks = (int)n_obs; num_residuals = (int)(nr * n_obs); num_residuals_reduced = 0;
  for (int64_t k = 0; k < n_obs; ++k) {
    const int i = obs_img[k], c = img_cam[i], p = obs_pt[k];
    const bool act = (q_off[i] >= 0) || (t_off[i] >= 0) || (k_off[c] >= 0) || (p_off[p] >= 0);
    obs_active[k] = act;
    if (act) num_residuals_reduced += nr;
  }
  // CSR by point
  pt_ptr.assign(n_pt + 1, 0);
  for (int64_t k = 0; k < n_obs; ++k) pt_ptr[obs_pt[k] + 1]++;
  for (int p = 0; p < n_pt; ++p) pt_ptr[p + 1] += pt_ptr[p];
  pt_obs.resize(n_obs);
  std::vector<int64_t> fill(pt_ptr.begin(), pt_ptr.end() - 1);
  for (int64_t k = 0; k < n_obs; ++k) pt_obs[fill[obs_pt[k]]++] = k;
}

int BAProblem::ObsCols(int64_t k, int* cols) const {
  const int i = obs_img[k], c = img_cam[i];
  int n = 0;
  if (q_off[i] >= 0) for (int j = 0; j < 3; ++j) cols[n++] = q_off[i] + j;
  if (t_off[i] >= 0) for (int j = 0; j < t_dim[i]; ++j) cols[n++] = t_off[i] + j;
  if (k_off[c] >= 0) for (int j = 0; j < k_dim[c]; ++j) cols[n++] = k_off[c] + j;
  return n;
}

// ---------------------------------------------------------------------------
// Functor evaluation (cost_functions.h:57-81) with Jet<4+3+3+P>, as
// AutoDiffCostFunction<..., 2, 4, 3, 3, kNumParams> does (cost_functions.h:50-55).
// ---------------------------------------------------------------------------
static inline double ScalarPart(double x) { return x; }
template <int N> static inline double ScalarPart(const Jet<N>& x) { return x.a; }

// Bitmap::InterpolateBilinear (src/util/bitmap.cc:238-269) over T (double or Jet) on a top-down ROW-major float
// map: the cell is chosen by the VALUE (floor acts on the scalar part, as every Jet comparison does), the
// interpolation weights carry the derivatives -- so the Jet result is the exact Jacobian of the interpolant
// inside its cell.  false: the sample is rejected (x0 < 0, x1 >= W, y0 < 0, y1 >= H).
template <typename T>
static bool BilinearT(const float* map, int H, int W, const T& x, const T& y, T* out) {
  const T inv_y = double(H - 1) - y;
  const double fx = std::floor(ScalarPart(x)), fy = std::floor(ScalarPart(inv_y));
  if (!(fx >= 0.0 && fx + 1.0 < (double)W && fy >= 0.0 && fy + 1.0 < (double)H)) return false;
  const int x0 = (int)fx, x1 = x0 + 1, y0 = (int)fy, y1 = y0 + 1;
  const T dx = x - double(x0), dy = inv_y - double(y0);
  const T dx_1 = 1.0 - dx, dy_1 = 1.0 - dy;
  const float* line0 = map + (size_t)(H - 1 - y0) * W;  // FreeImage scanline y0 (bottom-up) in the top-down map
  const float* line1 = map + (size_t)(H - 1 - y1) * W;
  const T v0 = dx_1 * double(line0[x0]) + dx * double(line0[x1]);
  const T v1 = dx_1 * double(line1[x0]) + dx * double(line1[x1]);
  *out = dy_1 * v0 + dy * v1;
  return true;
}

// rows = 2, or 3 when map != nullptr (the semantic row, extension X1); J blocks are row-major rows x cols
template <int P>
static void EvalObsJet(int model, const double* q, const double* t, const double* X,
                       const double* params, const double* xy, const float* map, int H, int W, double weight,
                       double* r, double* Jq, double* Jt, double* JX, double* Jk) {
  constexpr int N = 10 + P;
  typedef Jet<N> J;
  J jq[4], jt[3], jX[3], jp[P];
  for (int i = 0; i < 4; ++i) jq[i] = J(q[i], i);
  for (int i = 0; i < 3; ++i) jt[i] = J(t[i], 4 + i);
  for (int i = 0; i < 3; ++i) jX[i] = J(X[i], 7 + i);
  for (int i = 0; i < P; ++i) jp[i] = J(params[i], 10 + i);
  J proj[3];
  UnitQuaternionRotatePoint(jq, jX, proj);
  proj[0] += jt[0]; proj[1] += jt[1]; proj[2] += jt[2];
  proj[0] /= proj[2];
  proj[1] /= proj[2];
  J res[3];
  Project<J>(model, jp, proj[0], proj[1], &res[0], &res[1]);
  const int rows = map ? 3 : 2;
  if (map) {
    J m;
    if (BilinearT<J>(map, H, W, res[0], res[1], &m)) res[2] = weight * m; else res[2] = J(0.0);
  }
  res[0] = res[0] - xy[0];
  res[1] = res[1] - xy[1];
  for (int row = 0; row < rows; ++row) {
    r[row] = res[row].a;
    if (Jq) for (int i = 0; i < 4; ++i) Jq[row * 4 + i] = res[row].v[i];
    if (Jt) for (int i = 0; i < 3; ++i) Jt[row * 3 + i] = res[row].v[4 + i];
    if (JX) for (int i = 0; i < 3; ++i) JX[row * 3 + i] = res[row].v[7 + i];
    if (Jk) for (int i = 0; i < P; ++i) Jk[row * P + i] = res[row].v[10 + i];
  }
}

// Rig functor (cost_functions.h:173-209) with Jet<4+3+4+3+3+P>, as
// AutoDiffCostFunction<..., 2, 4, 3, 4, 3, 3, kNumParams> does (cost_functions.h:171-176).
// J: per residual row 17 + 12 columns (rig_q, rig_t, rel_q, rel_t, X, params padded to 12).
template <int P>
static void EvalRigJet(int model, const double* rig_q, const double* rig_t, const double* rel_q, const double* rel_t,
                       const double* X, const double* params, const double* xy, double* r, double* Jout) {
  constexpr int N = 17 + P;
  typedef Jet<N> J;
  J gq[4], gt[3], lq[4], lt[3], jX[3], jp[P];
  for (int i = 0; i < 4; ++i) gq[i] = J(rig_q[i], i);
  for (int i = 0; i < 3; ++i) gt[i] = J(rig_t[i], 4 + i);
  for (int i = 0; i < 4; ++i) lq[i] = J(rel_q[i], 7 + i);
  for (int i = 0; i < 3; ++i) lt[i] = J(rel_t[i], 11 + i);
  for (int i = 0; i < 3; ++i) jX[i] = J(X[i], 14 + i);
  for (int i = 0; i < P; ++i) jp[i] = J(params[i], 17 + i);
  J q[4], t[3], proj[3];
  QuaternionProduct(lq, gq, q);
  UnitQuaternionRotatePoint(lq, gt, t);
  t[0] += lt[0]; t[1] += lt[1]; t[2] += lt[2];
  UnitQuaternionRotatePoint(q, jX, proj);
  proj[0] += t[0]; proj[1] += t[1]; proj[2] += t[2];
  proj[0] /= proj[2];
  proj[1] /= proj[2];
  J res[2];
  Project<J>(model, jp, proj[0], proj[1], &res[0], &res[1]);
  res[0] = res[0] - xy[0];
  res[1] = res[1] - xy[1];
  for (int row = 0; row < 2; ++row) {
    r[row] = res[row].a;
    if (!Jout) continue;
    for (int i = 0; i < 29; ++i) Jout[row * 29 + i] = i < N ? res[row].v[i] : 0.0;
  }
}

void EvalRigObservation(int model, const double* rig_q, const double* rig_t, const double* rel_q, const double* rel_t,
                        const double* X, const double* params, const double* xy, double* r, double* J) {
  switch (NumParams(model)) {
    case 3: EvalRigJet<3>(model, rig_q, rig_t, rel_q, rel_t, X, params, xy, r, J); break;
    case 4: EvalRigJet<4>(model, rig_q, rig_t, rel_q, rel_t, X, params, xy, r, J); break;
    case 5: EvalRigJet<5>(model, rig_q, rig_t, rel_q, rel_t, X, params, xy, r, J); break;
    case 8: EvalRigJet<8>(model, rig_q, rig_t, rel_q, rel_t, X, params, xy, r, J); break;
    case 12: EvalRigJet<12>(model, rig_q, rig_t, rel_q, rel_t, X, params, xy, r, J); break;
    default: r[0] = r[1] = std::numeric_limits<double>::quiet_NaN();
  }
}

// residual-only (double) evaluation, used for candidate-step cost; r[2], or r[3] with a map
static inline void EvalObsValue(int model, const double* q, const double* t, const double* X,
                                const double* params, const double* xy, const float* map, int H, int W, double weight,
                                double* r) {
  double proj[3];
  UnitQuaternionRotatePoint(q, X, proj);
  proj[0] += t[0]; proj[1] += t[1]; proj[2] += t[2];
  proj[0] /= proj[2];
  proj[1] /= proj[2];
  double x, y;
  Project<double>(model, params, proj[0], proj[1], &x, &y);
  r[0] = x - xy[0];
  r[1] = y - xy[1];
  if (map) {
    double m;
    r[2] = BilinearT<double>(map, H, W, x, y, &m) ? weight * m : 0.0;
  }
}

void EvalObservationSem(int model, const double* q, const double* t, const double* X, const double* params,
                        const double* xy, const float* map, int H, int W, double weight, double* r, double* Jq,
                        double* Jt, double* JX, double* Jk) {
  switch (NumParams(model)) {
    case 3: EvalObsJet<3>(model, q, t, X, params, xy, map, H, W, weight, r, Jq, Jt, JX, Jk); break;
    case 4: EvalObsJet<4>(model, q, t, X, params, xy, map, H, W, weight, r, Jq, Jt, JX, Jk); break;
    case 5: EvalObsJet<5>(model, q, t, X, params, xy, map, H, W, weight, r, Jq, Jt, JX, Jk); break;
    case 8: EvalObsJet<8>(model, q, t, X, params, xy, map, H, W, weight, r, Jq, Jt, JX, Jk); break;
    case 12: EvalObsJet<12>(model, q, t, X, params, xy, map, H, W, weight, r, Jq, Jt, JX, Jk); break;
    default: r[0] = r[1] = std::numeric_limits<double>::quiet_NaN(); if (map) r[2] = r[0];
  }
}
void EvalObservation(int model, const double* q, const double* t, const double* X,
                     const double* params, const double* xy, double* r, double* Jq, double* Jt,
                     double* JX, double* Jk) {
  EvalObservationSem(model, q, t, X, params, xy, nullptr, 0, 0, 0.0, r, Jq, Jt, JX, Jk);
}

// ceres loss functions (TrivialLoss / SoftLOneLoss / CauchyLoss), App. B.2
static inline void LossEval(int type, double a, double s, double rho[3]) {
  if (type == SBA_LOSS_SOFT_L1) {
    const double b = a * a, c = 1.0 / b;
    const double sum = 1.0 + s * c, tmp = std::sqrt(sum);
    rho[0] = 2.0 * b * (tmp - 1.0);
    rho[1] = std::max(std::numeric_limits<double>::min(), 1.0 / tmp);
    rho[2] = -(c * rho[1]) / (2.0 * sum);
  } else if (type == SBA_LOSS_CAUCHY) {
    const double b = a * a, c = 1.0 / b;
    const double sum = 1.0 + s * c, inv = 1.0 / sum;
    rho[0] = b * std::log(sum);
    rho[1] = std::max(std::numeric_limits<double>::min(), inv);
    rho[2] = -c * (inv * inv);
  } else {
    rho[0] = s; rho[1] = 1.0; rho[2] = 0.0;
  }
}

static inline const float* MapOf(const BAProblem& p, int i) {
  return p.sem_maps ? p.sem_maps + (size_t)i * p.sem_h * p.sem_w : nullptr;
}

void Linearize(const BAProblem& p, int loss_type, double loss_scale, bool apply_loss,
               Linearization* lin, int num_threads) {
  const int dc = p.dc_max, nr = p.nr;
  lin->r.assign((size_t)nr * p.n_obs, 0.0);
  lin->F.assign((size_t)p.n_obs * nr * dc, 0.0);
  lin->E.assign((size_t)p.n_obs * nr * 3, 0.0);
  // (sums are kept per thread and combined in thread order: an OpenMP reduction clause combines in arrival order, which
  //  makes the last bits of the cost -- and through the trust-region radius the whole inexact-Newton trajectory --
  //  differ from run to run)
  const int nt = std::max(1, num_threads);
  std::vector<double> cost_t(nt * 8, 0.0), fixed_t(nt * 8, 0.0);
#pragma omp parallel for num_threads(nt) schedule(static)
  for (int64_t k = 0; k < p.n_obs; ++k) {
    double& cost = cost_t[8 * omp_get_thread_num()];
    double& fixed = fixed_t[8 * omp_get_thread_num()];
    const int i = p.obs_img[k], c = p.img_cam[i], pt = p.obs_pt[k];
    const int m = p.model[c], P = NumParams(m);
    double r[3] = {0, 0, 0}, Jq[12], Jt[9], JX[9], Jk[36];
    EvalObservationSem(m, &p.q[4 * i], &p.t[3 * i], &p.X[3 * pt], &p.cam[SBA_MAX_CAMERA_PARAMS * c],
                       &p.obs_xy[2 * k], MapOf(p, i), p.sem_h, p.sem_w, p.sem_weight, r, Jq, Jt, JX, Jk);
    double s = 0.0;
    for (int row = 0; row < nr; ++row) s += r[row] * r[row];
    double rho[3];
    LossEval(loss_type, loss_scale, s, rho);
    if (!p.obs_active[k]) {
      fixed += 0.5 * rho[0];
      for (int row = 0; row < nr; ++row) lin->r[(size_t)nr * k + row] = r[row];  // raw, parity surface only
      continue;
    }
    cost += 0.5 * rho[0];
    // local (tangent) Jacobian, compact camera-side layout [q(3)][t cols][k cols]
    double* F = &lin->F[(size_t)k * nr * dc];
    double* E = &lin->E[(size_t)k * nr * 3];
    int n = 0;
    if (p.q_off[i] >= 0) {
      double PJ[12];
      QuaternionPlusJacobian(&p.q[4 * i], PJ);
      for (int row = 0; row < nr; ++row)
        for (int j = 0; j < 3; ++j) {
          double acc = 0.0;
          for (int a = 0; a < 4; ++a) acc += Jq[row * 4 + a] * PJ[a * 3 + j];
          F[row * dc + n + j] = acc;
        }
      n += 3;
    }
    if (p.t_off[i] >= 0) {
      for (int j = 0; j < p.t_dim[i]; ++j) {
        const int a = p.t_cols[3 * i + j];
        for (int row = 0; row < nr; ++row) F[row * dc + n + j] = Jt[row * 3 + a];
      }
      n += p.t_dim[i];
    }
    if (p.k_off[c] >= 0) {
      for (int j = 0; j < p.k_dim[c]; ++j) {
        const int a = p.k_cols[SBA_MAX_CAMERA_PARAMS * c + j];
        for (int row = 0; row < nr; ++row) F[row * dc + n + j] = Jk[row * P + a];
      }
      n += p.k_dim[c];
    }
    if (p.p_off[pt] >= 0) for (int j = 0; j < 3 * nr; ++j) E[j] = JX[j];
    // Corrector (ceres/internal/corrector.cc): Jacobian first, then residuals
    if (apply_loss) {
      const double sqrt_rho1 = std::sqrt(rho[1]);
      double residual_scaling, alpha_sq_norm;
      if (s == 0.0 || rho[2] <= 0.0) {
        residual_scaling = sqrt_rho1; alpha_sq_norm = 0.0;
      } else {
        const double D = 1.0 + 2.0 * s * rho[2] / rho[1];
        const double alpha = 1.0 - std::sqrt(D);
        residual_scaling = sqrt_rho1 / (1 - alpha); alpha_sq_norm = alpha / s;
      }
      // one column: J[row * stride] for row < nr
      auto correct = [&](double* J, int stride) {
        if (alpha_sq_norm == 0.0) { for (int row = 0; row < nr; ++row) J[row * stride] *= sqrt_rho1; return; }
        double rtj = 0.0;
        for (int row = 0; row < nr; ++row) rtj += r[row] * J[row * stride];
        for (int row = 0; row < nr; ++row) J[row * stride] = sqrt_rho1 * (J[row * stride] - alpha_sq_norm * r[row] * rtj);
      };
      for (int j = 0; j < n; ++j) correct(&F[j], dc);
      for (int j = 0; j < 3; ++j) correct(&E[j], 3);
      for (int row = 0; row < nr; ++row) r[row] *= residual_scaling;
    }
    for (int row = 0; row < nr; ++row) lin->r[(size_t)nr * k + row] = r[row];
  }
  double cost = 0.0, fixed = 0.0;
  for (int t = 0; t < nt; ++t) { cost += cost_t[8 * t]; fixed += fixed_t[8 * t]; }
  lin->cost = cost; lin->fixed_cost = fixed;
}

double CostOnly(const BAProblem& p, int loss_type, double loss_scale, int num_threads) {
  const int nt = std::max(1, num_threads);
  std::vector<double> cost_t(nt * 8, 0.0);  // per thread, combined in thread order (see Linearize)
#pragma omp parallel for num_threads(nt) schedule(static)
  for (int64_t k = 0; k < p.n_obs; ++k) {
    if (!p.obs_active[k]) continue;
    double& cost = cost_t[8 * omp_get_thread_num()];
    const int i = p.obs_img[k], c = p.img_cam[i], pt = p.obs_pt[k];
    double r[3] = {0, 0, 0};
    EvalObsValue(p.model[c], &p.q[4 * i], &p.t[3 * i], &p.X[3 * pt],
                 &p.cam[SBA_MAX_CAMERA_PARAMS * c], &p.obs_xy[2 * k], MapOf(p, i), p.sem_h, p.sem_w, p.sem_weight, r);
    double rho[3];
    LossEval(loss_type, loss_scale, r[0] * r[0] + r[1] * r[1] + r[2] * r[2], rho);
    cost += 0.5 * rho[0];
  }
  double cost = 0.0;
  for (int t = 0; t < nt; ++t) cost += cost_t[8 * t];
  return cost;
}

// ---------------------------------------------------------------------------
// small dense helpers
// ---------------------------------------------------------------------------
// in-place lower Cholesky of row-major n x n; returns false if not SPD
static bool CholeskyLower(double* A, int n, int num_threads) {
  for (int j = 0; j < n; ++j) {
    double d = A[(size_t)j * n + j];
    for (int k = 0; k < j; ++k) d -= A[(size_t)j * n + k] * A[(size_t)j * n + k];
    if (!(d > 0.0) || !std::isfinite(d)) return false;
    d = std::sqrt(d);
    A[(size_t)j * n + j] = d;
    const double inv = 1.0 / d;
#pragma omp parallel for num_threads(num_threads) schedule(static) if (n - j > 256)
    for (int i = j + 1; i < n; ++i) {
      double s = A[(size_t)i * n + j];
      const double* ai = &A[(size_t)i * n];
      const double* aj = &A[(size_t)j * n];
      for (int k = 0; k < j; ++k) s -= ai[k] * aj[k];
      A[(size_t)i * n + j] = s * inv;
    }
  }
  return true;
}
static void CholeskySolve(const double* L, int n, double* b) {
  for (int i = 0; i < n; ++i) {
    double s = b[i];
    for (int k = 0; k < i; ++k) s -= L[(size_t)i * n + k] * b[k];
    b[i] = s / L[(size_t)i * n + i];
  }
  for (int i = n - 1; i >= 0; --i) {
    double s = b[i];
    for (int k = i + 1; k < n; ++k) s -= L[(size_t)k * n + i] * b[k];
    b[i] = s / L[(size_t)i * n + i];
  }
}
// inverse of SPD 3x3 (row-major full) via LLT solve of identity, as
// InvertPSDMatrix<3> does in Ceres for full-rank EtE.
static bool InvertSPD(const double* A, int n, double* Ainv) {
  double L[12 * 12];
  for (int i = 0; i < n * n; ++i) L[i] = A[i];
  if (!CholeskyLower(L, n, 1)) return false;
  for (int c = 0; c < n; ++c) {
    double e[12];
    for (int i = 0; i < n; ++i) e[i] = (i == c) ? 1.0 : 0.0;
    CholeskySolve(L, n, e);
    for (int i = 0; i < n; ++i) Ainv[i * n + c] = e[i];
  }
  return true;
}

// ---------------------------------------------------------------------------
// Linear solver state for one LM step: scaled Jacobian + D
// ---------------------------------------------------------------------------
struct StepContext {
  const BAProblem* p;
  const Linearization* lin;  // SCALED jacobian blocks, corrected residuals
  const double* Dc;          // [n_cdim]
  const double* Dp;          // [3*n_pvar]
  int threads;
  std::vector<double> Cinv;  // [n_pvar*9]
  std::vector<double> g;     // [n_pvar*3]  E^T r
};

// Row accessors of one observation's blocks (nr rows): F row-major nr x dc, E row-major nr x 3, r[nr].
#define OBS_F(lin, p, k) (&(lin).F[(size_t)(k) * (p).nr * (p).dc_max])
#define OBS_E(lin, p, k) (&(lin).E[(size_t)(k) * (p).nr * 3])
#define OBS_R(lin, p, k) (&(lin).r[(size_t)(k) * (p).nr])

// f = F_k x  over the observation's camera columns
static inline void RowsTimesX(const BAProblem& p, const double* F, const int* cols, int n, const double* x, double* f) {
  const int dc = p.dc_max;
  for (int row = 0; row < p.nr; ++row) {
    double a = 0.0;
    for (int j = 0; j < n; ++j) a += F[row * dc + j] * x[cols[j]];
    f[row] = a;
  }
}

static bool ComputePointBlocks(StepContext* c) {
  const BAProblem& p = *c->p;
  const int nr = p.nr;
  c->Cinv.assign((size_t)p.n_pvar * 9, 0.0);
  c->g.assign((size_t)p.n_pvar * 3, 0.0);
  bool ok = true;
#pragma omp parallel for num_threads(c->threads) schedule(static)
  for (int pt = 0; pt < p.n_pt; ++pt) {
    const int po = p.p_off[pt];
    if (po < 0) continue;
    double C[9] = {0}, g[3] = {0};
    for (int64_t idx = p.pt_ptr[pt]; idx < p.pt_ptr[pt + 1]; ++idx) {
      const int64_t k = p.pt_obs[idx];
      const double* E = OBS_E(*c->lin, p, k);
      const double* r = OBS_R(*c->lin, p, k);
      for (int a = 0; a < 3; ++a) {
        for (int b = 0; b < 3; ++b)
          for (int row = 0; row < nr; ++row) C[a * 3 + b] += E[3 * row + a] * E[3 * row + b];
        for (int row = 0; row < nr; ++row) g[a] += E[3 * row + a] * r[row];
      }
    }
    for (int a = 0; a < 3; ++a) C[a * 3 + a] += c->Dp[3 * po + a] * c->Dp[3 * po + a];
    if (!InvertSPD(C, 3, &c->Cinv[(size_t)po * 9])) {
#pragma omp critical
      ok = false;
    }
    for (int a = 0; a < 3; ++a) c->g[(size_t)po * 3 + a] = g[a];
  }
  return ok;
}

// per-thread accumulators over the camera columns, summed in thread order
struct ThreadAcc {
  std::vector<std::vector<double>> acc;
  ThreadAcc(int nt, size_t n) : acc(std::max(1, nt), std::vector<double>(n, 0.0)) {}
  void Sum(double* out, size_t n) const {
    for (size_t j = 0; j < n; ++j) { double s = 0.0; for (const auto& a : acc) s += a[j]; out[j] += s; }
  }
};

// rhs of the reduced system: F^T r - sum_p W_p Cinv_p g_p
static void ReducedRhs(const StepContext& c, std::vector<double>* rhs) {
  const BAProblem& p = *c.p;
  const int dc = p.dc_max, nr = p.nr;
  rhs->assign(p.n_cdim, 0.0);
  const int nt = std::max(1, c.threads);
  ThreadAcc ta(nt, p.n_cdim);
#pragma omp parallel num_threads(nt)
  {
    std::vector<double>& acc = ta.acc[omp_get_thread_num()];
#pragma omp for schedule(static)
    for (int64_t k = 0; k < p.n_obs; ++k) {
      if (!p.obs_active[k]) continue;
      int cols[24];
      const int n = p.ObsCols(k, cols);
      if (n == 0) continue;
      const double* F = OBS_F(*c.lin, p, k);
      const double* r = OBS_R(*c.lin, p, k);
      double e[3] = {0, 0, 0};
      for (int row = 0; row < nr; ++row) e[row] = r[row];
      const int po = p.p_off[p.obs_pt[k]];
      if (po >= 0) {
        const double* E = OBS_E(*c.lin, p, k);
        const double* Ci = &c.Cinv[(size_t)po * 9];
        const double* g = &c.g[(size_t)po * 3];
        double w[3];
        for (int a = 0; a < 3; ++a) w[a] = Ci[a * 3] * g[0] + Ci[a * 3 + 1] * g[1] + Ci[a * 3 + 2] * g[2];
        for (int row = 0; row < nr; ++row) e[row] -= E[3 * row] * w[0] + E[3 * row + 1] * w[1] + E[3 * row + 2] * w[2];
      }
      for (int j = 0; j < n; ++j) {
        double a = 0.0;
        for (int row = 0; row < nr; ++row) a += F[row * dc + j] * e[row];
        acc[cols[j]] += a;
      }
    }
  }
  ta.Sum(rhs->data(), p.n_cdim);
}

// y = S x, S = F^T F + Dc^2 - sum_p W_p Cinv_p W_p^T, evaluated implicitly
// (ImplicitSchurComplement::RightMultiply)
static void SchurMatVec(const StepContext& c, const double* x, double* y) {
  const BAProblem& p = *c.p;
  const int dc = p.dc_max, nr = p.nr;
  std::vector<double> v((size_t)p.n_pvar * 3, 0.0);
#pragma omp parallel for num_threads(c.threads) schedule(static)
  for (int pt = 0; pt < p.n_pt; ++pt) {
    const int po = p.p_off[pt];
    if (po < 0) continue;
    double u[3] = {0, 0, 0};
    for (int64_t idx = p.pt_ptr[pt]; idx < p.pt_ptr[pt + 1]; ++idx) {
      const int64_t k = p.pt_obs[idx];
      int cols[24];
      const int n = p.ObsCols(k, cols);
      const double* E = OBS_E(*c.lin, p, k);
      double f[3];
      RowsTimesX(p, OBS_F(*c.lin, p, k), cols, n, x, f);
      for (int a = 0; a < 3; ++a)
        for (int row = 0; row < nr; ++row) u[a] += E[3 * row + a] * f[row];
    }
    const double* Ci = &c.Cinv[(size_t)po * 9];
    for (int a = 0; a < 3; ++a) v[(size_t)po * 3 + a] = Ci[a * 3] * u[0] + Ci[a * 3 + 1] * u[1] + Ci[a * 3 + 2] * u[2];
  }
  const int nt = std::max(1, c.threads);
  ThreadAcc ta(nt, p.n_cdim);
#pragma omp parallel num_threads(nt)
  {
    std::vector<double>& a = ta.acc[omp_get_thread_num()];
#pragma omp for schedule(static)
    for (int64_t k = 0; k < p.n_obs; ++k) {
      if (!p.obs_active[k]) continue;
      int cols[24];
      const int n = p.ObsCols(k, cols);
      if (n == 0) continue;
      const double* F = OBS_F(*c.lin, p, k);
      double f[3];
      RowsTimesX(p, F, cols, n, x, f);
      const int po = p.p_off[p.obs_pt[k]];
      if (po >= 0) {
        const double* E = OBS_E(*c.lin, p, k);
        const double* w = &v[(size_t)po * 3];
        for (int row = 0; row < nr; ++row) f[row] -= E[3 * row] * w[0] + E[3 * row + 1] * w[1] + E[3 * row + 2] * w[2];
      }
      for (int j = 0; j < n; ++j) {
        double s = 0.0;
        for (int row = 0; row < nr; ++row) s += F[row * dc + j] * f[row];
        a[cols[j]] += s;
      }
    }
  }
  for (int j = 0; j < p.n_cdim; ++j) y[j] = c.Dc[j] * c.Dc[j] * x[j];
  ta.Sum(y, p.n_cdim);
}

// W = F^T E (n x 3) of one observation
static inline void ObsW(const BAProblem& p, const double* F, const double* E, int n, double* G /* [n*3] */) {
  const int dc = p.dc_max;
  for (int j = 0; j < n; ++j)
    for (int m = 0; m < 3; ++m) {
      double a = 0.0;
      for (int row = 0; row < p.nr; ++row) a += F[row * dc + j] * E[3 * row + m];
      G[j * 3 + m] = a;
    }
}

// Explicit reduced camera matrix (SchurEliminator::Eliminate), dense row-major.
static void BuildDenseSchur(const StepContext& c, std::vector<double>* Sout) {
  const BAProblem& p = *c.p;
  const int dc = p.dc_max, n = p.n_cdim, nr = p.nr;
  std::vector<double>& S = *Sout;
  S.assign((size_t)n * n, 0.0);
  for (int j = 0; j < n; ++j) S[(size_t)j * n + j] = c.Dc[j] * c.Dc[j];
  // per point: + F^T F of its observations, - W_p Cinv_p W_p^T  (constant points: only the first)
#pragma omp parallel for num_threads(c.threads) schedule(dynamic, 64)
  for (int pt = 0; pt < p.n_pt; ++pt) {
    const int po = p.p_off[pt];
    const int64_t b0 = p.pt_ptr[pt], b1 = p.pt_ptr[pt + 1];
    const int len = (int)(b1 - b0);
    if (len == 0) continue;
    std::vector<double> G((size_t)len * 24 * 3), H((size_t)len * 24 * 3);
    std::vector<int> cols((size_t)len * 24), ncs(len);
    for (int a = 0; a < len; ++a) {
      const int64_t k = p.pt_obs[b0 + a];
      ncs[a] = p.obs_active[k] ? p.ObsCols(k, &cols[(size_t)a * 24]) : 0;
      const double* F = OBS_F(*c.lin, p, k);
      for (int ja = 0; ja < ncs[a]; ++ja)
        for (int jb = 0; jb < ncs[a]; ++jb) {
          double v = 0.0;
          for (int row = 0; row < nr; ++row) v += F[row * dc + ja] * F[row * dc + jb];
#pragma omp atomic
          S[(size_t)cols[(size_t)a * 24 + ja] * n + cols[(size_t)a * 24 + jb]] += v;
        }
      if (po < 0) continue;
      const double* Ci = &c.Cinv[(size_t)po * 9];
      ObsW(p, F, OBS_E(*c.lin, p, k), ncs[a], &G[(size_t)a * 24 * 3]);
      for (int j = 0; j < ncs[a]; ++j)
        for (int m = 0; m < 3; ++m) {
          const double* g = &G[((size_t)a * 24 + j) * 3];
          H[((size_t)a * 24 + j) * 3 + m] = g[0] * Ci[m] + g[1] * Ci[3 + m] + g[2] * Ci[6 + m];
        }
    }
    if (po < 0) continue;
    for (int a = 0; a < len; ++a)
      for (int b = 0; b < len; ++b)
        for (int ja = 0; ja < ncs[a]; ++ja) {
          const double* h = &H[((size_t)a * 24 + ja) * 3];
          for (int jb = 0; jb < ncs[b]; ++jb) {
            const double* g = &G[((size_t)b * 24 + jb) * 3];
            const double val = h[0] * g[0] + h[1] * g[1] + h[2] * g[2];
#pragma omp atomic
            S[(size_t)cols[(size_t)a * 24 + ja] * n + cols[(size_t)b * 24 + jb]] -= val;
          }
        }
  }
}

// SCHUR_JACOBI: block diagonal of S at parameter-block granularity, inverted.
static bool BuildSchurJacobi(const StepContext& c, std::vector<double>* Minv) {
  const BAProblem& p = *c.p;
  const int dc = p.dc_max, nr = p.nr;
  const int nb = (int)p.blk_off.size();
  std::vector<int> col2blk(p.n_cdim, -1);
  std::vector<size_t> boff(nb + 1, 0);
  for (int b = 0; b < nb; ++b) {
    for (int j = 0; j < p.blk_dim[b]; ++j) col2blk[p.blk_off[b] + j] = b;
    boff[b + 1] = boff[b] + (size_t)p.blk_dim[b] * p.blk_dim[b];
  }
  std::vector<double> M(boff[nb], 0.0);
  for (int b = 0; b < nb; ++b)
    for (int j = 0; j < p.blk_dim[b]; ++j)
      M[boff[b] + (size_t)j * p.blk_dim[b] + j] = c.Dc[p.blk_off[b] + j] * c.Dc[p.blk_off[b] + j];
  const int nt = std::max(1, c.threads);
  ThreadAcc ta(nt, boff[nb]);
#pragma omp parallel num_threads(nt)
  {
    std::vector<double>& Mt = ta.acc[omp_get_thread_num()];
#pragma omp for schedule(static)
    for (int pt = 0; pt < p.n_pt; ++pt) {
      const int po = p.p_off[pt];
      // T_{p,B} = sum over this point's observations touching block B of F_B^T E
      std::vector<int> blks;
      std::vector<std::vector<double>> Ts;
      for (int64_t idx = p.pt_ptr[pt]; idx < p.pt_ptr[pt + 1]; ++idx) {
        const int64_t k = p.pt_obs[idx];
        if (!p.obs_active[k]) continue;
        int cols[24];
        const int nc = p.ObsCols(k, cols);
        const double* F = OBS_F(*c.lin, p, k);
        for (int a = 0; a < nc; ++a)
          for (int b = 0; b < nc; ++b) {
            const int ba = col2blk[cols[a]];
            if (ba != col2blk[cols[b]]) continue;
            const int d = p.blk_dim[ba];
            double v = 0.0;
            for (int row = 0; row < nr; ++row) v += F[row * dc + a] * F[row * dc + b];
            Mt[boff[ba] + (size_t)(cols[a] - p.blk_off[ba]) * d + (cols[b] - p.blk_off[ba])] += v;
          }
        if (po < 0) continue;
        double G[24 * 3];
        ObsW(p, F, OBS_E(*c.lin, p, k), nc, G);
        for (int j = 0; j < nc; ++j) {
          const int b = col2blk[cols[j]];
          size_t pos = std::find(blks.begin(), blks.end(), b) - blks.begin();
          if (pos == blks.size()) { blks.push_back(b); Ts.emplace_back((size_t)p.blk_dim[b] * 3, 0.0); }
          const int lj = cols[j] - p.blk_off[b];
          for (int m = 0; m < 3; ++m) Ts[pos][(size_t)lj * 3 + m] += G[j * 3 + m];
        }
      }
      if (po < 0) continue;
      const double* Ci = &c.Cinv[(size_t)po * 9];
      for (size_t bi = 0; bi < blks.size(); ++bi) {
        const int b = blks[bi], d = p.blk_dim[b];
        const std::vector<double>& T = Ts[bi];
        for (int a = 0; a < d; ++a) {
          double h[3];
          for (int m = 0; m < 3; ++m) h[m] = T[a * 3] * Ci[m] + T[a * 3 + 1] * Ci[3 + m] + T[a * 3 + 2] * Ci[6 + m];
          for (int e = 0; e < d; ++e)
            Mt[boff[b] + (size_t)a * d + e] -= h[0] * T[e * 3] + h[1] * T[e * 3 + 1] + h[2] * T[e * 3 + 2];
        }
      }
    }
  }
  ta.Sum(M.data(), boff[nb]);
  Minv->assign(boff[nb], 0.0);
  for (int b = 0; b < nb; ++b)
    if (!InvertSPD(&M[boff[b]], p.blk_dim[b], &(*Minv)[boff[b]])) return false;
  return true;
}
static void ApplySchurJacobi(const BAProblem& p, const std::vector<double>& Minv, const double* x, double* y) {
  size_t off = 0;
  for (size_t b = 0; b < p.blk_off.size(); ++b) {
    const int d = p.blk_dim[b], o = p.blk_off[b];
    for (int i = 0; i < d; ++i) {
      double s = 0;
      for (int j = 0; j < d; ++j) s += Minv[off + (size_t)i * d + j] * x[o + j];
      y[o + i] = s;
    }
    off += (size_t)d * d;
  }
}

static double Dot(const std::vector<double>& a, const std::vector<double>& b) {
  double s = 0; for (size_t i = 0; i < a.size(); ++i) s += a[i] * b[i]; return s;
}

// ConjugateGradientsSolver::Solve (Ceres 2.1.0), App. B.5.  Returns: 0 success
// (converged), 1 no-convergence (still usable), 2 failure.
static int SolveCG(const StepContext& c, const std::vector<double>& Minv, const std::vector<double>& b,
                   int min_it, int max_it, double q_tol, std::vector<double>* xout, int* iters) {
  const int n = c.p->n_cdim;
  std::vector<double>& x = *xout;
  x.assign(n, 0.0);
  *iters = 0;
  const double norm_b = std::sqrt(Dot(b, b));
  if (norm_b == 0.0) return 0;
  std::vector<double> r(b), z(n), pvec(n), qv(n), tmp(n);
  // r = b - A*0 = b
  double rho = 1.0;
  double Q0 = -1.0 * 0.0;  // x == 0
  {
    // Q0 = -x'(b + r)
    std::vector<double> br(n); for (int i = 0; i < n; ++i) br[i] = b[i] + r[i];
    Q0 = -1.0 * Dot(x, br);
  }
  for (int it = 1;; ++it) {
    *iters = it;
    ApplySchurJacobi(*c.p, Minv, r.data(), z.data());
    const double last_rho = rho;
    rho = Dot(r, z);
    if (rho == 0.0 || !std::isfinite(rho)) return 2;
    if (it == 1) {
      pvec = z;
    } else {
      const double beta = rho / last_rho;
      if (beta == 0.0 || !std::isfinite(beta)) return 2;
      for (int i = 0; i < n; ++i) pvec[i] = z[i] + beta * pvec[i];
    }
    SchurMatVec(c, pvec.data(), qv.data());
    const double pq = Dot(pvec, qv);
    if (pq <= 0.0 || std::isinf(pq)) return 1;
    const double alpha = rho / pq;
    if (std::isinf(alpha)) return 2;
    for (int i = 0; i < n; ++i) x[i] += alpha * pvec[i];
    if (it % 10 == 0) {
      SchurMatVec(c, x.data(), tmp.data());
      for (int i = 0; i < n; ++i) r[i] = b[i] - tmp[i];
    } else {
      for (int i = 0; i < n; ++i) r[i] -= alpha * qv[i];
    }
    double Q1 = 0.0;
    for (int i = 0; i < n; ++i) Q1 += x[i] * (b[i] + r[i]);
    Q1 = -1.0 * Q1;
    const double zeta = it * (Q1 - Q0) / Q1;
    if (zeta < q_tol && it >= min_it) return 0;
    Q0 = Q1;
    if (it >= max_it) break;
  }
  return 1;
}

// ---------------------------------------------------------------------------
// TrustRegionMinimizer + LevenbergMarquardtStrategy (Ceres 2.1.0), App. B.3-B.4
// ---------------------------------------------------------------------------
// mode 0: column square norms; mode 1: gradient J^T r.  Camera part through per-thread accumulators, point part
// per point (CSR), so both are parallel and the result does not depend on the schedule.
static void ColumnPass(const BAProblem& p, const Linearization& lin, int mode, std::vector<double>* nc,
                       std::vector<double>* np, int threads) {
  const int dc = p.dc_max, nr = p.nr;
  nc->assign(p.n_cdim, 0.0);
  np->assign((size_t)3 * p.n_pvar, 0.0);
  const int nt = std::max(1, threads);
  ThreadAcc ta(nt, p.n_cdim);
#pragma omp parallel num_threads(nt)
  {
    std::vector<double>& acc = ta.acc[omp_get_thread_num()];
#pragma omp for schedule(static)
    for (int64_t k = 0; k < p.n_obs; ++k) {
      if (!p.obs_active[k]) continue;
      int cols[24];
      const int n = p.ObsCols(k, cols);
      const double* F = OBS_F(lin, p, k);
      const double* r = OBS_R(lin, p, k);
      for (int j = 0; j < n; ++j) {
        double a = 0.0;
        for (int row = 0; row < nr; ++row) a += F[row * dc + j] * (mode ? r[row] : F[row * dc + j]);
        acc[cols[j]] += a;
      }
    }
#pragma omp for schedule(static)
    for (int pt = 0; pt < p.n_pt; ++pt) {
      const int po = p.p_off[pt];
      if (po < 0) continue;
      double v[3] = {0, 0, 0};
      for (int64_t idx = p.pt_ptr[pt]; idx < p.pt_ptr[pt + 1]; ++idx) {
        const int64_t k = p.pt_obs[idx];
        if (!p.obs_active[k]) continue;
        const double* E = OBS_E(lin, p, k);
        const double* r = OBS_R(lin, p, k);
        for (int a = 0; a < 3; ++a)
          for (int row = 0; row < nr; ++row) v[a] += E[3 * row + a] * (mode ? r[row] : E[3 * row + a]);
      }
      for (int a = 0; a < 3; ++a) (*np)[3 * (size_t)po + a] = v[a];
    }
  }
  ta.Sum(nc->data(), p.n_cdim);
}
static void ColumnSquaredNorms(const BAProblem& p, const Linearization& lin, std::vector<double>* nc,
                               std::vector<double>* np, int threads) {
  ColumnPass(p, lin, 0, nc, np, threads);
}

static void ScaleColumns(const BAProblem& p, const std::vector<double>& sc, const std::vector<double>& sp,
                         Linearization* lin, int threads) {
  const int dc = p.dc_max, nr = p.nr;
#pragma omp parallel for num_threads(threads) schedule(static)
  for (int64_t k = 0; k < p.n_obs; ++k) {
    if (!p.obs_active[k]) continue;
    int cols[24];
    const int n = p.ObsCols(k, cols);
    double* F = OBS_F(*lin, p, k);
    for (int j = 0; j < n; ++j)
      for (int row = 0; row < nr; ++row) F[row * dc + j] *= sc[cols[j]];
    const int po = p.p_off[p.obs_pt[k]];
    if (po >= 0) {
      double* E = OBS_E(*lin, p, k);
      for (int a = 0; a < 3; ++a)
        for (int row = 0; row < nr; ++row) E[3 * row + a] *= sp[3 * (size_t)po + a];
    }
  }
}

// x_plus = Plus(x, delta) on every variable block (manifolds, App. A.1)
struct State { std::vector<double> q, t, cam, X; };
static void ApplyPlus(const BAProblem& src, const std::vector<double>& dc, const std::vector<double>& dp,
                      State* dst) {
  dst->q = src.q; dst->t = src.t; dst->cam = src.cam; dst->X = src.X;
  for (int i = 0; i < src.n_img; ++i) {
    if (src.q_off[i] >= 0) QuaternionPlus(&src.q[4 * i], &dc[src.q_off[i]], &dst->q[4 * i]);
    if (src.t_off[i] >= 0)
      for (int j = 0; j < src.t_dim[i]; ++j) dst->t[3 * i + src.t_cols[3 * i + j]] += dc[src.t_off[i] + j];
  }
  for (int c = 0; c < src.n_cam; ++c)
    if (src.k_off[c] >= 0)
      for (int j = 0; j < src.k_dim[c]; ++j)
        dst->cam[SBA_MAX_CAMERA_PARAMS * c + src.k_cols[SBA_MAX_CAMERA_PARAMS * c + j]] += dc[src.k_off[c] + j];
  for (int p = 0; p < src.n_pt; ++p)
    if (src.p_off[p] >= 0)
      for (int a = 0; a < 3; ++a) dst->X[3 * p + a] += dp[3 * (size_t)src.p_off[p] + a];
}
static void SwapState(BAProblem* p, State* s) { p->q.swap(s->q); p->t.swap(s->t); p->cam.swap(s->cam); p->X.swap(s->X); }

// norm over the variable blocks (ambient coordinates) of the state, or of its difference to `other`
static double AmbientNorm(const BAProblem& p, const State* other) {
  double s = 0;
  auto acc = [&](const std::vector<double>& a, const std::vector<double>* b, size_t i) {
    const double d = b ? a[i] - (*b)[i] : a[i];
    s += d * d;
  };
  for (int i = 0; i < p.n_img; ++i) {
    if (p.q_off[i] >= 0) for (int j = 0; j < 4; ++j) acc(p.q, other ? &other->q : nullptr, 4 * i + j);
    if (p.t_off[i] >= 0) for (int j = 0; j < 3; ++j) acc(p.t, other ? &other->t : nullptr, 3 * i + j);
  }
  for (int c = 0; c < p.n_cam; ++c)
    if (p.k_off[c] >= 0)
      for (int j = 0; j < NumParams(p.model[c]); ++j) acc(p.cam, other ? &other->cam : nullptr, SBA_MAX_CAMERA_PARAMS * c + j);
  for (int pt = 0; pt < p.n_pt; ++pt)
    if (p.p_off[pt] >= 0) for (int j = 0; j < 3; ++j) acc(p.X, other ? &other->X : nullptr, 3 * pt + j);
  return std::sqrt(s);
}

int Solve(BAProblem* pp, const sba_solver_options& o, sba_summary* sum,
          std::vector<sba_iteration_summary>* trace, int threads) {
  BAProblem& p = *pp;
  const auto t_start = std::chrono::steady_clock::now();
  std::memset(sum, 0, sizeof(*sum));
  sum->num_residual_blocks = p.num_residual_blocks;
  sum->num_residuals = p.num_residuals;
  sum->num_residuals_reduced = p.num_residuals_reduced;
  sum->num_parameter_blocks_reduced = p.num_parameter_blocks_reduced;
  sum->num_parameters_reduced = p.num_parameters_reduced;
  sum->num_effective_parameters_reduced = p.num_effective_parameters_reduced;
  if (p.n_obs == 0) return SBA_ERR_EMPTY_PROBLEM;
  int solver = o.linear_solver_type;
  if (solver == SBA_SOLVER_AUTO) {
    const int nimg = o.num_config_images > 0 ? o.num_config_images : p.n_img;
    solver = nimg <= 50 ? SBA_DENSE_SCHUR : (nimg <= 1000 ? SBA_SPARSE_SCHUR : SBA_ITERATIVE_SCHUR);
  }
  sum->linear_solver_type_used = solver;
  trace->clear();

  Linearization lin;
  Linearize(p, o.loss_function_type, o.loss_function_scale, true, &lin, threads);
  double x_cost = lin.cost;
  const double fixed_cost = lin.fixed_cost;
  sum->fixed_cost = fixed_cost;
  sum->initial_cost = x_cost + fixed_cost;

  // gradient (unscaled J) for the max-norm column of the table
  auto gradient_max_norm = [&](const Linearization& l, const std::vector<double>* sc, const std::vector<double>* sp) {
    std::vector<double> gc, gp;
    ColumnPass(p, l, 1, &gc, &gp, threads);
    double m = 0;
    for (int j = 0; j < p.n_cdim; ++j) m = std::max(m, std::fabs(sc ? gc[j] / (*sc)[j] : gc[j]));
    for (size_t j = 0; j < gp.size(); ++j) m = std::max(m, std::fabs(sp ? gp[j] / (*sp)[j] : gp[j]));
    return m;
  };

  std::vector<double> scale_c(p.n_cdim, 1.0), scale_p((size_t)3 * p.n_pvar, 1.0);
  double gmax = gradient_max_norm(lin, nullptr, nullptr);
  if (o.jacobi_scaling) {
    std::vector<double> nc, np;
    ColumnSquaredNorms(p, lin, &nc, &np, threads);
    for (int j = 0; j < p.n_cdim; ++j) scale_c[j] = 1.0 / (1.0 + std::sqrt(nc[j]));
    for (size_t j = 0; j < np.size(); ++j) scale_p[j] = 1.0 / (1.0 + std::sqrt(np[j]));
    ScaleColumns(p, scale_c, scale_p, &lin, threads);
  }

  double radius = o.initial_trust_region_radius;
  double decrease_factor = 2.0;
  bool reuse_diagonal = false;
  std::vector<double> diag_c, diag_p;
  int num_consecutive_invalid = 0;
  double x_norm = AmbientNorm(p, nullptr);

  sba_iteration_summary it0;
  std::memset(&it0, 0, sizeof(it0));
  it0.iteration = 0; it0.step_is_valid = 0; it0.step_is_successful = 0;
  it0.cost = x_cost + fixed_cost; it0.gradient_max_norm = gmax; it0.trust_region_radius = radius; it0.eta = o.eta;
  trace->push_back(it0);

  int termination = SBA_NO_CONVERGENCE;
  int iteration = 0;
  while (true) {
    // FinalizeIterationAndCheckIfMinimizerCanContinue
    if (iteration >= o.max_num_iterations) { termination = SBA_NO_CONVERGENCE; break; }
    if (trace->back().gradient_max_norm <= o.gradient_tolerance) { termination = SBA_CONVERGENCE; break; }
    if (radius < o.min_trust_region_radius) { termination = SBA_CONVERGENCE; break; }
    ++iteration;
    sba_iteration_summary it;
    std::memset(&it, 0, sizeof(it));
    it.iteration = iteration; it.eta = o.eta;

    // ---- LevenbergMarquardtStrategy::ComputeStep
    if (!reuse_diagonal) {
      ColumnSquaredNorms(p, lin, &diag_c, &diag_p, threads);
      for (auto& d : diag_c) d = std::min(std::max(d, o.min_lm_diagonal), o.max_lm_diagonal);
      for (auto& d : diag_p) d = std::min(std::max(d, o.min_lm_diagonal), o.max_lm_diagonal);
    }
    std::vector<double> Dc(p.n_cdim), Dp(diag_p.size());
    for (int j = 0; j < p.n_cdim; ++j) Dc[j] = std::sqrt(diag_c[j] / radius);
    for (size_t j = 0; j < diag_p.size(); ++j) Dp[j] = std::sqrt(diag_p[j] / radius);

    StepContext ctx;
    ctx.p = &p; ctx.lin = &lin; ctx.Dc = Dc.data(); ctx.Dp = Dp.data(); ctx.threads = threads;
    bool solver_ok = ComputePointBlocks(&ctx);
    std::vector<double> yc(p.n_cdim, 0.0), yp((size_t)3 * p.n_pvar, 0.0);
    int lin_iters = 0;
    if (solver_ok && p.n_cdim > 0) {
      std::vector<double> rhs;
      ReducedRhs(ctx, &rhs);
      if (solver == SBA_ITERATIVE_SCHUR) {
        std::vector<double> Minv;
        solver_ok = BuildSchurJacobi(ctx, &Minv);
        if (solver_ok) {
          const int rc = SolveCG(ctx, Minv, rhs, o.min_linear_solver_iterations,
                                 o.max_linear_solver_iterations, o.eta, &yc, &lin_iters);
          solver_ok = (rc != 2);
        }
      } else {
        std::vector<double> S;
        BuildDenseSchur(ctx, &S);
        solver_ok = CholeskyLower(S.data(), p.n_cdim, threads);
        if (solver_ok) { yc = rhs; CholeskySolve(S.data(), p.n_cdim, yc.data()); }
        lin_iters = 1;
      }
    }
    if (solver_ok) {
      // back substitution: y_p = Cinv (g_p - sum_o E^T F y_c)
#pragma omp parallel for num_threads(threads) schedule(static)
      for (int pt = 0; pt < p.n_pt; ++pt) {
        const int po = p.p_off[pt];
        if (po < 0) continue;
        double u[3] = {ctx.g[3 * (size_t)po], ctx.g[3 * (size_t)po + 1], ctx.g[3 * (size_t)po + 2]};
        for (int64_t idx = p.pt_ptr[pt]; idx < p.pt_ptr[pt + 1]; ++idx) {
          const int64_t k = p.pt_obs[idx];
          int cols[24];
          const int n = p.ObsCols(k, cols);
          const double* E = OBS_E(lin, p, k);
          double f[3];
          RowsTimesX(p, OBS_F(lin, p, k), cols, n, yc.data(), f);
          for (int a = 0; a < 3; ++a)
            for (int row = 0; row < p.nr; ++row) u[a] -= E[3 * row + a] * f[row];
        }
        const double* Ci = &ctx.Cinv[(size_t)po * 9];
        for (int a = 0; a < 3; ++a) yp[3 * (size_t)po + a] = Ci[a * 3] * u[0] + Ci[a * 3 + 1] * u[1] + Ci[a * 3 + 2] * u[2];
      }
      for (double v : yc) if (!std::isfinite(v)) solver_ok = false;
      for (double v : yp) if (!std::isfinite(v)) solver_ok = false;
    }
    reuse_diagonal = true;
    it.linear_solver_iterations = lin_iters;
    sum->num_linear_solver_iterations += lin_iters;

    double model_cost_change = 0.0;
    bool step_valid = false;
    std::vector<double> step_c(p.n_cdim), step_p(yp.size());
    if (solver_ok) {
      for (int j = 0; j < p.n_cdim; ++j) step_c[j] = -yc[j];
      for (size_t j = 0; j < yp.size(); ++j) step_p[j] = -yp[j];
      // model_cost_change = -(J step)'(r + J step / 2)
      const int nt = std::max(1, threads);
      std::vector<double> acc_t(nt * 8, 0.0);  // per thread, combined in thread order (see Linearize)
#pragma omp parallel for num_threads(nt) schedule(static)
      for (int64_t k = 0; k < p.n_obs; ++k) {
        if (!p.obs_active[k]) continue;
        double& acc = acc_t[8 * omp_get_thread_num()];
        int cols[24];
        const int n = p.ObsCols(k, cols);
        double m[3];
        RowsTimesX(p, OBS_F(lin, p, k), cols, n, step_c.data(), m);
        const int po = p.p_off[p.obs_pt[k]];
        if (po >= 0) {
          const double* E = OBS_E(lin, p, k);
          for (int row = 0; row < p.nr; ++row)
            for (int a = 0; a < 3; ++a) m[row] += E[3 * row + a] * step_p[3 * (size_t)po + a];
        }
        const double* r = OBS_R(lin, p, k);
        for (int row = 0; row < p.nr; ++row) acc += m[row] * (r[row] + m[row] / 2.0);
      }
      double acc = 0.0;
      for (int t = 0; t < nt; ++t) acc += acc_t[8 * t];
      model_cost_change = -acc;
      step_valid = model_cost_change > 0.0;
    }
    it.step_is_valid = step_valid;
    if (!step_valid) {
      // HandleInvalidStep -> StepIsInvalid == StepRejected(0)
      ++num_consecutive_invalid;
      radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;
      it.cost = x_cost + fixed_cost; it.cost_change = 0.0; it.gradient_max_norm = trace->back().gradient_max_norm;
      it.step_norm = 0.0; it.relative_decrease = 0.0; it.trust_region_radius = radius;
      trace->push_back(it);
      sum->num_unsuccessful_steps++;
      if (num_consecutive_invalid >= o.max_num_consecutive_invalid_steps) { termination = SBA_FAILURE; break; }
      continue;
    }
    num_consecutive_invalid = 0;
    std::vector<double> delta_c(p.n_cdim), delta_p(step_p.size());
    for (int j = 0; j < p.n_cdim; ++j) delta_c[j] = step_c[j] * scale_c[j];
    for (size_t j = 0; j < step_p.size(); ++j) delta_p[j] = step_p[j] * scale_p[j];

    State cand;  // candidate state; the problem evaluates it in place (swapped in and out)
    ApplyPlus(p, delta_c, delta_p, &cand);
    SwapState(&p, &cand);
    double cand_cost = CostOnly(p, o.loss_function_type, o.loss_function_scale, threads);
    SwapState(&p, &cand);
    if (!std::isfinite(cand_cost)) cand_cost = std::numeric_limits<double>::max();
    const double step_norm = AmbientNorm(p, &cand);
    it.step_norm = step_norm;
    it.cost_change = x_cost - cand_cost;
    // ParameterToleranceReached
    // (Minimize() returns from both tests before the iteration's row is appended to Summary::iterations)
    if (step_norm <= o.parameter_tolerance * (x_norm + o.parameter_tolerance)) { termination = SBA_CONVERGENCE; break; }
    // FunctionToleranceReached
    if (std::fabs(it.cost_change) <= o.function_tolerance * x_cost) { termination = SBA_CONVERGENCE; break; }
    const double relative_decrease = (x_cost - cand_cost) / model_cost_change;
    it.relative_decrease = relative_decrease;
    if (relative_decrease > o.min_relative_decrease) {
      // HandleSuccessfulStep
      SwapState(&p, &cand);
      x_norm = AmbientNorm(p, nullptr);
      Linearize(p, o.loss_function_type, o.loss_function_scale, true, &lin, threads);
      x_cost = lin.cost;
      ScaleColumns(p, scale_c, scale_p, &lin, threads);
      it.gradient_max_norm = gradient_max_norm(lin, &scale_c, &scale_p);
      it.step_is_successful = 1;
      radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * relative_decrease - 1.0, 3));
      radius = std::min(o.max_trust_region_radius, radius);
      decrease_factor = 2.0; reuse_diagonal = false;
      sum->num_successful_steps++;
    } else {
      radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;
      it.gradient_max_norm = trace->back().gradient_max_norm;
      sum->num_unsuccessful_steps++;
    }
    it.cost = x_cost + fixed_cost; it.trust_region_radius = radius;
    trace->push_back(it);
  }
  sum->termination_type = termination;
  sum->final_cost = x_cost + fixed_cost;
  sum->num_iterations = (int)trace->size();
  sum->total_time_in_seconds =
      std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
  return SBA_OK;
}

}  // namespace oracle
