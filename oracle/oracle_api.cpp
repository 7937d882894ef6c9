// ORACLE (test infrastructure only -- never linked into the product library).
// extern "C" surface of the CPU restatement, loaded with ctypes by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
#include <limits>
#include <omp.h>

#include <cstring>
#include <vector>

#include "ba_oracle.h"
#include "projection.h"
#include "rotation.h"
#include "semantic_oracle.h"

using namespace oracle;

extern "C" {

int oracle_num_threads(void) { return omp_get_max_threads(); }

void oracle_world_to_image(int model, const double* params, double u, double v, double* x, double* y) {
  Project<double>(model, params, u, v, x, y);
}
void oracle_image_to_world(int model, const double* params, double x, double y, double* u, double* v) {
  Unproject(model, params, x, y, u, v);
}
int oracle_model_num_params(int model) { return NumParams(model); }

// CalculateSquaredReprojectionError (src/base/projection.cc:111-128) and the depth HasPointPositiveDepth tests
// (:191-195), with the rotation matrix of the normalised quaternion (src/base/pose.cc:63-91).
void oracle_observation_error(int model, const double* q, const double* t, const double* X, const double* params,
                              const double* xy, double* sq_error, double* depth) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  const double n = std::sqrt(w * w + x * x + y * y + z * z);
  if (n == 0.0) { w = 1.0; x = y = z = 0.0; } else { w /= n; x /= n; y /= n; z /= n; }
  const double R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z),     2 * (x * z + w * y),
                       2 * (x * y + w * z),     1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                       2 * (x * z - w * y),     2 * (y * z + w * x),     1 - 2 * (x * x + y * y)};
  const double px = R[0] * X[0] + R[1] * X[1] + R[2] * X[2] + t[0];
  const double py = R[3] * X[0] + R[4] * X[1] + R[5] * X[2] + t[1];
  const double pz = R[6] * X[0] + R[7] * X[1] + R[8] * X[2] + t[2];
  *depth = pz;
  if (pz < std::numeric_limits<double>::epsilon()) { *sq_error = std::numeric_limits<double>::max(); return; }
  double ix, iy;
  Project<double>(model, params, px / pz, py / pz, &ix, &iy);
  *sq_error = (ix - xy[0]) * (ix - xy[0]) + (iy - xy[1]) * (iy - xy[1]);
}

// CostFunction::Evaluate of one residual block (cost_functions.h:57-81 / 115-140)
void oracle_eval_observation(int model, const double* q, const double* t, const double* X,
                             const double* params, const double* xy, double* r, double* Jq,
                             double* Jt, double* JX, double* Jk) {
  EvalObservation(model, q, t, X, params, xy, r, Jq, Jt, JX, Jk);
}

// Rig functor (cost_functions.h:173-209): residual only, for the known answers at
// cost_functions_test.cc:101-132.
void oracle_eval_rig_observation(int model, const double* rig_q, const double* rig_t,
                                 const double* rel_q, const double* rel_t, const double* X,
                                 const double* params, const double* xy, double* r) {
  double q[4], t[3];
  QuaternionProduct(rel_q, rig_q, q);
  UnitQuaternionRotatePoint(rel_q, rig_t, t);
  t[0] += rel_t[0]; t[1] += rel_t[1]; t[2] += rel_t[2];
  EvalObservation(model, q, t, X, params, xy, r, nullptr, nullptr, nullptr, nullptr);
}

// Rig functor with its autodiff Jacobian (2 x 29: rig_q, rig_t, rel_q, rel_t, X, params padded to 12) -- the layout
// of sba_rig_evaluate (include/sba_b200.h).
void oracle_eval_rig_observation_jac(int model, const double* rig_q, const double* rig_t, const double* rel_q,
                                     const double* rel_t, const double* X, const double* params, const double* xy,
                                     double* r, double* J) {
  EvalRigObservation(model, rig_q, rig_t, rel_q, rel_t, X, params, xy, r, J);
}

// Problem::Evaluate equivalent in the sba_evaluate() / sba_get_semantic_jacobian() output format (include/sba_b200.h).
// The first two rows of every block go to residuals / jac_*; the third (semantic, extension X1) row -- present when the
// description carries maps -- to sem_residuals [n], sem_pose [n*6], sem_intr [n*12], sem_point [n*3].
int oracle_ba_evaluate_rows(const sba_problem_desc* d, int loss_type, double loss_scale, int apply_loss,
                            int threads, double* cost, double* residuals, double* jac_pose,
                            double* jac_intr, double* jac_point, double* jac_qvec_ambient,
                            double* sem_residuals, double* sem_pose, double* sem_intr, double* sem_point) {
  BAProblem p;
  p.Build(*d);
  if (threads <= 0) threads = omp_get_max_threads();
  Linearization lin;
  Linearize(p, loss_type, loss_scale, apply_loss != 0, &lin, threads);
  if (cost) *cost = lin.cost + lin.fixed_cost;
  const int dc = p.dc_max, nr = p.nr, KP = SBA_MAX_CAMERA_PARAMS;
  if ((sem_residuals || sem_pose || sem_intr || sem_point) && nr != 3) return 1;
  for (int64_t k = 0; k < p.n_obs; ++k) {
    const int i = p.obs_img[k], c = p.img_cam[i], pt = p.obs_pt[k];
    const double* F = &lin.F[(size_t)k * nr * dc];
    const double* E = &lin.E[(size_t)k * nr * 3];
    const double* r = &lin.r[(size_t)k * nr];  // inactive blocks: the raw functor output
    if (residuals) { residuals[2 * k] = r[0]; residuals[2 * k + 1] = r[1]; }
    if (sem_residuals) sem_residuals[k] = r[2];
    int n = 0;
    if (jac_pose) std::memset(&jac_pose[k * 12], 0, 12 * sizeof(double));
    if (jac_intr) std::memset(&jac_intr[k * 2 * KP], 0, 2 * KP * sizeof(double));
    if (jac_point) std::memset(&jac_point[k * 6], 0, 6 * sizeof(double));
    if (sem_pose) std::memset(&sem_pose[k * 6], 0, 6 * sizeof(double));
    if (sem_intr) std::memset(&sem_intr[k * KP], 0, KP * sizeof(double));
    if (sem_point) std::memset(&sem_point[k * 3], 0, 3 * sizeof(double));
    if (p.obs_active[k]) {
      if (p.q_off[i] >= 0) {
        for (int j = 0; j < 3; ++j) {
          if (jac_pose) { jac_pose[k * 12 + j] = F[n + j]; jac_pose[k * 12 + 6 + j] = F[dc + n + j]; }
          if (sem_pose) sem_pose[k * 6 + j] = F[2 * dc + n + j];
        }
        n += 3;
      }
      if (p.t_off[i] >= 0) {
        for (int j = 0; j < p.t_dim[i]; ++j) {
          const int a = p.t_cols[3 * i + j];
          if (jac_pose) { jac_pose[k * 12 + 3 + a] = F[n + j]; jac_pose[k * 12 + 9 + a] = F[dc + n + j]; }
          if (sem_pose) sem_pose[k * 6 + 3 + a] = F[2 * dc + n + j];
        }
        n += p.t_dim[i];
      }
      if (p.k_off[c] >= 0) {
        for (int j = 0; j < p.k_dim[c]; ++j) {
          const int a = p.k_cols[KP * c + j];
          if (jac_intr) { jac_intr[k * 2 * KP + a] = F[n + j]; jac_intr[k * 2 * KP + KP + a] = F[dc + n + j]; }
          if (sem_intr) sem_intr[k * KP + a] = F[2 * dc + n + j];
        }
        n += p.k_dim[c];
      }
      if (p.p_off[pt] >= 0) {
        if (jac_point) for (int j = 0; j < 6; ++j) jac_point[k * 6 + j] = E[j];
        if (sem_point) for (int j = 0; j < 3; ++j) sem_point[k * 3 + j] = E[6 + j];
      }
    }
    if (jac_qvec_ambient) {
      double rr[3], Jq[12];
      EvalObservation(p.model[c], &p.q[4 * i], &p.t[3 * i], &p.X[3 * pt], &p.cam[KP * c],
                      &p.obs_xy[2 * k], rr, Jq, nullptr, nullptr, nullptr);
      for (int j = 0; j < 8; ++j) jac_qvec_ambient[k * 8 + j] = Jq[j];
    }
  }
  return 0;
}
int oracle_ba_evaluate(const sba_problem_desc* d, int loss_type, double loss_scale, int apply_loss,
                       int threads, double* cost, double* residuals, double* jac_pose,
                       double* jac_intr, double* jac_point, double* jac_qvec_ambient) {
  return oracle_ba_evaluate_rows(d, loss_type, loss_scale, apply_loss, threads, cost, residuals, jac_pose, jac_intr,
                                 jac_point, jac_qvec_ambient, nullptr, nullptr, nullptr, nullptr);
}

// BundleAdjustmentOptions' solver options (src/optim/bundle_adjustment.h:78-93) on top of the ceres::Solver::Options
// 2.1.0 defaults, restated here so that the CPU arm of bench.py needs nothing from the product library.
void oracle_default_solver_options(sba_solver_options* o) {
  std::memset(o, 0, sizeof(*o));
  o->linear_solver_type = SBA_SOLVER_AUTO;
  o->max_num_iterations = 100;            // bundle_adjustment.h:85
  o->max_linear_solver_iterations = 200;  // :86
  o->min_linear_solver_iterations = 0;
  o->max_num_consecutive_invalid_steps = 10;  // :87
  o->function_tolerance = 0.0; o->gradient_tolerance = 0.0; o->parameter_tolerance = 0.0;  // :81-83
  o->loss_function_type = SBA_LOSS_TRIVIAL; o->loss_function_scale = 1.0;
  o->initial_trust_region_radius = 1e4; o->max_trust_region_radius = 1e16; o->min_trust_region_radius = 1e-32;
  o->min_relative_decrease = 1e-3; o->min_lm_diagonal = 1e-6; o->max_lm_diagonal = 1e32;
  o->eta = 1e-1; o->jacobi_scaling = 1; o->update_state_every_iteration = 0;
}

int oracle_ba_counts(const sba_problem_desc* d, sba_summary* s) {
  BAProblem p;
  p.Build(*d);
  std::memset(s, 0, sizeof(*s));
  s->num_residual_blocks = p.num_residual_blocks; s->num_residuals = p.num_residuals;
  s->num_residuals_reduced = p.num_residuals_reduced;
  s->num_parameter_blocks_reduced = p.num_parameter_blocks_reduced;
  s->num_parameters_reduced = p.num_parameters_reduced;
  s->num_effective_parameters_reduced = p.num_effective_parameters_reduced;
  return 0;
}

int oracle_ba_solve(const sba_problem_desc* d, const sba_solver_options* o, int threads,
                    sba_summary* summary, sba_iteration_summary* trace, int cap, int* n_trace,
                    double* qvec, double* tvec, double* cams, double* points) {
  BAProblem p;
  p.Build(*d);
  if (threads <= 0) threads = omp_get_max_threads();
  std::vector<sba_iteration_summary> tr;
  const int rc = Solve(&p, *o, summary, &tr, threads);
  if (trace && n_trace) {
    const int n = std::min<int>(cap, (int)tr.size());
    for (int i = 0; i < n; ++i) trace[i] = tr[i];
    *n_trace = n;
  }
  if (qvec) std::memcpy(qvec, p.q.data(), p.q.size() * sizeof(double));
  if (tvec) std::memcpy(tvec, p.t.data(), p.t.size() * sizeof(double));
  if (cams) std::memcpy(cams, p.cam.data(), p.cam.size() * sizeof(double));
  if (points) std::memcpy(points, p.X.data(), p.X.size() * sizeof(double));
  return rc;
}

// Timed pieces for bench.py's cpu_baseline: one residual+Jacobian evaluation
// (Evaluator::Evaluate equivalent) on all cores; returns seconds.
double oracle_ba_time_linearize(const sba_problem_desc* d, int loss_type, double loss_scale, int threads, int repeat) {
  BAProblem p;
  p.Build(*d);
  if (threads <= 0) threads = omp_get_max_threads();
  Linearization lin;
  Linearize(p, loss_type, loss_scale, true, &lin, threads);  // warm
  const double t0 = omp_get_wtime();
  for (int i = 0; i < repeat; ++i) Linearize(p, loss_type, loss_scale, true, &lin, threads);
  return (omp_get_wtime() - t0) / repeat;
}

int oracle_semantic_evaluate(const sba_semantic_desc* d, int threads, double* residuals, int32_t* status, double* jac) {
  if (threads <= 0) threads = omp_get_max_threads();
  SemanticEvaluate(*d, threads, residuals, status, jac);
  return 0;
}

int oracle_cylinder_evaluate(const sba_cylinder_desc* d, int threads, double* residuals, double* jac) {
  if (threads <= 0) threads = omp_get_max_threads();
  CylinderEvaluate(*d, threads, residuals, jac);
  return 0;
}

void oracle_draw_quadrilateral(const double* pts8, int H, int W, uint8_t* mask_rowmajor) {
  DrawQuadrilateralRowMajor(pts8, H, W, mask_rowmajor);
}

int oracle_cylinder_quadrilateral(const double* cam_q, const double* cam_t, const double* cam_params,
                                  const double* cyl_q, const double* cyl_t, double radius, double height,
                                  double* pts8) {
  return CylinderQuadrilateral(cam_q, cam_t, cam_params, cyl_q, cyl_t, radius, height, pts8) ? 1 : 0;
}

// extension X1 (no reference implementation): per observation,
// weight * bilinear(map_of_image, projection), 0 where the sample is rejected.
int oracle_ba_sem_rows(const sba_problem_desc* d, double* out) {
  if (!d->sem_maps) return 1;
  const size_t HW = (size_t)d->sem_height * d->sem_width;
  for (int64_t k = 0; k < d->num_observations; ++k) {
    const int i = d->obs_image[k], c = d->image_camera[i], pt = d->obs_point[k];
    double p[3];
    UnitQuaternionRotatePoint(&d->image_qvec[4 * i], &d->points[3 * pt], p);
    p[0] += d->image_tvec[3 * i]; p[1] += d->image_tvec[3 * i + 1]; p[2] += d->image_tvec[3 * i + 2];
    p[0] /= p[2]; p[1] /= p[2];
    double x, y;
    Project<double>(d->camera_model[c], &d->camera_params[SBA_MAX_CAMERA_PARAMS * c], p[0], p[1], &x, &y);
    int ok = 0;
    const double v = BilinearSample(d->sem_maps + i * HW, d->sem_height, d->sem_width, x, y, &ok);
    out[k] = ok ? d->sem_weight * v : 0.0;
  }
  return 0;
}

double oracle_bilinear(const float* map, int H, int W, double x, double y, int* ok) {
  return BilinearSample(map, H, W, x, y, ok);
}

}  // extern "C"
