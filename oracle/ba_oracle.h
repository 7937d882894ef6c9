// ORACLE (test infrastructure only -- never linked into the product library).
//
// CPU restatement of the reference's Ceres-backed bundle-adjustment path:
//   residual blocks      /root/reference/src/base/cost_functions.h:57-81, 115-140
//   problem construction /root/reference/src/optim/bundle_adjustment.cc:326-530
//   loss                 /root/reference/src/optim/bundle_adjustment.cc:53-68
//   solver policy        /root/reference/src/optim/bundle_adjustment.cc:271-286
//   ceres::Solve         /root/reference/src/optim/bundle_adjustment.cc:306
// The LM / Schur / CG arithmetic lives in Ceres Solver 2.1.0, an un-vendored
// dependency (CMakeLists.txt:91; pinned at scripts/python/build.py:388-392).  Its
// published algorithm (trust_region_minimizer, levenberg_marquardt_strategy,
// schur_eliminator, implicit_schur_complement, conjugate_gradients_solver,
// schur_jacobi_preconditioner, corrector, QuaternionManifold / SubsetManifold) is
// restated here per SURVEY.md Appendix B.
//
// PARITY PINNING: residual values are pinned by the reference's known answers
// (src/base/cost_functions_test.cc:41-99), projection by the camera-model
// round-trip vectors (src/base/camera_models_test.cc:133-218), problem structure
// by the counts in src/optim/bundle_adjustment_test.cc:186-645.  The numeric LM
// trajectory is NOT pinned by any reference test ("parity unpinned" for the
// solve: no Ceres binary exists in this environment).
#pragma once
#include <cstdint>
#include <vector>

#include "../include/sba_b200.h"

namespace oracle {

struct BAProblem {
  int n_img = 0, n_cam = 0, n_pt = 0;
  int64_t n_obs = 0;
  std::vector<double> q, t, cam, X;  // current state (ambient)
  std::vector<int> img_cam, model;
  std::vector<uint8_t> pose_const, tvec_mask, pt_const;
  std::vector<uint16_t> cam_mask;
  std::vector<int> obs_img, obs_pt;
  std::vector<double> obs_xy;

  // ---- reduced-program structure (tangent space) ----
  // camera-side vector: per image [q(3)] [t(3-k)], per camera [free intrinsics]
  std::vector<int> q_off, t_off, t_dim, k_off, k_dim;  // -1 offset == constant/unused
  std::vector<int> t_cols;                             // [n_img*3] ambient idx of each tangent col
  std::vector<int> k_cols;                             // [n_cam*12]
  int n_cdim = 0;                                      // camera-side tangent dims
  std::vector<int> p_off;                              // [n_pt] offset/3 into point-side, -1 const/unused
  int n_pvar = 0;                                      // variable points
  int dc_max = 0;                                      // max camera-side cols per observation
  std::vector<uint8_t> obs_active;                     // residual block kept in reduced program
  // parameter-block list for SCHUR_JACOBI (offset, size) on the camera side
  std::vector<int> blk_off, blk_dim;
  // CSR observation lists
  std::vector<int64_t> pt_ptr; std::vector<int64_t> pt_obs;

  // extension X1 (no reference implementation): with per-image maps every residual
  // block has a third row  r_s = sem_weight * bilinear(map_img, x, y)  at the projection (x, y), bilinear as
  // Bitmap::InterpolateBilinear (src/util/bitmap.cc:238-269); the Jet of the interpolant is its Jacobian oracle.
  int nr = 2;                       // residual rows per block
  const float* sem_maps = nullptr;  // caller-owned, [n_img * sem_h * sem_w]
  int sem_h = 0, sem_w = 0;
  double sem_weight = 0.0;

  // counts (ceres::Solver::Summary)
  int num_residual_blocks = 0, num_residuals = 0, num_residuals_reduced = 0;
  int num_parameter_blocks_reduced = 0, num_parameters_reduced = 0, num_effective_parameters_reduced = 0;

  void Build(const sba_problem_desc& d);
  int ObsCols(int64_t k, int* cols) const;  // camera-side global columns of obs k
};

struct Linearization {  // nr = BAProblem::nr rows per block
  std::vector<double> r;  // [nr*n_obs]   (loss-corrected, NOT scaled)
  std::vector<double> F;  // [n_obs*nr*dc_max] camera-side local Jacobian, row-major nr x dc_max
  std::vector<double> E;  // [n_obs*nr*3]
  double cost = 0.0, fixed_cost = 0.0;
};

// raw functor evaluation of observation k at explicit parameter values:
// r[2], Jq[2x4], Jt[2x3], JX[2x3], Jk[2xP] (P = NumParams(model)), row-major.
void EvalRigObservation(int model, const double* rig_q, const double* rig_t, const double* rel_q, const double* rel_t,
                        const double* X, const double* params, const double* xy, double* r, double* J /* 2x29 or null */);
void EvalObservation(int model, const double* q, const double* t, const double* X,
                     const double* params, const double* xy, double* r, double* Jq, double* Jt,
                     double* JX, double* Jk);
// the same with the semantic row: r[3], every J block gets a third row (row-major 3 x cols); map == nullptr: 2 rows
void EvalObservationSem(int model, const double* q, const double* t, const double* X, const double* params,
                        const double* xy, const float* map, int H, int W, double weight, double* r, double* Jq,
                        double* Jt, double* JX, double* Jk);

void Linearize(const BAProblem& p, int loss_type, double loss_scale, bool apply_loss,
               Linearization* lin, int num_threads);
double CostOnly(const BAProblem& p, int loss_type, double loss_scale, int num_threads);

int Solve(BAProblem* p, const sba_solver_options& o, sba_summary* s,
          std::vector<sba_iteration_summary>* trace, int num_threads);

}  // namespace oracle
