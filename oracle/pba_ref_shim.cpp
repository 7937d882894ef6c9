// ORACLE / test infrastructure only.  Thin extern "C" door into the REFERENCE's own lib/PBA CPU solver, compiled
// from the sources where they lie under /root/reference/lib/PBA (oracle/Makefile, target `ref`) into
// oracle/_ref/libpba_ref.so.  It is the one reference-owned bundle-adjustment solver that builds without
// Ceres / Eigen / glog / Boost.  Feeding PBA follows what the reference's wrapper does:
//   ParallelBundleAdjuster::Solve / AddImagesToProblem / AddPointsToProblem
//   /root/reference/src/optim/bundle_adjustment.cc:560-627, 699-783
// -- SIMPLE_RADIAL only, unshared intrinsics, measurements relative to the principal point (:770-771), row-major
// rotation matrix (:713-720), observations contiguous per point (:644-646), PBA_CPU_DOUBLE above 100 k residuals (:577-580).
// PBA is a different LM (multiplicative damping, float storage of cameras / points), so this is a second-opinion
// baseline and sanity check, not a bit-level oracle.
#include <chrono>
#include <cmath>
#include <cstdint>
#include <vector>

#include "PBA/pba.h"
#include "../include/sba_b200.h"

namespace {
void RotationRowMajor(const double* q, double* R) {  // QuaternionToRotationMatrix, src/base/pose.cc:63-80 (q = w,x,y,z)
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const double w = q[0] / n, x = q[1] / n, y = q[2] / n, z = q[3] / n;
  R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - w * z);     R[2] = 2 * (x * z + w * y);
  R[3] = 2 * (x * y + w * z);     R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - w * x);
  R[6] = 2 * (x * z - w * y);     R[7] = 2 * (y * z + w * x);     R[8] = 1 - 2 * (x * x + y * y);
}
}  // namespace

extern "C" {

// Returns 0 on success, 1 when the problem is outside what ParallelBundleAdjuster::IsSupported accepts
// (bundle_adjustment.cc:629-651).  out[0] initial cost, out[1] final cost (the wrapper's MSE * residuals / 4,
// :617-622), out[2] LM iterations, out[3] seconds inside RunBundleAdjustment, out[4] threads requested.
int pba_ref_solve(const sba_problem_desc* d, int max_num_iterations, int num_threads, int use_double, double* out,
                  double* points_out) {
  const int n_img = d->num_images, n_pt = d->num_points;
  const int64_t n_obs = d->num_observations;
  std::vector<uint8_t> cam_used(static_cast<size_t>(d->num_cameras), 0);
  for (int i = 0; i < n_img; ++i) {
    const int c = d->image_camera[i];
    if (d->camera_model[c] != SBA_SIMPLE_RADIAL || cam_used[c]) return 1;
    cam_used[c] = 1;
  }
  std::vector<pba::CameraT> cams(static_cast<size_t>(n_img));
  for (int i = 0; i < n_img; ++i) {
    const int c = d->image_camera[i];
    const double* prm = d->camera_params + static_cast<size_t>(c) * SBA_MAX_CAMERA_PARAMS;
    double R[9];
    RotationRowMajor(d->image_qvec + 4 * static_cast<size_t>(i), R);
    pba::CameraT& cam = cams[i];
    cam.SetFocalLength(prm[0]);
    cam.SetProjectionDistortion(prm[3]);
    cam.SetMatrixRotation(R);
    cam.SetTranslation(d->image_tvec + 3 * static_cast<size_t>(i));
    const bool intr_const = (d->camera_param_constant_mask[c] & 0x9) == 0x9;  // f and k held
    if (d->image_pose_constant && d->image_pose_constant[i]) cam.SetConstantCamera();
    else if (intr_const) cam.SetFixedIntrinsic();
    else cam.SetVariableCamera();
  }
  std::vector<pba::Point3D> pts(static_cast<size_t>(n_pt));
  for (int p = 0; p < n_pt; ++p) pts[p].SetPoint(d->points + 3 * static_cast<size_t>(p));
  // contiguous per point: stable counting sort of the observations by point index
  std::vector<int64_t> start(static_cast<size_t>(n_pt) + 1, 0);
  for (int64_t k = 0; k < n_obs; ++k) start[d->obs_point[k] + 1]++;
  for (int p = 0; p < n_pt; ++p) start[p + 1] += start[p];
  std::vector<pba::Point2D> meas(static_cast<size_t>(n_obs));
  std::vector<int> cam_idx(static_cast<size_t>(n_obs)), pt_idx(static_cast<size_t>(n_obs));
  std::vector<int64_t> at(start.begin(), start.end() - 1);
  for (int64_t k = 0; k < n_obs; ++k) {
    const int p = d->obs_point[k], i = d->obs_image[k];
    const double* prm = d->camera_params + static_cast<size_t>(d->image_camera[i]) * SBA_MAX_CAMERA_PARAMS;
    const int64_t s = at[p]++;
    meas[s].SetPoint2D(d->obs_xy[2 * k] - prm[1], d->obs_xy[2 * k + 1] - prm[2]);
    cam_idx[s] = i;
    pt_idx[s] = p;
  }
  const int num_residuals = static_cast<int>(2 * n_obs);
  pba::ParallelBA ba(use_double ? pba::ParallelBA::PBA_CPU_DOUBLE : pba::ParallelBA::PBA_CPU_FLOAT, num_threads);
  ba.SetNextBundleMode(pba::ParallelBA::BUNDLE_FULL);
  ba.EnableRadialDistortion(pba::ParallelBA::PBA_PROJECTION_DISTORTION);
  ba.SetFixedIntrinsics(false);
  pba::ConfigBA* cfg = ba.GetInternalConfig();
  cfg->__lm_delta_threshold /= 100.0f;
  cfg->__lm_gradient_threshold /= 100.0f;
  cfg->__lm_mse_threshold = 0.0f;
  cfg->__cg_min_iteration = 10;
  cfg->__verbose_level = 0;
  cfg->__lm_max_iteration = max_num_iterations;
  ba.SetCameraData(cams.size(), cams.data());
  ba.SetPointData(pts.size(), pts.data());
  ba.SetProjection(meas.size(), meas.data(), pt_idx.data(), cam_idx.data());
  const auto t0 = std::chrono::steady_clock::now();
  ba.RunBundleAdjustment();
  const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (out) {
    out[0] = cfg->GetInitialMSE() * num_residuals / 4.0;
    out[1] = cfg->GetFinalMSE() * num_residuals / 4.0;
    out[2] = cfg->GetIterationsLM();
    out[3] = sec;
    out[4] = num_threads;
  }
  if (points_out)
    for (int p = 0; p < n_pt; ++p) pts[p].GetPoint(points_out + 3 * static_cast<size_t>(p));
  return 0;
}

}  // extern "C"
