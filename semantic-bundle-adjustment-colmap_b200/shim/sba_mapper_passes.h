// sba_mapper_passes.h -- header-only C++17 shims for the mapper-side steps around global bundle adjustment
//  over include/sba_b200.h:
//
//   Normalize<Reconstruction>(rec, extent, p0, p1, use_images)
//       Reconstruction::Normalize                      /root/reference/src/base/reconstruction.cc:321-345
//       (the mapper's step after global BA,            /root/reference/src/sfm/incremental_mapper.cc:711)
//   RefineAbsolutePose<Camera, Vec2, Vec3, Vec4>(options, inlier_mask, points2D, points3D, qvec, tvec, camera)
//       RefineAbsolutePose                             /root/reference/src/estimators/pose.cc:198-322
//       AbsolutePoseRefinementOptions                  /root/reference/src/estimators/pose.h:81-105
//
// Accessor contract on top of what sba_bundle_adjuster.h needs (all present in COLMAP's classes):
//   Reconstruction::Images()       iterable of (image_t, Image)                 reconstruction.h:96
//   Reconstruction::RegImageIds()  std::vector<image_t>                         reconstruction.h:99
//   Reconstruction::Points3D()     iterable of (point3D_t, Point3D)             reconstruction.h:97
//   Image::Qvec() / Tvec(), Point3D::XYZ()  with .data()
// Both shims only flatten, call the library and scatter back: no arithmetic of the path runs on the host, and there is
// no CPU fallback (std::runtime_error when the device call fails).
#ifndef SBA_SHIM_MAPPER_PASSES_H_
#define SBA_SHIM_MAPPER_PASSES_H_

#include <cstdint>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <unordered_set>
#include <vector>

#include "sba_b200.h"

namespace sba_shim {

// Returns the scale applied (1 when fewer than two elements: nothing changes, reconstruction.cc:325-328).
template <class Reconstruction>
double Normalize(Reconstruction* reconstruction, double extent = 10.0, double p0 = 0.1, double p1 = 0.9, bool use_images = true) {
  if (!(extent > 0.0)) throw std::invalid_argument("Normalize: extent must be positive");  // CHECK_GT, :323
  std::vector<double> q, t, X;
  std::vector<uint8_t> reg;
  std::vector<std::decay_t<decltype(reconstruction->RegImageIds().front())>> image_ids;
  std::unordered_set<uint64_t> registered;
  for (const auto id : reconstruction->RegImageIds()) registered.insert((uint64_t)id);
  for (const auto& el : reconstruction->Images()) {
    image_ids.push_back(el.first);
    const auto& im = el.second;
    for (int j = 0; j < 4; ++j) q.push_back(im.Qvec().data()[j]);
    for (int j = 0; j < 3; ++j) t.push_back(im.Tvec().data()[j]);
    reg.push_back(registered.count((uint64_t)el.first) ? 1 : 0);
  }
  std::vector<uint64_t> point_ids;
  for (const auto& el : reconstruction->Points3D()) {
    point_ids.push_back((uint64_t)el.first);
    for (int j = 0; j < 3; ++j) X.push_back(el.second.XYZ().data()[j]);
  }
  double scale = 1.0, tr[3];
  const int rc = sba_normalize((int32_t)image_ids.size(), q.data(), t.data(), reg.data(), (int64_t)point_ids.size(), X.data(), extent, p0,
                               p1, use_images ? 1 : 0, &scale, tr);
  if (rc != SBA_OK) throw std::runtime_error(std::string("sba_normalize: ") + sba_last_error());
  for (size_t k = 0; k < image_ids.size(); ++k) {
    auto& im = reconstruction->Image(image_ids[k]);
    for (int j = 0; j < 4; ++j) im.Qvec().data()[j] = q[4 * k + j];
    for (int j = 0; j < 3; ++j) im.Tvec().data()[j] = t[3 * k + j];
  }
  for (size_t k = 0; k < point_ids.size(); ++k) {
    auto& p = reconstruction->Point3D(point_ids[k]);
    for (int j = 0; j < 3; ++j) p.XYZ().data()[j] = X[3 * k + j];
  }
  return scale;
}

struct AbsolutePoseRefinementOptions {  // pose.h:81-105
  double gradient_tolerance = 1.0;
  int max_num_iterations = 100;
  double loss_function_scale = 1.0;
  bool refine_focal_length = true;
  bool refine_extra_params = true;
  bool print_summary = true;
  void Check() const {
    if (gradient_tolerance < 0.0 || max_num_iterations < 0 || loss_function_scale < 0.0)
      throw std::invalid_argument("AbsolutePoseRefinementOptions::Check failed");
  }
};

// In place on *qvec, *tvec and camera->Params(); returns Summary::IsSolutionUsable() like the reference.  The
// covariance output of the reference (ceres::Covariance) is not provided.  summary_out: optional.
template <class Camera, class Vec2, class Vec3, class Vec4>
bool RefineAbsolutePose(const AbsolutePoseRefinementOptions& options, const std::vector<char>& inlier_mask,
                        const std::vector<Vec2>& points2D, const std::vector<Vec3>& points3D, Vec4* qvec, Vec3* tvec,
                        Camera* camera, sba_summary* summary_out = nullptr) {
  if (inlier_mask.size() != points2D.size() || points2D.size() != points3D.size())
    throw std::invalid_argument("RefineAbsolutePose: inlier_mask, points2D and points3D differ in length");  // CHECK_EQ, pose.cc:204-205
  options.Check();
  std::vector<double> X, xy;
  std::vector<int32_t> obs_img, obs_pt;
  for (size_t i = 0; i < points2D.size(); ++i) {
    if (!inlier_mask[i]) continue;  // pose.cc:221-224
    obs_img.push_back(0); obs_pt.push_back((int32_t)obs_pt.size());
    for (int j = 0; j < 3; ++j) X.push_back(points3D[i].data()[j]);
    for (int j = 0; j < 2; ++j) xy.push_back(points2D[i].data()[j]);
  }
  if (obs_pt.empty()) return true;  // an empty ceres::Problem converges at once: the solution is usable
  const size_t np = camera->NumParams();
  // constant camera parameters: principal point always; focal length / extra parameters by option (pose.cc:249-281)
  uint16_t mask = 0;
  for (size_t k : camera->PrincipalPointIdxs()) mask |= (uint16_t)(1u << k);
  if (!options.refine_focal_length) for (size_t k : camera->FocalLengthIdxs()) mask |= (uint16_t)(1u << k);
  if (!options.refine_extra_params) for (size_t k : camera->ExtraParamsIdxs()) mask |= (uint16_t)(1u << k);
  if (!options.refine_focal_length && !options.refine_extra_params) mask = (uint16_t)((1u << np) - 1);
  std::vector<double> params(SBA_MAX_CAMERA_PARAMS, 0.0);
  for (size_t k = 0; k < np; ++k) params[k] = camera->Params()[k];
  std::vector<uint8_t> point_constant(obs_pt.size(), 1);
  const int32_t image_camera = 0, model = camera->ModelId();
  sba_problem_desc d{};
  d.num_images = 1; d.num_cameras = 1; d.num_points = (int32_t)obs_pt.size(); d.num_observations = (int64_t)obs_pt.size();
  d.image_qvec = qvec->data(); d.image_tvec = tvec->data(); d.image_camera = &image_camera;
  d.camera_model = &model; d.camera_params = params.data(); d.camera_param_constant_mask = &mask;
  d.points = X.data(); d.point_constant = point_constant.data();
  d.obs_image = obs_img.data(); d.obs_point = obs_pt.data(); d.obs_xy = xy.data();
  d.shard_rank = 0; d.shard_count = 1; d.global_num_observations = d.num_observations;
  sba_solver_options o;
  sba_default_solver_options(&o);
  o.function_tolerance = 1e-6; o.parameter_tolerance = 1e-8;  // Ceres defaults: pose.cc sets only the next three (:283-286)
  o.gradient_tolerance = options.gradient_tolerance;
  o.max_num_iterations = options.max_num_iterations;
  o.linear_solver_type = SBA_DENSE_SCHUR;  // DENSE_QR: with no point block to eliminate the same normal equations
  o.loss_function_type = SBA_LOSS_CAUCHY; o.loss_function_scale = options.loss_function_scale;
  sba_handle* h = nullptr;
  if (sba_create(&d, &h) != SBA_OK) throw std::runtime_error(std::string("sba_create: ") + sba_last_error());
  struct Guard { sba_handle* h; ~Guard() { if (h) sba_destroy(h); } } guard{h};
  sba_summary s{};
  if (sba_solve(h, &o, nullptr, nullptr, &s) != SBA_OK) throw std::runtime_error(std::string("sba_solve: ") + sba_last_error());
  std::vector<double> Xout(X.size());
  if (sba_get_parameters(h, qvec->data(), tvec->data(), params.data(), Xout.data()) != SBA_OK)
    throw std::runtime_error(std::string("sba_get_parameters: ") + sba_last_error());
  for (size_t k = 0; k < np; ++k) camera->Params()[k] = params[k];
  if (summary_out) *summary_out = s;
  return s.termination_type == SBA_CONVERGENCE || s.termination_type == SBA_NO_CONVERGENCE || s.termination_type == SBA_USER_SUCCESS;
}

}  // namespace sba_shim

#endif  // SBA_SHIM_MAPPER_PASSES_H_
