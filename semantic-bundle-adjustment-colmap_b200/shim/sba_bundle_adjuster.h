// sba_bundle_adjuster.h -- header-only C++17 host side of the B200 bundle-adjustment path.
//
// Mirrors the front-end of the reference so that a caller keeps writing
//     BundleAdjuster ba(options, config);  ba.Solve(&reconstruction);  ba.Summary();
//   BundleAdjustmentOptions   /root/reference/src/optim/bundle_adjustment.h:48-100
//   BundleAdjustmentConfig    /root/reference/src/optim/bundle_adjustment.h:103-167
//   BundleAdjuster            /root/reference/src/optim/bundle_adjustment.h:171-203
// but hands the arithmetic to libsba_b200.so through the C ABI of include/sba_b200.h instead of
// building a ceres::Problem.  Everything below is host integer work: walking the containers in the
// reference's residual order (SetUp, bundle_adjustment.cc:326-342) and packing plain arrays.
//
// The adjuster is a template over the Reconstruction type.  It only uses the accessor contract
//   rec->Image(id)  : CameraId(), NormalizeQvec(), Qvec().data(), Tvec().data(), Points2D(), Point2D(i)
//   Point2D         : HasPoint3D(), Point3DId(), XY().data()
//   rec->Camera(id) : ModelId(), NumParams(), ParamsData(), FocalLengthIdxs(), PrincipalPointIdxs(), ExtraParamsIdxs()
//   rec->Point3D(id): XYZ().data(), Track().Length(), Track().Elements() -> {image_id, point2D_idx}
// (src/base/{reconstruction,image,camera,point2d,point3d,track}.h), so the same header compiles against
// COLMAP's own classes and against shim/mock_reconstruction.h (tests/shim_test.cpp).
//
// No CPU solve exists here: when libsba_b200.so reports no device, Solve() throws.
#ifndef SBA_SHIM_BUNDLE_ADJUSTER_H_
#define SBA_SHIM_BUNDLE_ADJUSTER_H_

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "sba_b200.h"

namespace sba_shim {

using camera_t = uint32_t;   // src/util/types.h:80-94
using image_t = uint32_t;
using point3D_t = uint64_t;

// ---- the slice of ceres::Solver::{Options,Summary} and ceres::IterationCallback the reference touches ----
using IterationSummary = sba_iteration_summary;

enum CallbackReturnType { SOLVER_CONTINUE = 0, SOLVER_TERMINATE_SUCCESSFULLY = 1, SOLVER_ABORT = 2 };

class IterationCallback {  // src/controllers/bundle_adjustment.cc:43-61 derives from ceres::IterationCallback
 public:
  virtual ~IterationCallback() = default;
  virtual CallbackReturnType operator()(const IterationSummary& summary) = 0;
};

enum LinearSolverType { POLICY = SBA_SOLVER_AUTO, DENSE_SCHUR = SBA_DENSE_SCHUR, SPARSE_SCHUR = SBA_SPARSE_SCHUR,
                        ITERATIVE_SCHUR = SBA_ITERATIVE_SCHUR };

struct SolverOptions {  // defaults: bundle_adjustment.h:80-93 over Ceres 2.1.0's
  double function_tolerance = 0.0;
  double gradient_tolerance = 0.0;
  double parameter_tolerance = 0.0;
  bool minimizer_progress_to_stdout = false;
  int max_num_iterations = 100;
  int max_linear_solver_iterations = 200;
  int max_num_consecutive_invalid_steps = 10;
  int max_consecutive_nonmonotonic_steps = 10;  // inert: the reference never enables nonmonotonic steps
  int num_threads = -1;                          // CPU notion; ignored by the device path
  LinearSolverType linear_solver_type = POLICY;  // Solve() overwrites it by image count, bundle_adjustment.cc:276-286
  bool update_state_every_iteration = false;
  std::vector<IterationCallback*> callbacks;
};

struct SolverSummary : sba_summary {
  SolverSummary() : sba_summary() {}
  std::vector<IterationSummary> iterations;
  bool IsSolutionUsable() const {
    return termination_type == SBA_CONVERGENCE || termination_type == SBA_NO_CONVERGENCE ||
           termination_type == SBA_USER_SUCCESS;
  }
  std::string BriefReport() const {
    char b[256];
    std::snprintf(b, sizeof b, "B200 BA: iterations %d, initial cost %e, final cost %e, termination %d", num_iterations,
                  initial_cost, final_cost, termination_type);
    return b;
  }
};

// ---- BundleAdjustmentOptions -----------------------------------------------------------------------
struct BundleAdjustmentOptions {
  enum class LossFunctionType { TRIVIAL, SOFT_L1, CAUCHY };
  LossFunctionType loss_function_type = LossFunctionType::TRIVIAL;
  double loss_function_scale = 1.0;
  bool refine_focal_length = true;
  bool refine_principal_point = false;
  bool refine_extra_params = true;
  bool refine_extrinsics = true;
  bool print_summary = true;
  int min_num_residuals_for_multi_threading = 50000;  // kept for source compatibility; no host threads here
  SolverOptions solver_options;

  // CreateLossFunction (bundle_adjustment.cc:53-68) returns a heap ceres::LossFunction in the reference; across
  // the C ABI a loss is the pair (type, scale) and is applied inside the evaluation kernel.
  int LossFunctionCode() const { return static_cast<int>(loss_function_type); }

  bool Check() const { return loss_function_scale >= 0; }  // CHECK_OPTION_GE, bundle_adjustment.cc:70-73
};

// ---- BundleAdjustmentConfig ------------------------------------------------------------------------
// Same six hash containers as the reference (bundle_adjustment.h:161-166): Images() is iterated by Solve(), and
// its iteration order *is* the residual order, so the container type must not change.
class BundleAdjustmentConfig {
 public:
  size_t NumImages() const { return image_ids_.size(); }
  size_t NumPoints() const { return variable_point3D_ids_.size() + constant_point3D_ids_.size(); }
  size_t NumConstantCameras() const { return constant_camera_ids_.size(); }
  size_t NumConstantPoses() const { return constant_poses_.size(); }
  size_t NumConstantTvecs() const { return constant_tvecs_.size(); }
  size_t NumVariablePoints() const { return variable_point3D_ids_.size(); }
  size_t NumConstantPoints() const { return constant_point3D_ids_.size(); }

  template <class Reconstruction>
  size_t NumResiduals(const Reconstruction& rec) const {  // bundle_adjustment.cc:103-139
    size_t n = 0;
    for (image_t id : image_ids_)
      for (const auto& p : rec.Image(id).Points2D()) n += p.HasPoint3D() ? 1 : 0;
    auto outside = [&](point3D_t pid) {
      size_t k = 0;
      for (const auto& el : rec.Point3D(pid).Track().Elements()) k += image_ids_.count(el.image_id) ? 0 : 1;
      return k;
    };
    for (point3D_t pid : variable_point3D_ids_) n += outside(pid);
    for (point3D_t pid : constant_point3D_ids_) n += outside(pid);
    return 2 * n;
  }

  void AddImage(image_t id) { image_ids_.insert(id); }
  bool HasImage(image_t id) const { return image_ids_.count(id) > 0; }
  void RemoveImage(image_t id) { image_ids_.erase(id); }

  void SetConstantCamera(camera_t id) { constant_camera_ids_.insert(id); }
  void SetVariableCamera(camera_t id) { constant_camera_ids_.erase(id); }
  bool IsConstantCamera(camera_t id) const { return constant_camera_ids_.count(id) > 0; }

  void SetConstantPose(image_t id) {
    Require(HasImage(id) && !HasConstantTvec(id), "SetConstantPose: image not added or has constant tvec");
    constant_poses_.insert(id);
  }
  void SetVariablePose(image_t id) { constant_poses_.erase(id); }
  bool HasConstantPose(image_t id) const { return constant_poses_.count(id) > 0; }

  void SetConstantTvec(image_t id, const std::vector<int>& idxs) {
    Require(!idxs.empty() && idxs.size() <= 3, "SetConstantTvec: 1..3 indices expected");
    Require(HasImage(id) && !HasConstantPose(id), "SetConstantTvec: image not added or has constant pose");
    for (size_t a = 0; a < idxs.size(); ++a)
      for (size_t b = a + 1; b < idxs.size(); ++b) Require(idxs[a] != idxs[b], "Tvec indices must not contain duplicates");
    constant_tvecs_[id] = idxs;
  }
  void RemoveConstantTvec(image_t id) { constant_tvecs_.erase(id); }
  bool HasConstantTvec(image_t id) const { return constant_tvecs_.count(id) > 0; }

  void AddVariablePoint(point3D_t id) {
    Require(!HasConstantPoint(id), "point is already constant");
    variable_point3D_ids_.insert(id);
  }
  void AddConstantPoint(point3D_t id) {
    Require(!HasVariablePoint(id), "point is already variable");
    constant_point3D_ids_.insert(id);
  }
  bool HasPoint(point3D_t id) const { return HasVariablePoint(id) || HasConstantPoint(id); }
  bool HasVariablePoint(point3D_t id) const { return variable_point3D_ids_.count(id) > 0; }
  bool HasConstantPoint(point3D_t id) const { return constant_point3D_ids_.count(id) > 0; }
  void RemoveVariablePoint(point3D_t id) { variable_point3D_ids_.erase(id); }
  void RemoveConstantPoint(point3D_t id) { constant_point3D_ids_.erase(id); }

  const std::unordered_set<image_t>& Images() const { return image_ids_; }
  const std::unordered_set<point3D_t>& VariablePoints() const { return variable_point3D_ids_; }
  const std::unordered_set<point3D_t>& ConstantPoints() const { return constant_point3D_ids_; }
  const std::vector<int>& ConstantTvec(image_t id) const { return constant_tvecs_.at(id); }

 private:
  static void Require(bool ok, const char* what) {
    if (!ok) throw std::logic_error(what);  // glog CHECK in the reference (process abort)
  }
  std::unordered_set<camera_t> constant_camera_ids_;
  std::unordered_set<image_t> image_ids_;
  std::unordered_set<point3D_t> variable_point3D_ids_;
  std::unordered_set<point3D_t> constant_point3D_ids_;
  std::unordered_set<image_t> constant_poses_;
  std::unordered_map<image_t, std::vector<int>> constant_tvecs_;
};

// ---- flattened problem: the arrays behind one sba_problem_desc ---------------------------------------
struct FlatProblem {
  std::vector<double> qvec, tvec, cam_params, points, obs_xy;
  std::vector<int32_t> image_camera, camera_model, obs_image, obs_point;
  std::vector<uint8_t> pose_constant, tvec_mask, point_constant;
  std::vector<uint16_t> cam_mask;
  // dense slot -> id of the Reconstruction (scatter-back), in first-seen order
  std::vector<image_t> image_ids;
  std::vector<camera_t> camera_ids;
  std::vector<point3D_t> point_ids;

  sba_problem_desc Desc() const {
    sba_problem_desc d{};
    d.num_images = static_cast<int32_t>(image_ids.size());
    d.num_cameras = static_cast<int32_t>(camera_ids.size());
    d.num_points = static_cast<int32_t>(point_ids.size());
    d.num_observations = static_cast<int64_t>(obs_image.size());
    d.image_qvec = qvec.data();
    d.image_tvec = tvec.data();
    d.image_camera = image_camera.data();
    d.image_pose_constant = pose_constant.data();
    d.image_tvec_constant_mask = tvec_mask.data();
    d.camera_model = camera_model.data();
    d.camera_params = cam_params.data();
    d.camera_param_constant_mask = cam_mask.data();
    d.points = points.data();
    d.point_constant = point_constant.data();
    d.obs_image = obs_image.data();
    d.obs_point = obs_point.data();
    d.obs_xy = obs_xy.data();
    d.shard_rank = 0;
    d.shard_count = 1;
    return d;
  }
};

// ---- BundleAdjuster --------------------------------------------------------------------------------
template <class Reconstruction>
class BundleAdjusterT {
 public:
  BundleAdjusterT(const BundleAdjustmentOptions& options, const BundleAdjustmentConfig& config)
      : options_(options), config_(config) {  // both copied by value, bundle_adjustment.cc:252-256
    if (!options_.Check()) throw std::invalid_argument("BundleAdjustmentOptions::Check failed");
  }

  // bundle_adjustment.cc:258-320.  true unless the problem is empty; convergence does not matter (:319).
  bool Solve(Reconstruction* reconstruction) {
    if (reconstruction == nullptr) throw std::invalid_argument("reconstruction is null");        // CHECK_NOTNULL :259
    if (used_) throw std::logic_error("Cannot use the same BundleAdjuster multiple times");       // CHECK :260
    used_ = true;
    Flatten(reconstruction);
    if (flat_.obs_image.empty()) return false;  // :267-269

    sba_solver_options o;
    sba_default_solver_options(&o);
    const SolverOptions& so = options_.solver_options;
    // the reference overwrites linear_solver_type by image count (:276-286); POLICY asks the library to do the same
    o.linear_solver_type = SBA_SOLVER_AUTO;
    o.num_config_images = static_cast<int32_t>(config_.NumImages());
    o.max_num_iterations = so.max_num_iterations;
    o.max_linear_solver_iterations = so.max_linear_solver_iterations;
    o.max_num_consecutive_invalid_steps = so.max_num_consecutive_invalid_steps;
    o.function_tolerance = so.function_tolerance;
    o.gradient_tolerance = so.gradient_tolerance;
    o.parameter_tolerance = so.parameter_tolerance;
    o.loss_function_type = options_.LossFunctionCode();
    o.loss_function_scale = options_.loss_function_scale;
    o.update_state_every_iteration = so.update_state_every_iteration ? 1 : 0;

    const sba_problem_desc d = flat_.Desc();
    sba_handle* h = nullptr;
    Check(sba_create(&d, &h));
    // the handle is released on every way out, including an exception thrown by a user callback or by ScatterBack
    struct HandleGuard { sba_handle* h; ~HandleGuard() { sba_destroy(h); } } guard{h};
    CallbackState st{this, reconstruction, h};
    const bool have_cb = !so.callbacks.empty();
    const int rc = sba_solve(h, &o, have_cb ? &BundleAdjusterT::Trampoline : nullptr, &st, &summary_);
    if (rc == SBA_OK) {
      int32_t n = 0;
      summary_.iterations.resize(static_cast<size_t>(summary_.num_iterations));
      sba_get_trace(h, summary_.iterations.data(), summary_.num_iterations, &n);
      summary_.iterations.resize(static_cast<size_t>(n));
      ScatterBack(reconstruction, h);
    }
    Check(rc);
    if (options_.print_summary) PrintSolverSummary(summary_);
    return true;
  }

  const SolverSummary& Summary() const { return summary_; }

  // Residual blocks in the reference order -> plain arrays (a7).  Public for the parity tests.
  const FlatProblem& Flatten(Reconstruction* rec) {
    flat_ = FlatProblem();
    image_slot_.clear(); camera_slot_.clear(); point_slot_.clear();
    seen_per_point_.clear(); problem_cameras_.clear();
    for (image_t id : config_.Images()) AddImage(id, rec);                 // :330-332
    for (point3D_t id : config_.VariablePoints()) AddPoint(id, rec);       // :333-335
    for (point3D_t id : config_.ConstantPoints()) AddPoint(id, rec);       // :336-338
    PackCameras(rec);                                                      // ParameterizeCameras :480-516
    PackPoints(rec);                                                       // ParameterizePoints  :518-530
    PackPoses(rec);
    return flat_;
  }

  const BundleAdjustmentConfig& Config() const { return config_; }

  static void PrintSolverSummary(const SolverSummary& s) {  // bundle_adjustment.cc:1142-1196
    static const char* kTerm[] = {"Convergence", "No convergence", "Failure", "User success", "User failure"};
    const double nres = s.num_residuals_reduced > 0 ? s.num_residuals_reduced : 1;
    std::printf("    Residuals : %d\n   Parameters : %d\n   Iterations : %d\n         Time : %g [s]\n"
                " Initial cost : %g [px]\n   Final cost : %g [px]\n  Termination : %s\n\n",
                s.num_residuals_reduced, s.num_effective_parameters_reduced,
                s.num_successful_steps + s.num_unsuccessful_steps, s.total_time_in_seconds,
                std::sqrt(s.initial_cost / nres), std::sqrt(s.final_cost / nres),
                (s.termination_type >= 0 && s.termination_type < 5) ? kTerm[s.termination_type] : "Unknown");
  }

 private:
  struct CallbackState { BundleAdjusterT* self; Reconstruction* rec; sba_handle* h; };

  static int Trampoline(const sba_iteration_summary* it, void* user) {
    auto* st = static_cast<CallbackState*>(user);
    const SolverOptions& so = st->self->options_.solver_options;
    if (so.update_state_every_iteration) st->self->ScatterBack(st->rec, st->h);
    for (IterationCallback* cb : so.callbacks) {
      const CallbackReturnType r = (*cb)(*it);
      if (r != SOLVER_CONTINUE) return static_cast<int>(r);
    }
    return 0;
  }

  static void Check(int rc) {
    if (rc == SBA_OK) return;
    const std::string msg = std::string("libsba_b200: ") + sba_last_error();
    if (rc == SBA_ERR_UNKNOWN_CAMERA_MODEL) throw std::domain_error(msg);  // camera_models.h:140-141
    throw std::runtime_error(msg);
  }

  int32_t ImageSlot(image_t id, bool constant_pose) {
    auto it = image_slot_.find(id);
    if (it != image_slot_.end()) return it->second;
    const int32_t s = static_cast<int32_t>(flat_.image_ids.size());
    image_slot_.emplace(id, s);
    flat_.image_ids.push_back(id);
    flat_.pose_constant.push_back(constant_pose ? 1 : 0);
    flat_.tvec_mask.push_back(0);
    return s;
  }
  int32_t CameraSlot(camera_t id) {
    auto it = camera_slot_.find(id);
    if (it != camera_slot_.end()) return it->second;
    const int32_t s = static_cast<int32_t>(flat_.camera_ids.size());
    camera_slot_.emplace(id, s);
    flat_.camera_ids.push_back(id);
    return s;
  }
  int32_t PointSlot(point3D_t id) {
    auto it = point_slot_.find(id);
    if (it != point_slot_.end()) return it->second;
    const int32_t s = static_cast<int32_t>(flat_.point_ids.size());
    point_slot_.emplace(id, s);
    flat_.point_ids.push_back(id);
    return s;
  }
  void PushObservation(int32_t img, int32_t pt, const double* xy) {
    flat_.obs_image.push_back(img);
    flat_.obs_point.push_back(pt);
    flat_.obs_xy.push_back(xy[0]);
    flat_.obs_xy.push_back(xy[1]);
  }

  void AddImage(image_t image_id, Reconstruction* rec) {  // AddImageToProblem :348-427
    auto& image = rec->Image(image_id);
    image.NormalizeQvec();  // side effect kept (:355)
    const bool constant_pose = !options_.refine_extrinsics || config_.HasConstantPose(image_id);
    size_t n = 0;
    int32_t slot = -1;
    for (const auto& p2 : image.Points2D()) {
      if (!p2.HasPoint3D()) continue;
      if (slot < 0) slot = ImageSlot(image_id, constant_pose);
      ++n;
      seen_per_point_[p2.Point3DId()] += 1;
      PushObservation(slot, PointSlot(p2.Point3DId()), p2.XY().data());
    }
    if (n == 0) return;
    problem_cameras_.insert(image.CameraId());
    CameraSlot(image.CameraId());
    if (!constant_pose && config_.HasConstantTvec(image_id)) {
      uint8_t m = 0;
      for (int k : config_.ConstantTvec(image_id)) m |= static_cast<uint8_t>(1u << k);
      flat_.tvec_mask[slot] = m;
    }
  }

  void AddPoint(point3D_t pid, Reconstruction* rec) {  // AddPointToProblem :429-478
    auto& p3 = rec->Point3D(pid);
    size_t& seen = seen_per_point_[pid];
    if (seen == p3.Track().Length()) return;
    for (const auto& el : p3.Track().Elements()) {
      if (config_.HasImage(el.image_id)) continue;
      ++seen;
      auto& image = rec->Image(el.image_id);
      if (problem_cameras_.count(image.CameraId()) == 0) {  // :456-459
        problem_cameras_.insert(image.CameraId());
        config_.SetConstantCamera(image.CameraId());
      }
      CameraSlot(image.CameraId());
      PushObservation(ImageSlot(el.image_id, true), PointSlot(pid), image.Point2D(el.point2D_idx).XY().data());
    }
  }

  void PackCameras(Reconstruction* rec) {
    const bool all_constant =
        !options_.refine_focal_length && !options_.refine_principal_point && !options_.refine_extra_params;
    for (camera_t id : flat_.camera_ids) {
      auto& cam = rec->Camera(id);
      const size_t P = cam.NumParams();
      uint16_t mask = 0;
      auto hold = [&mask](const std::vector<size_t>& idxs) {
        for (size_t j : idxs) mask |= static_cast<uint16_t>(1u << j);
      };
      if (all_constant || config_.IsConstantCamera(id)) {
        mask = static_cast<uint16_t>((1u << P) - 1u);
      } else {
        if (!options_.refine_focal_length) hold(cam.FocalLengthIdxs());
        if (!options_.refine_principal_point) hold(cam.PrincipalPointIdxs());
        if (!options_.refine_extra_params) hold(cam.ExtraParamsIdxs());
      }
      flat_.cam_mask.push_back(mask);
      flat_.camera_model.push_back(static_cast<int32_t>(cam.ModelId()));
      const double* prm = cam.ParamsData();
      for (size_t j = 0; j < SBA_MAX_CAMERA_PARAMS; ++j) flat_.cam_params.push_back(j < P ? prm[j] : 0.0);
    }
  }

  void PackPoints(Reconstruction* rec) {
    flat_.point_constant.assign(flat_.point_ids.size(), 0);
    flat_.points.resize(3 * flat_.point_ids.size());
    for (size_t s = 0; s < flat_.point_ids.size(); ++s) {
      const point3D_t pid = flat_.point_ids[s];
      auto& p3 = rec->Point3D(pid);
      const double* X = p3.XYZ().data();
      for (int k = 0; k < 3; ++k) flat_.points[3 * s + k] = X[k];
      if (p3.Track().Length() > seen_per_point_[pid] || config_.HasConstantPoint(pid)) flat_.point_constant[s] = 1;
    }
  }

  void PackPoses(Reconstruction* rec) {
    const size_t n = flat_.image_ids.size();
    flat_.qvec.resize(4 * n); flat_.tvec.resize(3 * n); flat_.image_camera.resize(n);
    for (size_t s = 0; s < n; ++s) {
      auto& image = rec->Image(flat_.image_ids[s]);
      const double* q = image.Qvec().data();
      const double* t = image.Tvec().data();
      for (int k = 0; k < 4; ++k) flat_.qvec[4 * s + k] = q[k];
      for (int k = 0; k < 3; ++k) flat_.tvec[3 * s + k] = t[k];
      flat_.image_camera[s] = camera_slot_.at(image.CameraId());
    }
  }

  // The reference lets Ceres write through the raw pointers (:357-359); here the optimised blocks come back
  // in one D2H copy and are stored through the same accessors.
  void ScatterBack(Reconstruction* rec, sba_handle* h) {
    std::vector<double> q(flat_.qvec.size()), t(flat_.tvec.size()), c(flat_.cam_params.size()), X(flat_.points.size());
    Check(sba_get_parameters(h, q.data(), t.data(), c.data(), X.data()));
    for (size_t s = 0; s < flat_.image_ids.size(); ++s) {
      auto& image = rec->Image(flat_.image_ids[s]);
      double* qd = image.Qvec().data();
      double* td = image.Tvec().data();
      for (int k = 0; k < 4; ++k) qd[k] = q[4 * s + k];
      for (int k = 0; k < 3; ++k) td[k] = t[3 * s + k];
    }
    for (size_t s = 0; s < flat_.camera_ids.size(); ++s) {
      auto& cam = rec->Camera(flat_.camera_ids[s]);
      double* prm = cam.ParamsData();
      for (size_t j = 0; j < cam.NumParams(); ++j) prm[j] = c[SBA_MAX_CAMERA_PARAMS * s + j];
    }
    for (size_t s = 0; s < flat_.point_ids.size(); ++s) {
      double* Xd = rec->Point3D(flat_.point_ids[s]).XYZ().data();
      for (int k = 0; k < 3; ++k) Xd[k] = X[3 * s + k];
    }
  }

  const BundleAdjustmentOptions options_;
  BundleAdjustmentConfig config_;
  SolverSummary summary_;
  bool used_ = false;
  FlatProblem flat_;
  std::unordered_map<image_t, int32_t> image_slot_;
  std::unordered_map<camera_t, int32_t> camera_slot_;
  std::unordered_map<point3D_t, int32_t> point_slot_;
  std::unordered_map<point3D_t, size_t> seen_per_point_;   // point3D_num_observations_, bundle_adjustment.h:202
  std::unordered_set<camera_t> problem_cameras_;           // camera_ids_, bundle_adjustment.h:201
};

}  // namespace sba_shim
#endif  // SBA_SHIM_BUNDLE_ADJUSTER_H_
