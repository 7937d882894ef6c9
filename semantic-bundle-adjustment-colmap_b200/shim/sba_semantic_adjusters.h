// sba_semantic_adjusters.h -- header-only C++17 host side of the fork's two semantic adjusters.
//
// Mirrors the front-ends
//   SemanticBundleAdjustmentOptions / ...Config / SemanticBundleAdjuster
//       /root/reference/src/optim/semantic_bundle_adjustment.h:53-272
//   GeometricSemanticBundleAdjustmentOptions / ...Config / GeometricSemanticBundleAdjuster<CylinderClass>
//       /root/reference/src/optim/geometric_semantic_bundle_adjustment.h:51-299
// (both Config classes are verbatim clones of BundleAdjustmentConfig and are aliases here) and hands the arithmetic to
// sba_semantic_solve / sba_cylinder_solve (include/sba_b200.h).  What stays on the host: the precondition checks
// (Assert, semantic_bundle_adjustment.cc:604-644 / geometric_semantic_bundle_adjustment.cc:667-711), walking
// config.Images() in its iteration order (= residual order), and copying poses / cylinders in and out.
// The reference reads the per-image float32 TIFF maps from <data_path>/{depth_tiff,semantic_tiff}/<stem>_{depth,semantic}.tiff
// (semantic_bundle_adjustment.cc:1021-1068); file IO is not part of the hot path, so the maps are handed over with
// SetMaps(image name, ...) -- model_io.py reads and writes the same TIFFs for the Python mirror.
#ifndef SBA_SHIM_SEMANTIC_ADJUSTERS_H_
#define SBA_SHIM_SEMANTIC_ADJUSTERS_H_

#include <array>
#include <string>
#include <unordered_map>
#include <vector>

#include "sba_bundle_adjuster.h"

namespace sba_shim {

using SemanticBundleAdjustmentConfig = BundleAdjustmentConfig;
using GeometricSemanticBundleAdjustmentConfig = BundleAdjustmentConfig;

inline SolverOptions SemanticSolverDefaults() {  // semantic_bundle_adjustment.h:115-131 (the GSBA options are the same)
  SolverOptions o;
  o.function_tolerance = o.gradient_tolerance = o.parameter_tolerance = 1e-8;
  o.update_state_every_iteration = true;
  return o;
}

struct SemanticBundleAdjustmentOptions {  // semantic_bundle_adjustment.h:53-136
  enum class LossFunctionType { TRIVIAL, SOFT_L1, CAUCHY };
  LossFunctionType loss_function_type = LossFunctionType::TRIVIAL;
  std::string data_path, output_path, visualization_path = "{output_path}/run";
  bool export_csv = false;
  double depth_error_threshold = 2;
  int error_computation_pixel_step = 10;
  double numeric_relative_step_size = 1e-3;
  double loss_function_scale = 1.0;
  bool refine_focal_length = true, refine_principal_point = false, refine_extra_params = true, refine_extrinsics = true;
  bool print_summary = true;
  int min_num_residuals_for_multi_threading = 50000;
  SolverOptions solver_options = SemanticSolverDefaults();
  bool Check() const { return loss_function_scale >= 0; }
};

struct GeometricSemanticBundleAdjustmentOptions {  // geometric_semantic_bundle_adjustment.h:51-160
  enum class LossFunctionType { TRIVIAL, SOFT_L1, CAUCHY };
  LossFunctionType loss_function_type = LossFunctionType::TRIVIAL;
  std::string data_path, output_path, input_geometry_path, visualization_path = "{output_path}/run";
  double trunk_semantic_class = 250.;
  bool refine_geometry = true;
  bool use_cylinder_by_2_points = false;  // the CylinderClass template argument of the reference (:223, cylinder_by_2_points.h)
  bool include_landmark_error = false;    // not available on the device path (include/sba_b200.h): Solve() throws when set
  double landmark_error_weight = 1;
  double numeric_relative_step_size = 1e-3;
  double loss_function_scale = 1.0;
  bool refine_focal_length = true, refine_principal_point = false, refine_extra_params = true, refine_extrinsics = true;
  bool print_summary = true;
  int min_num_residuals_for_multi_threading = 50000;
  SolverOptions solver_options = SemanticSolverDefaults();
  bool Check() const { return loss_function_scale >= 0; }
};

struct FloatMap { int height = 0, width = 0; std::vector<float> data; };  // ROW-major (y * width + x)

// Cylinder (src/util/cylinder.h:123-347: qvec, tvec = bottom centre, radius, height) or CylinderBy2Points
// (src/util/cylinder_by_2_points.h: tvec_1, tvec_2, radius), as the nine doubles of sba_cylinder_problem::cylinder_params
struct CylinderParams { std::array<double, 9> v{}; };

namespace detail {
inline void FillSolverOptions(const SolverOptions& so, int loss_type, double loss_scale, size_t num_images, sba_solver_options* o) {
  sba_default_solver_options(o);
  o->linear_solver_type = SBA_SOLVER_AUTO;
  o->num_config_images = static_cast<int32_t>(num_images);
  o->max_num_iterations = so.max_num_iterations;
  o->max_linear_solver_iterations = so.max_linear_solver_iterations;
  o->max_num_consecutive_invalid_steps = so.max_num_consecutive_invalid_steps;
  o->function_tolerance = so.function_tolerance;
  o->gradient_tolerance = so.gradient_tolerance;
  o->parameter_tolerance = so.parameter_tolerance;
  o->loss_function_type = loss_type;
  o->loss_function_scale = loss_scale;
  o->update_state_every_iteration = so.update_state_every_iteration ? 1 : 0;
}
inline void CheckRc(int rc) {
  if (rc != SBA_OK) throw std::runtime_error(std::string("libsba_b200: ") + sba_last_error());
}
// the checks both adjusters make before building the problem (semantic_bundle_adjustment.cc:604-644)
template <class Reconstruction>
void AssertSimplePinholeConstant(const BundleAdjustmentConfig& config, const Reconstruction& rec, bool refine_extrinsics) {
  for (image_t id : config.Images())
    if (!config.IsConstantCamera(rec.Image(id).CameraId()))
      throw std::runtime_error("ERROR: camera intrinsics of image '" + rec.Image(id).Name() + "' are not set to constant. This is not supported.");
  for (image_t id : config.Images())
    if (rec.Camera(rec.Image(id).CameraId()).ModelId() != 0 /* SimplePinholeCameraModel::kModelId */)
      throw std::runtime_error("ERROR: the only supported camera model is SimplePinholeCameraModel.");
  if (!refine_extrinsics) throw std::runtime_error("ERROR: the argument 'refine_extrinsics' must be set to true.");
}
// poses / constness of config.Images() in iteration order
template <class Reconstruction>
struct PoseArrays {
  std::vector<image_t> ids;
  std::vector<double> q, t, cam;
  std::vector<uint8_t> pose_const, tvec_mask;
  void Gather(const BundleAdjustmentConfig& config, Reconstruction* rec, bool refine_extrinsics) {
    for (image_t id : config.Images()) {
      auto& im = rec->Image(id);
      ids.push_back(id);
      for (int k = 0; k < 4; ++k) q.push_back(im.Qvec().data()[k]);
      for (int k = 0; k < 3; ++k) t.push_back(im.Tvec().data()[k]);
      const double* prm = rec->Camera(im.CameraId()).ParamsData();
      for (int k = 0; k < 3; ++k) cam.push_back(prm[k]);
      pose_const.push_back((!refine_extrinsics || config.HasConstantPose(id)) ? 1 : 0);
      uint8_t m = 0;
      if (config.HasConstantTvec(id)) for (int idx : config.ConstantTvec(id)) m |= static_cast<uint8_t>(1u << idx);
      tvec_mask.push_back(m);
    }
  }
  void Scatter(Reconstruction* rec) const {
    for (size_t s = 0; s < ids.size(); ++s) {
      auto& im = rec->Image(ids[s]);
      for (int k = 0; k < 4; ++k) im.Qvec().data()[k] = q[4 * s + k];
      for (int k = 0; k < 3; ++k) im.Tvec().data()[k] = t[3 * s + k];
    }
  }
};
}  // namespace detail

// ---- SemanticBundleAdjuster (semantic_bundle_adjustment.h:217-272) ------------------------------------------------
template <class Reconstruction>
class SemanticBundleAdjusterT {
 public:
  SemanticBundleAdjusterT(const SemanticBundleAdjustmentOptions& options, const SemanticBundleAdjustmentConfig& config)
      : options_(options), config_(config) {
    if (!options_.Check()) throw std::invalid_argument("SemanticBundleAdjustmentOptions::Check failed");
  }
  // replaces ReadDepthAndSemanticMaps (:1021-1068): maps are keyed by image name, as depth_maps_ / semantic_maps_ are
  void SetMaps(const std::string& image_name, const FloatMap& depth, const FloatMap& semantic) {
    depth_maps_[image_name] = depth; semantic_maps_[image_name] = semantic;
  }
  // semantic_bundle_adjustment.cc:430-544: false when the problem has no residuals, true otherwise
  bool Solve(Reconstruction* reconstruction) {
    if (reconstruction == nullptr) throw std::invalid_argument("reconstruction is null");
    if (used_) throw std::logic_error("Cannot use the same BundleAdjuster multiple times");
    used_ = true;
    detail::AssertSimplePinholeConstant(config_, *reconstruction, options_.refine_extrinsics);
    detail::PoseArrays<Reconstruction> P;
    P.Gather(config_, reconstruction, options_.refine_extrinsics);
    const size_t n = P.ids.size();
    int H = 0, W = 0;
    std::vector<float> depth, label;
    for (size_t s = 0; s < n; ++s) {
      const std::string& name = reconstruction->Image(P.ids[s]).Name();
      auto d = depth_maps_.find(name), l = semantic_maps_.find(name);
      if (d == depth_maps_.end() || l == semantic_maps_.end()) throw std::runtime_error("ERROR: no depth / semantic map for image '" + name + "'");
      if (s == 0) { H = d->second.height; W = d->second.width; }
      if (d->second.height != H || d->second.width != W || l->second.height != H || l->second.width != W)
        throw std::runtime_error("ERROR: depth / semantic maps of different sizes");
      depth.insert(depth.end(), d->second.data.begin(), d->second.data.end());
      label.insert(label.end(), l->second.data.begin(), l->second.data.end());
    }
    if (n == 0) return false;
    sba_semantic_problem p{};
    p.num_images = static_cast<int32_t>(n); p.height = H; p.width = W;
    p.image_qvec = P.q.data(); p.image_tvec = P.t.data(); p.camera_params = P.cam.data();
    p.depth_maps = depth.data(); p.semantic_maps = label.data();
    p.image_pose_constant = P.pose_const.data(); p.image_tvec_constant_mask = P.tvec_mask.data();
    p.pixel_step = options_.error_computation_pixel_step;
    p.depth_error_threshold = options_.depth_error_threshold;
    p.relative_step_size = options_.numeric_relative_step_size;
    sba_solver_options o;
    detail::FillSolverOptions(options_.solver_options, static_cast<int>(options_.loss_function_type), options_.loss_function_scale,
                              config_.NumImages(), &o);
    State st{this, reconstruction, &P};
    summary_.iterations.assign(static_cast<size_t>(o.max_num_iterations) + 2, IterationSummary());
    int32_t n_rows = 0;
    const bool have_cb = !options_.solver_options.callbacks.empty();
    const int rc = sba_semantic_solve(&p, &o, have_cb ? &SemanticBundleAdjusterT::Trampoline : nullptr, &st, &summary_,
                                      summary_.iterations.data(), static_cast<int32_t>(summary_.iterations.size()), &n_rows);
    summary_.iterations.resize(static_cast<size_t>(n_rows));
    if (rc == SBA_ERR_EMPTY_PROBLEM) return false;  // `if (problem_->NumResiduals() == 0) return false` (:484-486)
    detail::CheckRc(rc);
    P.Scatter(reconstruction);
    return true;
  }
  const SolverSummary& Summary() const { return summary_; }

 private:
  struct State { SemanticBundleAdjusterT* self; Reconstruction* rec; detail::PoseArrays<Reconstruction>* P; };
  static int Trampoline(const sba_iteration_summary* it, void* user) {
    auto* st = static_cast<State*>(user);
    const SolverOptions& so = st->self->options_.solver_options;
    if (so.update_state_every_iteration) st->P->Scatter(st->rec);  // the library has refreshed the arrays (update_state_every_iteration)
    for (IterationCallback* cb : so.callbacks) {
      const CallbackReturnType r = (*cb)(*it);
      if (r != SOLVER_CONTINUE) return static_cast<int>(r);
    }
    return 0;
  }
  const SemanticBundleAdjustmentOptions options_;
  SemanticBundleAdjustmentConfig config_;
  SolverSummary summary_;
  bool used_ = false;
  std::unordered_map<std::string, FloatMap> depth_maps_, semantic_maps_;
};

// ---- GeometricSemanticBundleAdjuster<CylinderClass> (geometric_semantic_bundle_adjustment.h:223-299) ---------------
template <class Reconstruction>
class GeometricSemanticBundleAdjusterT {
 public:
  GeometricSemanticBundleAdjusterT(const GeometricSemanticBundleAdjustmentOptions& options,
                                   const GeometricSemanticBundleAdjustmentConfig& config)
      : options_(options), config_(config) {
    if (!options_.Check()) throw std::invalid_argument("GeometricSemanticBundleAdjustmentOptions::Check failed");
  }
  // semantic map of an image (ReadDepthAndSemanticMaps); the bool mask label == trunk_semantic_class is formed in Solve
  // (geometric_semantic_bundle_adjustment.cc:1328-1333)
  void SetSemanticMap(const std::string& image_name, const FloatMap& semantic) { semantic_maps_[image_name] = semantic; }
  // the cylinders read from input_geometry_path (cylinders.txt, src/util/cylinder.h:289-329)
  std::vector<CylinderParams>& Cylinders() { return cylinders_; }
  const std::vector<CylinderParams>& Cylinders() const { return cylinders_; }

  bool Solve(Reconstruction* reconstruction) {
    if (reconstruction == nullptr) throw std::invalid_argument("reconstruction is null");
    if (used_) throw std::logic_error("Cannot use the same BundleAdjuster multiple times");
    used_ = true;
    detail::AssertSimplePinholeConstant(config_, *reconstruction, true);
    if (options_.loss_function_type != GeometricSemanticBundleAdjustmentOptions::LossFunctionType::TRIVIAL)
      throw std::runtime_error("ERROR: the only supported loss function is 'LossFunctionType::TRIVIAL'.");  // :699-706
    if (options_.include_landmark_error)
      throw std::runtime_error("include_landmark_error is not available on the device path (include/sba_b200.h)");
    detail::PoseArrays<Reconstruction> P;
    P.Gather(config_, reconstruction, options_.refine_extrinsics);
    const size_t n = P.ids.size();
    if (n == 0) return false;
    int H = 0, W = 0;
    std::vector<uint8_t> masks;
    for (size_t s = 0; s < n; ++s) {
      const std::string& name = reconstruction->Image(P.ids[s]).Name();
      auto l = semantic_maps_.find(name);
      if (l == semantic_maps_.end()) throw std::runtime_error("ERROR: no semantic map for image '" + name + "'");
      if (s == 0) { H = l->second.height; W = l->second.width; }
      if (l->second.height != H || l->second.width != W) throw std::runtime_error("ERROR: semantic maps of different sizes");
      for (float v : l->second.data) masks.push_back(static_cast<double>(v) == options_.trunk_semantic_class ? 1 : 0);
    }
    std::vector<double> cyl;
    for (const CylinderParams& c : cylinders_) cyl.insert(cyl.end(), c.v.begin(), c.v.end());
    sba_cylinder_problem p{};
    p.num_images = static_cast<int32_t>(n); p.height = H; p.width = W;
    p.image_qvec = P.q.data(); p.image_tvec = P.t.data(); p.camera_params = P.cam.data();
    p.masks = masks.data();
    p.image_pose_constant = P.pose_const.data(); p.image_tvec_constant_mask = P.tvec_mask.data();
    p.num_cylinders = static_cast<int32_t>(cylinders_.size());
    p.cylinder_kind = options_.use_cylinder_by_2_points ? 1 : 0;
    p.cylinder_params = cyl.data();
    p.refine_geometry = options_.refine_geometry ? 1 : 0;
    p.relative_step_size = options_.numeric_relative_step_size;
    sba_solver_options o;
    detail::FillSolverOptions(options_.solver_options, SBA_LOSS_TRIVIAL, options_.loss_function_scale, config_.NumImages(), &o);
    summary_.iterations.assign(static_cast<size_t>(o.max_num_iterations) + 2, IterationSummary());
    int32_t n_rows = 0;
    State st{this, reconstruction, &P, &cyl};
    const bool have_cb = !options_.solver_options.callbacks.empty();
    const int rc = sba_cylinder_solve(&p, &o, have_cb ? &GeometricSemanticBundleAdjusterT::Trampoline : nullptr, &st, &summary_,
                                      summary_.iterations.data(), static_cast<int32_t>(summary_.iterations.size()), &n_rows);
    summary_.iterations.resize(static_cast<size_t>(n_rows));
    if (rc == SBA_ERR_EMPTY_PROBLEM) return false;
    detail::CheckRc(rc);
    P.Scatter(reconstruction);
    ScatterCylinders(cyl);
    return true;
  }
  const SolverSummary& Summary() const { return summary_; }

 private:
  struct State { GeometricSemanticBundleAdjusterT* self; Reconstruction* rec; detail::PoseArrays<Reconstruction>* P; std::vector<double>* cyl; };
  void ScatterCylinders(const std::vector<double>& cyl) {
    for (size_t c = 0; c < cylinders_.size(); ++c) for (int k = 0; k < 9; ++k) cylinders_[c].v[k] = cyl[9 * c + k];
  }
  static int Trampoline(const sba_iteration_summary* it, void* user) {
    auto* st = static_cast<State*>(user);
    const SolverOptions& so = st->self->options_.solver_options;
    if (so.update_state_every_iteration) { st->P->Scatter(st->rec); st->self->ScatterCylinders(*st->cyl); }
    for (IterationCallback* cb : so.callbacks) {
      const CallbackReturnType r = (*cb)(*it);
      if (r != SOLVER_CONTINUE) return static_cast<int>(r);
    }
    return 0;
  }
  const GeometricSemanticBundleAdjustmentOptions options_;
  GeometricSemanticBundleAdjustmentConfig config_;
  SolverSummary summary_;
  bool used_ = false;
  std::unordered_map<std::string, FloatMap> semantic_maps_;
  std::vector<CylinderParams> cylinders_;
};

// ParallelBundleAdjuster (src/optim/bundle_adjustment.h:208-268) accepts a strict subset of BundleAdjuster's problems
// (SIMPLE_RADIAL, unshared intrinsics, no constant sub-blocks: bundle_adjustment.cc:541-558, 643-663) and exists in the
// reference only to reach lib/PBA's GPU path; on this side the same solver serves it.
template <class Reconstruction>
using ParallelBundleAdjusterT = BundleAdjusterT<Reconstruction>;

}  // namespace sba_shim
#endif  // SBA_SHIM_SEMANTIC_ADJUSTERS_H_
