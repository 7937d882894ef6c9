// sba_bundle_adjuster.h -- header-only C++17 host side of the B200 bundle-adjustment path.
//
// Mirrors the front-end of the reference so that a caller keeps writing
//     BundleAdjuster ba(options, config);  ba.Solve(&reconstruction);  ba.Summary();
//   BundleAdjustmentOptions   /root/reference/src/optim/bundle_adjustment.h:48-100
//   BundleAdjustmentConfig    /root/reference/src/optim/bundle_adjustment.h:103-167
//   BundleAdjuster            /root/reference/src/optim/bundle_adjustment.h:171-203
// but hands the arithmetic to libsba_b200.so through the C ABI of include/sba_b200.h instead of
// building a ceres::Problem.  Everything below is host integer work: walking the containers in the
// reference's residual order (SetUp, bundle_adjustment.cc:326-342) and packing plain arrays.  At the headline size
// (1778 images / 1 M points / 5 M observations) Flatten takes 0.17-0.35 s on one host core (tools/shim_flatten_bench.cpp;
// 1.1 s with two hash lookups per observation), against 33 ms for everything from sba_create to the read-back.
//
// The adjuster is a template over the Reconstruction type.  It only uses the accessor contract
//   rec->Image(id)  : CameraId(), NormalizeQvec(), Qvec().data(), Tvec().data(), Points2D(), Point2D(i)
//   Point2D         : HasPoint3D(), Point3DId(), XY().data()
//   rec->Camera(id) : ModelId(), NumParams(), ParamsData(), FocalLengthIdxs(), PrincipalPointIdxs(), ExtraParamsIdxs()
//   rec->Point3D(id): XYZ().data(), Track().Length(), Track().Elements() -> {image_id, point2D_idx}
//   rec->Points3D() : iterable of (point3D_t, Point3D) with size()   (reconstruction.h:97; walked instead of one hash
//                     lookup per point when most of the reconstruction's points take part)
// (src/base/{reconstruction,image,camera,point2d,point3d,track}.h), so the same header compiles against
// COLMAP's own classes and against shim/mock_reconstruction.h (tests/shim_test.cpp).
//
// No CPU solve exists here: when libsba_b200.so reports no device, Solve() throws.
#ifndef SBA_SHIM_BUNDLE_ADJUSTER_H_
#define SBA_SHIM_BUNDLE_ADJUSTER_H_

#include <cmath>
#include <chrono>
#include <cstdint>
#include <mutex>
#include <type_traits>
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
// point3D id -> dense slot.  COLMAP hands point ids out from a counter (Reconstruction::AddPoint3D,
// src/base/reconstruction.cc:299-304), so the ids a problem meets are dense up to the gaps deleted points leave: a direct
// table when the largest id is within a small factor of the number of observations, a hash map otherwise.  (Two hash
// lookups per observation were 0.8 of the 1.1 s Flatten took at C4's size.)
class PointSlots {
 public:
  struct Entry { int32_t slot; uint32_t seen; };  // seen: point3D_num_observations_ of the reference (bundle_adjustment.h:202)
  static bool& ForceHash() { static bool f = false; return f; }  // tests: take the fallback for sparse ids
  void Reset(point3D_t max_id, size_t expected) {
    dense_ = max_id < 16 * static_cast<point3D_t>(expected) + (1u << 20);
    if (ForceHash()) dense_ = false;
    table_.clear(); map_.clear();
    if (dense_) table_.assign(static_cast<size_t>(max_id) + 1, Entry{-1, 0});
    else map_.reserve(expected);
  }
  Entry* Find(point3D_t id) {
    if (dense_) return (id < table_.size() && table_[static_cast<size_t>(id)].slot >= 0) ? &table_[static_cast<size_t>(id)] : nullptr;
    auto it = map_.find(id);
    return it == map_.end() ? nullptr : &it->second;
  }
  Entry* Insert(point3D_t id, int32_t slot) {
    if (dense_) {
      if (id >= table_.size()) table_.resize(static_cast<size_t>(id) + 1, Entry{-1, 0});  // (an id the size pass did not see)
      table_[static_cast<size_t>(id)] = Entry{slot, 0};
      return &table_[static_cast<size_t>(id)];
    }
    return &map_.emplace(id, Entry{slot, 0}).first->second;
  }
  void Prefetch(point3D_t id) const {
#if defined(__GNUC__) || defined(__clang__)
    if (dense_ && id < table_.size()) __builtin_prefetch(&table_[static_cast<size_t>(id)]);
#else
    (void)id;
#endif
  }

 private:
  bool dense_ = false;
  std::vector<Entry> table_;
  std::unordered_map<point3D_t, Entry> map_;
};

struct FlatProblem {
  std::vector<double> qvec, tvec, cam_params, points, obs_xy;
  std::vector<int32_t> image_camera, camera_model, obs_image, obs_point;
  std::vector<uint8_t> pose_constant, tvec_mask, point_constant;
  std::vector<uint16_t> cam_mask;
  // dense slot -> id of the Reconstruction (scatter-back), in first-seen order
  std::vector<image_t> image_ids;
  std::vector<camera_t> camera_ids;
  std::vector<point3D_t> point_ids;

  void Clear() {  // keeps the capacity (and with it the mapped pages) of every array
    qvec.clear(); tvec.clear(); cam_params.clear(); points.clear(); obs_xy.clear();
    image_camera.clear(); camera_model.clear(); obs_image.clear(); obs_point.clear();
    pose_constant.clear(); tvec_mask.clear(); point_constant.clear(); cam_mask.clear();
    image_ids.clear(); camera_ids.clear(); point_ids.clear();
  }

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

// The arrays of the last adjusters of this process, kept for the next ones: a mapper adjusts again and again with problems
// of growing size, and 150 MB of freshly mapped vectors cost ~45 ms of page faults per Flatten at C4's size.
class FlatPool {
 public:
  static FlatProblem Take() {
    std::lock_guard<std::mutex> lk(Mu());
    auto& f = Free();
    if (f.empty()) return FlatProblem();
    FlatProblem p = std::move(f.back());
    f.pop_back();
    p.Clear();
    return p;
  }
  static void Give(FlatProblem&& p) {
    std::lock_guard<std::mutex> lk(Mu());
    if (Free().size() < 2) Free().push_back(std::move(p));
  }

 private:
  static std::mutex& Mu() { static std::mutex m; return m; }
  static std::vector<FlatProblem>& Free() { static std::vector<FlatProblem> f; return f; }
};

// ---- BundleAdjuster --------------------------------------------------------------------------------
template <class Reconstruction>
class BundleAdjusterT {
 public:
  BundleAdjusterT(const BundleAdjustmentOptions& options, const BundleAdjustmentConfig& config)
      : options_(options), config_(config) {  // both copied by value, bundle_adjustment.cc:252-256
    if (!options_.Check()) throw std::invalid_argument("BundleAdjustmentOptions::Check failed");
  }
  ~BundleAdjusterT() { FlatPool::Give(std::move(flat_)); }
  BundleAdjusterT(const BundleAdjusterT&) = delete;
  BundleAdjusterT& operator=(const BundleAdjusterT&) = delete;

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
    FlatPool::Give(std::move(flat_));
    flat_ = FlatPool::Take();
    image_slot_.clear(); camera_slot_.clear();
    problem_cameras_.clear();
    {  // size pass: observation count and the largest point id (sequential reads of the images' Points2D)
      size_t n_obs = 0;
      point3D_t max_id = 0;
      for (image_t id : config_.Images())
        for (const auto& p2 : rec->Image(id).Points2D())
          if (p2.HasPoint3D()) { ++n_obs; if (p2.Point3DId() > max_id) max_id = p2.Point3DId(); }
      for (point3D_t id : config_.VariablePoints()) if (id > max_id) max_id = id;
      for (point3D_t id : config_.ConstantPoints()) if (id > max_id) max_id = id;
      point_slot_.Reset(max_id, n_obs + config_.NumPoints());
      flat_.obs_image.reserve(n_obs); flat_.obs_point.reserve(n_obs); flat_.obs_xy.reserve(2 * n_obs);
    }
    const auto t0 = std::chrono::steady_clock::now();
    for (image_t id : config_.Images()) AddImage(id, rec);                 // :330-332
    const auto t1 = std::chrono::steady_clock::now();
    for (point3D_t id : config_.VariablePoints()) AddPoint(id, rec);       // :333-335
    for (point3D_t id : config_.ConstantPoints()) AddPoint(id, rec);       // :336-338
    const auto t2 = std::chrono::steady_clock::now();
    PackCameras(rec);                                                      // ParameterizeCameras :480-516
    PackPoints(rec);                                                       // ParameterizePoints  :518-530
    PackPoses(rec);
    const auto t3 = std::chrono::steady_clock::now();
    flatten_ms_[0] = std::chrono::duration<double, std::milli>(t1 - t0).count();
    flatten_ms_[1] = std::chrono::duration<double, std::milli>(t2 - t1).count();
    flatten_ms_[2] = std::chrono::duration<double, std::milli>(t3 - t2).count();
    return flat_;
  }

  const BundleAdjustmentConfig& Config() const { return config_; }
  // host time of the last Flatten: images' observations, extra track elements of the config's points, parameter packing (ms)
  const double* FlattenPhaseMs() const { return flatten_ms_; }

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
  PointSlots::Entry* PointSlot(point3D_t id) {
    if (PointSlots::Entry* e = point_slot_.Find(id)) return e;
    const int32_t s = static_cast<int32_t>(flat_.point_ids.size());
    flat_.point_ids.push_back(id);
    return point_slot_.Insert(id, s);
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
    const auto& points2D = image.Points2D();
    const size_t n2 = points2D.size();
    constexpr size_t kAhead = 12;  // the id -> slot table is read at random: ask for the entries a few features ahead
    for (size_t i = 0; i < n2; ++i) {
      if (i + kAhead < n2 && points2D[i + kAhead].HasPoint3D()) point_slot_.Prefetch(points2D[i + kAhead].Point3DId());
      const auto& p2 = points2D[i];
      if (!p2.HasPoint3D()) continue;
      if (slot < 0) slot = ImageSlot(image_id, constant_pose);
      ++n;
      PointSlots::Entry* e = PointSlot(p2.Point3DId());
      e->seen += 1;  // point3D_num_observations_ (:371)
      PushObservation(slot, e->slot, p2.XY().data());
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
    const PointSlots::Entry* known = point_slot_.Find(pid);
    if ((known ? known->seen : 0u) == p3.Track().Length()) return;
    for (const auto& el : p3.Track().Elements()) {
      if (config_.HasImage(el.image_id)) continue;
      PointSlot(pid)->seen += 1;
      auto& image = rec->Image(el.image_id);
      if (problem_cameras_.count(image.CameraId()) == 0) {  // :456-459
        problem_cameras_.insert(image.CameraId());
        config_.SetConstantCamera(image.CameraId());
      }
      CameraSlot(image.CameraId());
      PushObservation(ImageSlot(el.image_id, true), PointSlot(pid)->slot, image.Point2D(el.point2D_idx).XY().data());
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
    const size_t n = flat_.point_ids.size();
    flat_.point_constant.assign(n, 0);
    flat_.points.resize(3 * n);
    auto pack = [this](size_t s, point3D_t pid, const auto& p3, uint32_t seen) {
      const double* X = p3.XYZ().data();
      for (int k = 0; k < 3; ++k) flat_.points[3 * s + k] = X[k];
      if (p3.Track().Length() > seen || config_.HasConstantPoint(pid)) flat_.point_constant[s] = 1;
    };
    if (2 * n >= rec->Points3D().size()) {
      // most of the reconstruction's points take part (global bundle adjustment): one walk over the container instead of
      // a hash lookup per point (reconstruction.h:97)
      for (const auto& el : rec->Points3D())
        if (const PointSlots::Entry* e = point_slot_.Find(el.first)) pack(static_cast<size_t>(e->slot), el.first, el.second, e->seen);
    } else {
      for (size_t s = 0; s < n; ++s) {
        const point3D_t pid = flat_.point_ids[s];
        pack(s, pid, rec->Point3D(pid), point_slot_.Find(pid)->seen);
      }
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
    if (2 * flat_.point_ids.size() >= rec->Points3D().size()) {
      // (the container is walked through its const accessor, reconstruction.h:97; the Reconstruction itself is ours to write)
      for (const auto& el : rec->Points3D())
        if (const PointSlots::Entry* e = point_slot_.Find(el.first)) {
          using P3 = std::remove_const_t<std::remove_reference_t<decltype(el.second)>>;
          double* Xd = const_cast<P3&>(el.second).XYZ().data();
          for (int k = 0; k < 3; ++k) Xd[k] = X[3 * static_cast<size_t>(e->slot) + k];
        }
    } else {
      for (size_t s = 0; s < flat_.point_ids.size(); ++s) {
        double* Xd = rec->Point3D(flat_.point_ids[s]).XYZ().data();
        for (int k = 0; k < 3; ++k) Xd[k] = X[3 * s + k];
      }
    }
  }

  const BundleAdjustmentOptions options_;
  BundleAdjustmentConfig config_;
  SolverSummary summary_;
  bool used_ = false;
  FlatProblem flat_;
  std::unordered_map<image_t, int32_t> image_slot_;
  std::unordered_map<camera_t, int32_t> camera_slot_;
  double flatten_ms_[3] = {0, 0, 0};
  PointSlots point_slot_;
  std::unordered_set<camera_t> problem_cameras_;           // camera_ids_, bundle_adjustment.h:201
};

}  // namespace sba_shim
#endif  // SBA_SHIM_BUNDLE_ADJUSTER_H_
