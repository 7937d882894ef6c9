// sba_rig_bundle_adjuster.h -- header-only C++17 shim of RigBundleAdjuster over sba_rig_solve (include/sba_b200.h).
//
//   RigBundleAdjuster            /root/reference/src/optim/bundle_adjustment.h:270-330, bundle_adjustment.cc:789-1140
//
// Templated on the Reconstruction AND the CameraRig type, so that it compiles against COLMAP's classes -- whose
// CameraRig::ComputeAbsolutePose / Check (src/base/camera_rig.cc:86-110, 230-258) it calls as they are -- and against the
// mocks of mock_reconstruction.h.  Accessor contract of CameraRig: GetCameraIds(), Snapshots(), NumSnapshots(),
// RefCameraId(), RelativeQvec(camera_id) / RelativeTvec(camera_id) (mutable, .data()), Check(reconstruction),
// ComputeAbsolutePose(snapshot_idx, reconstruction, &qvec, &tvec).
// The shim flattens the residual blocks in the reference's order, gates the rig observations by max_reproj_error
// through the device pass sba_observation_errors (:943-975), calls sba_rig_solve (dense small-problem path, see
// csrc/rig_solve.cu) and writes the concatenated poses back (TearDown, :889-903).  No arithmetic of the solve runs on the
// host and there is no CPU fallback.
#ifndef SBA_SHIM_RIG_BUNDLE_ADJUSTER_H_
#define SBA_SHIM_RIG_BUNDLE_ADJUSTER_H_

#include <array>
#include <cmath>
#include <map>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

#include "sba_bundle_adjuster.h"

namespace sba_shim {

namespace rig_detail {
using Q4 = std::array<double, 4>;
using V3 = std::array<double, 3>;
inline Q4 Normalized(const double* q) {  // NormalizeQuaternion, src/base/pose.cc:82-91
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n == 0.0) return Q4{1.0, 0.0, 0.0, 0.0};
  return Q4{q[0] / n, q[1] / n, q[2] / n, q[3] / n};
}
inline V3 Rotate(const double* qvec, const double* p) {  // QuaternionRotatePoint, src/base/pose.cc:110-116
  const Q4 q = Normalized(qvec);
  const double w = q[0], x = q[1], y = q[2], z = q[3];
  const double R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z),     2 * (x * z + w * y),
                       2 * (x * y + w * z),     1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                       2 * (x * z - w * y),     2 * (y * z + w * x),     1 - 2 * (x * x + y * y)};
  return V3{R[0] * p[0] + R[1] * p[1] + R[2] * p[2], R[3] * p[0] + R[4] * p[1] + R[5] * p[2], R[6] * p[0] + R[7] * p[1] + R[8] * p[2]};
}
// ConcatenatePoses, src/base/pose.cc:183-190: q12 = normalise(q2 * q1), t12 = t2 + R(q2) t1
inline void Concatenate(const double* q1, const double* t1, const double* q2, const double* t2, double* q12, double* t12) {
  const Q4 a = Normalized(q2), b = Normalized(q1);
  const double c[4] = {a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                       a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]};
  const Q4 n = Normalized(c);
  const V3 r = Rotate(q2, t1);
  for (int j = 0; j < 4; ++j) q12[j] = n[j];
  for (int j = 0; j < 3; ++j) t12[j] = t2[j] + r[j];
}
}  // namespace rig_detail

struct RigBundleAdjusterOptions {  // RigBundleAdjuster::Options, bundle_adjustment.h:272-282
  bool refine_relative_poses = true;
  double max_reproj_error = 1000.0;
};

template <class Reconstruction, class CameraRig>
class RigBundleAdjusterT {
 public:
  RigBundleAdjusterT(const BundleAdjustmentOptions& options, const RigBundleAdjusterOptions& rig_options, const BundleAdjustmentConfig& config)
      : options_(options), rig_options_(rig_options), config_(config) {
    if (!options_.Check()) throw std::invalid_argument("BundleAdjustmentOptions::Check failed");
  }

  bool Solve(Reconstruction* rec, std::vector<CameraRig>* camera_rigs) {
    if (!rec || !camera_rigs) throw std::invalid_argument("reconstruction / camera_rigs must not be null");  // CHECK_NOTNULL, :796-797
    if (used_) throw std::logic_error("Cannot use the same BundleAdjuster multiple times");
    used_ = true;
    Flatten(rec, camera_rigs);
    if (b_rig_.empty()) return false;  // :824-826
    sba_rig_problem p{};
    p.num_rig_poses = (int32_t)rig_const_.size(); p.rig_qvec = rig_q_.data(); p.rig_tvec = rig_t_.data(); p.rig_pose_constant = rig_const_.data();
    p.num_rel_poses = (int32_t)rel_const_.size(); p.rel_qvec = rel_q_.data(); p.rel_tvec = rel_t_.data();
    p.rel_pose_constant = rel_const_.data(); p.rel_tvec_constant_mask = rel_mask_.data();
    p.num_cameras = (int32_t)cam_model_.size(); p.camera_model = cam_model_.data(); p.camera_params = cam_params_.data();
    p.camera_param_constant_mask = cam_mask_.data();
    p.num_points = (int32_t)pt_const_.size(); p.points = points_.data(); p.point_constant = pt_const_.data();
    p.num_blocks = (int64_t)b_rig_.size();
    p.block_rig = b_rig_.data(); p.block_rel = b_rel_.data(); p.block_camera = b_cam_.data(); p.block_point = b_pt_.data(); p.block_xy = b_xy_.data();
    sba_solver_options o;
    sba_default_solver_options(&o);
    const SolverOptions& so = options_.solver_options;
    o.num_config_images = (int32_t)config_.NumImages();
    o.max_num_iterations = so.max_num_iterations;
    o.max_num_consecutive_invalid_steps = so.max_num_consecutive_invalid_steps;
    o.function_tolerance = so.function_tolerance; o.gradient_tolerance = so.gradient_tolerance; o.parameter_tolerance = so.parameter_tolerance;
    o.loss_function_type = options_.LossFunctionCode(); o.loss_function_scale = options_.loss_function_scale;
    if (sba_rig_solve(&p, &o, nullptr, nullptr, &summary_, nullptr, 0, nullptr) != SBA_OK)
      throw std::runtime_error(std::string("sba_rig_solve: ") + sba_last_error());
    ScatterBack(rec, camera_rigs);
    return true;
  }
  const SolverSummary& Summary() const { return summary_; }
  size_t NumResidualBlocks() const { return b_rig_.size(); }

  // SetUp (:869-887): public so that tests can look at the flat problem
  void Flatten(Reconstruction* rec, std::vector<CameraRig>* rigs) {
    using namespace rig_detail;
    // validity of the rigs (:800-817)
    std::unordered_set<uint64_t> rig_camera_ids;
    image_rig_.clear();
    for (size_t r = 0; r < rigs->size(); ++r) {
      CameraRig& rig = (*rigs)[r];
      rig.Check(*rec);
      for (const auto camera_id : rig.GetCameraIds()) {
        if (!rig_camera_ids.insert((uint64_t)camera_id).second) throw std::invalid_argument("Camera must not be part of multiple camera rigs");
      }
      for (const auto& snapshot : rig.Snapshots())
        for (const auto image_id : snapshot)
          if (!image_rig_.emplace((uint64_t)image_id, r).second) throw std::invalid_argument("Image must not be part of multiple camera rigs");
    }
    // slot 0: the constant identity pose the images outside rigs hang on; then ComputeCameraRigPoses (:1104-1129)
    rig_q_ = {1.0, 0.0, 0.0, 0.0}; rig_t_ = {0.0, 0.0, 0.0}; rig_const_ = {1};
    image_rig_slot_.clear();
    for (CameraRig& rig : *rigs)
      for (size_t s = 0; s < rig.NumSnapshots(); ++s) {
        std::decay_t<decltype(rig.RelativeQvec(rig.RefCameraId()))> q{};
        std::decay_t<decltype(rig.RelativeTvec(rig.RefCameraId()))> t{};
        rig.ComputeAbsolutePose(s, *rec, &q, &t);
        for (int j = 0; j < 4; ++j) rig_q_.push_back(q.data()[j]);
        for (int j = 0; j < 3; ++j) rig_t_.push_back(t.data()[j]);
        rig_const_.push_back(0);
        for (const auto image_id : rig.Snapshots()[s]) image_rig_slot_[(uint64_t)image_id] = (int32_t)rig_const_.size() - 1;
      }
    // the outlier gate (:943-975): squared reprojection error of every rig observation under the concatenated pose
    std::map<std::pair<uint64_t, uint32_t>, double> gate;
    {
      std::vector<double> gq, gt, gxy, gX, gparams;
      std::vector<int32_t> gcam, gmodel, go_i, go_p;
      std::unordered_map<uint64_t, int32_t> cam_l, pt_l;
      std::vector<std::pair<uint64_t, uint32_t>> keys;
      for (const auto image_id : config_.Images()) {
        if (!image_rig_.count((uint64_t)image_id)) continue;
        auto& im = rec->Image(image_id);
        CameraRig& rig = (*rigs)[image_rig_.at((uint64_t)image_id)];
        const int32_t s = image_rig_slot_.at((uint64_t)image_id);
        double q[4], t[3];
        Concatenate(&rig_q_[4 * (size_t)s], &rig_t_[3 * (size_t)s], rig.RelativeQvec(im.CameraId()).data(), rig.RelativeTvec(im.CameraId()).data(), q, t);
        gq.insert(gq.end(), q, q + 4); gt.insert(gt.end(), t, t + 3);
        auto cit = cam_l.find((uint64_t)im.CameraId());
        if (cit == cam_l.end()) {
          cit = cam_l.emplace((uint64_t)im.CameraId(), (int32_t)gmodel.size()).first;
          const auto& cam = rec->Camera(im.CameraId());
          gmodel.push_back(cam.ModelId());
          for (size_t k = 0; k < SBA_MAX_CAMERA_PARAMS; ++k) gparams.push_back(k < cam.NumParams() ? cam.Params()[k] : 0.0);
        }
        gcam.push_back(cit->second);
        uint32_t idx = 0;
        for (const auto& p2 : im.Points2D()) {
          if (p2.HasPoint3D()) {
            auto pit = pt_l.find((uint64_t)p2.Point3DId());
            if (pit == pt_l.end()) {
              pit = pt_l.emplace((uint64_t)p2.Point3DId(), (int32_t)(gX.size() / 3)).first;
              const auto& X = rec->Point3D(p2.Point3DId()).XYZ();
              gX.insert(gX.end(), X.data(), X.data() + 3);
            }
            go_i.push_back((int32_t)gcam.size() - 1); go_p.push_back(pit->second);
            gxy.push_back(p2.XY().data()[0]); gxy.push_back(p2.XY().data()[1]);
            keys.emplace_back((uint64_t)image_id, idx);
          }
          ++idx;
        }
      }
      if (!keys.empty()) {
        sba_problem_desc d{};
        d.num_images = (int32_t)gcam.size(); d.num_cameras = (int32_t)gmodel.size(); d.num_points = (int32_t)(gX.size() / 3);
        d.num_observations = (int64_t)keys.size();
        d.image_qvec = gq.data(); d.image_tvec = gt.data(); d.image_camera = gcam.data(); d.camera_model = gmodel.data();
        d.camera_params = gparams.data(); d.points = gX.data(); d.obs_image = go_i.data(); d.obs_point = go_p.data(); d.obs_xy = gxy.data();
        d.shard_count = 1; d.global_num_observations = d.num_observations;
        std::vector<double> err(keys.size());
        if (sba_observation_errors(&d, err.data(), nullptr) != SBA_OK) throw std::runtime_error(std::string("sba_observation_errors: ") + sba_last_error());
        for (size_t k = 0; k < keys.size(); ++k) gate[keys[k]] = err[k];
      }
    }
    const double max_sq = rig_options_.max_reproj_error * rig_options_.max_reproj_error;
    rel_index_.clear(); cam_index_.clear(); pt_index_.clear();
    rel_q_.clear(); rel_t_.clear(); rel_const_.clear(); rel_mask_.clear();
    b_rig_.clear(); b_rel_.clear(); b_cam_.clear(); b_pt_.clear(); b_xy_.clear();
    std::unordered_map<uint64_t, size_t> num_obs;
    std::unordered_set<uint64_t> camera_ids;
    auto cam_slot = [&](uint64_t id) { return cam_index_.emplace(id, (int32_t)cam_index_.size()).first->second; };
    auto pt_slot = [&](uint64_t id) { return pt_index_.emplace(id, (int32_t)pt_index_.size()).first->second; };
    auto rel_slot = [&](const std::pair<uint64_t, uint64_t>& key, const double* q, const double* t, bool konst, uint8_t mask) {
      auto it = rel_index_.find(key);
      if (it == rel_index_.end()) {
        it = rel_index_.emplace(key, (int32_t)rel_const_.size()).first;
        rel_q_.insert(rel_q_.end(), q, q + 4); rel_t_.insert(rel_t_.end(), t, t + 3);
        rel_const_.push_back(konst ? 1 : 0); rel_mask_.push_back(mask);
      }
      return it->second;
    };
    for (const auto image_id : config_.Images()) {  // AddImageToProblem, :905-1055
      auto& im = rec->Image(image_id);
      const bool constant_pose = config_.HasConstantPose(image_id), constant_tvec = config_.HasConstantTvec(image_id);
      const bool in_rig = image_rig_.count((uint64_t)image_id) > 0;
      if (in_rig && (constant_pose || constant_tvec))
        throw std::invalid_argument("Images contained in a camera rig must not have constant pose / tvec");  // CHECK, :929-932
      if (!in_rig) im.NormalizeQvec();
      camera_ids.insert((uint64_t)im.CameraId());
      uint32_t idx = 0;
      for (const auto& p2 : im.Points2D()) {
        const uint32_t this_idx = idx++;
        if (!p2.HasPoint3D()) continue;
        if (in_rig && gate.at({(uint64_t)image_id, this_idx}) > max_sq) continue;
        num_obs[(uint64_t)p2.Point3DId()] += 1;
        int32_t r;
        if (in_rig) {
          CameraRig& rig = (*rigs)[image_rig_.at((uint64_t)image_id)];
          const bool konst = !rig_options_.refine_relative_poses || im.CameraId() == rig.RefCameraId();  // :1040-1046
          r = rel_slot({1 + image_rig_.at((uint64_t)image_id), (uint64_t)im.CameraId()}, rig.RelativeQvec(im.CameraId()).data(),
                       rig.RelativeTvec(im.CameraId()).data(), konst, 0);
          b_rig_.push_back(image_rig_slot_.at((uint64_t)image_id));
        } else {
          uint8_t mask = 0;
          if (!constant_pose && constant_tvec) for (int k : config_.ConstantTvec(image_id)) mask |= (uint8_t)(1u << k);
          r = rel_slot({0, (uint64_t)image_id}, im.Qvec().data(), im.Tvec().data(), constant_pose, mask);
          b_rig_.push_back(0);
        }
        b_rel_.push_back(r); b_cam_.push_back(cam_slot((uint64_t)im.CameraId())); b_pt_.push_back(pt_slot((uint64_t)p2.Point3DId()));
        b_xy_.push_back(p2.XY().data()[0]); b_xy_.push_back(p2.XY().data()[1]);
      }
    }
    auto add_point = [&](uint64_t pid) {  // AddPointToProblem, :1057-1102
      auto& p3 = rec->Point3D(pid);
      if (num_obs[pid] == p3.Track().Length()) return;
      for (const auto& el : p3.Track().Elements()) {
        if (config_.HasImage(el.image_id)) continue;
        num_obs[pid] += 1;
        auto& im = rec->Image(el.image_id);
        if (camera_ids.insert((uint64_t)im.CameraId()).second) config_.SetConstantCamera(im.CameraId());
        const int32_t r = rel_slot({0, (uint64_t)el.image_id}, im.Qvec().data(), im.Tvec().data(), true, 0);
        b_rig_.push_back(0); b_rel_.push_back(r); b_cam_.push_back(cam_slot((uint64_t)im.CameraId())); b_pt_.push_back(pt_slot(pid));
        const auto& xy = im.Point2D(el.point2D_idx).XY();
        b_xy_.push_back(xy.data()[0]); b_xy_.push_back(xy.data()[1]);
      }
    };
    for (const auto pid : config_.VariablePoints()) add_point((uint64_t)pid);
    for (const auto pid : config_.ConstantPoints()) add_point((uint64_t)pid);
    // ParameterizeCameras / ParameterizePoints as BundleAdjuster (:480-530)
    const bool constant_camera = !(options_.refine_focal_length || options_.refine_principal_point || options_.refine_extra_params);
    cam_model_.assign(cam_index_.size(), 0); cam_mask_.assign(cam_index_.size(), 0);
    cam_params_.assign(cam_index_.size() * SBA_MAX_CAMERA_PARAMS, 0.0);
    for (const auto& kv : cam_index_) {
      const auto& cam = rec->Camera((decltype(rec->Image(0).CameraId()))kv.first);
      uint16_t m = 0;
      if (constant_camera || config_.IsConstantCamera((decltype(rec->Image(0).CameraId()))kv.first)) {
        m = (uint16_t)((1u << cam.NumParams()) - 1);
      } else {
        if (!options_.refine_focal_length) for (size_t k : cam.FocalLengthIdxs()) m |= (uint16_t)(1u << k);
        if (!options_.refine_principal_point) for (size_t k : cam.PrincipalPointIdxs()) m |= (uint16_t)(1u << k);
        if (!options_.refine_extra_params) for (size_t k : cam.ExtraParamsIdxs()) m |= (uint16_t)(1u << k);
      }
      cam_mask_[kv.second] = m; cam_model_[kv.second] = cam.ModelId();
      for (size_t k = 0; k < cam.NumParams(); ++k) cam_params_[(size_t)kv.second * SBA_MAX_CAMERA_PARAMS + k] = cam.Params()[k];
    }
    points_.assign(pt_index_.size() * 3, 0.0); pt_const_.assign(pt_index_.size(), 0);
    for (const auto& kv : pt_index_) {
      const auto& p3 = rec->Point3D(kv.first);
      for (int j = 0; j < 3; ++j) points_[3 * (size_t)kv.second + j] = p3.XYZ().data()[j];
      if (p3.Track().Length() > num_obs[kv.first] || config_.HasConstantPoint(kv.first)) pt_const_[kv.second] = 1;
    }
  }

 private:
  void ScatterBack(Reconstruction* rec, std::vector<CameraRig>* rigs) {
    for (const auto& kv : rel_index_) {
      const double* q = &rel_q_[4 * (size_t)kv.second];
      const double* t = &rel_t_[3 * (size_t)kv.second];
      if (kv.first.first == 0) {
        auto& im = rec->Image((decltype(rec->Image(0).ImageId()))kv.first.second);
        for (int j = 0; j < 4; ++j) im.Qvec().data()[j] = q[j];
        for (int j = 0; j < 3; ++j) im.Tvec().data()[j] = t[j];
      } else {
        CameraRig& rig = (*rigs)[kv.first.first - 1];
        const auto cid = (decltype(rig.RefCameraId()))kv.first.second;
        for (int j = 0; j < 4; ++j) rig.RelativeQvec(cid).data()[j] = q[j];
        for (int j = 0; j < 3; ++j) rig.RelativeTvec(cid).data()[j] = t[j];
      }
    }
    for (const auto& kv : cam_index_) {
      auto& cam = rec->Camera((decltype(rec->Image(0).CameraId()))kv.first);
      for (size_t k = 0; k < cam.NumParams(); ++k) cam.Params()[k] = cam_params_[(size_t)kv.second * SBA_MAX_CAMERA_PARAMS + k];
    }
    for (const auto& kv : pt_index_) {
      auto& p3 = rec->Point3D(kv.first);
      for (int j = 0; j < 3; ++j) p3.XYZ().data()[j] = points_[3 * (size_t)kv.second + j];
    }
    for (const auto& kv : image_rig_) {  // TearDown, :889-903
      auto& im = rec->Image((decltype(rec->Image(0).ImageId()))kv.first);
      CameraRig& rig = (*rigs)[kv.second];
      const int32_t s = image_rig_slot_.at(kv.first);
      rig_detail::Concatenate(&rig_q_[4 * (size_t)s], &rig_t_[3 * (size_t)s], rig.RelativeQvec(im.CameraId()).data(),
                              rig.RelativeTvec(im.CameraId()).data(), im.Qvec().data(), im.Tvec().data());
    }
  }

  BundleAdjustmentOptions options_;
  RigBundleAdjusterOptions rig_options_;
  BundleAdjustmentConfig config_;
  SolverSummary summary_;
  bool used_ = false;
  std::unordered_map<uint64_t, size_t> image_rig_;          // image -> index of its rig
  std::unordered_map<uint64_t, int32_t> image_rig_slot_;    // image -> rig pose slot
  std::map<std::pair<uint64_t, uint64_t>, int32_t> rel_index_;  // (0, image) | (1 + rig, camera) -> relative pose slot
  std::unordered_map<uint64_t, int32_t> cam_index_, pt_index_;
  std::vector<double> rig_q_, rig_t_, rel_q_, rel_t_, cam_params_, points_, b_xy_;
  std::vector<uint8_t> rig_const_, rel_const_, rel_mask_, pt_const_;
  std::vector<uint16_t> cam_mask_;
  std::vector<int32_t> cam_model_, b_rig_, b_rel_, b_cam_, b_pt_;
};

}  // namespace sba_shim

#endif  // SBA_SHIM_RIG_BUNDLE_ADJUSTER_H_
