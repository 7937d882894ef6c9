// mock_reconstruction.h -- the smallest Reconstruction that honours the accessor contract
// sba_bundle_adjuster.h relies on (see its header comment).  It stands in for COLMAP's
//   Reconstruction / Image / Camera / Point2D / Point3D / Track
//   (/root/reference/src/base/reconstruction.h:83-102, image.h:114-148,229-237, camera.h:54,86-105,162,
//    point2d.h:51-69, point3d.h:53-89, track.h:43-61)
// in builds where Eigen / Boost / glog are not installed (this container).  Test infrastructure + example only.
#ifndef SBA_SHIM_MOCK_RECONSTRUCTION_H_
#define SBA_SHIM_MOCK_RECONSTRUCTION_H_

#include <array>
#include <cmath>
#include <cstdint>
#include <limits>
#include <stdexcept>
#include <unordered_map>
#include <string>
#include <vector>

#include "sba_b200.h"

namespace sba_mock {

using camera_t = uint32_t;
using image_t = uint32_t;
using point3D_t = uint64_t;
constexpr point3D_t kInvalidPoint3DId = std::numeric_limits<point3D_t>::max();  // src/util/types.h:103

using Vector2d = std::array<double, 2>;
using Vector3d = std::array<double, 3>;
using Vector4d = std::array<double, 4>;

struct TrackElement {
  image_t image_id;
  uint32_t point2D_idx;
};

class Track {
 public:
  size_t Length() const { return elements_.size(); }
  const std::vector<TrackElement>& Elements() const { return elements_; }
  void AddElement(image_t image_id, uint32_t point2D_idx) { elements_.push_back({image_id, point2D_idx}); }
  void DeleteElement(image_t image_id, uint32_t point2D_idx) {
    for (size_t i = 0; i < elements_.size(); ++i)
      if (elements_[i].image_id == image_id && elements_[i].point2D_idx == point2D_idx) {
        elements_.erase(elements_.begin() + static_cast<long>(i));
        return;
      }
  }

 private:
  std::vector<TrackElement> elements_;
};

class Point2D {
 public:
  const Vector2d& XY() const { return xy_; }
  void SetXY(const Vector2d& xy) { xy_ = xy; }
  bool HasPoint3D() const { return point3D_id_ != kInvalidPoint3DId; }
  point3D_t Point3DId() const { return point3D_id_; }
  void SetPoint3DId(point3D_t id) { point3D_id_ = id; }

 private:
  Vector2d xy_{{0, 0}};
  point3D_t point3D_id_ = kInvalidPoint3DId;
};

class Point3D {
 public:
  Vector3d& XYZ() { return xyz_; }
  const Vector3d& XYZ() const { return xyz_; }
  class Track& Track() { return track_; }
  const class Track& Track() const { return track_; }

 private:
  Vector3d xyz_{{0, 0, 0}};
  class Track track_;
};

// parameter-group index tables of the 11 models, src/base/camera_models.h:189-349 (ids) and the
// InitializeFocalLengthIdxs / PrincipalPointIdxs / ExtraParamsIdxs bodies (:597-607 ...)
struct ModelInfo {
  int num_params;
  std::vector<size_t> focal, pp, extra;
};
inline const ModelInfo& Model(int model_id) {
  static const std::vector<ModelInfo> kModels = {
      {3, {0}, {1, 2}, {}},                                  // SIMPLE_PINHOLE
      {4, {0, 1}, {2, 3}, {}},                               // PINHOLE
      {4, {0}, {1, 2}, {3}},                                 // SIMPLE_RADIAL
      {5, {0}, {1, 2}, {3, 4}},                              // RADIAL
      {8, {0, 1}, {2, 3}, {4, 5, 6, 7}},                     // OPENCV
      {8, {0, 1}, {2, 3}, {4, 5, 6, 7}},                     // OPENCV_FISHEYE
      {12, {0, 1}, {2, 3}, {4, 5, 6, 7, 8, 9, 10, 11}},      // FULL_OPENCV
      {5, {0, 1}, {2, 3}, {4}},                              // FOV
      {4, {0}, {1, 2}, {3}},                                 // SIMPLE_RADIAL_FISHEYE
      {5, {0}, {1, 2}, {3, 4}},                              // RADIAL_FISHEYE
      {12, {0, 1}, {2, 3}, {4, 5, 6, 7, 8, 9, 10, 11}}};     // THIN_PRISM_FISHEYE
  if (model_id < 0 || model_id >= static_cast<int>(kModels.size()))
    throw std::domain_error("Camera model does not exist");  // camera_models.h:140-141
  return kModels[static_cast<size_t>(model_id)];
}

class Camera {
 public:
  camera_t CameraId() const { return camera_id_; }
  void SetCameraId(camera_t id) { camera_id_ = id; }
  int ModelId() const { return model_id_; }
  void SetModelId(int model_id) {
    params_.assign(static_cast<size_t>(Model(model_id).num_params), 0.0);
    model_id_ = model_id;
  }
  size_t NumParams() const { return params_.size(); }
  std::vector<double>& Params() { return params_; }
  const std::vector<double>& Params() const { return params_; }
  double* ParamsData() { return params_.data(); }
  const std::vector<size_t>& FocalLengthIdxs() const { return Model(model_id_).focal; }
  const std::vector<size_t>& PrincipalPointIdxs() const { return Model(model_id_).pp; }
  const std::vector<size_t>& ExtraParamsIdxs() const { return Model(model_id_).extra; }

 private:
  camera_t camera_id_ = 0;
  int model_id_ = -1;
  std::vector<double> params_;
};

class Image {
 public:
  image_t ImageId() const { return image_id_; }
  void SetImageId(image_t id) { image_id_ = id; }
  camera_t CameraId() const { return camera_id_; }
  void SetCameraId(camera_t id) { camera_id_ = id; }
  const std::string& Name() const { return name_; }   // src/base/image.h:150-152: keys the map tables of the semantic adjusters
  void SetName(const std::string& name) { name_ = name; }
  Vector4d& Qvec() { return qvec_; }
  const Vector4d& Qvec() const { return qvec_; }
  Vector3d& Tvec() { return tvec_; }
  const Vector3d& Tvec() const { return tvec_; }
  void NormalizeQvec() {  // src/base/pose.cc:82-91
    const double n = std::sqrt(qvec_[0] * qvec_[0] + qvec_[1] * qvec_[1] + qvec_[2] * qvec_[2] + qvec_[3] * qvec_[3]);
    if (n == 0.0) {
      qvec_ = {{1, 0, 0, 0}};
    } else {
      for (double& v : qvec_) v /= n;
    }
  }
  std::vector<class Point2D>& Points2D() { return points2D_; }
  const std::vector<class Point2D>& Points2D() const { return points2D_; }
  class Point2D& Point2D(uint32_t idx) { return points2D_.at(idx); }
  const class Point2D& Point2D(uint32_t idx) const { return points2D_.at(idx); }
  void SetPoints2D(const std::vector<Vector2d>& xy) {
    points2D_.assign(xy.size(), sba_mock::Point2D());
    for (size_t i = 0; i < xy.size(); ++i) points2D_[i].SetXY(xy[i]);
  }

 private:
  image_t image_id_ = 0;
  camera_t camera_id_ = 0;
  std::string name_;
  Vector4d qvec_{{1, 0, 0, 0}};
  Vector3d tvec_{{0, 0, 0}};
  std::vector<class Point2D> points2D_;
};

class Reconstruction {
 public:
  void AddCamera(const class Camera& c) { cameras_[c.CameraId()] = c; }
  void AddImage(const class Image& im) { images_[im.ImageId()] = im; reg_image_ids_.push_back(im.ImageId()); }  // (every mock image is registered)
  point3D_t AddPoint3D(const Vector3d& xyz, const Track& track) {  // reconstruction.cc:299-324
    const point3D_t id = ++num_added_points3D_;
    class Point3D& p = points3D_[id];
    p.XYZ() = xyz;
    p.Track() = track;
    for (const TrackElement& el : track.Elements()) images_.at(el.image_id).Point2D(el.point2D_idx).SetPoint3DId(id);
    return id;
  }
  void DeleteObservation(image_t image_id, uint32_t point2D_idx) {  // reconstruction.cc:368-388 (track stays >= 2 here)
    class Point2D& p2 = images_.at(image_id).Point2D(point2D_idx);
    points3D_.at(p2.Point3DId()).Track().DeleteElement(image_id, point2D_idx);
    p2.SetPoint3DId(kInvalidPoint3DId);
  }
  class Camera& Camera(camera_t id) { return cameras_.at(id); }
  const class Camera& Camera(camera_t id) const { return cameras_.at(id); }
  class Image& Image(image_t id) { return images_.at(id); }
  const class Image& Image(image_t id) const { return images_.at(id); }
  class Point3D& Point3D(point3D_t id) { return points3D_.at(id); }
  const class Point3D& Point3D(point3D_t id) const { return points3D_.at(id); }
  size_t NumImages() const { return images_.size(); }
  size_t NumPoints3D() const { return points3D_.size(); }
  const std::unordered_map<point3D_t, class Point3D>& Points3D() const { return points3D_; }
  const std::unordered_map<image_t, class Image>& Images() const { return images_; }   // reconstruction.h:96
  const std::vector<image_t>& RegImageIds() const { return reg_image_ids_; }           // reconstruction.h:99

 private:
  std::unordered_map<camera_t, class Camera> cameras_;
  std::unordered_map<image_t, class Image> images_;
  std::unordered_map<point3D_t, class Point3D> points3D_;
  std::vector<image_t> reg_image_ids_;
  point3D_t num_added_points3D_ = 0;
};

// ---- CameraRig: the part the rig bundle adjuster uses (/root/reference/src/base/camera_rig.h:27-107, camera_rig.cc:86-110,
// 230-258; pose helpers /root/reference/src/base/pose.cc:82-143, 192-196) ----------------------------------------------
class CameraRig {
 public:
  void AddCamera(camera_t camera_id, const Vector4d& rel_qvec, const Vector3d& rel_tvec) { rig_cameras_[camera_id] = {rel_qvec, rel_tvec}; order_.push_back(camera_id); }
  void AddSnapshot(const std::vector<image_t>& image_ids) { snapshots_.push_back(image_ids); }
  void SetRefCameraId(camera_t id) { ref_camera_id_ = id; }
  camera_t RefCameraId() const { return ref_camera_id_; }
  bool HasCamera(camera_t id) const { return rig_cameras_.count(id) > 0; }
  size_t NumCameras() const { return rig_cameras_.size(); }
  size_t NumSnapshots() const { return snapshots_.size(); }
  const std::vector<camera_t>& GetCameraIds() const { return order_; }
  const std::vector<std::vector<image_t>>& Snapshots() const { return snapshots_; }
  Vector4d& RelativeQvec(camera_t id) { return rig_cameras_.at(id).first; }
  const Vector4d& RelativeQvec(camera_t id) const { return rig_cameras_.at(id).first; }
  Vector3d& RelativeTvec(camera_t id) { return rig_cameras_.at(id).second; }
  const Vector3d& RelativeTvec(camera_t id) const { return rig_cameras_.at(id).second; }
  void Check(const Reconstruction& rec) const {  // camera_rig.cc:86-110
    if (!HasCamera(ref_camera_id_)) throw std::invalid_argument("reference camera is not part of the rig");
    for (const auto& kv : rig_cameras_) rec.Camera(kv.first);
    for (const auto& snap : snapshots_) {
      if (snap.empty() || snap.size() > NumCameras()) throw std::invalid_argument("invalid snapshot");
      bool has_ref = false;
      for (image_t id : snap) {
        if (!HasCamera(rec.Image(id).CameraId())) throw std::invalid_argument("snapshot image's camera is not part of the rig");
        has_ref = has_ref || rec.Image(id).CameraId() == ref_camera_id_;
      }
      if (!has_ref) throw std::invalid_argument("snapshot without the reference camera");
    }
  }
  // average of the rig poses the snapshot's images imply (camera_rig.cc:230-258)
  void ComputeAbsolutePose(size_t snapshot_idx, const Reconstruction& rec, Vector4d* abs_qvec, Vector3d* abs_tvec) const {
    const auto& snap = snapshots_.at(snapshot_idx);
    double A[4][4] = {};
    *abs_tvec = {0.0, 0.0, 0.0};
    Vector4d last{};
    for (image_t id : snap) {
      const Image& im = rec.Image(id);
      const Vector4d rq = Normalized(RelativeQvec(im.CameraId()));
      const Vector4d inv{rq[0], -rq[1], -rq[2], -rq[3]};                      // InvertQuaternion
      const Vector4d q = Normalized(Mul(inv, Normalized(im.Qvec())));         // ConcatenateQuaternions(image, inv_rel) = inv_rel * image
      const Vector3d d{im.Tvec()[0] - RelativeTvec(im.CameraId())[0], im.Tvec()[1] - RelativeTvec(im.CameraId())[1], im.Tvec()[2] - RelativeTvec(im.CameraId())[2]};
      const Vector3d t = Rot(inv, d);
      for (int a = 0; a < 4; ++a) for (int b = 0; b < 4; ++b) A[a][b] += q[a] * q[b];
      for (int a = 0; a < 3; ++a) (*abs_tvec)[a] += t[a];
      last = q;
    }
    for (int a = 0; a < 3; ++a) (*abs_tvec)[a] /= (double)snap.size();
    if (snap.size() == 1) { *abs_qvec = last; return; }
    // AverageQuaternions (pose.cc:118-143): eigenvector of the largest eigenvalue of sum q q^T -- cyclic Jacobi on the 4 x 4
    double V[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
    for (int sweep = 0; sweep < 50; ++sweep) {
      double off = 0.0;
      for (int p = 0; p < 4; ++p) for (int q = p + 1; q < 4; ++q) off += A[p][q] * A[p][q];
      if (off < 1e-30) break;
      for (int p = 0; p < 4; ++p)
        for (int q = p + 1; q < 4; ++q) {
          if (std::fabs(A[p][q]) < 1e-300) continue;
          const double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
          const double tt = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
          const double c = 1.0 / std::sqrt(tt * tt + 1.0), s = tt * c;
          for (int k = 0; k < 4; ++k) { const double akp = A[k][p], akq = A[k][q]; A[k][p] = c * akp - s * akq; A[k][q] = s * akp + c * akq; }
          for (int k = 0; k < 4; ++k) { const double apk = A[p][k], aqk = A[q][k]; A[p][k] = c * apk - s * aqk; A[q][k] = s * apk + c * aqk; }
          for (int k = 0; k < 4; ++k) { const double vkp = V[k][p], vkq = V[k][q]; V[k][p] = c * vkp - s * vkq; V[k][q] = s * vkp + c * vkq; }
        }
    }
    int best = 0;
    for (int k = 1; k < 4; ++k) if (A[k][k] > A[best][best]) best = k;
    *abs_qvec = {V[0][best], V[1][best], V[2][best], V[3][best]};
  }

 private:
  static Vector4d Normalized(const Vector4d& q) {
    const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    return n == 0.0 ? Vector4d{1.0, 0.0, 0.0, 0.0} : Vector4d{q[0] / n, q[1] / n, q[2] / n, q[3] / n};
  }
  static Vector4d Mul(const Vector4d& a, const Vector4d& b) {
    return {a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
            a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]};
  }
  static Vector3d Rot(const Vector4d& qv, const Vector3d& p) {
    const Vector4d q = Normalized(qv);
    const double w = q[0], x = q[1], y = q[2], z = q[3];
    return {(1 - 2 * (y * y + z * z)) * p[0] + 2 * (x * y - w * z) * p[1] + 2 * (x * z + w * y) * p[2],
            2 * (x * y + w * z) * p[0] + (1 - 2 * (x * x + z * z)) * p[1] + 2 * (y * z - w * x) * p[2],
            2 * (x * z - w * y) * p[0] + 2 * (y * z + w * x) * p[1] + (1 - 2 * (x * x + y * y)) * p[2]};
  }
  camera_t ref_camera_id_ = std::numeric_limits<camera_t>::max();
  std::unordered_map<camera_t, std::pair<Vector4d, Vector3d>> rig_cameras_;
  std::vector<camera_t> order_;
  std::vector<std::vector<image_t>> snapshots_;
};

}  // namespace sba_mock
#endif  // SBA_SHIM_MOCK_RECONSTRUCTION_H_
