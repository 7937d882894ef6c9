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

}  // namespace sba_mock
#endif  // SBA_SHIM_MOCK_RECONSTRUCTION_H_
