// Mapper-side per-observation passes around global bundle adjustment: the squared
// reprojection error and the depth of every observation, in one launch over the caller's observation list.
//
//   CalculateSquaredReprojectionError      /root/reference/src/base/projection.cc:111-128
//     (behind Reconstruction::FilterPoints3DWithLargeReprojectionError, src/base/reconstruction.cc:1472-1523)
//   HasPointPositiveDepth                  /root/reference/src/base/projection.cc:191-195
//     (behind Reconstruction::FilterObservationsWithNegativeDepth, src/base/reconstruction.cc:647-665)
//
// Both take the rotation of the NORMALISED quaternion (QuaternionRotatePoint / ComposeProjectionMatrix,
// src/base/pose.cc:63-80, 110-116), unlike the bundle-adjustment functor, and the error is
// std::numeric_limits<double>::max() for a point that is not in front of the camera.  The filters' bookkeeping
// (which track elements / points to delete, in which order) is host logic in the package's `filters` module.
#include <cfloat>
#include <cmath>

#include <cub/cub.cuh>

#include "device_math.cuh"
#include "solver.h"

namespace sba {

__global__ void __launch_bounds__(256) observation_error_kernel(int64_t n_obs, const double* __restrict__ qvec,
                                                                const double* __restrict__ tvec,
                                                                const int* __restrict__ image_camera,
                                                                const int* __restrict__ camera_model,
                                                                const double* __restrict__ camera_params,
                                                                const double* __restrict__ points,
                                                                const int* __restrict__ obs_image, const int* __restrict__ obs_point,
                                                                const double2* __restrict__ obs_xy,
                                                                double* __restrict__ sq_error, double* __restrict__ depth) {
  const int64_t k = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (k >= n_obs) return;
  const int i = obs_image[k], c = image_camera[i];
  const double* q = qvec + 4 * (size_t)i;
  const double* t = tvec + 3 * (size_t)i;
  const double* X = points + 3 * (size_t)obs_point[k];
  double w = q[0], x = q[1], y = q[2], z = q[3];
  const double norm = sqrt(w * w + x * x + y * y + z * z);
  if (norm == 0.0) { w = 1.0; x = y = z = 0.0; }  // NormalizeQuaternion, src/base/pose.cc:82-91
  else { w /= norm; x /= norm; y /= norm; z /= norm; }
  const double R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z),     2 * (x * z + w * y),
                       2 * (x * y + w * z),     1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                       2 * (x * z - w * y),     2 * (y * z + w * x),     1 - 2 * (x * x + y * y)};
  const double px = R[0] * X[0] + R[1] * X[1] + R[2] * X[2] + t[0];
  const double py = R[3] * X[0] + R[4] * X[1] + R[5] * X[2] + t[1];
  const double pz = R[6] * X[0] + R[7] * X[1] + R[8] * X[2] + t[2];
  if (depth) depth[k] = pz;
  if (!sq_error) return;
  if (pz < DBL_EPSILON) { sq_error[k] = DBL_MAX; return; }
  double ix, iy;
  WorldToImageValDyn(camera_model[c], camera_params + SBA_MAX_CAMERA_PARAMS * (size_t)c, px / pz, py / pz, &ix, &iy);
  const double2 m = obs_xy[k];
  sq_error[k] = (ix - m.x) * (ix - m.x) + (iy - m.y) * (iy - m.y);
}


// ---- Reconstruction::Normalize (/root/reference/src/base/reconstruction.cc:321-345) ---------------------------------
// bounds and centroid from the float-cast coordinates of the registered images' projection centres (or of the points),
// sorted per axis (ComputeBoundsAndCentroid, :358-422); then Transform (:424-431) with the similarity
// (scale, identity rotation, -scale * centroid): points through SimilarityTransform3::TransformPoint, poses through
// ::TransformPose (/root/reference/src/base/similarity_transform.cc:211-240).

// Eigen::Quaterniond::toRotationMatrix of the NORMALISED quaternion (QuaternionToRotationMatrix, src/base/pose.cc:63-80)
__device__ inline void QuatToRot(const double* qv, double* R) {
  double w = qv[0], x = qv[1], y = qv[2], z = qv[3];
  const double norm = sqrt(w * w + x * x + y * y + z * z);
  if (norm == 0.0) { w = 1.0; x = y = z = 0.0; }
  else { w /= norm; x /= norm; y /= norm; z /= norm; }
  const double tx = 2.0 * x, ty = 2.0 * y, tz = 2.0 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1.0 - (tyy + tzz); R[1] = txy - twz;         R[2] = txz + twy;
  R[3] = txy + twz;         R[4] = 1.0 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy;         R[7] = tyz + twx;         R[8] = 1.0 - (txx + tyy);
}

// Eigen::Quaterniond(rotation matrix) (RotationMatrixToQuaternion, src/base/pose.cc:72-80): Shepperd's branches as Eigen orders them
__device__ inline void RotToQuat(const double* m, double* q) {
  double t = m[0] + m[4] + m[8];
  if (t > 0.0) {
    t = sqrt(t + 1.0);
    q[0] = 0.5 * t;
    t = 0.5 / t;
    q[1] = (m[7] - m[5]) * t; q[2] = (m[2] - m[6]) * t; q[3] = (m[3] - m[1]) * t;
  } else {
    int i = 0;
    if (m[4] > m[0]) i = 1;
    if (m[8] > m[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = sqrt(m[4 * i] - m[4 * j] - m[4 * k] + 1.0);
    q[1 + i] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (m[3 * k + j] - m[3 * j + k]) * t;
    q[1 + j] = (m[3 * j + i] + m[3 * i + j]) * t;
    q[1 + k] = (m[3 * k + i] + m[3 * i + k]) * t;
  }
}

// float coordinates of element e: projection centre of the e-th REGISTERED image (-R^T t, ProjectionCenterFromPose,
// src/base/pose.cc:164-171) or the e-th point
__global__ void normalize_coords_kernel(int64_t n, int use_images, const int* __restrict__ reg, const double* __restrict__ qvec,
                                        const double* __restrict__ tvec, const double* __restrict__ points,
                                        float* __restrict__ cx, float* __restrict__ cy, float* __restrict__ cz) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  double c[3];
  if (use_images) {
    const int i = reg[e];
    double R[9];
    QuatToRot(qvec + 4 * (size_t)i, R);
    const double* t = tvec + 3 * (size_t)i;
    // conjugate quaternion applied to -t  ==  R^T (-t)
    c[0] = -(R[0] * t[0] + R[3] * t[1] + R[6] * t[2]);
    c[1] = -(R[1] * t[0] + R[4] * t[1] + R[7] * t[2]);
    c[2] = -(R[2] * t[0] + R[5] * t[1] + R[8] * t[2]);
  } else {
    c[0] = points[3 * e]; c[1] = points[3 * e + 1]; c[2] = points[3 * e + 2];
  }
  cx[e] = (float)c[0]; cy[e] = (float)c[1]; cz[e] = (float)c[2];
}

// out[a] = sum of sorted axis a over [P0, P1] in double (one CTA per axis, fixed order)
__global__ void __launch_bounds__(256) normalize_mean_kernel(const float* __restrict__ sx, const float* __restrict__ sy,
                                                             const float* __restrict__ sz, int64_t P0, int64_t P1, double* out) {
  const float* s = blockIdx.x == 0 ? sx : (blockIdx.x == 1 ? sy : sz);
  __shared__ double sm[256];
  double acc = 0.0;
  for (int64_t i = P0 + threadIdx.x; i <= P1; i += 256) acc += (double)s[i];
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) out[blockIdx.x] = sm[0];
}

// X <- s X + tr   (transform_ * xyz with the matrix [s I | tr])
__global__ void normalize_points_kernel(int64_t n, double s, double tx, double ty, double tz, double* __restrict__ X) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  X[3 * e] = s * X[3 * e] + tx; X[3 * e + 1] = s * X[3 * e + 1] + ty; X[3 * e + 2] = s * X[3 * e + 2] + tz;
}

// [R | t] <- ([R | t] * S^-1) * s,  S^-1 = [I / s | -tr / s]; qvec from the rotation block
__global__ void normalize_poses_kernel(int n, double s, double tx, double ty, double tz, double* __restrict__ qvec, double* __restrict__ tvec) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double R[9], D[9], q[4];
  QuatToRot(qvec + 4 * (size_t)i, R);
  const double inv = 1.0 / s;
  const double it[3] = {-(inv * tx), -(inv * ty), -(inv * tz)};  // translation of the inverse transform
  double* t = tvec + 3 * (size_t)i;
  double nt[3];
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) D[3 * r + c] = (R[3 * r + c] * inv) * s;
    nt[r] = (R[3 * r] * it[0] + R[3 * r + 1] * it[1] + R[3 * r + 2] * it[2] + t[r]) * s;
  }
  RotToQuat(D, q);
  for (int c = 0; c < 4; ++c) qvec[4 * (size_t)i + c] = q[c];
  for (int c = 0; c < 3; ++c) t[c] = nt[c];
}

}  // namespace sba

using namespace sba;

extern "C" int sba_observation_errors(const sba_problem_desc* d, double* sq_error, double* depth) {
  if (!d || (!sq_error && !depth)) { SetError(SBA_ERR_INVALID_ARGUMENT, "null problem description / outputs"); return SBA_ERR_INVALID_ARGUMENT; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    return SBA_ERR_NO_DEVICE;
  }
  const int64_t n = d->num_observations;
  if (n == 0) return SBA_OK;
  for (int c = 0; c < d->num_cameras; ++c)
    if (ModelNumParams(d->camera_model[c]) < 0) { SetError(SBA_ERR_UNKNOWN_CAMERA_MODEL, "camera model does not exist"); return SBA_ERR_UNKNOWN_CAMERA_MODEL; }
  for (int i = 0; i < d->num_images; ++i)
    if (d->image_camera[i] < 0 || d->image_camera[i] >= d->num_cameras) { SetError(SBA_ERR_INVALID_ARGUMENT, "image camera index out of range"); return SBA_ERR_INVALID_ARGUMENT; }
  for (int64_t k = 0; k < n; ++k)
    if (d->obs_image[k] < 0 || d->obs_image[k] >= d->num_images || d->obs_point[k] < 0 || d->obs_point[k] >= d->num_points) {
      SetError(SBA_ERR_INVALID_ARGUMENT, "observation index out of range");
      return SBA_ERR_INVALID_ARGUMENT;
    }
  cudaStream_t st;
  SBA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  DevBuf<double> q, t, cam, pts, xy, err, dep;
  DevBuf<int> ic, model, oi, op;
  SBA_CUDA(q.Alloc(4 * (size_t)d->num_images)); SBA_CUDA(t.Alloc(3 * (size_t)d->num_images));
  SBA_CUDA(cam.Alloc(SBA_MAX_CAMERA_PARAMS * (size_t)d->num_cameras)); SBA_CUDA(pts.Alloc(3 * (size_t)std::max(1, d->num_points)));
  SBA_CUDA(xy.Alloc(2 * (size_t)n)); SBA_CUDA(ic.Alloc(d->num_images)); SBA_CUDA(model.Alloc(d->num_cameras));
  SBA_CUDA(oi.Alloc(n)); SBA_CUDA(op.Alloc(n));
  SBA_CUDA(UploadLarge(q.p, d->image_qvec, 4 * (size_t)d->num_images * sizeof(double), st));
  SBA_CUDA(UploadLarge(t.p, d->image_tvec, 3 * (size_t)d->num_images * sizeof(double), st));
  SBA_CUDA(UploadLarge(cam.p, d->camera_params, SBA_MAX_CAMERA_PARAMS * (size_t)d->num_cameras * sizeof(double), st));
  SBA_CUDA(UploadLarge(pts.p, d->points, 3 * (size_t)d->num_points * sizeof(double), st));
  SBA_CUDA(UploadLarge(xy.p, d->obs_xy, 2 * (size_t)n * sizeof(double), st));
  SBA_CUDA(UploadLarge(ic.p, d->image_camera, (size_t)d->num_images * sizeof(int), st));
  SBA_CUDA(UploadLarge(model.p, d->camera_model, (size_t)d->num_cameras * sizeof(int), st));
  SBA_CUDA(UploadLarge(oi.p, d->obs_image, (size_t)n * sizeof(int), st));
  SBA_CUDA(UploadLarge(op.p, d->obs_point, (size_t)n * sizeof(int), st));
  if (sq_error) SBA_CUDA(err.Alloc(n));
  if (depth) SBA_CUDA(dep.Alloc(n));
  observation_error_kernel<<<DivUp(n, 256), 256, 0, st>>>(n, q.p, t.p, ic.p, model.p, cam.p, pts.p, oi.p, op.p,
                                                         reinterpret_cast<const double2*>(xy.p), sq_error ? err.p : nullptr,
                                                         depth ? dep.p : nullptr);
  SBA_CUDA(cudaGetLastError());
  if (sq_error) SBA_CUDA(cudaMemcpyAsync(sq_error, err.p, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (depth) SBA_CUDA(cudaMemcpyAsync(depth, dep.p, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st));
  SBA_CUDA(cudaStreamSynchronize(st));
  cudaStreamDestroy(st);
  return SBA_OK;
}

extern "C" int sba_normalize(int32_t num_images, double* image_qvec, double* image_tvec, const uint8_t* image_registered,
                             int64_t num_points, double* points, double extent, double p0, double p1, int32_t use_images,
                             double* scale_out, double* translation_out) {
  if (!(extent > 0.0) || !(p0 >= 0.0 && p0 <= 1.0 && p1 >= 0.0 && p1 <= 1.0 && p0 <= p1)) {  // CHECK_GT / CHECK_GE / CHECK_LE, :323, 361-365
    SetError(SBA_ERR_INVALID_ARGUMENT, "sba_normalize: extent must be > 0 and 0 <= p0 <= p1 <= 1");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  if (num_images < 0 || num_points < 0 || (num_images > 0 && (!image_qvec || !image_tvec)) || (num_points > 0 && !points)) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "sba_normalize: null arrays"); return SBA_ERR_INVALID_ARGUMENT;
  }
  if (scale_out) *scale_out = 1.0;
  if (translation_out) translation_out[0] = translation_out[1] = translation_out[2] = 0.0;
  std::vector<int> reg;
  for (int i = 0; i < num_images; ++i) if (!image_registered || image_registered[i]) reg.push_back(i);
  const int64_t n_el = use_images ? (int64_t)reg.size() : num_points;
  if (n_el < 2) return SBA_OK;  // :325-328: nothing to do
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    return SBA_ERR_NO_DEVICE;
  }
  cudaStream_t st;
  SBA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamDestroy(s); } } guard{st};
  DevBuf<double> q, t, pts, sums;
  DevBuf<int> d_reg;
  DevBuf<float> c[3], sorted[3];
  DevBuf<char> tmp;
  if (num_images > 0) {
    SBA_CUDA(q.Alloc(4 * (size_t)num_images)); SBA_CUDA(t.Alloc(3 * (size_t)num_images));
    SBA_CUDA(UploadLarge(q.p, image_qvec, 4 * (size_t)num_images * sizeof(double), st));
    SBA_CUDA(UploadLarge(t.p, image_tvec, 3 * (size_t)num_images * sizeof(double), st));
  }
  if (num_points > 0) {
    SBA_CUDA(pts.Alloc(3 * (size_t)num_points));
    SBA_CUDA(UploadLarge(pts.p, points, 3 * (size_t)num_points * sizeof(double), st));
  }
  if (use_images) SBA_CUDA(d_reg.Upload(reg, st));
  for (int a = 0; a < 3; ++a) { SBA_CUDA(c[a].Alloc(n_el)); SBA_CUDA(sorted[a].Alloc(n_el)); }
  SBA_CUDA(sums.Alloc(3));
  normalize_coords_kernel<<<(unsigned)DivUp(n_el, 256), 256, 0, st>>>(n_el, use_images ? 1 : 0, d_reg.p, q.p, t.p, pts.p, c[0].p, c[1].p, c[2].p);
  SBA_CUDA(cudaGetLastError());
  size_t tb = 0;
  SBA_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tb, c[0].p, sorted[0].p, (int)n_el, 0, 32, st));
  SBA_CUDA(tmp.Alloc(tb));
  for (int a = 0; a < 3; ++a) SBA_CUDA(cub::DeviceRadixSort::SortKeys(tmp.p, tb, c[a].p, sorted[a].p, (int)n_el, 0, 32, st));
  // percentile positions (:405-408)
  const int64_t P0 = (int64_t)((n_el > 3) ? p0 * (double)(n_el - 1) : 0.0);
  const int64_t P1 = (int64_t)((n_el > 3) ? p1 * (double)(n_el - 1) : (double)(n_el - 1));
  normalize_mean_kernel<<<3, 256, 0, st>>>(sorted[0].p, sorted[1].p, sorted[2].p, P0, P1, sums.p);
  SBA_CUDA(cudaGetLastError());
  float lo[3], hi[3];
  double hs[3];
  for (int a = 0; a < 3; ++a) {
    SBA_CUDA(cudaMemcpyAsync(&lo[a], sorted[a].p + P0, sizeof(float), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaMemcpyAsync(&hi[a], sorted[a].p + P1, sizeof(float), cudaMemcpyDeviceToHost, st));
  }
  SBA_CUDA(cudaMemcpyAsync(hs, sums.p, sizeof(hs), cudaMemcpyDeviceToHost, st));
  SBA_CUDA(cudaStreamSynchronize(st));
  double ext2 = 0.0, mean[3];
  for (int a = 0; a < 3; ++a) {
    const double d = (double)hi[a] - (double)lo[a];
    ext2 += d * d;
    mean[a] = hs[a] / (double)(P1 - P0 + 1);
  }
  const double old_extent = std::sqrt(ext2);
  const double scale = old_extent < DBL_EPSILON ? 1.0 : extent / old_extent;  // :334-340
  const double tr[3] = {-scale * mean[0], -scale * mean[1], -scale * mean[2]};
  if (num_points > 0) normalize_points_kernel<<<(unsigned)DivUp(num_points, 256), 256, 0, st>>>(num_points, scale, tr[0], tr[1], tr[2], pts.p);
  if (num_images > 0) normalize_poses_kernel<<<DivUp(num_images, 128), 128, 0, st>>>(num_images, scale, tr[0], tr[1], tr[2], q.p, t.p);
  SBA_CUDA(cudaGetLastError());
  if (num_points > 0) SBA_CUDA(DownloadLarge(points, pts.p, 3 * (size_t)num_points * sizeof(double), st));
  if (num_images > 0) {
    SBA_CUDA(cudaMemcpyAsync(image_qvec, q.p, 4 * (size_t)num_images * sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaMemcpyAsync(image_tvec, t.p, 3 * (size_t)num_images * sizeof(double), cudaMemcpyDeviceToHost, st));
  }
  SBA_CUDA(cudaStreamSynchronize(st));
  if (scale_out) *scale_out = scale;
  if (translation_out) { translation_out[0] = tr[0]; translation_out[1] = tr[1]; translation_out[2] = tr[2]; }
  return SBA_OK;
}
