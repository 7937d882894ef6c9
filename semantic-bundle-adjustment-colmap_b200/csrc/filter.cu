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
