// Camera-rig reprojection functor:
//   RigBundleAdjustmentCostFunction<CameraModel>::operator()   /root/reference/src/base/cost_functions.h:160-214
//   q = rel_q (x) rig_q (ceres::QuaternionProduct, :180), t = R(rel_q) rig_t + rel_t (:184-187), then the
//   ordinary reprojection residual (:189-206).  The reference differentiates it with
//   AutoDiffCostFunction<..., 2, 4, 3, 4, 3, 3, kNumParams> (:171-176); here the six ambient Jacobian blocks are
//   analytic: the chain rule through the quaternion product and the two rotation polynomials.
// One thread per residual block over SoA arrays (rig poses, relative poses, points, cameras).  This is the functor
// level of the rig path -- what RigBundleAdjuster (src/optim/bundle_adjustment.cc:789-1140) adds to the Ceres
// problem per observation; its solve on the device is listed as next in DESIGN.md section 10.
#include "rig_device.cuh"

namespace sba {

__global__ void __launch_bounds__(128) rig_eval_kernel(sba_rig_desc d, const double* __restrict__ rig_q,
                                                       const double* __restrict__ rig_t, const double* __restrict__ rel_q,
                                                       const double* __restrict__ rel_t, const int* __restrict__ model,
                                                       const double* __restrict__ cam, const double* __restrict__ pts,
                                                       const int* __restrict__ b_rig, const int* __restrict__ b_rel,
                                                       const int* __restrict__ b_cam, const int* __restrict__ b_pt,
                                                       const double* __restrict__ b_xy, double* __restrict__ res,
                                                       double* __restrict__ jac) {
  const int64_t k = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (k >= d.num_blocks) return;
  const int ig = b_rig[k], il = b_rel[k], ic = b_cam[k], ip = b_pt[k];
  double w[4], z[4], tg[3], tl[3], X[3], prm[12];
#pragma unroll
  for (int j = 0; j < 4; ++j) { w[j] = rig_q[4 * (size_t)ig + j]; z[j] = rel_q[4 * (size_t)il + j]; }
#pragma unroll
  for (int j = 0; j < 3; ++j) { tg[j] = rig_t[3 * (size_t)ig + j]; tl[j] = rel_t[3 * (size_t)il + j]; X[j] = pts[3 * (size_t)ip + j]; }
#pragma unroll
  for (int j = 0; j < 12; ++j) prm[j] = cam[12 * (size_t)ic + j];
  double r[2];
  const double xy[2] = {b_xy[2 * k], b_xy[2 * k + 1]};
  RigBlock(w, tg, z, tl, X, prm, model[ic], xy, r, jac ? jac + (size_t)k * 2 * kRigJac : nullptr);
  res[2 * k] = r[0];
  res[2 * k + 1] = r[1];
}

}  // namespace sba

using namespace sba;

extern "C" int sba_rig_evaluate(const sba_rig_desc* d, double* residuals, double* jac, double* ms_device) {
  if (!d || !residuals || d->num_blocks < 0) { SetError(SBA_ERR_INVALID_ARGUMENT, "null rig description / output"); return SBA_ERR_INVALID_ARGUMENT; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    return SBA_ERR_NO_DEVICE;
  }
  const int64_t n = d->num_blocks;
  if (n == 0) return SBA_OK;
  for (int c = 0; c < d->num_cameras; ++c)
    if (ModelNumParams(d->camera_model[c]) < 0) { SetError(SBA_ERR_UNKNOWN_CAMERA_MODEL, "camera model does not exist"); return SBA_ERR_UNKNOWN_CAMERA_MODEL; }
  for (int64_t k = 0; k < n; ++k)
    if (d->block_rig[k] < 0 || d->block_rig[k] >= d->num_rig_poses || d->block_rel[k] < 0 || d->block_rel[k] >= d->num_rel_poses ||
        d->block_camera[k] < 0 || d->block_camera[k] >= d->num_cameras || d->block_point[k] < 0 || d->block_point[k] >= d->num_points) {
      SetError(SBA_ERR_INVALID_ARGUMENT, "rig block index out of range");
      return SBA_ERR_INVALID_ARGUMENT;
    }
  cudaStream_t st;
  SBA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  DevBuf<double> gq, gt, lq, lt, cam, pts, xy, res, J;
  DevBuf<int> model, bg, bl, bc, bp;
  auto up_d = [&](DevBuf<double>& b, const double* src, size_t cnt) {
    cudaError_t e = b.Alloc(std::max<size_t>(cnt, 1));
    return (e == cudaSuccess && cnt) ? cudaMemcpyAsync(b.p, src, cnt * sizeof(double), cudaMemcpyHostToDevice, st) : e;
  };
  auto up_i = [&](DevBuf<int>& b, const int32_t* src, size_t cnt) {
    cudaError_t e = b.Alloc(std::max<size_t>(cnt, 1));
    return (e == cudaSuccess && cnt) ? cudaMemcpyAsync(b.p, src, cnt * sizeof(int), cudaMemcpyHostToDevice, st) : e;
  };
  SBA_CUDA(up_d(gq, d->rig_qvec, 4 * (size_t)d->num_rig_poses)); SBA_CUDA(up_d(gt, d->rig_tvec, 3 * (size_t)d->num_rig_poses));
  SBA_CUDA(up_d(lq, d->rel_qvec, 4 * (size_t)d->num_rel_poses)); SBA_CUDA(up_d(lt, d->rel_tvec, 3 * (size_t)d->num_rel_poses));
  SBA_CUDA(up_d(cam, d->camera_params, 12 * (size_t)d->num_cameras)); SBA_CUDA(up_i(model, d->camera_model, d->num_cameras));
  SBA_CUDA(up_d(pts, d->points, 3 * (size_t)d->num_points)); SBA_CUDA(up_d(xy, d->block_xy, 2 * (size_t)n));
  SBA_CUDA(up_i(bg, d->block_rig, n)); SBA_CUDA(up_i(bl, d->block_rel, n));
  SBA_CUDA(up_i(bc, d->block_camera, n)); SBA_CUDA(up_i(bp, d->block_point, n));
  SBA_CUDA(res.Alloc(2 * (size_t)n));
  if (jac) SBA_CUDA(J.Alloc((size_t)n * 2 * kRigJac));
  cudaEvent_t e0, e1;
  SBA_CUDA(cudaEventCreate(&e0)); SBA_CUDA(cudaEventCreate(&e1));
  SBA_CUDA(cudaEventRecord(e0, st));
  rig_eval_kernel<<<DivUp(n, 128), 128, 0, st>>>(*d, gq.p, gt.p, lq.p, lt.p, model.p, cam.p, pts.p, bg.p, bl.p, bc.p, bp.p,
                                                 xy.p, res.p, jac ? J.p : nullptr);
  SBA_CUDA(cudaEventRecord(e1, st));
  SBA_CUDA(cudaGetLastError());
  SBA_CUDA(cudaMemcpyAsync(residuals, res.p, 2 * (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (jac) SBA_CUDA(cudaMemcpyAsync(jac, J.p, (size_t)n * 2 * kRigJac * sizeof(double), cudaMemcpyDeviceToHost, st));
  SBA_CUDA(cudaStreamSynchronize(st));
  if (ms_device) { float ms = 0.f; cudaEventElapsedTime(&ms, e0, e1); *ms_device = ms; }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaStreamDestroy(st);
  return SBA_OK;
}
