// Camera-rig reprojection functor:
//   RigBundleAdjustmentCostFunction<CameraModel>::operator()   /root/reference/src/base/cost_functions.h:160-214
//   q = rel_q (x) rig_q (ceres::QuaternionProduct, :180), t = R(rel_q) rig_t + rel_t (:184-187), then the
//   ordinary reprojection residual (:189-206).  The reference differentiates it with
//   AutoDiffCostFunction<..., 2, 4, 3, 4, 3, 3, kNumParams> (:171-176); here the six ambient Jacobian blocks are
//   analytic: the chain rule through the quaternion product and the two rotation polynomials.
// One thread per residual block over SoA arrays (rig poses, relative poses, points, cameras).  This is the functor
// level of the rig path -- what RigBundleAdjuster (src/optim/bundle_adjustment.cc:789-1140) adds to the Ceres
// problem per observation; its solve on the device is listed as next in DESIGN.md section 10.
#include "device_math.cuh"
#include "solver.h"

namespace sba {

// ceres::UnitQuaternionRotatePoint polynomial: rotation matrix part (no normalisation)
__device__ inline void RotPoly(const double q[4], double R[9]) {
  const double q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
  const double t2 = q0 * q1, t3 = q0 * q2, t4 = q0 * q3, t5 = -q1 * q1, t6 = q1 * q2, t7 = q1 * q3, t8 = -q2 * q2,
               t9 = q2 * q3, t1 = -q3 * q3;
  R[0] = 1.0 + 2.0 * (t8 + t1); R[1] = 2.0 * (t6 - t4);       R[2] = 2.0 * (t3 + t7);
  R[3] = 2.0 * (t4 + t6);       R[4] = 1.0 + 2.0 * (t5 + t1); R[5] = 2.0 * (t9 - t2);
  R[6] = 2.0 * (t7 - t3);       R[7] = 2.0 * (t2 + t9);       R[8] = 1.0 + 2.0 * (t5 + t8);
}
// d (RotPoly(q) X) / d q, 3x4 row-major, columns (w, x, y, z)
__device__ inline void RotPolyDq(const double q[4], const double X[3], double D[12]) {
  const double q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3], X0 = X[0], X1 = X[1], X2 = X[2];
  D[0] = 2.0 * (-q3 * X1 + q2 * X2); D[1] = 2.0 * (q2 * X1 + q3 * X2);
  D[2] = 2.0 * (-2.0 * q2 * X0 + q1 * X1 + q0 * X2); D[3] = 2.0 * (-2.0 * q3 * X0 - q0 * X1 + q1 * X2);
  D[4] = 2.0 * (q3 * X0 - q1 * X2); D[5] = 2.0 * (q2 * X0 - 2.0 * q1 * X1 - q0 * X2);
  D[6] = 2.0 * (q1 * X0 + q3 * X2); D[7] = 2.0 * (q0 * X0 - 2.0 * q3 * X1 + q2 * X2);
  D[8] = 2.0 * (-q2 * X0 + q1 * X1); D[9] = 2.0 * (q3 * X0 + q0 * X1 - 2.0 * q1 * X2);
  D[10] = 2.0 * (-q0 * X0 + q3 * X1 - 2.0 * q2 * X2); D[11] = 2.0 * (q1 * X0 + q2 * X1);
}
__device__ inline void Mat3Vec(const double R[9], const double v[3], double o[3]) {
  o[0] = R[0] * v[0] + R[1] * v[1] + R[2] * v[2];
  o[1] = R[3] * v[0] + R[4] * v[1] + R[5] * v[2];
  o[2] = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
}

constexpr int kRigJac = 4 + 3 + 4 + 3 + 3 + SBA_MAX_CAMERA_PARAMS;  // 29 columns per residual row

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
  // q = z (x) w   (ceres::QuaternionProduct)
  const double q[4] = {z[0] * w[0] - z[1] * w[1] - z[2] * w[2] - z[3] * w[3], z[0] * w[1] + z[1] * w[0] + z[2] * w[3] - z[3] * w[2],
                       z[0] * w[2] - z[1] * w[3] + z[2] * w[0] + z[3] * w[1], z[0] * w[3] + z[1] * w[2] - z[2] * w[1] + z[3] * w[0]};
  double Rl[9], Rq[9], t[3], p[3];
  RotPoly(z, Rl);
  Mat3Vec(Rl, tg, t);
  t[0] += tl[0]; t[1] += tl[1]; t[2] += tl[2];
  RotPoly(q, Rq);
  Mat3Vec(Rq, X, p);
  p[0] += t[0]; p[1] += t[1]; p[2] += t[2];
  const double iz = 1.0 / p[2], u = p[0] * iz, v = p[1] * iz;
  double x, y, A[4], Jk[24];
  WorldToImageJacDyn(model[ic], prm, u, v, &x, &y, A, Jk);
  res[2 * k] = x - b_xy[2 * k];
  res[2 * k + 1] = y - b_xy[2 * k + 1];
  if (!jac) return;
  // AB = d r / d p (2x3)
  const double AB[6] = {A[0] * iz, A[1] * iz, -(A[0] * u + A[1] * v) * iz, A[2] * iz, A[3] * iz, -(A[2] * u + A[3] * v) * iz};
  double Dq[12], Dl[12];
  RotPolyDq(q, X, Dq);    // d p / d q
  RotPolyDq(z, tg, Dl);   // d t / d rel_q
  // d q / d w (rig) and d q / d z (rel), 4x4 row-major
  const double Qw[16] = {z[0], -z[1], -z[2], -z[3], z[1], z[0], -z[3], z[2], z[2], z[3], z[0], -z[1], z[3], -z[2], z[1], z[0]};
  const double Qz[16] = {w[0], -w[1], -w[2], -w[3], w[1], w[0], w[3], -w[2], w[2], -w[3], w[0], w[1], w[3], w[2], -w[1], w[0]};
  double* J = jac + (size_t)k * 2 * kRigJac;
#pragma unroll
  for (int row = 0; row < 2; ++row) {
    const double a0 = AB[3 * row], a1 = AB[3 * row + 1], a2 = AB[3 * row + 2];
    double rq[4];  // d r / d q (composed quaternion)
#pragma unroll
    for (int j = 0; j < 4; ++j) rq[j] = a0 * Dq[j] + a1 * Dq[4 + j] + a2 * Dq[8 + j];
    double* o = J + row * kRigJac;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      o[j] = rq[0] * Qw[j] + rq[1] * Qw[4 + j] + rq[2] * Qw[8 + j] + rq[3] * Qw[12 + j];                    // rig_q
      o[7 + j] = rq[0] * Qz[j] + rq[1] * Qz[4 + j] + rq[2] * Qz[8 + j] + rq[3] * Qz[12 + j] +
                 a0 * Dl[j] + a1 * Dl[4 + j] + a2 * Dl[8 + j];                                              // rel_q
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      o[4 + j] = a0 * Rl[j] + a1 * Rl[3 + j] + a2 * Rl[6 + j];                                              // rig_t
      o[11 + j] = AB[3 * row + j];                                                                          // rel_t
      o[14 + j] = a0 * Rq[j] + a1 * Rq[3 + j] + a2 * Rq[6 + j];                                             // point
    }
#pragma unroll
    for (int j = 0; j < SBA_MAX_CAMERA_PARAMS; ++j) o[17 + j] = Jk[12 * row + j];                           // camera params
  }
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
