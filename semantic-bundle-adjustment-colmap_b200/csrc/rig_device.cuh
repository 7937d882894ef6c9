// Device side of the camera-rig reprojection functor, shared by the evaluation surface (rig.cu) and the rig solve
// (rig_solve.cu):  RigBundleAdjustmentCostFunction<CameraModel>::operator()   /root/reference/src/base/cost_functions.h:160-214
#pragma once
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

// One residual block: w = rig q, tg = rig t, z = relative q, tl = relative t.  r[2]; J (nullable) = 2 x kRigJac ambient
// blocks d r / d (rig_q(4), rig_t(3), rel_q(4), rel_t(3), point(3), camera params(12)), row-major.
__device__ inline void RigBlock(const double w[4], const double tg[3], const double z[4], const double tl[3], const double X[3],
                                const double prm[12], int model, const double xy[2], double r[2], double* J) {
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
  WorldToImageJacDyn(model, prm, u, v, &x, &y, A, Jk);
  r[0] = x - xy[0];
  r[1] = y - xy[1];
  if (!J) return;
  // AB = d r / d p (2x3)
  const double AB[6] = {A[0] * iz, A[1] * iz, -(A[0] * u + A[1] * v) * iz, A[2] * iz, A[3] * iz, -(A[2] * u + A[3] * v) * iz};
  double Dq[12], Dl[12];
  RotPolyDq(q, X, Dq);    // d p / d q
  RotPolyDq(z, tg, Dl);   // d t / d rel_q
  // d q / d w (rig) and d q / d z (rel), 4x4 row-major
  const double Qw[16] = {z[0], -z[1], -z[2], -z[3], z[1], z[0], -z[3], z[2], z[2], z[3], z[0], -z[1], z[3], -z[2], z[1], z[0]};
  const double Qz[16] = {w[0], -w[1], -w[2], -w[3], w[1], w[0], w[3], -w[2], w[2], -w[3], w[0], w[1], w[3], w[2], -w[1], w[0]};
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
