// ORACLE (test infrastructure only -- never linked into the product library).
//
// Restatement of the Ceres 2.1.0 `ceres/rotation.h` helpers the reference calls
// (Ceres is NOT under /root/reference: un-vendored dependency pinned to 2.1.0 at
// /root/reference/scripts/python/build.py:388-392).  Call sites in the reference:
//   UnitQuaternionRotatePoint  src/base/cost_functions.h:63,122,184,191
//   QuaternionProduct          src/base/cost_functions.h:180
//   QuaternionRotatePoint      src/util/rotation_extension.h:80-86 (via PoseTransformPoint)
//   AngleAxisRotatePoint / AngleAxisToQuaternion   src/util/cylinder.h:403-409,
//                              src/util/cylinder_by_2_points.h:104
// Quaternion order is (w, x, y, z).
#pragma once
#include <cmath>

#include "jet.h"

namespace oracle {

// Published 2.1.0 algorithm: the rotation-matrix polynomial with temporaries
// t1..t9, NO normalisation of q.
template <typename T>
inline void UnitQuaternionRotatePoint(const T q[4], const T pt[3], T out[3]) {
  const T t2 = q[0] * q[1];
  const T t3 = q[0] * q[2];
  const T t4 = q[0] * q[3];
  const T t5 = -q[1] * q[1];
  const T t6 = q[1] * q[2];
  const T t7 = q[1] * q[3];
  const T t8 = -q[2] * q[2];
  const T t9 = q[2] * q[3];
  const T t1 = -q[3] * q[3];
  out[0] = T(2.0) * ((t8 + t1) * pt[0] + (t6 - t4) * pt[1] + (t3 + t7) * pt[2]) + pt[0];
  out[1] = T(2.0) * ((t4 + t6) * pt[0] + (t5 + t1) * pt[1] + (t9 - t2) * pt[2]) + pt[1];
  out[2] = T(2.0) * ((t7 - t3) * pt[0] + (t2 + t9) * pt[1] + (t5 + t8) * pt[2]) + pt[2];
}

template <typename T>
inline void QuaternionRotatePoint(const T q[4], const T pt[3], T out[3]) {
  const T scale = T(1.0) / oracle::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const T unit[4] = {scale * q[0], scale * q[1], scale * q[2], scale * q[3]};
  UnitQuaternionRotatePoint(unit, pt, out);
}

// zw = z * w (Hamilton product)
template <typename T>
inline void QuaternionProduct(const T z[4], const T w[4], T zw[4]) {
  zw[0] = z[0] * w[0] - z[1] * w[1] - z[2] * w[2] - z[3] * w[3];
  zw[1] = z[0] * w[1] + z[1] * w[0] + z[2] * w[3] - z[3] * w[2];
  zw[2] = z[0] * w[2] - z[1] * w[3] + z[2] * w[0] + z[3] * w[1];
  zw[3] = z[0] * w[3] + z[1] * w[2] - z[2] * w[1] + z[3] * w[0];
}

// QuaternionManifold::Plus (Ceres 2.1.0): x_plus = [cos|d|, sin|d|/|d| d] * x
inline void QuaternionPlus(const double x[4], const double d[3], double out[4]) {
  const double n = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  if (n == 0.0) {
    for (int i = 0; i < 4; ++i) out[i] = x[i];
    return;
  }
  const double s = std::sin(n) / n;
  const double qd[4] = {std::cos(n), s * d[0], s * d[1], s * d[2]};
  QuaternionProduct(qd, x, out);
}

// QuaternionManifold::PlusJacobian, 4x3 row-major (rows w,x,y,z)
inline void QuaternionPlusJacobian(const double x[4], double J[12]) {
  J[0] = -x[1]; J[1] = -x[2]; J[2] = -x[3];
  J[3] = x[0];  J[4] = x[3];  J[5] = -x[2];
  J[6] = -x[3]; J[7] = x[0];  J[8] = x[1];
  J[9] = x[2];  J[10] = -x[1]; J[11] = x[0];
}

inline void AngleAxisToQuaternion(const double aa[3], double q[4]) {
  const double th2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
  if (th2 > 0.0) {
    const double th = std::sqrt(th2), half = th * 0.5, k = std::sin(half) / th;
    q[0] = std::cos(half); q[1] = aa[0] * k; q[2] = aa[1] * k; q[3] = aa[2] * k;
  } else {
    q[0] = 1.0; q[1] = aa[0] * 0.5; q[2] = aa[1] * 0.5; q[3] = aa[2] * 0.5;
  }
}

inline void AngleAxisRotatePoint(const double aa[3], const double pt[3], double out[3]) {
  const double th2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
  if (th2 > 2.220446049250313e-16) {
    const double th = std::sqrt(th2), c = std::cos(th), s = std::sin(th), it = 1.0 / th;
    const double w[3] = {aa[0] * it, aa[1] * it, aa[2] * it};
    const double wxp[3] = {w[1] * pt[2] - w[2] * pt[1], w[2] * pt[0] - w[0] * pt[2],
                           w[0] * pt[1] - w[1] * pt[0]};
    const double tmp = (w[0] * pt[0] + w[1] * pt[1] + w[2] * pt[2]) * (1.0 - c);
    for (int i = 0; i < 3; ++i) out[i] = pt[i] * c + wxp[i] * s + w[i] * tmp;
  } else {
    const double wxp[3] = {aa[1] * pt[2] - aa[2] * pt[1], aa[2] * pt[0] - aa[0] * pt[2],
                           aa[0] * pt[1] - aa[1] * pt[0]};
    for (int i = 0; i < 3; ++i) out[i] = pt[i] + wxp[i];
  }
}

}  // namespace oracle
