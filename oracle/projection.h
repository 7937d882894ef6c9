// ORACLE (test infrastructure only -- never linked into the product library).
//
// CPU restatement of the 11 COLMAP camera models' normalised-plane -> pixel map
// (`WorldToImage`) and its inverse, written as one generic function over a
// scalar type T so it can be evaluated with doubles or with oracle::Jet<N>.
//
// Follows /root/reference/src/base/camera_models.h:
//   SIMPLE_PINHOLE :614-626   PINHOLE :663-676        SIMPLE_RADIAL :714-757
//   RADIAL :783-827           OPENCV :853-902         OPENCV_FISHEYE :928-986
//   FULL_OPENCV :1023-1079    FOV :1104-1173          SIMPLE_RADIAL_FISHEYE :1239-1289
//   RADIAL_FISHEYE :1315-1366 THIN_PRISM_FISHEYE :1404-1480
// Parameter-group index tables follow :597-607 (and the per-model equivalents).
// Branches (fisheye r > eps, FOV omega^2 / r^2 < 1e-4) are taken on the VALUE
// part, exactly as Jet comparisons do under Ceres autodiff.
#pragma once
#include <cfloat>
#include <cmath>
#include <vector>

#include "jet.h"

namespace oracle {

enum ModelId {
  kSimplePinhole = 0, kPinhole = 1, kSimpleRadial = 2, kRadial = 3, kOpenCV = 4,
  kOpenCVFisheye = 5, kFullOpenCV = 6, kFOV = 7, kSimpleRadialFisheye = 8,
  kRadialFisheye = 9, kThinPrismFisheye = 10, kNumModels = 11
};

inline int NumParams(int model) {
  static const int n[kNumModels] = {3, 4, 4, 5, 8, 8, 12, 5, 4, 5, 12};
  return (model >= 0 && model < kNumModels) ? n[model] : -1;
}
// number of focal-length params (1 for "simple" f, 2 for fx,fy); principal point follows.
inline int NumFocal(int model) {
  switch (model) {
    case kSimplePinhole: case kSimpleRadial: case kRadial:
    case kSimpleRadialFisheye: case kRadialFisheye: return 1;
    default: return 2;
  }
}
inline std::vector<int> FocalIdxs(int model) {
  return NumFocal(model) == 1 ? std::vector<int>{0} : std::vector<int>{0, 1};
}
inline std::vector<int> PrincipalIdxs(int model) {
  const int nf = NumFocal(model);
  return {nf, nf + 1};
}
inline std::vector<int> ExtraIdxs(int model) {
  std::vector<int> r;
  for (int i = NumFocal(model) + 2; i < NumParams(model); ++i) r.push_back(i);
  return r;
}

// "value part" for branch decisions
inline double Val(double x) { return x; }
template <int N> inline double Val(const Jet<N>& x) { return x.a; }

// Distortion (du, dv) of the radial/tangential families. `e` points at the
// extra params.  FOV is special-cased in Project (it returns u*factor directly).
template <typename T>
inline void Distort(int model, const T* e, const T& u, const T& v, T* du, T* dv) {
  switch (model) {
    case kSimplePinhole:
    case kPinhole:
      *du = T(0.0); *dv = T(0.0);
      return;
    case kSimpleRadial: {  // :746-757
      const T r2 = u * u + v * v;
      const T radial = e[0] * r2;
      *du = u * radial; *dv = v * radial;
      return;
    }
    case kRadial: {  // :815-827
      const T r2 = u * u + v * v;
      const T radial = e[0] * r2 + e[1] * r2 * r2;
      *du = u * radial; *dv = v * radial;
      return;
    }
    case kOpenCV: {  // :887-902
      const T u2 = u * u, uv = u * v, v2 = v * v;
      const T r2 = u2 + v2;
      const T radial = e[0] * r2 + e[1] * r2 * r2;
      *du = u * radial + T(2.0) * e[2] * uv + e[3] * (r2 + T(2.0) * u2);
      *dv = v * radial + T(2.0) * e[3] * uv + e[2] * (r2 + T(2.0) * v2);
      return;
    }
    case kFullOpenCV: {  // :1057-1079
      const T u2 = u * u, uv = u * v, v2 = v * v;
      const T r2 = u2 + v2, r4 = r2 * r2, r6 = r4 * r2;
      const T radial = (T(1.0) + e[0] * r2 + e[1] * r4 + e[4] * r6) /
                       (T(1.0) + e[5] * r2 + e[6] * r4 + e[7] * r6);
      *du = u * radial + T(2.0) * e[2] * uv + e[3] * (r2 + T(2.0) * u2) - u;
      *dv = v * radial + T(2.0) * e[3] * uv + e[2] * (r2 + T(2.0) * v2) - v;
      return;
    }
    case kOpenCVFisheye:        // :962-986
    case kSimpleRadialFisheye:  // :1271-1289
    case kRadialFisheye: {      // :1347-1366
      const T r = oracle::sqrt(u * u + v * v);
      if (Val(r) > DBL_EPSILON) {
        const T th = oracle::atan(r);
        const T th2 = th * th;
        T poly;
        if (model == kSimpleRadialFisheye) {
          poly = T(1.0) + e[0] * th2;
        } else if (model == kRadialFisheye) {
          const T th4 = th2 * th2;
          poly = T(1.0) + e[0] * th2 + e[1] * th4;
        } else {
          const T th4 = th2 * th2, th6 = th4 * th2, th8 = th4 * th4;
          poly = T(1.0) + e[0] * th2 + e[1] * th4 + e[2] * th6 + e[3] * th8;
        }
        const T thd = th * poly;
        *du = u * thd / r - u;
        *dv = v * thd / r - v;
      } else {
        *du = T(0.0); *dv = T(0.0);
      }
      return;
    }
    case kThinPrismFisheye: {  // :1459-1480 (applied to the theta-scaled uu, vv)
      const T u2 = u * u, uv = u * v, v2 = v * v;
      const T r2 = u2 + v2, r4 = r2 * r2, r6 = r4 * r2, r8 = r6 * r2;
      const T radial = e[0] * r2 + e[1] * r4 + e[4] * r6 + e[5] * r8;
      *du = u * radial + T(2.0) * e[2] * uv + e[3] * (r2 + T(2.0) * u2) + e[6] * r2;
      *dv = v * radial + T(2.0) * e[3] * uv + e[2] * (r2 + T(2.0) * v2) + e[7] * r2;
      return;
    }
    default:
      *du = T(0.0); *dv = T(0.0);
  }
}

// FOV distortion factor, :1137-1173
template <typename T>
inline T FovFactor(const T& omega, const T& u, const T& v) {
  const double kEps = 1e-4;
  const T radius2 = u * u + v * v;
  const T omega2 = omega * omega;
  if (Val(omega2) < kEps) {
    return (omega2 * radius2) / T(3.0) - omega2 / T(12.0) + T(1.0);
  } else if (Val(radius2) < kEps) {
    const T t = oracle::tan(omega / T(2.0));
    return (T(-2.0) * t * (T(4.0) * radius2 * t * t - T(3.0))) / (T(3.0) * omega);
  }
  const T radius = oracle::sqrt(radius2);
  const T num = oracle::atan(radius * T(2.0) * oracle::tan(omega / T(2.0)));
  return num / (radius * omega);
}

// WorldToImage for any model.
template <typename T>
inline void Project(int model, const T* p, const T& u, const T& v, T* x, T* y) {
  const int nf = NumFocal(model);
  const T& f1 = p[0];
  const T& f2 = p[nf - 1];
  const T& c1 = p[nf];
  const T& c2 = p[nf + 1];
  const T* e = p + nf + 2;
  if (model == kFOV) {  // :1104-1119, x = f*(u*factor)+c
    const T fac = FovFactor(e[0], u, v);
    *x = f1 * (u * fac) + c1;
    *y = f2 * (v * fac) + c2;
    return;
  }
  T uu = u, vv = v;
  if (model == kThinPrismFisheye) {  // :1412-1422
    const T r = oracle::sqrt(u * u + v * v);
    if (Val(r) > DBL_EPSILON) {
      const T th = oracle::atan(r);
      uu = th * u / r;
      vv = th * v / r;
    }
  }
  T du, dv;
  Distort(model, e, uu, vv, &du, &dv);
  *x = f1 * (uu + du) + c1;
  *y = f2 * (vv + dv) + c2;
}

// ---- inverse (only for pinning Project against the reference's round-trip
// tests, src/base/camera_models_test.cc:39-62, 110-130) -------------------------

// Newton iteration with central differences, camera_models.h:543-590
inline void IterativeUndistort(int model, const double* e, double* u, double* v) {
  const int kIters = 100;
  const double kMaxStepNorm = 1e-10, kRel = 1e-6;
  const double x0 = *u, y0 = *v;
  double x = *u, y = *v;
  for (int i = 0; i < kIters; ++i) {
    const double s0 = std::max(DBL_EPSILON, std::fabs(kRel * x));
    const double s1 = std::max(DBL_EPSILON, std::fabs(kRel * y));
    double dx, dy, a0, a1, b0, b1, c0, c1, d0, d1;
    Distort<double>(model, e, x, y, &dx, &dy);
    Distort<double>(model, e, x - s0, y, &a0, &a1);
    Distort<double>(model, e, x + s0, y, &b0, &b1);
    Distort<double>(model, e, x, y - s1, &c0, &c1);
    Distort<double>(model, e, x, y + s1, &d0, &d1);
    const double j00 = 1 + (b0 - a0) / (2 * s0), j01 = (d0 - c0) / (2 * s1);
    const double j10 = (b1 - a1) / (2 * s0), j11 = 1 + (d1 - c1) / (2 * s1);
    const double det = j00 * j11 - j01 * j10;
    const double rx = x + dx - x0, ry = y + dy - y0;
    const double sx = (j11 * rx - j01 * ry) / det;
    const double sy = (-j10 * rx + j00 * ry) / det;
    x -= sx; y -= sy;
    if (sx * sx + sy * sy < kMaxStepNorm) break;
  }
  *u = x; *v = y;
}

inline void Unproject(int model, const double* p, double x, double y, double* u, double* v) {
  const int nf = NumFocal(model);
  const double f1 = p[0], f2 = p[nf - 1], c1 = p[nf], c2 = p[nf + 1];
  const double* e = p + nf + 2;
  *u = (x - c1) / f1;
  *v = (y - c2) / f2;
  if (model == kSimplePinhole || model == kPinhole) return;
  if (model == kFOV) {  // closed-form Undistortion :1177-1213
    const double uu = *u, vv = *v, om = e[0];
    const double r2 = uu * uu + vv * vv, om2 = om * om;
    double fac;
    if (om2 < 1e-4) {
      fac = (om2 * r2) / 3.0 - om2 / 12.0 + 1.0;
    } else if (r2 < 1e-4) {
      fac = (om * (om * om * r2 + 3.0)) / (6.0 * std::tan(om / 2.0));
    } else {
      const double r = std::sqrt(r2);
      fac = std::tan(r * om) / (r * 2.0 * std::tan(om / 2.0));
    }
    *u = uu * fac; *v = vv * fac;
    return;
  }
  IterativeUndistort(model, e, u, v);
  if (model == kThinPrismFisheye) {  // :1440-1456
    const double th = std::sqrt(*u * *u + *v * *v);
    const double tc = th * std::cos(th);
    if (tc > DBL_EPSILON) {
      const double s = std::sin(th) / tc;
      *u *= s; *v *= s;
    }
  }
}

}  // namespace oracle
