// Device-side projection math with analytic Jacobians.
//
// What is computed follows the reference's templated camera models
//   /root/reference/src/base/camera_models.h:614-1480 (WorldToImage / Distortion)
// and the reprojection functor /root/reference/src/base/cost_functions.h:57-81.
// Polynomial models (SIMPLE_PINHOLE .. OPENCV) carry hand-derived closed-form
// derivatives; the transcendental / rational models are differentiated with a
// small forward-mode dual number over (u, v, extra params) so that the
// derivative is that of the branch the VALUE selects, which is what Ceres'
// autodiff yields on the reference side (Jet comparisons act on the value part).
#pragma once
#include <cfloat>

#include "common.cuh"

namespace sba {

__host__ __device__ constexpr inline int ModelNumParams(int m) {
  switch (m) {
    case SBA_SIMPLE_PINHOLE: return 3;
    case SBA_PINHOLE: return 4;
    case SBA_SIMPLE_RADIAL: return 4;
    case SBA_RADIAL: return 5;
    case SBA_OPENCV: return 8;
    case SBA_OPENCV_FISHEYE: return 8;
    case SBA_FULL_OPENCV: return 12;
    case SBA_FOV: return 5;
    case SBA_SIMPLE_RADIAL_FISHEYE: return 4;
    case SBA_RADIAL_FISHEYE: return 5;
    case SBA_THIN_PRISM_FISHEYE: return 12;
    default: return -1;
  }
}
// 1 focal parameter for the "simple"/radial families, else fx, fy
__host__ __device__ inline int ModelNumFocal(int m) {
  return (m == SBA_SIMPLE_PINHOLE || m == SBA_SIMPLE_RADIAL || m == SBA_RADIAL ||
          m == SBA_SIMPLE_RADIAL_FISHEYE || m == SBA_RADIAL_FISHEYE) ? 1 : 2;
}

// ---- forward-mode dual number ------------------------------------------------------
template <int N>
struct DJet {
  double a;
  double v[N];
  __device__ DJet() {}
  __device__ DJet(double s) : a(s) {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = 0.0;
  }
  __device__ static DJet Seed(double s, int k) {
    DJet j(s);
#pragma unroll
    for (int i = 0; i < N; ++i) if (i == k) j.v[i] = 1.0;
    return j;
  }
};
#define SBA_JET_BIN(op, aexpr, vexpr)                                              \
  template <int N>                                                                 \
  __device__ inline DJet<N> operator op(const DJet<N>& f, const DJet<N>& g) {      \
    DJet<N> h;                                                                     \
    h.a = aexpr;                                                                   \
    _Pragma("unroll") for (int i = 0; i < N; ++i) h.v[i] = vexpr;                  \
    return h;                                                                      \
  }
SBA_JET_BIN(+, f.a + g.a, f.v[i] + g.v[i])
SBA_JET_BIN(-, f.a - g.a, f.v[i] - g.v[i])
SBA_JET_BIN(*, f.a * g.a, f.a * g.v[i] + f.v[i] * g.a)
#undef SBA_JET_BIN
template <int N>
__device__ inline DJet<N> operator/(const DJet<N>& f, const DJet<N>& g) {
  DJet<N> h;
  const double inv = 1.0 / g.a;
  h.a = f.a * inv;
#pragma unroll
  for (int i = 0; i < N; ++i) h.v[i] = (f.v[i] - h.a * g.v[i]) * inv;
  return h;
}
template <int N>
__device__ inline DJet<N> jsqrt(const DJet<N>& f) {
  DJet<N> h;
  h.a = sqrt(f.a);
  const double d = 1.0 / (2.0 * h.a);
#pragma unroll
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * d;
  return h;
}
template <int N>
__device__ inline DJet<N> jatan(const DJet<N>& f) {
  DJet<N> h;
  h.a = atan(f.a);
  const double d = 1.0 / (1.0 + f.a * f.a);
#pragma unroll
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * d;
  return h;
}
template <int N>
__device__ inline DJet<N> jtan(const DJet<N>& f) {
  DJet<N> h;
  h.a = tan(f.a);
  const double d = 1.0 + h.a * h.a;
#pragma unroll
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * d;
  return h;
}

// (uu+du, vv+dv) for the non-polynomial models, as duals over (u, v, e[0..NE)).
// Returns the "distorted normalised" coordinates; focal / principal point are
// applied analytically by the caller.
template <int MODEL, int NE>
__device__ inline void DistortedDual(const double* e, double u, double v, DJet<2 + NE>* ou,
                                     DJet<2 + NE>* ov) {
  typedef DJet<2 + NE> J;
  J ju = J::Seed(u, 0), jv = J::Seed(v, 1);
  J je[NE];
#pragma unroll
  for (int i = 0; i < NE; ++i) je[i] = J::Seed(e[i], 2 + i);
  if constexpr (MODEL == SBA_FULL_OPENCV) {  // camera_models.h:1057-1079
    const J u2 = ju * ju, uv = ju * jv, v2 = jv * jv;
    const J r2 = u2 + v2, r4 = r2 * r2, r6 = r4 * r2;
    const J radial = (J(1.0) + je[0] * r2 + je[1] * r4 + je[4] * r6) /
                     (J(1.0) + je[5] * r2 + je[6] * r4 + je[7] * r6);
    *ou = ju * radial + J(2.0) * je[2] * uv + je[3] * (r2 + J(2.0) * u2);
    *ov = jv * radial + J(2.0) * je[3] * uv + je[2] * (r2 + J(2.0) * v2);
  } else if constexpr (MODEL == SBA_OPENCV_FISHEYE || MODEL == SBA_SIMPLE_RADIAL_FISHEYE ||
             MODEL == SBA_RADIAL_FISHEYE) {  // :962-986, :1271-1289, :1347-1366
    const J r = jsqrt(ju * ju + jv * jv);
    if (r.a > DBL_EPSILON) {
      const J th = jatan(r);
      const J th2 = th * th;
      J poly = J(1.0) + je[0] * th2;
      if constexpr (MODEL == SBA_RADIAL_FISHEYE) {
        poly = poly + je[1] * (th2 * th2);
      } else if constexpr (MODEL == SBA_OPENCV_FISHEYE) {
        const J th4 = th2 * th2;
        poly = poly + je[1 % NE] * th4 + je[2 % NE] * (th4 * th2) + je[3 % NE] * (th4 * th4);
      }
      const J thd = th * poly;
      // reference: du = u*thd/r - u, x = f*(u+du)+c  ==> u + du = u*thd/r
      *ou = ju * thd / r;
      *ov = jv * thd / r;
    } else {
      *ou = ju;
      *ov = jv;
    }
  } else if constexpr (MODEL == SBA_FOV) {  // :1137-1173 ; x = f*(u*factor)+c
    const J omega = je[0];
    const J radius2 = ju * ju + jv * jv;
    const J omega2 = omega * omega;
    J factor;
    if (omega2.a < 1e-4) {
      factor = (omega2 * radius2) / J(3.0) - omega2 / J(12.0) + J(1.0);
    } else if (radius2.a < 1e-4) {
      const J t = jtan(omega / J(2.0));
      factor = (J(-2.0) * t * (J(4.0) * radius2 * t * t - J(3.0))) / (J(3.0) * omega);
    } else {
      const J radius = jsqrt(radius2);
      const J num = jatan(radius * J(2.0) * jtan(omega / J(2.0)));
      factor = num / (radius * omega);
    }
    *ou = ju * factor;
    *ov = jv * factor;
  } else if constexpr (MODEL == SBA_THIN_PRISM_FISHEYE) {  // :1404-1480
    J uu = ju, vv = jv;
    const J r = jsqrt(ju * ju + jv * jv);
    if (r.a > DBL_EPSILON) {
      const J th = jatan(r);
      uu = th * ju / r;
      vv = th * jv / r;
    }
    const J u2 = uu * uu, uv = uu * vv, v2 = vv * vv;
    const J r2 = u2 + v2, r4 = r2 * r2, r6 = r4 * r2, r8 = r6 * r2;
    const J radial = je[0] * r2 + je[1] * r4 + je[4 % NE] * r6 + je[5 % NE] * r8;
    *ou = uu + uu * radial + J(2.0) * je[2 % NE] * uv + je[3 % NE] * (r2 + J(2.0) * u2) + je[6 % NE] * r2;
    *ov = vv + vv * radial + J(2.0) * je[3 % NE] * uv + je[2 % NE] * (r2 + J(2.0) * v2) + je[7 % NE] * r2;
  }
}

// ---- WorldToImage with Jacobians ----------------------------------------------------
// x,y pixel; A = d(x,y)/d(u,v) row-major 2x2; Jk[0..P) = d x/d params, Jk[12..12+P) = d y/d params.
template <int MODEL, int NE>
__device__ inline void DualModelJac(const double* p, double u, double v, double* x, double* y,
                                    double* A, double* Jk) {
  DJet<2 + NE> du, dv;
  DistortedDual<MODEL, NE>(p + ModelNumFocal(MODEL) + 2, u, v, &du, &dv);
  const int nf = ModelNumFocal(MODEL);
  const double f1 = p[0], f2 = p[nf - 1];
  *x = f1 * du.a + p[nf];
  *y = f2 * dv.a + p[nf + 1];
  A[0] = f1 * du.v[0]; A[1] = f1 * du.v[1];
  A[2] = f2 * dv.v[0]; A[3] = f2 * dv.v[1];
#pragma unroll
  for (int i = 0; i < 12; ++i) { Jk[i] = 0.0; Jk[12 + i] = 0.0; }
  if (nf == 1) { Jk[0] = du.a; Jk[12] = dv.a; } else { Jk[0] = du.a; Jk[12 + 1] = dv.a; }
  Jk[nf] = 1.0; Jk[12 + nf + 1] = 1.0;
#pragma unroll
  for (int i = 0; i < NE; ++i) { Jk[nf + 2 + i] = f1 * du.v[2 + i]; Jk[12 + nf + 2 + i] = f2 * dv.v[2 + i]; }
}

template <int MODEL>
__device__ inline void WorldToImageJac(const double* p, double u, double v, double* x, double* y,
                                       double* A, double* Jk) {
  if constexpr (MODEL == SBA_SIMPLE_PINHOLE) {  // camera_models.h:614-626
    const double f = p[0];
    *x = f * u + p[1]; *y = f * v + p[2];
    A[0] = f; A[1] = 0.0; A[2] = 0.0; A[3] = f;
#pragma unroll
    for (int i = 0; i < 12; ++i) { Jk[i] = 0.0; Jk[12 + i] = 0.0; }
    Jk[0] = u; Jk[1] = 1.0; Jk[12] = v; Jk[12 + 2] = 1.0;
  } else if constexpr (MODEL == SBA_PINHOLE) {  // :663-676
    *x = p[0] * u + p[2]; *y = p[1] * v + p[3];
    A[0] = p[0]; A[1] = 0.0; A[2] = 0.0; A[3] = p[1];
#pragma unroll
    for (int i = 0; i < 12; ++i) { Jk[i] = 0.0; Jk[12 + i] = 0.0; }
    Jk[0] = u; Jk[2] = 1.0; Jk[12 + 1] = v; Jk[12 + 3] = 1.0;
  } else if constexpr (MODEL == SBA_SIMPLE_RADIAL) {  // :714-757
    const double f = p[0], k = p[3];
    const double u2 = u * u, v2 = v * v, r2 = u2 + v2, radial = k * r2;
    const double xd = u + u * radial, yd = v + v * radial;
    *x = f * xd + p[1]; *y = f * yd + p[2];
    const double kuv2 = 2.0 * k * u * v;
    A[0] = f * (1.0 + radial + 2.0 * k * u2); A[1] = f * kuv2;
    A[2] = f * kuv2;                          A[3] = f * (1.0 + radial + 2.0 * k * v2);
#pragma unroll
    for (int i = 0; i < 12; ++i) { Jk[i] = 0.0; Jk[12 + i] = 0.0; }
    Jk[0] = xd; Jk[1] = 1.0; Jk[3] = f * u * r2;
    Jk[12] = yd; Jk[12 + 2] = 1.0; Jk[12 + 3] = f * v * r2;
  } else if constexpr (MODEL == SBA_RADIAL) {  // :783-827
    const double f = p[0], k1 = p[3], k2 = p[4];
    const double u2 = u * u, v2 = v * v, r2 = u2 + v2, r4 = r2 * r2;
    const double radial = k1 * r2 + k2 * r4;
    const double g = k1 + 2.0 * k2 * r2;  // d radial / d r2
    const double xd = u + u * radial, yd = v + v * radial;
    *x = f * xd + p[1]; *y = f * yd + p[2];
    const double guv2 = 2.0 * g * u * v;
    A[0] = f * (1.0 + radial + 2.0 * g * u2); A[1] = f * guv2;
    A[2] = f * guv2;                          A[3] = f * (1.0 + radial + 2.0 * g * v2);
#pragma unroll
    for (int i = 0; i < 12; ++i) { Jk[i] = 0.0; Jk[12 + i] = 0.0; }
    Jk[0] = xd; Jk[1] = 1.0; Jk[3] = f * u * r2; Jk[4] = f * u * r4;
    Jk[12] = yd; Jk[12 + 2] = 1.0; Jk[12 + 3] = f * v * r2; Jk[12 + 4] = f * v * r4;
  } else if constexpr (MODEL == SBA_OPENCV) {  // :853-902
    const double fx = p[0], fy = p[1], k1 = p[4], k2 = p[5], p1 = p[6], p2 = p[7];
    const double u2 = u * u, uv = u * v, v2 = v * v, r2 = u2 + v2, r4 = r2 * r2;
    const double radial = k1 * r2 + k2 * r4;
    const double g = k1 + 2.0 * k2 * r2;
    const double xd = u + u * radial + 2.0 * p1 * uv + p2 * (r2 + 2.0 * u2);
    const double yd = v + v * radial + 2.0 * p2 * uv + p1 * (r2 + 2.0 * v2);
    *x = fx * xd + p[2]; *y = fy * yd + p[3];
    A[0] = fx * (1.0 + radial + 2.0 * g * u2 + 2.0 * p1 * v + 6.0 * p2 * u);
    A[1] = fx * (2.0 * g * uv + 2.0 * p1 * u + 2.0 * p2 * v);
    A[2] = fy * (2.0 * g * uv + 2.0 * p2 * v + 2.0 * p1 * u);
    A[3] = fy * (1.0 + radial + 2.0 * g * v2 + 2.0 * p2 * u + 6.0 * p1 * v);
#pragma unroll
    for (int i = 0; i < 12; ++i) { Jk[i] = 0.0; Jk[12 + i] = 0.0; }
    Jk[0] = xd; Jk[2] = 1.0;
    Jk[4] = fx * u * r2; Jk[5] = fx * u * r4; Jk[6] = fx * 2.0 * uv; Jk[7] = fx * (r2 + 2.0 * u2);
    Jk[12 + 1] = yd; Jk[12 + 3] = 1.0;
    Jk[12 + 4] = fy * v * r2; Jk[12 + 5] = fy * v * r4; Jk[12 + 6] = fy * (r2 + 2.0 * v2); Jk[12 + 7] = fy * 2.0 * uv;
  } else if constexpr (MODEL == SBA_OPENCV_FISHEYE) {
    DualModelJac<SBA_OPENCV_FISHEYE, 4>(p, u, v, x, y, A, Jk);
  } else if constexpr (MODEL == SBA_FULL_OPENCV) {
    DualModelJac<SBA_FULL_OPENCV, 8>(p, u, v, x, y, A, Jk);
  } else if constexpr (MODEL == SBA_FOV) {
    DualModelJac<SBA_FOV, 1>(p, u, v, x, y, A, Jk);
  } else if constexpr (MODEL == SBA_SIMPLE_RADIAL_FISHEYE) {
    DualModelJac<SBA_SIMPLE_RADIAL_FISHEYE, 1>(p, u, v, x, y, A, Jk);
  } else if constexpr (MODEL == SBA_RADIAL_FISHEYE) {
    DualModelJac<SBA_RADIAL_FISHEYE, 2>(p, u, v, x, y, A, Jk);
  } else if constexpr (MODEL == SBA_THIN_PRISM_FISHEYE) {
    DualModelJac<SBA_THIN_PRISM_FISHEYE, 8>(p, u, v, x, y, A, Jk);
  }
}

// value-only WorldToImage (candidate-step cost evaluation)
template <int MODEL>
__device__ inline void WorldToImageVal(const double* p, double u, double v, double* x, double* y) {
  if constexpr (MODEL == SBA_SIMPLE_PINHOLE) {
    *x = p[0] * u + p[1]; *y = p[0] * v + p[2];
  } else if constexpr (MODEL == SBA_PINHOLE) {
    *x = p[0] * u + p[2]; *y = p[1] * v + p[3];
  } else if constexpr (MODEL == SBA_SIMPLE_RADIAL) {
    const double radial = p[3] * (u * u + v * v);
    *x = p[0] * (u + u * radial) + p[1]; *y = p[0] * (v + v * radial) + p[2];
  } else if constexpr (MODEL == SBA_RADIAL) {
    const double r2 = u * u + v * v, radial = p[3] * r2 + p[4] * r2 * r2;
    *x = p[0] * (u + u * radial) + p[1]; *y = p[0] * (v + v * radial) + p[2];
  } else {
    double A[4], Jk[24];
    WorldToImageJac<MODEL>(p, u, v, x, y, A, Jk);  // dead derivative code is eliminated by the compiler
  }
}

// runtime dispatch (mixed-model problems)
__device__ inline void WorldToImageJacDyn(int model, const double* p, double u, double v, double* x,
                                          double* y, double* A, double* Jk) {
  switch (model) {
    case 0: WorldToImageJac<0>(p, u, v, x, y, A, Jk); break;
    case 1: WorldToImageJac<1>(p, u, v, x, y, A, Jk); break;
    case 2: WorldToImageJac<2>(p, u, v, x, y, A, Jk); break;
    case 3: WorldToImageJac<3>(p, u, v, x, y, A, Jk); break;
    case 4: WorldToImageJac<4>(p, u, v, x, y, A, Jk); break;
    case 5: WorldToImageJac<5>(p, u, v, x, y, A, Jk); break;
    case 6: WorldToImageJac<6>(p, u, v, x, y, A, Jk); break;
    case 7: WorldToImageJac<7>(p, u, v, x, y, A, Jk); break;
    case 8: WorldToImageJac<8>(p, u, v, x, y, A, Jk); break;
    case 9: WorldToImageJac<9>(p, u, v, x, y, A, Jk); break;
    default: WorldToImageJac<10>(p, u, v, x, y, A, Jk); break;
  }
}
__device__ inline void WorldToImageValDyn(int model, const double* p, double u, double v, double* x,
                                          double* y) {
  switch (model) {
    case 0: WorldToImageVal<0>(p, u, v, x, y); break;
    case 1: WorldToImageVal<1>(p, u, v, x, y); break;
    case 2: WorldToImageVal<2>(p, u, v, x, y); break;
    case 3: WorldToImageVal<3>(p, u, v, x, y); break;
    case 4: WorldToImageVal<4>(p, u, v, x, y); break;
    case 5: WorldToImageVal<5>(p, u, v, x, y); break;
    case 6: WorldToImageVal<6>(p, u, v, x, y); break;
    case 7: WorldToImageVal<7>(p, u, v, x, y); break;
    case 8: WorldToImageVal<8>(p, u, v, x, y); break;
    case 9: WorldToImageVal<9>(p, u, v, x, y); break;
    default: WorldToImageVal<10>(p, u, v, x, y); break;
  }
}

// Loss functions of BundleAdjustmentOptions::CreateLossFunction
// (/root/reference/src/optim/bundle_adjustment.cc:53-68): rho, rho', rho'' at s.
__device__ inline void LossEval(int type, double a, double s, double rho[3]) {
  if (type == SBA_LOSS_SOFT_L1) {
    const double b = a * a, c = 1.0 / b;
    const double sum = 1.0 + s * c, tmp = sqrt(sum);
    rho[0] = 2.0 * b * (tmp - 1.0);
    rho[1] = fmax(DBL_MIN, 1.0 / tmp);
    rho[2] = -(c * rho[1]) / (2.0 * sum);
  } else if (type == SBA_LOSS_CAUCHY) {
    const double b = a * a, c = 1.0 / b;
    const double sum = 1.0 + s * c, inv = 1.0 / sum;
    rho[0] = b * log(sum);
    rho[1] = fmax(DBL_MIN, inv);
    rho[2] = -c * (inv * inv);
  } else {
    rho[0] = s; rho[1] = 1.0; rho[2] = 0.0;
  }
}

}  // namespace sba
