// Device functions of the fork's two semantic residuals, shared by the evaluators (semantic.cu) and the two solves
// (semantic_solve.cu, cylinder_solve.cu).  Reference locations: see semantic.cu.  Every file that includes this is
// compiled with -fmad=false (csrc/Makefile): the residuals are integer-valued functions of rounded pixel coordinates, and
// IEEE mul / add / div / sqrt without contraction reproduce the CPU evaluation order bit for bit.
#pragma once
#include <cfloat>
#include <climits>

#include "common.cuh"

namespace sba {

// ---- small pose algebra (ceres/rotation.h semantics) ----------------------------------------
__device__ inline void UnitQuatRotate(const double q[4], const double pt[3], double out[3]) {
  const double t2 = q[0] * q[1], t3 = q[0] * q[2], t4 = q[0] * q[3], t5 = -q[1] * q[1], t6 = q[1] * q[2],
               t7 = q[1] * q[3], t8 = -q[2] * q[2], t9 = q[2] * q[3], t1 = -q[3] * q[3];
  out[0] = 2.0 * ((t8 + t1) * pt[0] + (t6 - t4) * pt[1] + (t3 + t7) * pt[2]) + pt[0];
  out[1] = 2.0 * ((t4 + t6) * pt[0] + (t5 + t1) * pt[1] + (t9 - t2) * pt[2]) + pt[1];
  out[2] = 2.0 * ((t7 - t3) * pt[0] + (t2 + t9) * pt[1] + (t5 + t8) * pt[2]) + pt[2];
}
__device__ inline void QuatRotate(const double q[4], const double pt[3], double out[3]) {
  const double scale = 1.0 / sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const double unit[4] = {scale * q[0], scale * q[1], scale * q[2], scale * q[3]};
  UnitQuatRotate(unit, pt, out);
}
__device__ inline void QuatToRotation(const double q[4], double R[9]) {
  const double a = q[0], b = q[1], c = q[2], d = q[3];
  const double aa = a * a, ab = a * b, ac = a * c, ad = a * d, bb = b * b, bc = b * c, bd = b * d, cc = c * c,
               cd = c * d, dd = d * d;
  R[0] = aa + bb - cc - dd; R[1] = 2.0 * (bc - ad);   R[2] = 2.0 * (ac + bd);
  R[3] = 2.0 * (ad + bc);   R[4] = aa - bb + cc - dd; R[5] = 2.0 * (cd - ab);
  R[6] = 2.0 * (bd - ac);   R[7] = 2.0 * (ab + cd);   R[8] = aa - bb - cc + dd;
  const double normalizer = 1.0 / (aa + bb + cc + dd);
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] *= normalizer;
}
__device__ inline void PoseInverse(const double q[4], const double t[3], double qi[4], double ti[3]) {
  const double scale = 1.0 / sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  qi[0] = scale * q[0]; qi[1] = -(scale * q[1]); qi[2] = -(scale * q[2]); qi[3] = -(scale * q[3]);
  double R[9];
  QuatToRotation(qi, R);
  ti[0] = -(R[0] * t[0] + R[1] * t[1] + R[2] * t[2]);
  ti[1] = -(R[3] * t[0] + R[4] * t[1] + R[5] * t[2]);
  ti[2] = -(R[6] * t[0] + R[7] * t[1] + R[8] * t[2]);
}
__device__ inline void PoseTransform(const double q[4], const double t[3], const double pt[3], double out[3]) {
  QuatRotate(q, pt, out);
  out[0] += t[0]; out[1] += t[1]; out[2] += t[2];
}
// double -> int the way the host (x86 cvttsd2si) does it: out-of-range / NaN give INT_MIN
__device__ inline int CvtInt(double v) {
  if (!(v > -2147483649.0 && v < 2147483648.0)) return INT_MIN;
  return (int)v;
}
__device__ inline double NumDiffDelta(double x, double rel) { return fmax(sqrt(DBL_EPSILON), fabs(x) * rel); }

// =================================== K-SEM-PAIR ===============================================
struct SemView {
  int H, W;
  const double* q;      // [n_img*4]
  const double* t;      // [n_img*3]
  const double* cam;    // [n_img*3]
  const float* depth;   // [n_img*H*W] row-major
  const float* label;   // [n_img*H*W]
  const int* i1; const int* i2; const int* x1; const int* y1;
  double thr, rel;
};

__device__ inline double PairResidual(const SemView& V, int x1, int y1, float depth1, float sem1, const float* depth2,
                                      const float* sem2, const double* x /*q1 t1 q2 t2*/, const double* cam1,
                                      const double* cam2, int* status) {
  const double u1 = ((double)x1 - cam1[1]) / cam1[0];
  const double v1 = ((double)y1 - cam1[2]) / cam1[0];
  const double P[3] = {u1 * depth1, v1 * depth1, (double)depth1};
  double qi[4], ti[3], Pw[3], Pc[3];
  PoseInverse(x, x + 4, qi, ti);
  PoseTransform(qi, ti, P, Pw);
  PoseTransform(x + 7, x + 11, Pw, Pc);
  const double u2 = Pc[0] / Pc[2], v2 = Pc[1] / Pc[2];
  const double x2 = cam2[0] * u2 + cam2[1], y2 = cam2[0] * v2 + cam2[2];
  const int px = CvtInt(round(x2)), py = CvtInt(round(y2));
  if (px < 0 || px >= V.W || py < 0 || py >= V.H) { *status = -1; return 0.0; }
  const size_t o = (size_t)py * V.W + px;
  const double d2 = (double)__ldg(depth2 + o);
  if (fabs(d2 - Pc[2]) > V.thr) { *status = -2; return 0.0; }
  *status = 10;
  return (sem1 == __ldg(sem2 + o)) ? 0.0 : 1.0;
}

// One warp per residual block: lane 0 evaluates the block, lanes 1..28 the +/- perturbations
// of the 14 ambient coordinates; the central differences are formed with shuffles.

struct IBox { int x, y, w, h; };
__device__ inline IBox BoundPts(const double* pts, int npts) {
  double mnx = pts[0], mxx = pts[0], mny = pts[1], mxy = pts[1];
  for (int i = 0; i < npts; ++i) {
    mnx = fmin(mnx, pts[2 * i]); mxx = fmax(mxx, pts[2 * i]);
    mny = fmin(mny, pts[2 * i + 1]); mxy = fmax(mxy, pts[2 * i + 1]);
  }
  IBox b;
  b.x = CvtInt(floor(mnx)); b.y = CvtInt(floor(mny));
  b.w = CvtInt(ceil(mxx)) - b.x + 1; b.h = CvtInt(ceil(mxy)) - b.y + 1;
  return b;
}
__device__ inline IBox ClipBox(const IBox& b, int W, int H) {
  const int tlx = max(b.x, 0), tly = max(b.y, 0);
  const int brx = min(b.x + b.w - 1, W - 1), bry = min(b.y + b.h - 1, H - 1);
  IBox o = {0, 0, 0, 0};
  if (brx < tlx || bry < tly) return o;
  o.x = tlx; o.y = tly; o.w = brx - tlx + 1; o.h = bry - tly + 1;
  return o;
}
__device__ inline bool InBox(const IBox& b, int x, int y) { return x >= b.x && x < b.x + b.w && y >= b.y && y < b.y + b.h; }

__device__ inline void AngleAxisRotateZ(double beta, const double pt[3], double out[3]) {
  // ceres::AngleAxisRotatePoint with angle_axis = (0, 0, beta)
  const double aa[3] = {0.0, 0.0, beta};
  const double th2 = beta * beta;
  if (th2 > 2.220446049250313e-16) {
    const double th = sqrt(th2), c = cos(th), s = sin(th), it = 1.0 / th;
    const double w[3] = {aa[0] * it, aa[1] * it, aa[2] * it};
    const double wxp[3] = {w[1] * pt[2] - w[2] * pt[1], w[2] * pt[0] - w[0] * pt[2], w[0] * pt[1] - w[1] * pt[0]};
    const double tmp = (w[0] * pt[0] + w[1] * pt[1] + w[2] * pt[2]) * (1.0 - c);
    for (int i = 0; i < 3; ++i) out[i] = pt[i] * c + wxp[i] * s + w[i] * tmp;
  } else {
    const double wxp[3] = {aa[1] * pt[2] - aa[2] * pt[1], aa[2] * pt[0] - aa[0] * pt[2], aa[0] * pt[1] - aa[1] * pt[0]};
    for (int i = 0; i < 3; ++i) out[i] = pt[i] + wxp[i];
  }
}

// Cylinder::ProjectToQuadrilateral (cylinder.h:352-474); false where the reference throws
__device__ inline bool CylinderQuad(const double* x16, const double* cam, double* pts8) {
  const double* cam_q = x16; const double* cam_t = x16 + 4; const double* cyl_q = x16 + 7; const double* cyl_t = x16 + 11;
  double radius = x16[14], height = x16[15];
  if (radius <= 0) radius = 1e-4;
  if (height <= 0) height = 1e-4;
  double cwq[4], cwt[3], qi[4], ti[3], c[3];
  PoseInverse(cam_q, cam_t, cwq, cwt);
  PoseInverse(cyl_q, cyl_t, qi, ti);
  PoseTransform(qi, ti, cwt, c);
  c[2] = 0;
  const double dist = sqrt(c[0] * c[0] + c[1] * c[1]);
  if (!(dist > radius)) return false;
  const double dir[3] = {c[0] / dist * radius, c[1] / dist * radius, 0.0};
  const double beta = acos(radius / dist);
  double p[4][3];
  AngleAxisRotateZ(beta, dir, p[0]);
  AngleAxisRotateZ(-beta, dir, p[1]);
  p[2][0] = p[1][0]; p[2][1] = p[1][1]; p[2][2] = p[1][2] + height;
  p[3][0] = p[0][0]; p[3][1] = p[0][1]; p[3][2] = p[0][2] + height;
  for (int i = 0; i < 4; ++i) {
    double pw[3], pc[3];
    PoseTransform(cyl_q, cyl_t, p[i], pw);
    PoseTransform(cam_q, cam_t, pw, pc);
    if (!(pc[2] > 0)) return false;
    pc[0] /= pc[2]; pc[1] /= pc[2];
    pts8[2 * i] = cam[0] * pc[0] + cam[1];
    pts8[2 * i + 1] = cam[0] * pc[1] + cam[2];
  }
  const double v0x = pts8[2] - pts8[0], v0y = pts8[3] - pts8[1];
  const double v1x = pts8[4] - pts8[0], v1y = pts8[5] - pts8[1];
  if (v0x * v1y - v0y * v1x < 0) {
    double t = pts8[2]; pts8[2] = pts8[6]; pts8[6] = t;
    t = pts8[3]; pts8[3] = pts8[7]; pts8[7] = t;
  }
  return true;
}

// drawQuadrilateral as a per-pixel predicate: start from "inside the clipped bbox", clear by
// every edge inside that edge's own clipped bbox, then clear the corner rectangles.
struct QuadRaster {
  IBox bb;
  IBox eb[4];
  IBox cb[4];
  bool corner_on[4];
  double p[8];
};
__device__ inline void BuildRaster(const double* p, int H, int W, QuadRaster* R) {
  for (int i = 0; i < 8; ++i) R->p[i] = p[i];
  R->bb = ClipBox(BoundPts(p, 4), W, H);
  for (int e = 0; e < 4; ++e) {
    const double two[4] = {p[2 * e], p[2 * e + 1], p[2 * ((e + 1) & 3)], p[2 * ((e + 1) & 3) + 1]};
    R->eb[e] = ClipBox(BoundPts(two, 2), W, H);
  }
  const IBox& bb = R->bb;
  const int xe = bb.x + bb.w - 1, ye = bb.y + bb.h - 1;
  for (int v = 0; v < 4; ++v) {
    const double vx = p[2 * v], vy = p[2 * v + 1];
    R->corner_on[v] = !(vx - bb.x < 1 || xe - vx < 1 || vy - bb.y < 1 || ye - vy < 1);
    R->cb[v] = IBox{0, 0, 0, 0};
    if (!R->corner_on[v]) continue;
    const int cx[4] = {bb.x, xe, xe, bb.x};
    const int cy[4] = {bb.y, bb.y, ye, ye};
    int best = 0; double bd = 1e300;
    for (int i = 0; i < 4; ++i) {
      const double dx = vx - cx[i], dy = vy - cy[i];
      const double dist = sqrt(dx * dx + dy * dy);
      if (dist < bd) { bd = dist; best = i; }
    }
    const double two[4] = {(double)cx[best], (double)cy[best], vx, vy};
    R->cb[v] = ClipBox(BoundPts(two, 2), W, H);
  }
}
__device__ inline bool RasterInside(const QuadRaster& R, int x, int y) {
  bool m = true;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    if (R.eb[e].w * R.eb[e].h == 0 || !InBox(R.eb[e], x, y)) continue;
    const double ax = R.p[2 * e], ay = R.p[2 * e + 1], bx = R.p[2 * ((e + 1) & 3)], by = R.p[2 * ((e + 1) & 3) + 1];
    const double cross = ((double)x - ax) * (by - ay) - ((double)y - ay) * (bx - ax);
    if (cross > 0) m = false;
  }
#pragma unroll
  for (int v = 0; v < 4; ++v)
    if (R.corner_on[v] && InBox(R.cb[v], x, y)) m = false;
  return m;
}

}  // namespace sba
