// ORACLE (test infrastructure only -- never linked into the product library).
//
// CPU restatement of the fork's two semantic residuals and their Ceres-style
// CENTRAL numeric differentiation:
//   semantic pair residual   /root/reference/src/base/semantic_cost_functions.h:88-208
//   pose algebra             /root/reference/src/util/rotation_extension.h:44-101
//   cylinder -> quadrilateral /root/reference/src/util/cylinder.h:352-474
//   simplePinholeProject     /root/reference/src/util/utils.h:23-55
//   drawQuadrilateral        /root/reference/src/util/cylinder.h:27-121
//   XYWH boxes               /root/reference/src/util/xywh.h:76-131
//   IoU                      /root/reference/src/util/cylinder.h:497-540
//   GSBA functor             /root/reference/src/base/geometric_semantic_cost_functions.h:55-93
//   bilinear (extension X1)  /root/reference/src/util/bitmap.cc:238-269
// NumericDiff CENTRAL follows Ceres 2.1.0 (un-vendored).
// PARITY UNPINNED: the reference has no tests for any semantic code (SURVEY.md
// section 0 fact 7); only the bilinear rule has a known-answer test
// (src/util/bitmap_test.cc:277-288).
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../include/sba_b200.h"
#include "rotation.h"

namespace oracle {

// ceres::QuaternionToRotation (row-major), scaled rotation divided by |q|^2
inline void QuaternionToRotation(const double q[4], double R[9]) {
  const double a = q[0], b = q[1], c = q[2], d = q[3];
  const double aa = a * a, ab = a * b, ac = a * c, ad = a * d, bb = b * b, bc = b * c, bd = b * d,
               cc = c * c, cd = c * d, dd = d * d;
  R[0] = aa + bb - cc - dd; R[1] = 2.0 * (bc - ad);    R[2] = 2.0 * (ac + bd);
  R[3] = 2.0 * (ad + bc);   R[4] = aa - bb + cc - dd;  R[5] = 2.0 * (cd - ab);
  R[6] = 2.0 * (bd - ac);   R[7] = 2.0 * (ab + cd);    R[8] = aa - bb - cc + dd;
  const double normalizer = 1.0 / (aa + bb + cc + dd);
  for (int i = 0; i < 9; ++i) R[i] *= normalizer;
}

// rotation_extension.h:44-78
inline void PoseInverse(const double q[4], const double t[3], double qi[4], double ti[3]) {
  const double scale = 1.0 / std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  qi[0] = scale * q[0]; qi[1] = -(scale * q[1]); qi[2] = -(scale * q[2]); qi[3] = -(scale * q[3]);
  double R[9];
  QuaternionToRotation(qi, R);
  ti[0] = -(R[0] * t[0] + R[1] * t[1] + R[2] * t[2]);
  ti[1] = -(R[3] * t[0] + R[4] * t[1] + R[5] * t[2]);
  ti[2] = -(R[6] * t[0] + R[7] * t[1] + R[8] * t[2]);
}
// rotation_extension.h:80-86
inline void PoseTransformPoint(const double q[4], const double t[3], const double pt[3], double out[3]) {
  QuaternionRotatePoint(q, pt, out);
  out[0] += t[0]; out[1] += t[1]; out[2] += t[2];
}

// ---- semantic pair residual (semantic_cost_functions.h:88-208) -----------------
struct SemanticBlock {
  int x1, y1;
  float depth1, sem1;
  const float* depth2;  // row-major H x W
  const float* sem2;
  int H, W;
  double thr;
};
inline double SemanticPairResidual(const SemanticBlock& b, const double* q1, const double* t1,
                                   const double* q2, const double* t2, const double* cam1,
                                   const double* cam2, int* status) {
  // ImageToWorld of SIMPLE_PINHOLE (camera_models.h:629-637)
  const double u1 = ((double)b.x1 - cam1[1]) / cam1[0];
  const double v1 = ((double)b.y1 - cam1[2]) / cam1[0];
  const double P[3] = {u1 * b.depth1, v1 * b.depth1, (double)b.depth1};
  double qi[4], ti[3], Pw[3], Pc[3];
  PoseInverse(q1, t1, qi, ti);
  PoseTransformPoint(qi, ti, P, Pw);
  PoseTransformPoint(q2, t2, Pw, Pc);
  const double u2 = Pc[0] / Pc[2], v2 = Pc[1] / Pc[2];
  const double x2 = cam2[0] * u2 + cam2[1], y2 = cam2[0] * v2 + cam2[2];
  const int px = (int)std::round(x2), py = (int)std::round(y2);
  if (px < 0 || px >= b.W || py < 0 || py >= b.H) { *status = -1; return 0.0; }
  const double depth2 = (double)b.depth2[(size_t)py * b.W + px];
  if (std::fabs(depth2 - Pc[2]) > b.thr) { *status = -2; return 0.0; }
  *status = 10;
  return (b.sem1 == b.sem2[(size_t)py * b.W + px]) ? 0.0 : 1.0;
}

// Ceres NumericDiff step: max(sqrt(eps), |x| * relative_step_size)
inline double NumDiffDelta(double x, double rel) {
  return std::max(std::sqrt(DBL_EPSILON), std::fabs(x) * rel);
}

inline void SemanticEvaluate(const sba_semantic_desc& d, int threads, double* residuals,
                             int32_t* status, double* jac) {
  const size_t HW = (size_t)d.height * d.width;
#pragma omp parallel for num_threads(threads) schedule(static)
  for (int64_t k = 0; k < d.num_blocks; ++k) {
    const int i1 = d.block_image1[k], i2 = d.block_image2[k];
    SemanticBlock b;
    b.x1 = d.block_x1[k]; b.y1 = d.block_y1[k];
    b.depth1 = d.depth_maps[i1 * HW + (size_t)b.y1 * d.width + b.x1];
    b.sem1 = d.semantic_maps[i1 * HW + (size_t)b.y1 * d.width + b.x1];
    b.depth2 = d.depth_maps + i2 * HW; b.sem2 = d.semantic_maps + i2 * HW;
    b.H = d.height; b.W = d.width; b.thr = d.depth_error_threshold;
    double x[14];
    for (int j = 0; j < 4; ++j) { x[j] = d.image_qvec[4 * i1 + j]; x[7 + j] = d.image_qvec[4 * i2 + j]; }
    for (int j = 0; j < 3; ++j) { x[4 + j] = d.image_tvec[3 * i1 + j]; x[11 + j] = d.image_tvec[3 * i2 + j]; }
    const double* c1 = d.camera_params + 3 * i1;
    const double* c2 = d.camera_params + 3 * i2;
    int st;
    const double r = SemanticPairResidual(b, x, x + 4, x + 7, x + 11, c1, c2, &st);
    if (residuals) residuals[k] = r;
    if (status) status[k] = st;
    if (jac) {
      for (int j = 0; j < 14; ++j) {
        const double xj = x[j], delta = NumDiffDelta(xj, d.relative_step_size);
        int s2;
        x[j] = xj + delta;
        const double fp = SemanticPairResidual(b, x, x + 4, x + 7, x + 11, c1, c2, &s2);
        x[j] = xj - delta;
        const double fm = SemanticPairResidual(b, x, x + 4, x + 7, x + 11, c1, c2, &s2);
        x[j] = xj;
        jac[k * 14 + j] = (fp - fm) * ((1.0 / delta) / 2.0);
      }
    }
  }
}

// ---- boxes (xywh.h) --------------------------------------------------------------
struct Box {
  int x = 0, y = 0, w = 0, h = 0;
  int x_end() const { return x + w - 1; }
  int y_end() const { return y + h - 1; }
  int n() const { return w * h; }
};
inline Box BoundPoints(const double* pts, int npts) {  // xywh.h:85-111
  double mnx = pts[0], mxx = pts[0], mny = pts[1], mxy = pts[1];
  for (int i = 0; i < npts; ++i) {
    mnx = std::min(mnx, pts[2 * i]); mxx = std::max(mxx, pts[2 * i]);
    mny = std::min(mny, pts[2 * i + 1]); mxy = std::max(mxy, pts[2 * i + 1]);
  }
  Box b;
  b.x = (int)std::floor(mnx); b.y = (int)std::floor(mny);
  b.w = (int)std::ceil(mxx) - b.x + 1; b.h = (int)std::ceil(mxy) - b.y + 1;
  return b;
}
inline Box ClipToImage(const Box& b, int W, int H) {  // xywh.h:113-131
  const int tlx = std::max(b.x, 0), tly = std::max(b.y, 0);
  const int brx = std::min(b.x_end(), W - 1), bry = std::min(b.y_end(), H - 1);
  Box o;
  if (brx < tlx || bry < tly) return o;
  o.x = tlx; o.y = tly; o.w = brx - tlx + 1; o.h = bry - tly + 1;
  return o;
}

// drawQuadrilateral (cylinder.h:27-121) into a ROW-major uint8 mask; returns the clipped bbox.
inline Box DrawQuadrilateralRowMajorBox(const double* p, int H, int W, uint8_t* m) {
  std::memset(m, 0, (size_t)H * W);
  const Box bb = ClipToImage(BoundPoints(p, 4), W, H);
  for (int y = bb.y; y < bb.y + bb.h; ++y)
    for (int x = bb.x; x < bb.x + bb.w; ++x) m[(size_t)y * W + x] = 1;
  auto edge = [&](const double* a, const double* b) {
    const double two[4] = {a[0], a[1], b[0], b[1]};
    const Box e = ClipToImage(BoundPoints(two, 2), W, H);
    if (e.n() == 0) return;
    for (int y = e.y; y <= e.y_end(); ++y)
      for (int x = e.x; x <= e.x_end(); ++x) {
        const double cross = ((double)x - a[0]) * (b[1] - a[1]) - ((double)y - a[1]) * (b[0] - a[0]);
        if (cross > 0) m[(size_t)y * W + x] = 0;
      }
  };
  edge(p + 0, p + 2); edge(p + 2, p + 4); edge(p + 4, p + 6); edge(p + 6, p + 0);
  auto corner = [&](const double* v) {
    if (v[0] - bb.x < 1 || bb.x_end() - v[0] < 1 || v[1] - bb.y < 1 || bb.y_end() - v[1] < 1) return;
    const int cx[4] = {bb.x, bb.x_end(), bb.x_end(), bb.x};
    const int cy[4] = {bb.y, bb.y, bb.y_end(), bb.y_end()};
    int best = 0; double bd = 1e300;
    for (int i = 0; i < 4; ++i) {
      const double dx = v[0] - cx[i], dy = v[1] - cy[i];
      const double dist = std::sqrt(dx * dx + dy * dy);
      if (dist < bd) { bd = dist; best = i; }
    }
    const double two[4] = {(double)cx[best], (double)cy[best], v[0], v[1]};
    const Box c = ClipToImage(BoundPoints(two, 2), W, H);
    for (int y = c.y; y < c.y + c.h; ++y)
      for (int x = c.x; x < c.x + c.w; ++x) m[(size_t)y * W + x] = 0;
  };
  corner(p + 0); corner(p + 2); corner(p + 4); corner(p + 6);
  return bb;
}
inline void DrawQuadrilateralRowMajor(const double* p, int H, int W, uint8_t* m) {
  DrawQuadrilateralRowMajorBox(p, H, W, m);
}

// Cylinder::ProjectToQuadrilateral (cylinder.h:352-474); false == the reference throws.
inline bool CylinderQuadrilateral(const double* cam_q, const double* cam_t, const double* cam_params,
                                  const double* cyl_q, const double* cyl_t, double radius,
                                  double height, double* pts8) {
  if (radius <= 0) radius = 1e-4;  // Cylinder::Check, cylinder.h:268-281
  if (height <= 0) height = 1e-4;
  double cwq[4], cwt[3], qi[4], ti[3], c[3];
  PoseInverse(cam_q, cam_t, cwq, cwt);
  PoseInverse(cyl_q, cyl_t, qi, ti);
  PoseTransformPoint(qi, ti, cwt, c);
  c[2] = 0;
  const double dist = std::sqrt(c[0] * c[0] + c[1] * c[1]);
  if (dist <= radius) return false;
  const double dir[3] = {c[0] / dist * radius, c[1] / dist * radius, 0.0};
  const double beta = std::acos(radius / dist);
  const double aap[3] = {0, 0, beta}, aan[3] = {0, 0, -beta};
  double p[4][3];
  AngleAxisRotatePoint(aap, dir, p[0]);
  AngleAxisRotatePoint(aan, dir, p[1]);
  p[2][0] = p[1][0]; p[2][1] = p[1][1]; p[2][2] = p[1][2] + height;
  p[3][0] = p[0][0]; p[3][1] = p[0][1]; p[3][2] = p[0][2] + height;
  for (int i = 0; i < 4; ++i) {
    double pw[3], pc[3];
    PoseTransformPoint(cyl_q, cyl_t, p[i], pw);
    PoseTransformPoint(cam_q, cam_t, pw, pc);
    if (pc[2] <= 0) return false;
    pc[0] /= pc[2]; pc[1] /= pc[2];
    pts8[2 * i] = cam_params[0] * pc[0] + cam_params[1];
    pts8[2 * i + 1] = cam_params[0] * pc[1] + cam_params[2];
  }
  const double v0x = pts8[2] - pts8[0], v0y = pts8[3] - pts8[1];
  const double v1x = pts8[4] - pts8[0], v1y = pts8[5] - pts8[1];
  if (v0x * v1y - v0y * v1x < 0) { std::swap(pts8[2], pts8[6]); std::swap(pts8[3], pts8[7]); }
  return true;
}

// 1 - IoU (cylinder.h:497-540 + geometric_semantic_cost_functions.h:55-93)
inline double CylinderResidual(const double* x16, const double* cam_params, const uint8_t* sem,
                               int sem_pos_count, int H, int W, std::vector<uint8_t>* scratch) {
  double pts[8];
  if (!CylinderQuadrilateral(x16, x16 + 4, cam_params, x16 + 7, x16 + 11, x16[14], x16[15], pts))
    return 1.0 - 0.0;
  scratch->resize((size_t)H * W);
  const Box bb = DrawQuadrilateralRowMajorBox(pts, H, W, scratch->data());
  int inside_pos = 0, tp = 0, fp = 0, sem_and_not_mask = 0;
  for (int y = bb.y; y < bb.y + bb.h; ++y)
    for (int x = bb.x; x < bb.x + bb.w; ++x) {
      const bool s = sem[(size_t)y * W + x] != 0, m = (*scratch)[(size_t)y * W + x] != 0;
      inside_pos += s; tp += (s && m); fp += (!s && m); sem_and_not_mask += (s && !m);
    }
  const int fn = (sem_pos_count - inside_pos) + sem_and_not_mask;
  const double iou = (double)tp / (tp * 1. + fp + fn);
  return 1.0 - iou;
}

inline void CylinderEvaluate(const sba_cylinder_desc& d, int threads, double* residuals, double* jac) {
  const size_t HW = (size_t)d.height * d.width;
  std::vector<int> pos(d.num_images, 0);
  for (int i = 0; i < d.num_images; ++i)
    for (size_t j = 0; j < HW; ++j) pos[i] += d.masks[i * HW + j] != 0;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1)
  for (int64_t k = 0; k < d.num_blocks; ++k) {
    std::vector<uint8_t> scratch;
    const int im = d.block_image[k], cy = d.block_cylinder[k];
    double x[16];
    for (int j = 0; j < 4; ++j) { x[j] = d.image_qvec[4 * im + j]; x[7 + j] = d.cyl_qvec[4 * cy + j]; }
    for (int j = 0; j < 3; ++j) { x[4 + j] = d.image_tvec[3 * im + j]; x[11 + j] = d.cyl_tvec[3 * cy + j]; }
    x[14] = d.cyl_radius[cy]; x[15] = d.cyl_height[cy];
    const double* cp = d.camera_params + 3 * im;
    const uint8_t* sem = d.masks + im * HW;
    if (residuals) residuals[k] = CylinderResidual(x, cp, sem, pos[im], d.height, d.width, &scratch);
    if (jac) {
      for (int j = 0; j < 16; ++j) {
        const double xj = x[j], delta = NumDiffDelta(xj, d.relative_step_size);
        x[j] = xj + delta;
        const double fp = CylinderResidual(x, cp, sem, pos[im], d.height, d.width, &scratch);
        x[j] = xj - delta;
        const double fm = CylinderResidual(x, cp, sem, pos[im], d.height, d.width, &scratch);
        x[j] = xj;
        jac[k * 16 + j] = (fp - fm) * ((1.0 / delta) / 2.0);
      }
    }
  }
}

// Bitmap::InterpolateBilinear (bitmap.cc:238-269) on a top-down ROW-major float
// map; FreeImage's bottom-up scanlines are reproduced through inv_y.
inline double BilinearSample(const float* map, int H, int W, double x, double y, int* ok) {
  const double inv_y = H - 1 - y;
  const int x0 = (int)std::floor(x), x1 = x0 + 1;
  const int y0 = (int)std::floor(inv_y), y1 = y0 + 1;
  if (x0 < 0 || x1 >= W || y0 < 0 || y1 >= H) { if (ok) *ok = 0; return 0.0; }
  const double dx = x - x0, dy = inv_y - y0, dx_1 = 1 - dx, dy_1 = 1 - dy;
  const float* line0 = map + (size_t)(H - 1 - y0) * W;
  const float* line1 = map + (size_t)(H - 1 - y1) * W;
  const double v0 = dx_1 * line0[x0] + dx * line0[x1];
  const double v1 = dx_1 * line1[x0] + dx * line1[x1];
  if (ok) *ok = 1;
  return dy_1 * v0 + dy * v1;
}

}  // namespace oracle
