// GeometricSemanticBundleAdjuster as a solve on the device (SURVEY.md section 8 rows a12 + a13).
//
// What the reference does (/root/reference/src/optim/geometric_semantic_bundle_adjustment.cc):
//   SetUpGeometryError :714-726   ScaledLoss(nullptr, 1 / N_images) on every IoU block; for image in Images: AddImageToProblem
//   AddImageToProblem  :835-996   skip a constant camera when the geometry is constant too; per cylinder ONE block:
//                                   constant camera  -> ConstantPoseGSBACostFunction[By2Points]   (cylinder blocks only)
//                                   refine_geometry  -> GSBACostFunction[By2Points]                (pose + cylinder blocks)
//                                   else             -> ConstantCylinderGSBACostFunction           (pose blocks only)
//                                 all five are 1 - Cylinder::ComputeSemanticIoU (geometric_semantic_cost_functions.h:33-348,
//                                 src/util/cylinder.h:497-540) under NumericDiffCostFunction<CENTRAL>; a CylinderBy2Points is
//                                 converted with ToCylinder first (src/util/cylinder_by_2_points.h:84-107)
//   SetUpCylinderManifolds :1126-1213  QuaternionManifold on the cylinder quaternion; lower bound 0 on the RADIUS (the call
//                                 meant for the height names the radius block again, :1205-1211: the height stays unbounded)
//   SetUpManifolds                pose manifolds as in BundleAdjuster
// Here: one CTA per (block, evaluation) rasterises the quadrilateral of the perturbed parameters against the image's mask
// (the evaluator's own device functions, semantic_device.cuh); a second kernel turns the central differences of a block
// into its tangent row and its record for the dense LM (dense_lm.cu).
#include <cfloat>
#include <climits>

#include "dense_lm.h"
#include "rig_lin.cuh"
#include "semantic_device.cuh"

namespace sba {
namespace {

constexpr int kCylAmb = 16;  // ambient coordinates of a block: camera q(4) t(3) + cylinder (9: q t r h | 7: t1 t2 r)

struct CylSolveView {
  int n_img, n_cyl, H, W, kind, refine_geometry;
  const double* q; const double* t;    // poses (current or candidate)
  const double* cyl;                   // [n_cyl * 9]
  const double* cam; const uint8_t* masks; const int* pos_count;
  const int* g_img; const int* g_cyl;  // groups = blocks
  const uint8_t* pose_const; const uint8_t* tvec_mask;
  double rel, loss_a;                  // ScaledLoss(nullptr, a)
};

// ceres::AngleAxisToQuaternion
__device__ inline void AngleAxisToQuat(const double a[3], double q[4]) {
  const double th2 = a[0] * a[0] + a[1] * a[1] + a[2] * a[2];
  if (th2 > 0.0) {
    const double th = sqrt(th2), half = th * 0.5, k = sin(half) / th;
    q[0] = cos(half); q[1] = a[0] * k; q[2] = a[1] * k; q[3] = a[2] * k;
  } else {
    q[0] = 1.0; q[1] = a[0] * 0.5; q[2] = a[1] * 0.5; q[3] = a[2] * 0.5;
  }
}
// CylinderBy2Points(t1, t2, r).ToCylinder() (cylinder_by_2_points.h:35-42, 84-107) -> q(4) t(3) r h
__device__ inline void TwoPointsToCylinder(const double* p7, double* c9) {
  double radius = p7[6];
  if (radius <= 0) radius = 1e-4;                      // Check()
  double d[3] = {p7[3] - p7[0], p7[4] - p7[1], p7[5] - p7[2]};
  const double dn = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  d[0] /= dn; d[1] /= dn; d[2] /= dn;
  double axis[3] = {-d[1], d[0], 0.0};                 // z x d
  const double an = sqrt(axis[0] * axis[0] + axis[1] * axis[1] + axis[2] * axis[2]);
  if (fabs(an) < 1e-10) { axis[0] = 1.0; axis[1] = 0.0; axis[2] = 0.0; }
  else { axis[0] /= an; axis[1] /= an; axis[2] /= an; }
  const double angle = acos(d[2]);
  const double aa[3] = {angle * axis[0], angle * axis[1], angle * axis[2]};
  AngleAxisToQuat(aa, c9);
  c9[4] = p7[0]; c9[5] = p7[1]; c9[6] = p7[2];
  c9[7] = radius;
  const double h[3] = {p7[0] - p7[3], p7[1] - p7[4], p7[2] - p7[5]};
  c9[8] = sqrt(h[0] * h[0] + h[1] * h[1] + h[2] * h[2]);  // Height()
}

__device__ inline int CylCoords(int kind) { return kind == 0 ? 16 : 14; }

// One CTA per (block, evaluation): evaluation 0 is the block itself, 1 + 2j / 2 + 2j the +/- perturbations of ambient
// coordinate j.  A coordinate of a constant block is not a parameter of the functor: its evaluations are skipped (their
// column is zero).  vals[k * 33 + e] = 1 - IoU, deltas[k * 16 + j] = the step of coordinate j.
__global__ void __launch_bounds__(256) cyl_eval_kernel(CylSolveView V, int n_eval, double* __restrict__ vals,
                                                       double* __restrict__ deltas) {
  const int k = blockIdx.x / n_eval, e = blockIdx.x - k * n_eval;
  const int im = V.g_img[k], cy = V.g_cyl[k];
  __shared__ QuadRaster R;
  __shared__ int ok;
  __shared__ int red[8][4];
  const int nc = CylCoords(V.kind);
  if (e > 0) {
    const int j = (e - 1) >> 1;
    const bool is_param = j < nc && (j < 7 ? !V.pose_const[im] : V.refine_geometry != 0);
    if (!is_param) { if (threadIdx.x == 0) vals[(size_t)k * n_eval + e] = 0.0; return; }
  }
  if (threadIdx.x == 0) {
    double x[kCylAmb];
    for (int j = 0; j < 4; ++j) x[j] = V.q[4 * im + j];
    for (int j = 0; j < 3; ++j) x[4 + j] = V.t[3 * im + j];
    for (int j = 0; j < 9; ++j) x[7 + j] = V.cyl[9 * cy + j];
    if (e > 0) {
      const int j = (e - 1) >> 1;
      const double delta = NumDiffDelta(x[j], V.rel);
      x[j] = ((e - 1) & 1) ? x[j] - delta : x[j] + delta;
      if (!((e - 1) & 1)) deltas[(size_t)k * kCylAmb + j] = delta;
    }
    if (V.kind == 1) { double c9[9]; TwoPointsToCylinder(x + 7, c9); for (int j = 0; j < 9; ++j) x[7 + j] = c9[j]; }
    double pts[8];
    ok = CylinderQuad(x, V.cam + 3 * im, pts) ? 1 : 0;
    if (ok) BuildRaster(pts, V.H, V.W, &R);
  }
  __syncthreads();
  if (!ok) {  // the reference catches the exception and returns IoU 0 (cylinder.h:506-511)
    if (threadIdx.x == 0) vals[(size_t)k * n_eval + e] = 1.0;
    return;
  }
  const uint8_t* sem = V.masks + (size_t)im * V.H * V.W;
  int inside_pos = 0, tp = 0, fp = 0, snm = 0;
  const int n_px = R.bb.w * R.bb.h;
  for (int idx = threadIdx.x; idx < n_px; idx += 256) {
    const int yy = idx / R.bb.w, xx = idx - yy * R.bb.w;
    const int x = R.bb.x + xx, y = R.bb.y + yy;
    const bool s = sem[(size_t)y * V.W + x] != 0;
    const bool m = RasterInside(R, x, y);
    inside_pos += s; tp += (s && m); fp += (!s && m); snm += (s && !m);
  }
  int v[4] = {inside_pos, tp, fp, snm};
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    for (int o = 16; o > 0; o >>= 1) v[a] += __shfl_down_sync(0xffffffffu, v[a], o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][a] = v[a];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int t[4] = {0, 0, 0, 0};
    for (int w = 0; w < 8; ++w) for (int a = 0; a < 4; ++a) t[a] += red[w][a];
    const int fn = (V.pos_count[im] - t[0]) + t[3];
    const double iou = (double)t[1] / (t[1] * 1. + t[2] + fn);
    vals[(size_t)k * n_eval + e] = 1.0 - iou;
  }
}

__global__ void __launch_bounds__(256) cyl_mask_count_kernel(const uint8_t* __restrict__ masks, size_t HW, int* __restrict__ count) {
  const uint8_t* m = masks + (size_t)blockIdx.x * HW;
  int c = 0;
  for (size_t j = threadIdx.x; j < HW; j += 256) c += m[j] != 0;
  __shared__ int sm[8];
  for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) { int t = 0; for (int i = 0; i < 8; ++i) t += sm[i]; count[blockIdx.x] = t; }
}

// local tangent layout of a block: [pose: q (3) | t free] [cylinder: kind 0: q (3) t (3) r h | kind 1: t1 (3) t2 (3) r]
// One thread per block: ambient central differences -> tangent row -> record (J^T J lower, J^T r, cost), or the cost only
__global__ void cyl_record_kernel(CylSolveView V, int n_groups, int n_eval, const double* __restrict__ vals,
                                  const double* __restrict__ deltas, double* __restrict__ rec, double* __restrict__ gcost) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_groups) return;
  const double sa = sqrt(V.loss_a);
  const double r = vals[(size_t)k * n_eval];
  if (!rec) { gcost[k] = 0.5 * V.loss_a * r * r; return; }
  const int im = V.g_img[k], cy = V.g_cyl[k], nc = CylCoords(V.kind);
  double ja[kCylAmb];
  for (int j = 0; j < kCylAmb; ++j) {
    const bool is_param = j < nc && (j < 7 ? !V.pose_const[im] : V.refine_geometry != 0);
    ja[j] = is_param ? (vals[(size_t)k * n_eval + 1 + 2 * j] - vals[(size_t)k * n_eval + 2 + 2 * j]) * ((1.0 / deltas[(size_t)k * kCylAmb + j]) / 2.0) : 0.0;
  }
  double jl[kGroupDim];
  for (int j = 0; j < kGroupDim; ++j) jl[j] = 0.0;
  int nd = 0;
  if (!V.pose_const[im]) {
    QuatTangent(V.q + 4 * im, ja, jl);
    nd = 3;
    for (int a = 0; a < 3; ++a) if (!((V.tvec_mask[im] >> a) & 1)) jl[nd++] = ja[4 + a];
  }
  if (V.refine_geometry) {
    if (V.kind == 0) {
      QuatTangent(V.cyl + 9 * cy, ja + 7, jl + nd);
      for (int a = 0; a < 5; ++a) jl[nd + 3 + a] = ja[11 + a];  // t (3), radius, height
      nd += 8;
    } else {
      for (int a = 0; a < 7; ++a) jl[nd + a] = ja[7 + a];       // t1 (3), t2 (3), radius
      nd += 7;
    }
  }
  // ScaledLoss(nullptr, a): rho = a s, rho' = a, rho'' = 0 -> J <- sqrt(a) J, r <- sqrt(a) r
  double* out = rec + (size_t)k * kGroupRec;
  for (int e = 0; e < kGroupRec; ++e) out[e] = 0.0;
  const double rc = sa * r;
  for (int a = 0; a < nd; ++a) {
    const double va = sa * jl[a];
    for (int b = 0; b <= a; ++b) out[a * kGroupDim + b] = va * (sa * jl[b]);
    out[kGroupDim * kGroupDim + a] = va * rc;
  }
  out[kGroupDim * kGroupDim + kGroupDim] = 0.5 * V.loss_a * r * r;
}

// candidate = Plus(current, delta): poses (quaternion manifold / tvec subset), cylinders (quaternion manifold for kind 0;
// radius clamped to its lower bound 0); squared norms of the ambient step and of the current point over variable blocks
__global__ void cyl_plus_kernel(CylSolveView V, const int* __restrict__ ent_off, const double* __restrict__ delta,
                                double* __restrict__ q2, double* __restrict__ t2, double* __restrict__ cyl2,
                                double* __restrict__ out) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;  // tens of entities: one thread, fixed order
  double step2 = 0.0, x2 = 0.0;
  for (int i = 0; i < V.n_img; ++i) {
    double qn[4] = {V.q[4 * i], V.q[4 * i + 1], V.q[4 * i + 2], V.q[4 * i + 3]};
    double tn[3] = {V.t[3 * i], V.t[3 * i + 1], V.t[3 * i + 2]};
    if (!V.pose_const[i]) {
      const double* d = delta + ent_off[i];
      QuatPlus(V.q + 4 * i, d, qn);
      int col = 3, free_t = 0;
      for (int a = 0; a < 3; ++a) if (!((V.tvec_mask[i] >> a) & 1)) { tn[a] += d[col++]; free_t++; }
      for (int a = 0; a < 4; ++a) { step2 += (qn[a] - V.q[4 * i + a]) * (qn[a] - V.q[4 * i + a]); x2 += V.q[4 * i + a] * V.q[4 * i + a]; }
      if (free_t > 0) for (int a = 0; a < 3; ++a) { step2 += (tn[a] - V.t[3 * i + a]) * (tn[a] - V.t[3 * i + a]); x2 += V.t[3 * i + a] * V.t[3 * i + a]; }
    }
    for (int a = 0; a < 4; ++a) q2[4 * i + a] = qn[a];
    for (int a = 0; a < 3; ++a) t2[3 * i + a] = tn[a];
  }
  for (int c = 0; c < V.n_cyl; ++c) {
    double x[9], xn[9];
    for (int a = 0; a < 9; ++a) { x[a] = V.cyl[9 * c + a]; xn[a] = x[a]; }
    if (V.refine_geometry) {
      const double* d = delta + ent_off[V.n_img + c];
      if (V.kind == 0) {
        QuatPlus(x, d, xn);
        for (int a = 0; a < 5; ++a) xn[4 + a] = x[4 + a] + d[3 + a];
        xn[7] = fmax(xn[7], 0.0);  // SetParameterLowerBound(radius, 0, 0): the bound is enforced inside Plus
        for (int a = 0; a < 9; ++a) { step2 += (xn[a] - x[a]) * (xn[a] - x[a]); x2 += x[a] * x[a]; }
      } else {
        for (int a = 0; a < 7; ++a) xn[a] = x[a] + d[a];
        xn[6] = fmax(xn[6], 0.0);
        for (int a = 0; a < 7; ++a) { step2 += (xn[a] - x[a]) * (xn[a] - x[a]); x2 += x[a] * x[a]; }
      }
    }
    for (int a = 0; a < 9; ++a) cyl2[9 * c + a] = xn[a];
  }
  out[0] = step2; out[1] = x2;
}

// landmark blocks: X+ = X + delta on the variable points (constant ones copied); adds their part of |x+ - x|^2, |x|^2 to out[0..1]
__global__ void __launch_bounds__(256) lm_points_plus_kernel(int n_pt, const double* __restrict__ X, const int* __restrict__ xcol,
                                                             const double* __restrict__ delta, double* __restrict__ X2,
                                                             double* __restrict__ out) {
  __shared__ double sa[256], sb[256];
  double s2 = 0.0, x2 = 0.0;
  for (int p = threadIdx.x; p < n_pt; p += 256)
    for (int j = 0; j < 3; ++j) {
      const double x = X[3 * (size_t)p + j];
      const double v = xcol[p] >= 0 ? x + delta[xcol[p] + j] : x;
      X2[3 * (size_t)p + j] = v;
      if (xcol[p] >= 0) { s2 += (v - x) * (v - x); x2 += x * x; }
    }
  sa[threadIdx.x] = s2; sb[threadIdx.x] = x2;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { sa[threadIdx.x] += sa[threadIdx.x + o]; sb[threadIdx.x] += sb[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { out[0] += sa[0]; out[1] += sb[0]; }
}
__global__ void add_scalar_kernel(double* __restrict__ dst, const double* __restrict__ src) { *dst += *src; }

__global__ void cyl_sum_kernel(int n, const double* __restrict__ v, double* __restrict__ out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) { double s = 0.0; for (int i = 0; i < n; ++i) s += v[i]; *out = s; }
}

}  // namespace
}  // namespace sba

using namespace sba;

extern "C" int sba_cylinder_solve(const sba_cylinder_problem* d, const sba_solver_options* opts, sba_iteration_callback cb,
                                  void* user, sba_summary* summary, sba_iteration_summary* trace_out, int32_t trace_cap,
                                  int32_t* trace_n) {
  if (!d || !summary || d->num_images <= 0 || d->height <= 0 || d->width <= 0 || d->num_cylinders < 0 || !d->image_qvec ||
      !d->image_tvec || !d->camera_params || !d->masks || (d->num_cylinders > 0 && !d->cylinder_params) ||
      d->cylinder_kind < 0 || d->cylinder_kind > 1) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "bad cylinder problem");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  const int64_t n_lm = d->num_landmarks;
  const int n_pts = d->num_points;
  if (n_lm < 0 || n_pts < 0 || (n_lm > 0 && (!d->points || !d->landmark_image || !d->landmark_point || !d->landmark_xy || n_pts == 0 ||
                                              !(d->landmark_weight > 0.0)))) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "bad landmark blocks"); return SBA_ERR_INVALID_ARGUMENT;
  }
  for (int64_t k = 0; k < n_lm; ++k)
    if (d->landmark_image[k] < 0 || d->landmark_image[k] >= d->num_images || d->landmark_point[k] < 0 || d->landmark_point[k] >= n_pts) {
      SetError(SBA_ERR_INVALID_ARGUMENT, "landmark index out of range"); return SBA_ERR_INVALID_ARGUMENT;
    }
  sba_solver_options o;
  if (opts) o = *opts; else sba_default_solver_options(&o);
  if (o.loss_function_type != SBA_LOSS_TRIVIAL) {  // geometric_semantic_bundle_adjustment.cc:699-706
    SetError(SBA_ERR_INVALID_ARGUMENT, "the only supported loss function is TRIVIAL");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    return SBA_ERR_NO_DEVICE;
  }
  std::memset(summary, 0, sizeof(*summary));
  const int n_img = d->num_images, n_cyl = d->num_cylinders, H = d->height, W = d->width, kind = d->cylinder_kind;
  const size_t HW = (size_t)H * W;
  const bool refine_geometry = d->refine_geometry != 0;
  cudaStream_t st;
  SBA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamSynchronize(s); cudaStreamDestroy(s); } } guard{st};
  std::vector<uint8_t> pc(n_img, 0), tm(n_img, 0);
  for (int i = 0; i < n_img; ++i) {
    pc[i] = d->image_pose_constant ? d->image_pose_constant[i] : 0;
    tm[i] = d->image_tvec_constant_mask ? d->image_tvec_constant_mask[i] : 0;
  }
  DenseGroups G;
  int off = 0;
  std::vector<uint8_t> img_in_problem(n_img, 0), img_has_lm(n_img, 0);
  for (int64_t k = 0; k < n_lm; ++k) img_has_lm[d->landmark_image[k]] = 1;
  for (int i = 0; i < n_img; ++i) {
    int dim = 0;
    if (!pc[i] && (n_cyl > 0 || img_has_lm[i])) { dim = 3; for (int a = 0; a < 3; ++a) dim += !((tm[i] >> a) & 1); }
    G.ent_off.push_back(off); G.ent_dim.push_back(dim);
    off += dim;
    if (dim) {
      summary->num_parameter_blocks_reduced += 1 + (dim > 3);
      summary->num_parameters_reduced += 4 + (dim > 3 ? 3 : 0);
      summary->num_effective_parameters_reduced += dim;
    }
  }
  // groups in the reference order: for image in Images, for cylinder in cylinders_
  for (int i = 0; i < n_img; ++i) {
    if (pc[i] && !refine_geometry) continue;  // :857-861
    for (int c = 0; c < n_cyl; ++c) { G.g_e1.push_back(i); G.g_e2.push_back(n_img + c); img_in_problem[i] = 1; }
  }
  const bool any_block = !G.g_e1.empty();
  for (int c = 0; c < n_cyl; ++c) {
    const int dim = (refine_geometry && any_block) ? (kind == 0 ? 8 : 7) : 0;
    G.ent_off.push_back(off); G.ent_dim.push_back(dim);
    off += dim;
    if (dim) {
      summary->num_parameter_blocks_reduced += kind == 0 ? 4 : 3;
      summary->num_parameters_reduced += kind == 0 ? 9 : 7;
      summary->num_effective_parameters_reduced += dim;
    }
  }
  // (tangent columns of poses and cylinders end here; the variable points of the landmark blocks follow)
  std::vector<uint8_t> pt_used(std::max(1, n_pts), 0);
  for (int64_t k = 0; k < n_lm; ++k) pt_used[d->landmark_point[k]] = 1;
  std::vector<int> xcol(std::max(1, n_pts), -1);
  for (int p = 0; p < n_pts && n_lm > 0; ++p) {
    if (!pt_used[p] || (d->point_constant && d->point_constant[p])) continue;
    xcol[p] = off; off += 3;
    summary->num_parameter_blocks_reduced += 1; summary->num_parameters_reduced += 3; summary->num_effective_parameters_reduced += 3;
  }
  G.n = off;
  if (n_lm > 0 && 2 * n_lm * (int64_t)std::max(1, off) > kRigMaxJacDoubles) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "sba_cylinder_solve: the landmark blocks hold their Jacobian dense; 2 * landmarks * tangent dimension exceeds 2^28");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  for (int i = 0; i < n_img; ++i) if (img_has_lm[i]) img_in_problem[i] = 1;  // image.NormalizeQvec(), :1371
  // a pose only owns columns when one of its blocks is in the problem (always, when there are cylinders)
  std::vector<double> q(d->image_qvec, d->image_qvec + 4 * (size_t)n_img), t(d->image_tvec, d->image_tvec + 3 * (size_t)n_img);
  for (int i = 0; i < n_img; ++i) {
    if (!img_in_problem[i]) continue;  // image.NormalizeQvec() (:849) runs before the skip; harmless either way
    const double nrm = std::sqrt(q[4 * i] * q[4 * i] + q[4 * i + 1] * q[4 * i + 1] + q[4 * i + 2] * q[4 * i + 2] + q[4 * i + 3] * q[4 * i + 3]);
    if (nrm == 0.0) { q[4 * i] = 1.0; q[4 * i + 1] = q[4 * i + 2] = q[4 * i + 3] = 0.0; }
    else for (int a = 0; a < 4; ++a) q[4 * i + a] /= nrm;
  }
  // one more pose slot behind the images: the constant identity the landmark blocks use as their "rig pose" (rig_lin.cuh)
  q.insert(q.end(), {1.0, 0.0, 0.0, 0.0}); t.insert(t.end(), {0.0, 0.0, 0.0});
  std::vector<double> cyl((size_t)std::max(1, n_cyl) * 9, 0.0);
  for (int c = 0; c < n_cyl; ++c) for (int a = 0; a < 9; ++a) cyl[9 * (size_t)c + a] = d->cylinder_params[9 * (size_t)c + a];
  int rc = G.Upload(st);
  if (rc) return rc;
  summary->num_residual_blocks = G.n_groups + (int)n_lm;
  summary->num_residuals = summary->num_residuals_reduced = G.n_groups + 2 * (int)n_lm;
  summary->linear_solver_type_used = n_img <= 50 ? SBA_DENSE_SCHUR : (n_img <= 1000 ? SBA_SPARSE_SCHUR : SBA_ITERATIVE_SCHUR);
  if (G.n_groups == 0 && n_lm == 0) {
    SetError(SBA_ERR_EMPTY_PROBLEM, "no residual blocks");
    return SBA_ERR_EMPTY_PROBLEM;
  }
  // the group arrays of the kernels: image and cylinder INDEX (not entity) per block
  std::vector<int> g_img(G.g_e1), g_cyl(G.n_groups);
  for (int g = 0; g < G.n_groups; ++g) g_cyl[g] = G.g_e2[g] - n_img;
  DevBuf<double> dq, dt, dq2, dt2, dcyl, dcyl2, dcam, vals, deltas, gcost, scal;
  DevBuf<uint8_t> dmask, dpc, dtm;
  DevBuf<int> dpos, dgi, dgc;
  SBA_CUDA(dq.Upload(q, st)); SBA_CUDA(dt.Upload(t, st)); SBA_CUDA(dcyl.Upload(cyl, st));
  SBA_CUDA(dq2.Upload(q, st)); SBA_CUDA(dt2.Upload(t, st)); SBA_CUDA(dcyl2.Alloc(cyl.size()));  // (both pose buffers hold the identity slot)
  SBA_CUDA(dcam.Alloc((size_t)n_img * 3));
  SBA_CUDA(cudaMemcpyAsync(dcam.p, d->camera_params, (size_t)n_img * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
  SBA_CUDA(dmask.Alloc(n_img * HW));
  SBA_CUDA(UploadLarge(dmask.p, d->masks, n_img * HW, st));
  SBA_CUDA(dpc.Upload(pc, st)); SBA_CUDA(dtm.Upload(tm, st)); SBA_CUDA(dgi.Upload(g_img, st)); SBA_CUDA(dgc.Upload(g_cyl, st));
  SBA_CUDA(dpos.Alloc(n_img));
  const int n_eval = 1 + 2 * kCylAmb;
  SBA_CUDA(vals.Alloc((size_t)G.n_groups * n_eval)); SBA_CUDA(deltas.Alloc((size_t)G.n_groups * kCylAmb));
  SBA_CUDA(gcost.Alloc(G.n_groups)); SBA_CUDA(scal.Alloc(4));
  cyl_mask_count_kernel<<<n_img, 256, 0, st>>>(dmask.p, HW, dpos.p);
  SBA_CUDA(cudaStreamSynchronize(st));  // the host vectors above are read by the asynchronous copies
  // ---- landmark blocks: reprojection blocks (identity "rig" pose, image pose, point, constant SIMPLE_PINHOLE camera) ----
  DevBuf<double> dX, dX2, dcam12, dlxy, dJ, drf, dlpart;
  DevBuf<int> dmodel, dqcol, dtcol, dpcol, dxcol, dla, dlb, dlp;
  const int64_t lm_rows = 2 * n_lm;
  const int lm_blocks = (int)DivUp(std::max<int64_t>(1, n_lm), 128);
  if (n_lm > 0) {
    std::vector<double> hX(d->points, d->points + 3 * (size_t)n_pts), cam12(12 * (size_t)n_img, 0.0);
    for (int i = 0; i < n_img; ++i) for (int a = 0; a < 3; ++a) cam12[12 * (size_t)i + a] = d->camera_params[3 * (size_t)i + a];
    std::vector<int> model(n_img, SBA_SIMPLE_PINHOLE), qcol(n_img + 1, -1), tcol(3 * (size_t)(n_img + 1), -1), pcol(12 * (size_t)n_img, -1);
    for (int i = 0; i < n_img; ++i) {
      if (!G.ent_dim[i]) continue;
      qcol[i] = G.ent_off[i];
      int c = 3;
      for (int a = 0; a < 3; ++a) if (!((tm[i] >> a) & 1)) tcol[3 * (size_t)i + a] = G.ent_off[i] + c++;
    }
    std::vector<int> la(n_lm, n_img), lb(d->landmark_image, d->landmark_image + n_lm), lp(d->landmark_point, d->landmark_point + n_lm);
    std::vector<double> lxy(d->landmark_xy, d->landmark_xy + 2 * n_lm);
    SBA_CUDA(dX.Upload(hX, st)); SBA_CUDA(dX2.Upload(hX, st)); SBA_CUDA(dcam12.Upload(cam12, st)); SBA_CUDA(dmodel.Upload(model, st));
    SBA_CUDA(dqcol.Upload(qcol, st)); SBA_CUDA(dtcol.Upload(tcol, st)); SBA_CUDA(dpcol.Upload(pcol, st)); SBA_CUDA(dxcol.Upload(xcol, st));
    SBA_CUDA(dla.Upload(la, st)); SBA_CUDA(dlb.Upload(lb, st)); SBA_CUDA(dlp.Upload(lp, st)); SBA_CUDA(dlxy.Upload(lxy, st));
    SBA_CUDA(dJ.Alloc((size_t)lm_rows * std::max(1, G.n))); SBA_CUDA(dJ.Zero(st));
    SBA_CUDA(drf.Alloc(lm_rows)); SBA_CUDA(dlpart.Alloc(lm_blocks));
    SBA_CUDA(cudaStreamSynchronize(st));
  }
  const LossView lm_loss = {SBA_LOSS_TRIVIAL, 1.0, 1};  // ScaledLoss(nullptr, weight): the weight rides in the view
  CylSolveView V;
  V.n_img = n_img; V.n_cyl = n_cyl; V.H = H; V.W = W; V.kind = kind; V.refine_geometry = refine_geometry ? 1 : 0;
  V.q = dq.p; V.t = dt.p; V.cyl = dcyl.p; V.cam = dcam.p; V.masks = dmask.p; V.pos_count = dpos.p;
  V.g_img = dgi.p; V.g_cyl = dgc.p; V.pose_const = dpc.p; V.tvec_mask = dtm.p;
  V.rel = d->relative_step_size; V.loss_a = 1.0 / n_img;  // 1. / config_.Images().size() (:721-722)
  bool cand_is_2 = true;
  auto view = [&](bool candidate) {
    CylSolveView v = V;
    const bool second = candidate == cand_is_2;
    v.q = second ? dq2.p : dq.p; v.t = second ? dt2.p : dt.p; v.cyl = second ? dcyl2.p : dcyl.p;
    return v;
  };
  auto lm_view = [&](bool candidate) {
    const CylSolveView v = view(candidate);
    const bool second = candidate == cand_is_2;
    RigView R;
    R.n_pose = n_img + 1; R.n_cam = n_img; R.n_pt = n_pts; R.n = G.n; R.n_obs = n_lm;
    R.pq = v.q; R.pt = v.t; R.cam = dcam12.p; R.X = second ? dX2.p : dX.p; R.model = dmodel.p;
    R.qcol = dqcol.p; R.tcol = dtcol.p; R.pcol = dpcol.p; R.xcol = dxcol.p;
    R.b_a = dla.p; R.b_b = dlb.p; R.b_c = dlb.p; R.b_p = dlp.p; R.b_xy = dlxy.p; R.weight = d->landmark_weight;
    return R;
  };
  auto read_back = [&]() -> int {
    if (n_lm > 0) SBA_CUDA(cudaMemcpyAsync(d->points, lm_view(false).X, 3 * (size_t)n_pts * sizeof(double), cudaMemcpyDeviceToHost, st));
    const CylSolveView v = view(false);
    SBA_CUDA(cudaMemcpyAsync(d->image_qvec, v.q, 4 * (size_t)n_img * sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaMemcpyAsync(d->image_tvec, v.t, 3 * (size_t)n_img * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (n_cyl > 0) SBA_CUDA(cudaMemcpyAsync(d->cylinder_params, v.cyl, 9 * (size_t)n_cyl * sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    return SBA_OK;
  };
  DenseLMCallbacks F;
  F.cb = cb; F.user = user;
  if (o.update_state_every_iteration) F.before_callback = [&]() { read_back(); };
  F.linearize = [&]() -> int {
    const CylSolveView v = view(false);
    if (G.n_groups > 0) {
      cyl_eval_kernel<<<G.n_groups * n_eval, 256, 0, st>>>(v, n_eval, vals.p, deltas.p);
      cyl_record_kernel<<<DivUp(G.n_groups, 128), 128, 0, st>>>(v, G.n_groups, n_eval, vals.p, deltas.p, G.d_rec.p, nullptr);
    }
    SBA_CUDA(cudaGetLastError());
    return SBA_OK;
  };
  if (n_lm > 0)  // H, g, cost = the group records of the IoU blocks + J^T J, J^T r, cost of the landmark blocks
    F.assemble = [&](double* d_H, double* d_g, double* d_cost) -> int {
      SBA_CUDA(cudaMemsetAsync(d_H, 0, (size_t)G.n * G.n * sizeof(double), st));
      const int rc2 = AssembleDense(G, d_H, d_g, d_cost, st, nullptr);
      if (rc2) return rc2;
      rig_lin_kernel<true><<<lm_blocks, 128, 0, st>>>(lm_view(false), lm_loss, dJ.p, drf.p, dlpart.p);
      rig_sum_kernel<<<1, 256, 0, st>>>(dlpart.p, lm_blocks, 1, 1, scal.p + 3);
      const int tiles = DivUp(G.n, 32);
      rig_jtj_kernel<<<dim3(tiles, tiles), dim3(32, 32), 0, st>>>(G.n, lm_rows, dJ.p, d_H, 1);
      rig_jtr_kernel<<<DivUp(G.n, 128), 128, 0, st>>>(G.n, lm_rows, dJ.p, drf.p, d_g, 1);
      add_scalar_kernel<<<1, 1, 0, st>>>(d_cost, scal.p + 3);
      SBA_CUDA(cudaGetLastError());
      return SBA_OK;
    };
  F.candidate = [&](const double* d_delta, double* cost, double* step2, double* x2) -> int {
    const CylSolveView cur = view(false), cand = view(true);
    cyl_plus_kernel<<<1, 32, 0, st>>>(cur, G.d_ent_off.p, d_delta, const_cast<double*>(cand.q), const_cast<double*>(cand.t),
                                      const_cast<double*>(cand.cyl), scal.p);
    if (G.n_groups > 0) {
      cyl_eval_kernel<<<G.n_groups, 256, 0, st>>>(cand, 1, vals.p, deltas.p);
      cyl_record_kernel<<<DivUp(G.n_groups, 128), 128, 0, st>>>(cand, G.n_groups, 1, vals.p, deltas.p, nullptr, gcost.p);
    }
    cyl_sum_kernel<<<1, 32, 0, st>>>(G.n_groups, gcost.p, scal.p + 2);
    if (n_lm > 0) {
      const RigView rc_cur = lm_view(false), rc_cand = lm_view(true);
      lm_points_plus_kernel<<<1, 256, 0, st>>>(n_pts, rc_cur.X, dxcol.p, d_delta, const_cast<double*>(rc_cand.X), scal.p);
      rig_lin_kernel<false><<<lm_blocks, 128, 0, st>>>(rc_cand, lm_loss, nullptr, nullptr, dlpart.p);
      rig_sum_kernel<<<1, 256, 0, st>>>(dlpart.p, lm_blocks, 1, 1, scal.p + 3);
    }
    SBA_CUDA(cudaGetLastError());
    double h[4] = {0.0, 0.0, 0.0, 0.0};
    SBA_CUDA(cudaMemcpyAsync(h, scal.p, (n_lm > 0 ? 4 : 3) * sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    *step2 = h[0]; *x2 = h[1]; *cost = h[2] + h[3];
    return SBA_OK;
  };
  F.accept = [&]() { cand_is_2 = !cand_is_2; };
  std::vector<sba_iteration_summary> trace;
  const sba_summary counts = *summary;
  rc = DenseLMSolve(G, o, F, st, summary, &trace);
  if (rc) return rc;
  summary->num_residual_blocks = counts.num_residual_blocks; summary->num_residuals = counts.num_residuals;
  summary->num_residuals_reduced = counts.num_residuals_reduced;
  summary->num_parameter_blocks_reduced = counts.num_parameter_blocks_reduced;
  summary->num_parameters_reduced = counts.num_parameters_reduced;
  summary->num_effective_parameters_reduced = counts.num_effective_parameters_reduced;
  summary->linear_solver_type_used = counts.linear_solver_type_used;
  if (trace_out && trace_n) {
    const int m = std::min<int>(trace_cap, (int)trace.size());
    for (int i = 0; i < m; ++i) trace_out[i] = trace[i];
    *trace_n = m;
  }
  return read_back();
}
