// SemanticBundleAdjuster as a solve on the device.
//
// What the reference does (/root/reference/src/optim/semantic_bundle_adjustment.cc):
//   SetUp :646-668           for id1 in Images, for id2 in Images: AddImagePairToProblem
//   AddImagePairToProblem    :699-906  skip id1 == id2, skip when both poses are constant (:783-788); grid y outer / x inner
//                            with error_computation_pixel_step (:796-799); skip depth < 1e-4 (:809-814); one residual
//                            block per remaining pixel: ConstantFirstPose / ConstantSecondPose / both-variable functor
//                            (:829-877) -- all three are compute_semantic_error (semantic_cost_functions.h:88-208) under
//                            NumericDiffCostFunction<CENTRAL> (:248-258), differing only in which pose is a parameter
//   SetUpManifolds :670-693  QuaternionManifold on variable poses, SubsetManifold on tvec components
//   Solve :430-544           solver policy by image count, ceres::Solve
// Here: the grid pixels with depth are listed once per image on the device (the list of pair (i1, i2) is the list of i1,
// in the reference order); one CTA per ordered pair evaluates its blocks -- a warp per block, lane 0 the residual, lanes
// 1..28 the +/- perturbations of the 14 ambient coordinates -- turns the central differences into the tangent row
// (PlusJacobian, subset mask, loss corrector) and reduces J^T J, J^T r and the cost of the pair in registers.  The dense
// LM of dense_lm.cu does the rest.  No per-block residual or Jacobian is ever stored.
#include <cfloat>
#include <climits>
#include <memory>

#include "dense_lm.h"
#include "device_math.cuh"
#include "semantic_device.cuh"

namespace sba {

namespace {

struct PairView {
  int n_img, H, W;
  const double* q; const double* t;   // [n_img*4], [n_img*3] current or candidate poses
  const double* cam;                  // [n_img*3]
  const float* depth; const float* label;
  const int* pix_ptr; const int* pix; // per image: valid grid pixels (y * W + x), reference order
  const int* g_i1; const int* g_i2;   // groups = ordered pairs
  const uint8_t* pose_const; const uint8_t* tvec_mask;
  double thr, rel;
  int loss_type; double loss_scale;
};

// compute_semantic_error (semantic_cost_functions.h:88-208) for pixel (x1, y1) of image 1 and the poses x = (q1 t1 q2 t2):
// the evaluator's own PairResidual (semantic_device.cuh), so that the solve and sba_semantic_evaluate cannot differ
__device__ inline double PairRes(const PairView& V, int x1, int y1, float depth1, float sem1, const float* depth2, const float* sem2,
                                 const double* x, const double* cam1, const double* cam2) {
  SemView S;
  S.H = V.H; S.W = V.W; S.thr = V.thr; S.rel = V.rel;
  int status;
  return PairResidual(S, x1, y1, depth1, sem1, depth2, sem2, x, cam1, cam2, &status);
}

__device__ inline double NumDelta(double x, double rel) { return NumDiffDelta(x, rel); }

// valid grid pixels of every image, y outer / x inner (one warp per image keeps the order with ballots)
__global__ void sem_grid_kernel(int n_img, int H, int W, int step, const float* __restrict__ depth, int* __restrict__ count,
                                int* __restrict__ pix /* null: count only */, const int* __restrict__ pix_ptr) {
  const int i = blockIdx.x, lane = threadIdx.x;
  const int gx = (W + step - 1) / step, gy = (H + step - 1) / step;
  int n = 0;
  for (int base = 0; base < gx * gy; base += 32) {
    const int idx = base + lane;
    bool ok = false; int p = 0;
    if (idx < gx * gy) {
      const int y = (idx / gx) * step, x = (idx % gx) * step;
      p = y * W + x;
      ok = !((double)depth[(size_t)i * H * W + p] < 1e-4);  // `if (depth < 1e-4) skip`, float promoted to double
    }
    const unsigned m = __ballot_sync(0xffffffffu, ok);
    if (ok && pix) pix[pix_ptr[i] + n + __popc(m & ((1u << lane) - 1))] = p;
    n += __popc(m);
  }
  if (lane == 0 && !pix) count[i] = n;
}

// One CTA per ordered pair.  JAC: a warp per block, records into rec; else: a thread per block, cost only into gcost.
template <bool JAC>
__global__ void __launch_bounds__(128) sem_group_kernel(PairView V, double* __restrict__ rec, double* __restrict__ gcost) {
  const int g = blockIdx.x, i1 = V.g_i1[g], i2 = V.g_i2[g];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const size_t HW = (size_t)V.H * V.W;
  const int l0 = V.pix_ptr[i1], cnt = V.pix_ptr[i1 + 1] - l0;
  const float* depth1 = V.depth + i1 * HW; const float* sem1 = V.label + i1 * HW;
  const float* depth2 = V.depth + i2 * HW; const float* sem2 = V.label + i2 * HW;
  double x0[14];
#pragma unroll
  for (int j = 0; j < 4; ++j) { x0[j] = V.q[4 * i1 + j]; x0[7 + j] = V.q[4 * i2 + j]; }
#pragma unroll
  for (int j = 0; j < 3; ++j) { x0[4 + j] = V.t[3 * i1 + j]; x0[11 + j] = V.t[3 * i2 + j]; }
  __shared__ double red[4][96];
  if (!JAC) {
    double c = 0.0;
    for (int l = threadIdx.x; l < cnt; l += 128) {
      const int p = V.pix[l0 + l];
      const double r = PairRes(V, p % V.W, p / V.W, depth1[p], sem1[p], depth2, sem2, x0, V.cam + 3 * i1, V.cam + 3 * i2);
      double rho[3];
      LossEval(V.loss_type, V.loss_scale, r * r, rho);
      c += 0.5 * rho[0];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    if (lane == 0) red[warp][0] = c;
    __syncthreads();
    if (threadIdx.x == 0) gcost[g] = (red[0][0] + red[1][0]) + (red[2][0] + red[3][0]);
    return;
  }
  // local tangent layout [q1 (3) | t1 free | q2 (3) | t2 free]; constant poses own no columns and are not perturbed
  const bool c1 = V.pose_const[i1] != 0, c2 = V.pose_const[i2] != 0;
  int tcol1[3], tcol2[3];
  int d1 = 0, d2 = 0;
  if (!c1) { d1 = 3; for (int a = 0; a < 3; ++a) tcol1[a] = ((V.tvec_mask[i1] >> a) & 1) ? -1 : d1++; } else { tcol1[0] = tcol1[1] = tcol1[2] = -1; }
  if (!c2) { d2 = 3; for (int a = 0; a < 3; ++a) tcol2[a] = ((V.tvec_mask[i2] >> a) & 1) ? -1 : d2++; } else { tcol2[0] = tcol2[1] = tcol2[2] = -1; }
  const int nd = d1 + d2, ntri = nd * (nd + 1) / 2;  // <= 12, 78
  // my perturbation: lanes 1..28 -> coordinate (lane - 1) / 2, sign + for odd lanes
  const int jc = (lane - 1) >> 1;
  const bool perturbs = lane >= 1 && lane <= 28 && !((jc < 7) ? c1 : c2);
  double x[14];
#pragma unroll
  for (int j = 0; j < 14; ++j) x[j] = x0[j];
  double my_delta = 1.0;
  if (lane < 14) my_delta = NumDelta(x0[lane], V.rel);
  if (perturbs) {
    double dj = 0.0;
#pragma unroll
    for (int j = 0; j < 14; ++j) if (j == jc) dj = NumDelta(x0[j], V.rel);
    const double sgn = (lane & 1) ? 1.0 : -1.0;
#pragma unroll
    for (int j = 0; j < 14; ++j) if (j == jc) x[j] = x0[j] + sgn * dj;
  }
  double acc[3] = {0.0, 0.0, 0.0};  // entries lane, lane + 32, lane + 64 of [lower triangle (ntri) | J^T r (nd) | cost]
  int ea[3], eb[3];                 // (a, b) of a triangle entry; (a, -1) of a gradient entry; (-1, -1) cost; (-2, .) none
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int e = lane + 32 * k;
    if (e < ntri) { int a = 0; while ((a + 1) * (a + 2) / 2 <= e) ++a; ea[k] = a; eb[k] = e - a * (a + 1) / 2; }
    else if (e < ntri + nd) { ea[k] = e - ntri; eb[k] = -1; }
    else if (e == ntri + nd) { ea[k] = -1; eb[k] = -1; }
    else { ea[k] = -2; eb[k] = -2; }
  }
  for (int l = warp; l < cnt; l += 4) {
    const int p = V.pix[l0 + l];
    double r = 0.0;
    if (lane == 0 || perturbs) r = PairRes(V, p % V.W, p / V.W, depth1[p], sem1[p], depth2, sem2, x, V.cam + 3 * i1, V.cam + 3 * i2);
    const double fp = __shfl_sync(0xffffffffu, r, (1 + 2 * lane) & 31), fm = __shfl_sync(0xffffffffu, r, (2 + 2 * lane) & 31);
    const double ja_mine = (lane < 14) ? (fp - fm) * ((1.0 / my_delta) / 2.0) : 0.0;  // Ceres: (f(x+d) - f(x-d)) * (1 / (2 d))
    const double r0 = __shfl_sync(0xffffffffu, r, 0);
    if (!__any_sync(0xffffffffu, ja_mine != 0.0) && r0 == 0.0) continue;  // a zero block adds nothing
    double ja[14];
#pragma unroll
    for (int j = 0; j < 14; ++j) ja[j] = __shfl_sync(0xffffffffu, ja_mine, j);
    double jl[12];
#pragma unroll
    for (int j = 0; j < 12; ++j) jl[j] = 0.0;
    if (!c1) {
      QuatTangent(x0, ja, jl);
#pragma unroll
      for (int a = 0; a < 3; ++a) if (tcol1[a] >= 0) { for (int j = 0; j < 12; ++j) if (j == tcol1[a]) jl[j] = ja[4 + a]; }
    }
    if (!c2) {
      double t3[3];
      QuatTangent(x0 + 7, ja + 7, t3);
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        for (int j = 0; j < 12; ++j) if (j == d1 + a) jl[j] = t3[a];
        if (tcol2[a] >= 0) { for (int j = 0; j < 12; ++j) if (j == d1 + tcol2[a]) jl[j] = ja[11 + a]; }
      }
    }
    // loss and corrector of a one-residual block (ceres corrector.cc, App. B.2)
    double rho[3];
    const double s = r0 * r0;
    LossEval(V.loss_type, V.loss_scale, s, rho);
    const double sq = sqrt(rho[1]);
    double jscale = sq, rscale = sq;
    if (!(s == 0.0 || rho[2] <= 0.0)) {
      const double alpha = 1.0 - sqrt(1.0 + 2.0 * s * rho[2] / rho[1]);
      jscale = sq * (1.0 - alpha); rscale = sq / (1.0 - alpha);
    }
    const double rc = r0 * rscale;
#pragma unroll
    for (int j = 0; j < 12; ++j) jl[j] *= jscale;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      double ja_ = 0.0, jb_ = 0.0;
#pragma unroll
      for (int j = 0; j < 12; ++j) { ja_ = (j == ea[k]) ? jl[j] : ja_; jb_ = (j == eb[k]) ? jl[j] : jb_; }
      acc[k] += ea[k] == -2 ? 0.0 : (ea[k] == -1 ? 0.5 * rho[0] : (eb[k] == -1 ? ja_ * rc : ja_ * jb_));
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) red[warp][lane + 32 * k] = acc[k];
  __syncthreads();
  double* out = rec + (size_t)g * kGroupRec;
  for (int e = threadIdx.x; e < kGroupRec; e += 128) out[e] = 0.0;
  __syncthreads();
  for (int e = threadIdx.x; e <= ntri + nd; e += 128) {
    const double v = (red[0][e] + red[1][e]) + (red[2][e] + red[3][e]);
    if (e < ntri) {
      int a = 0;
      while ((a + 1) * (a + 2) / 2 <= e) ++a;
      out[a * kGroupDim + (e - a * (a + 1) / 2)] = v;
    } else if (e < ntri + nd) {
      out[kGroupDim * kGroupDim + (e - ntri)] = v;
    } else {
      out[kGroupDim * kGroupDim + kGroupDim] = v;
    }
  }
}

// candidate = Plus(current, delta); squared norms of the ambient step / current point over the variable blocks
__global__ void sem_plus_kernel(int n_img, const double* __restrict__ q, const double* __restrict__ t, const int* __restrict__ ent_off,
                                const uint8_t* __restrict__ pose_const, const uint8_t* __restrict__ tvec_mask,
                                const double* __restrict__ delta, double* __restrict__ q2, double* __restrict__ t2,
                                double* __restrict__ out /* [0] step^2, [1] x^2 */) {
  __shared__ double s_step, s_x;
  if (threadIdx.x == 0) { s_step = 0.0; s_x = 0.0; }
  __syncthreads();
  double step2 = 0.0, x2 = 0.0;
  for (int i = threadIdx.x; i < n_img; i += blockDim.x) {
    double qn[4] = {q[4 * i], q[4 * i + 1], q[4 * i + 2], q[4 * i + 3]};
    double tn[3] = {t[3 * i], t[3 * i + 1], t[3 * i + 2]};
    if (!pose_const[i]) {
      const double* d = delta + ent_off[i];
      QuatPlus(q + 4 * i, d, qn);
      int col = 3, free_t = 0;
      for (int a = 0; a < 3; ++a) if (!((tvec_mask[i] >> a) & 1)) { tn[a] += d[col++]; free_t++; }
      for (int a = 0; a < 4; ++a) { step2 += (qn[a] - q[4 * i + a]) * (qn[a] - q[4 * i + a]); x2 += q[4 * i + a] * q[4 * i + a]; }
      if (free_t > 0) for (int a = 0; a < 3; ++a) { step2 += (tn[a] - t[3 * i + a]) * (tn[a] - t[3 * i + a]); x2 += t[3 * i + a] * t[3 * i + a]; }
    }
    for (int a = 0; a < 4; ++a) q2[4 * i + a] = qn[a];
    for (int a = 0; a < 3; ++a) t2[3 * i + a] = tn[a];
  }
  // (one CTA, a handful of images per thread: plain shared atomics are fine -- but order-dependent; serialise instead)
  for (int k = 0; k < (int)blockDim.x; ++k) {
    if ((int)threadIdx.x == k) { s_step += step2; s_x += x2; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { out[0] = s_step; out[1] = s_x; }
}

__global__ void sem_sum_kernel(int n, const double* __restrict__ v, double* __restrict__ out) {  // fixed order
  __shared__ double sm[8];
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) s += v[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) { double t = 0.0; for (int i = 0; i < 8; ++i) t += sm[i]; *out = t; }
}

}  // namespace
}  // namespace sba

using namespace sba;

extern "C" int sba_semantic_solve(const sba_semantic_problem* d, const sba_solver_options* opts, sba_iteration_callback cb,
                                  void* user, sba_summary* summary, sba_iteration_summary* trace_out, int32_t trace_cap,
                                  int32_t* trace_n) {
  if (!d || !summary || d->num_images <= 0 || d->height <= 0 || d->width <= 0 || !d->image_qvec || !d->image_tvec ||
      !d->camera_params || !d->depth_maps || !d->semantic_maps || d->pixel_step <= 0) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "bad semantic problem");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  sba_solver_options o;
  if (opts) o = *opts; else sba_default_solver_options(&o);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    return SBA_ERR_NO_DEVICE;
  }
  std::memset(summary, 0, sizeof(*summary));
  const int n_img = d->num_images, H = d->height, W = d->width;
  const size_t HW = (size_t)H * W;
  cudaStream_t st;
  SBA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamSynchronize(s); cudaStreamDestroy(s); } } guard{st};
  // ---- entities (poses) and groups (ordered pairs) in the reference order ----
  std::vector<uint8_t> pc(n_img, 0), tm(n_img, 0);
  for (int i = 0; i < n_img; ++i) {
    pc[i] = d->image_pose_constant ? d->image_pose_constant[i] : 0;
    tm[i] = d->image_tvec_constant_mask ? d->image_tvec_constant_mask[i] : 0;
  }
  DenseGroups G;
  int off = 0;
  for (int i = 0; i < n_img; ++i) {
    int dim = 0;
    if (!pc[i]) { dim = 3; for (int a = 0; a < 3; ++a) dim += !((tm[i] >> a) & 1); }
    G.ent_off.push_back(off); G.ent_dim.push_back(dim);
    off += dim;
    if (dim) {
      summary->num_parameter_blocks_reduced += 1 + (dim > 3);
      summary->num_parameters_reduced += 4 + (dim > 3 ? 3 : 0);
      summary->num_effective_parameters_reduced += dim;
    }
  }
  G.n = off;
  for (int i1 = 0; i1 < n_img; ++i1)
    for (int i2 = 0; i2 < n_img; ++i2) {
      if (i1 == i2 || (pc[i1] && pc[i2])) continue;  // :703, :783-788
      G.g_e1.push_back(i1); G.g_e2.push_back(i2);
    }
  // AddImagePairToProblem normalises the quaternion of both images of every pair it visits (:729-730) -- every image
  // as soon as there are two
  std::vector<double> q(d->image_qvec, d->image_qvec + 4 * (size_t)n_img), t(d->image_tvec, d->image_tvec + 3 * (size_t)n_img);
  if (n_img >= 2)
    for (int i = 0; i < n_img; ++i) {
      const double nrm = std::sqrt(q[4 * i] * q[4 * i] + q[4 * i + 1] * q[4 * i + 1] + q[4 * i + 2] * q[4 * i + 2] + q[4 * i + 3] * q[4 * i + 3]);
      if (nrm == 0.0) { q[4 * i] = 1.0; q[4 * i + 1] = q[4 * i + 2] = q[4 * i + 3] = 0.0; }  // NormalizeQvec of a zero quaternion
      else for (int a = 0; a < 4; ++a) q[4 * i + a] /= nrm;
    }
  int rc = G.Upload(st);
  if (rc) return rc;
  // ---- device data ----
  DevBuf<double> dq, dt, dq2, dt2, dcam, gcost, scal;
  DevBuf<float> ddepth, dlabel;
  DevBuf<uint8_t> dpc, dtm;
  DevBuf<int> dcount, dptr, dpix;
  SBA_CUDA(dq.Upload(q, st)); SBA_CUDA(dt.Upload(t, st));
  SBA_CUDA(dq2.Alloc(q.size())); SBA_CUDA(dt2.Alloc(t.size()));
  SBA_CUDA(dcam.Alloc((size_t)n_img * 3));
  SBA_CUDA(cudaMemcpyAsync(dcam.p, d->camera_params, (size_t)n_img * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
  SBA_CUDA(ddepth.Alloc(n_img * HW)); SBA_CUDA(dlabel.Alloc(n_img * HW));
  SBA_CUDA(UploadLarge(ddepth.p, d->depth_maps, n_img * HW * sizeof(float), st));
  SBA_CUDA(UploadLarge(dlabel.p, d->semantic_maps, n_img * HW * sizeof(float), st));
  SBA_CUDA(dpc.Upload(pc, st)); SBA_CUDA(dtm.Upload(tm, st));
  SBA_CUDA(dcount.Alloc(n_img)); SBA_CUDA(scal.Alloc(4)); SBA_CUDA(gcost.Alloc(std::max<size_t>(1, G.n_groups)));
  sem_grid_kernel<<<n_img, 32, 0, st>>>(n_img, H, W, d->pixel_step, ddepth.p, dcount.p, nullptr, nullptr);
  std::vector<int> count(n_img), ptr(n_img + 1, 0);
  SBA_CUDA(cudaMemcpyAsync(count.data(), dcount.p, n_img * sizeof(int), cudaMemcpyDeviceToHost, st));
  SBA_CUDA(cudaStreamSynchronize(st));
  for (int i = 0; i < n_img; ++i) ptr[i + 1] = ptr[i] + count[i];
  SBA_CUDA(dptr.Upload(ptr, st)); SBA_CUDA(dpix.Alloc(std::max(1, ptr[n_img])));
  sem_grid_kernel<<<n_img, 32, 0, st>>>(n_img, H, W, d->pixel_step, ddepth.p, nullptr, dpix.p, dptr.p);
  SBA_CUDA(cudaGetLastError());
  int64_t n_blocks = 0;
  for (int g = 0; g < G.n_groups; ++g) n_blocks += count[G.g_e1[g]];
  summary->num_residual_blocks = summary->num_residuals = summary->num_residuals_reduced = (int)std::min<int64_t>(n_blocks, INT_MAX);
  // BundleAdjuster's solver policy (:491-503); every type is the same dense factorisation here
  {
    int solver = o.linear_solver_type;
    if (solver == SBA_SOLVER_AUTO) {
      const int nimg = o.num_config_images > 0 ? o.num_config_images : n_img;
      solver = nimg <= 50 ? SBA_DENSE_SCHUR : (nimg <= 1000 ? SBA_SPARSE_SCHUR : SBA_ITERATIVE_SCHUR);
    }
    summary->linear_solver_type_used = solver;
  }
  if (n_blocks == 0) {  // `if (problem_->NumResiduals() == 0) return false` (:484-486)
    SetError(SBA_ERR_EMPTY_PROBLEM, "no residual blocks");
    return SBA_ERR_EMPTY_PROBLEM;
  }
  PairView V;
  V.n_img = n_img; V.H = H; V.W = W; V.q = dq.p; V.t = dt.p; V.cam = dcam.p; V.depth = ddepth.p; V.label = dlabel.p;
  V.pix_ptr = dptr.p; V.pix = dpix.p; V.g_i1 = G.d_g_e1.p; V.g_i2 = G.d_g_e2.p; V.pose_const = dpc.p; V.tvec_mask = dtm.p;
  V.thr = d->depth_error_threshold; V.rel = d->relative_step_size; V.loss_type = o.loss_function_type; V.loss_scale = o.loss_function_scale;
  bool cand_is_2 = true;  // which buffer pair holds the candidate
  auto cur_q = [&]() { return cand_is_2 ? dq.p : dq2.p; };
  auto cur_t = [&]() { return cand_is_2 ? dt.p : dt2.p; };
  auto cand_q = [&]() { return cand_is_2 ? dq2.p : dq.p; };
  auto cand_t = [&]() { return cand_is_2 ? dt2.p : dt.p; };
  auto read_back = [&]() -> int {
    SBA_CUDA(cudaMemcpyAsync(d->image_qvec, cur_q(), 4 * (size_t)n_img * sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaMemcpyAsync(d->image_tvec, cur_t(), 3 * (size_t)n_img * sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    return SBA_OK;
  };
  DenseLMCallbacks F;
  F.cb = cb; F.user = user;
  if (o.update_state_every_iteration) F.before_callback = [&]() { read_back(); };
  F.linearize = [&]() -> int {
    PairView v = V; v.q = cur_q(); v.t = cur_t();
    sem_group_kernel<true><<<G.n_groups, 128, 0, st>>>(v, G.d_rec.p, nullptr);
    SBA_CUDA(cudaGetLastError());
    return SBA_OK;
  };
  F.candidate = [&](const double* d_delta, double* cost, double* step2, double* x2) -> int {
    sem_plus_kernel<<<1, 128, 0, st>>>(n_img, cur_q(), cur_t(), G.d_ent_off.p, dpc.p, dtm.p, d_delta, cand_q(), cand_t(), scal.p);
    PairView v = V; v.q = cand_q(); v.t = cand_t();
    sem_group_kernel<false><<<G.n_groups, 128, 0, st>>>(v, nullptr, gcost.p);
    sem_sum_kernel<<<1, 256, 0, st>>>(G.n_groups, gcost.p, scal.p + 2);
    SBA_CUDA(cudaGetLastError());
    double h[3];
    SBA_CUDA(cudaMemcpyAsync(h, scal.p, 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    *step2 = h[0]; *x2 = h[1]; *cost = h[2];
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
