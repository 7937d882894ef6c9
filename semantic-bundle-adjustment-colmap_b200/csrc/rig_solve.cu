// RigBundleAdjuster as a solve: /root/reference/src/optim/bundle_adjustment.cc:789-1140.
//
// What the reference hands to Ceres: one RigBundleAdjustmentCostFunction block per kept observation of an image that
// belongs to a rig -- parameter blocks (rig q, rig t) of its snapshot, (relative q, relative t) of its camera in the
// rig, the point, the intrinsics (:1017-1031) -- and ordinary reprojection blocks for images outside rigs (:981-1014);
// QuaternionManifold on every q block in use (:1135-1139), the relative pose of the reference camera (or all of them,
// refine_relative_poses = false) constant (:1040-1046), SubsetManifold on tvec / intrinsics, then ceres::Solve.
//
// An ordinary block IS a rig block whose rig pose is the constant identity (q = z (x) (1,0,0,0) = z and
// t = R(z) 0 + t_rel = t_rel exactly, and its Jacobian w.r.t. the relative pose is the ordinary one), so the flat
// problem has one kind of block: (pose a, pose b, camera, point).
//
// SCOPE: this is the small-problem path.  The Schur pipeline of this library gives every observation ONE camera-side
// pose block; a rig observation has two, shared across images, and that generalisation is not built (DESIGN.md section
// 10).  Here the LM step solves the DENSE normal equations over ALL blocks, points included -- what Ceres' Schur solvers
// compute after eliminating the points, to rounding -- through csrc/dense_lm.cu: J is held as a dense
// (2 n_obs) x n matrix, H = J^T J and g = J^T r are formed by fixed-order kernels (bit-reproducible), and problems
// whose J would exceed kRigMaxJacDoubles are refused.  It covers the reference's rig tests and calibration-sized rigs.
#include <algorithm>
#include <cmath>
#include <vector>

#include "rig_lin.cuh"

namespace sba {

// candidate = Plus(current, delta) for every block (constant blocks are copied); partial[b*2 + {0, 1}] = this CTA's part
// of |x+ - x|^2 and |x|^2 over the ambient coordinates of the VARIABLE blocks (TrustRegionMinimizer's parameter tolerance)
__global__ void __launch_bounds__(256) rig_plus_kernel(RigView V, const double* __restrict__ delta, double* __restrict__ nq,
                                                       double* __restrict__ nt, double* __restrict__ ncam, double* __restrict__ nX,
                                                       double* __restrict__ partial) {
  double s2 = 0.0, x2 = 0.0;
  const int stride = gridDim.x * 256, t0 = blockIdx.x * 256 + threadIdx.x;
  for (int i = t0; i < V.n_pose; i += stride) {
    const double* q = V.pq + 4 * (size_t)i;
    double o[4] = {q[0], q[1], q[2], q[3]};
    if (V.qcol[i] >= 0) {
      QuatPlus(q, delta + V.qcol[i], o);
      for (int j = 0; j < 4; ++j) { s2 += (o[j] - q[j]) * (o[j] - q[j]); x2 += q[j] * q[j]; }
    }
    for (int j = 0; j < 4; ++j) nq[4 * (size_t)i + j] = o[j];
    const double* t = V.pt + 3 * (size_t)i;
    const bool var = V.tcol[3 * i] >= 0 || V.tcol[3 * i + 1] >= 0 || V.tcol[3 * i + 2] >= 0;
    for (int j = 0; j < 3; ++j) {
      const int c = V.tcol[3 * i + j];
      const double v = c >= 0 ? t[j] + delta[c] : t[j];
      nt[3 * (size_t)i + j] = v;
      if (var) { s2 += (v - t[j]) * (v - t[j]); x2 += t[j] * t[j]; }
    }
  }
  for (int c = t0; c < V.n_cam; c += stride) {
    const double* p = V.cam + 12 * (size_t)c;
    bool var = false;
    for (int j = 0; j < 12; ++j) var = var || V.pcol[12 * c + j] >= 0;
    const int np = ModelNumParams(V.model[c]);
    for (int j = 0; j < 12; ++j) {
      const int col = V.pcol[12 * c + j];
      const double v = col >= 0 ? p[j] + delta[col] : p[j];
      ncam[12 * (size_t)c + j] = v;
      if (var && j < np) { s2 += (v - p[j]) * (v - p[j]); x2 += p[j] * p[j]; }
    }
  }
  for (int p = t0; p < V.n_pt; p += stride) {
    const double* X = V.X + 3 * (size_t)p;
    for (int j = 0; j < 3; ++j) {
      const double v = V.xcol[p] >= 0 ? X[j] + delta[V.xcol[p] + j] : X[j];
      nX[3 * (size_t)p + j] = v;
      if (V.xcol[p] >= 0) { s2 += (v - X[j]) * (v - X[j]); x2 += X[j] * X[j]; }
    }
  }
  __shared__ double sa[256], sb[256];
  sa[threadIdx.x] = s2; sb[threadIdx.x] = x2;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { sa[threadIdx.x] += sa[threadIdx.x + o]; sb[threadIdx.x] += sb[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { partial[2 * blockIdx.x] = sa[0]; partial[2 * blockIdx.x + 1] = sb[0]; }
}

}  // namespace sba

using namespace sba;

extern "C" int sba_rig_solve(sba_rig_problem* d, const sba_solver_options* opts, sba_iteration_callback cb, void* user,
                             sba_summary* summary, sba_iteration_summary* trace, int32_t trace_cap, int32_t* trace_n) {
  if (!d || !opts || !summary) { SetError(SBA_ERR_INVALID_ARGUMENT, "sba_rig_solve: null problem / options / summary"); return SBA_ERR_INVALID_ARGUMENT; }
  std::memset(summary, 0, sizeof(*summary));
  if (trace_n) *trace_n = 0;
  const int n_rig = d->num_rig_poses, n_rel = d->num_rel_poses, n_pose = n_rig + n_rel, n_cam = d->num_cameras, n_pt = d->num_points;
  const int64_t n_obs = d->num_blocks;
  if (n_rig < 0 || n_rel < 0 || n_cam < 0 || n_pt < 0 || n_obs < 0 || (n_obs > 0 && (!d->block_rig || !d->block_rel || !d->block_camera || !d->block_point || !d->block_xy)) ||
      (n_rig > 0 && (!d->rig_qvec || !d->rig_tvec)) || (n_rel > 0 && (!d->rel_qvec || !d->rel_tvec)) || (n_cam > 0 && (!d->camera_model || !d->camera_params)) ||
      (n_pt > 0 && !d->points)) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "sba_rig_solve: null arrays / negative counts"); return SBA_ERR_INVALID_ARGUMENT;
  }
  for (int c = 0; c < n_cam; ++c)
    if (ModelNumParams(d->camera_model[c]) < 0) { SetError(SBA_ERR_UNKNOWN_CAMERA_MODEL, "camera model does not exist"); return SBA_ERR_UNKNOWN_CAMERA_MODEL; }
  // blocks in use (a parameter block exists in the Ceres problem only through a residual block)
  std::vector<uint8_t> use_pose(std::max(1, n_pose), 0), use_cam(std::max(1, n_cam), 0), use_pt(std::max(1, n_pt), 0);
  std::vector<int> b_a(std::max<int64_t>(1, n_obs)), b_b(std::max<int64_t>(1, n_obs));
  for (int64_t k = 0; k < n_obs; ++k) {
    const int a = d->block_rig[k], b = d->block_rel[k], c = d->block_camera[k], p = d->block_point[k];
    if (a < 0 || a >= n_rig || b < 0 || b >= n_rel || c < 0 || c >= n_cam || p < 0 || p >= n_pt) {
      SetError(SBA_ERR_INVALID_ARGUMENT, "rig block index out of range"); return SBA_ERR_INVALID_ARGUMENT;
    }
    b_a[k] = a; b_b[k] = n_rig + b;
    use_pose[a] = use_pose[n_rig + b] = use_cam[c] = use_pt[p] = 1;
  }
  if (n_obs == 0) { SetError(SBA_ERR_EMPTY_PROBLEM, "no residual block"); return SBA_ERR_EMPTY_PROBLEM; }  // Solve returns false, :824-826
  // tangent columns, in the order Ceres' Program holds the blocks of this build: poses (q then t), cameras, points
  std::vector<int> qcol(n_pose, -1), tcol(3 * (size_t)n_pose, -1), pcol(12 * (size_t)std::max(1, n_cam), -1), xcol(std::max(1, n_pt), -1);
  int n = 0, n_par = 0, n_par_blocks = 0;
  for (int i = 0; i < n_pose; ++i) {
    const bool rig = i < n_rig;
    const int j = rig ? i : i - n_rig;
    const bool konst = rig ? (d->rig_pose_constant && d->rig_pose_constant[j]) : (d->rel_pose_constant && d->rel_pose_constant[j]);
    if (!use_pose[i] || konst) continue;
    qcol[i] = n; n += 3; n_par += 4; n_par_blocks++;
    const unsigned mask = (!rig && d->rel_tvec_constant_mask) ? d->rel_tvec_constant_mask[j] & 7u : 0u;
    if (mask != 7u) {
      for (int c = 0; c < 3; ++c) if (!(mask >> c & 1u)) tcol[3 * (size_t)i + c] = n++;
      n_par += 3; n_par_blocks++;
    }
  }
  for (int c = 0; c < n_cam; ++c) {
    if (!use_cam[c]) continue;
    const int np = ModelNumParams(d->camera_model[c]);
    const unsigned mask = d->camera_param_constant_mask ? d->camera_param_constant_mask[c] : 0u;
    bool any = false;
    for (int j = 0; j < np; ++j) if (!(mask >> j & 1u)) { pcol[12 * (size_t)c + j] = n++; any = true; }
    if (any) { n_par += np; n_par_blocks++; }
  }
  for (int p = 0; p < n_pt; ++p) {
    if (!use_pt[p] || (d->point_constant && d->point_constant[p])) continue;
    xcol[p] = n; n += 3; n_par += 3; n_par_blocks++;
  }
  summary->num_residual_blocks = (int)n_obs; summary->num_residuals = 2 * (int)n_obs; summary->num_residuals_reduced = 2 * (int)n_obs;
  summary->num_parameter_blocks_reduced = n_par_blocks; summary->num_parameters_reduced = n_par;
  summary->num_effective_parameters_reduced = n;
  summary->linear_solver_type_used = SBA_DENSE_SCHUR;
  if (2 * n_obs * (int64_t)std::max(1, n) > kRigMaxJacDoubles) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "sba_rig_solve: the dense rig path holds J as (2 * blocks) x (tangent dimension) doubles; this problem exceeds 2^28 (DESIGN.md section 10)");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    return SBA_ERR_NO_DEVICE;
  }
  cudaStream_t st;
  SBA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamDestroy(s); } } guard{st};
  // ---- device state ----
  std::vector<double> hq(4 * (size_t)std::max(1, n_pose)), ht(3 * (size_t)std::max(1, n_pose));
  for (int i = 0; i < n_rig; ++i) { std::memcpy(&hq[4 * (size_t)i], d->rig_qvec + 4 * (size_t)i, 32); std::memcpy(&ht[3 * (size_t)i], d->rig_tvec + 3 * (size_t)i, 24); }
  for (int i = 0; i < n_rel; ++i) { std::memcpy(&hq[4 * (size_t)(n_rig + i)], d->rel_qvec + 4 * (size_t)i, 32); std::memcpy(&ht[3 * (size_t)(n_rig + i)], d->rel_tvec + 3 * (size_t)i, 24); }
  std::vector<double> hcam(d->camera_params, d->camera_params + 12 * (size_t)n_cam), hX(d->points, d->points + 3 * (size_t)n_pt);
  std::vector<int> hmodel(d->camera_model, d->camera_model + n_cam), hbc(d->block_camera, d->block_camera + n_obs), hbp(d->block_point, d->block_point + n_obs);
  std::vector<double> hxy(d->block_xy, d->block_xy + 2 * n_obs);
  DevBuf<double> q[2], t[2], cam[2], X[2], xy, Jf, rf, cpart, ppart, scal;
  DevBuf<int> model, dq, dt, dp, dx, da, db, dc, dpt;
  for (int s = 0; s < 2; ++s) { SBA_CUDA(q[s].Upload(hq, st)); SBA_CUDA(t[s].Upload(ht, st)); SBA_CUDA(cam[s].Upload(hcam, st)); SBA_CUDA(X[s].Upload(hX, st)); }
  SBA_CUDA(xy.Upload(hxy, st)); SBA_CUDA(model.Upload(hmodel, st));
  SBA_CUDA(dq.Upload(qcol, st)); SBA_CUDA(dt.Upload(tcol, st)); SBA_CUDA(dp.Upload(pcol, st)); SBA_CUDA(dx.Upload(xcol, st));
  SBA_CUDA(da.Upload(b_a, st)); SBA_CUDA(db.Upload(b_b, st)); SBA_CUDA(dc.Upload(hbc, st)); SBA_CUDA(dpt.Upload(hbp, st));
  const int64_t m = 2 * n_obs;
  const int lin_blocks = (int)DivUp(n_obs, 128), plus_blocks = 64;
  SBA_CUDA(Jf.Alloc((size_t)m * std::max(1, n))); SBA_CUDA(Jf.Zero(st));
  SBA_CUDA(rf.Alloc(m)); SBA_CUDA(cpart.Alloc(lin_blocks)); SBA_CUDA(ppart.Alloc(2 * plus_blocks)); SBA_CUDA(scal.Alloc(4));
  int cur = 0;
  auto view = [&](int s) {
    RigView V;
    V.n_pose = n_pose; V.n_cam = n_cam; V.n_pt = n_pt; V.n = n; V.n_obs = n_obs;
    V.pq = q[s].p; V.pt = t[s].p; V.cam = cam[s].p; V.X = X[s].p; V.model = model.p;
    V.qcol = dq.p; V.tcol = dt.p; V.pcol = dp.p; V.xcol = dx.p;
    V.b_a = da.p; V.b_b = db.p; V.b_c = dc.p; V.b_p = dpt.p; V.b_xy = xy.p; V.weight = 1.0;
    return V;
  };
  const LossView loss = {opts->loss_function_type, opts->loss_function_scale, 1};
  auto read_back = [&]() -> int {
    SBA_CUDA(cudaMemcpyAsync(hq.data(), q[cur].p, hq.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaMemcpyAsync(ht.data(), t[cur].p, ht.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (n_cam) SBA_CUDA(cudaMemcpyAsync(d->camera_params, cam[cur].p, 12 * (size_t)n_cam * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (n_pt) SBA_CUDA(cudaMemcpyAsync(d->points, X[cur].p, 3 * (size_t)n_pt * sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    for (int i = 0; i < n_rig; ++i) { std::memcpy(d->rig_qvec + 4 * (size_t)i, &hq[4 * (size_t)i], 32); std::memcpy(d->rig_tvec + 3 * (size_t)i, &ht[3 * (size_t)i], 24); }
    for (int i = 0; i < n_rel; ++i) { std::memcpy(d->rel_qvec + 4 * (size_t)i, &hq[4 * (size_t)(n_rig + i)], 32); std::memcpy(d->rel_tvec + 3 * (size_t)i, &ht[3 * (size_t)(n_rig + i)], 24); }
    return SBA_OK;
  };
  DenseGroups G;
  G.n = n;
  DenseLMCallbacks F;
  F.cb = cb; F.user = user;
  if (opts->update_state_every_iteration) F.before_callback = [&]() { read_back(); };
  F.linearize = [&]() -> int { return SBA_OK; };  // (the work is in assemble: it needs the dense system's buffers)
  F.assemble = [&](double* d_H, double* d_g, double* d_cost) -> int {
    rig_lin_kernel<true><<<lin_blocks, 128, 0, st>>>(view(cur), loss, Jf.p, rf.p, cpart.p);
    rig_sum_kernel<<<1, 256, 0, st>>>(cpart.p, lin_blocks, 1, 1, d_cost);
    if (n > 0) {
      const int tiles = DivUp(n, 32);
      rig_jtj_kernel<<<dim3(tiles, tiles), dim3(32, 32), 0, st>>>(n, m, Jf.p, d_H, 0);
      rig_jtr_kernel<<<DivUp(n, 128), 128, 0, st>>>(n, m, Jf.p, rf.p, d_g, 0);
    }
    SBA_CUDA(cudaGetLastError());
    return SBA_OK;
  };
  F.candidate = [&](const double* d_delta, double* cost, double* step2, double* x2) -> int {
    const int o = cur ^ 1;
    rig_plus_kernel<<<plus_blocks, 256, 0, st>>>(view(cur), d_delta, q[o].p, t[o].p, cam[o].p, X[o].p, ppart.p);
    rig_sum_kernel<<<1, 256, 0, st>>>(ppart.p, plus_blocks, 2, 2, scal.p + 1);
    rig_lin_kernel<false><<<lin_blocks, 128, 0, st>>>(view(o), loss, nullptr, nullptr, cpart.p);
    rig_sum_kernel<<<1, 256, 0, st>>>(cpart.p, lin_blocks, 1, 1, scal.p);
    SBA_CUDA(cudaGetLastError());
    double h[3];
    SBA_CUDA(cudaMemcpyAsync(h, scal.p, sizeof(h), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    *cost = h[0]; *step2 = h[1]; *x2 = h[2];
    return SBA_OK;
  };
  F.accept = [&]() { cur ^= 1; };
  std::vector<sba_iteration_summary> rows;
  int rc = DenseLMSolve(G, *opts, F, st, summary, &rows);
  if (rc) return rc;
  if ((rc = read_back())) return rc;
  if (trace && trace_n) {
    const int k = std::min<int>((int)rows.size(), trace_cap);
    for (int i = 0; i < k; ++i) trace[i] = rows[i];
    *trace_n = k;
  }
  return SBA_OK;
}
