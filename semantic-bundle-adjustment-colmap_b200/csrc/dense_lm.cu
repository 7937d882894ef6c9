// Dense-normal-equations Levenberg-Marquardt shared by the two semantic adjusters (dense_lm.h).
// Control flow, constants and stopping rules: Ceres 2.1.0 TrustRegionMinimizer + LevenbergMarquardtStrategy as restated
// in SURVEY.md App. B.3/B.4 -- the same rules csrc/lm.cu follows for the point-eliminating solver.
#include "dense_lm.h"

#include <chrono>
#include <cmath>
#include <cstring>

namespace sba {

int DenseGroups::Upload(cudaStream_t st) {
  n_ent = (int)ent_off.size();
  n_groups = (int)g_e1.size();
  std::vector<int> table((size_t)std::max(1, n_ent) * std::max(1, n_ent), -1), single(std::max(1, n_ent), -1);
  for (int g = 0; g < n_groups; ++g) {
    if (g_e2[g] < 0) single[g_e1[g]] = g;
    else table[(size_t)g_e1[g] * n_ent + g_e2[g]] = g;
  }
  SBA_CUDA(d_ent_off.Upload(ent_off, st)); SBA_CUDA(d_ent_dim.Upload(ent_dim, st));
  SBA_CUDA(d_g_e1.Upload(g_e1, st)); SBA_CUDA(d_g_e2.Upload(g_e2, st));
  SBA_CUDA(d_table.Upload(table, st)); SBA_CUDA(d_single.Upload(single, st));
  SBA_CUDA(d_rec.Alloc((size_t)std::max(1, n_groups) * kGroupRec));
  SBA_CUDA(cudaStreamSynchronize(st));  // the host vectors above are locals of this call
  return SBA_OK;
}

__device__ __forceinline__ double RecJtJ(const double* rec, int r, int c) {  // lower triangle stored
  return r >= c ? rec[r * kGroupDim + c] : rec[c * kGroupDim + r];
}

// One CTA per entity tile (a, b), b <= a: the tile of H and (on the diagonal) the run of g.
__global__ void __launch_bounds__(128) dense_assemble_kernel(int n_ent, int n, const int* __restrict__ ent_off,
                                                             const int* __restrict__ ent_dim, const int* __restrict__ table,
                                                             const int* __restrict__ single, const int* __restrict__ g_e1,
                                                             const double* __restrict__ rec, double* __restrict__ H,
                                                             double* __restrict__ g) {
  const int a = blockIdx.x, b = blockIdx.y;
  if (b > a) return;
  const int da = ent_dim[a], db = ent_dim[b], oa = ent_off[a], ob = ent_off[b];
  if (da == 0 || db == 0) return;
  if (a != b) {
    const int g1 = table[(size_t)a * n_ent + b], g2 = table[(size_t)b * n_ent + a];
    for (int t = threadIdx.x; t < da * db; t += 128) {
      const int i = t / db, j = t - i * db;
      double v = 0.0;
      if (g1 >= 0) v += RecJtJ(rec + (size_t)g1 * kGroupRec, i, da + j);       // a is entity 1: rows [0, da), b's columns behind
      if (g2 >= 0) v += RecJtJ(rec + (size_t)g2 * kGroupRec, db + i, j);       // b is entity 1, a's rows behind
      H[(size_t)(oa + i) * n + ob + j] = v;
    }
    return;
  }
  for (int t = threadIdx.x; t < da * da + da; t += 128) {
    const bool grad = t >= da * da;
    const int i = grad ? t - da * da : t / da, j = grad ? 0 : t - i * da;
    if (!grad && j > i) continue;
    double v = 0.0;
    for (int c = 0; c < n_ent; ++c) {  // fixed order: bit-reproducible
      if (c == a) continue;
      const int g1 = table[(size_t)a * n_ent + c], g2 = table[(size_t)c * n_ent + a];
      if (g1 >= 0) { const double* r = rec + (size_t)g1 * kGroupRec; v += grad ? r[kGroupDim * kGroupDim + i] : RecJtJ(r, i, j); }
      if (g2 >= 0) {
        const double* r = rec + (size_t)g2 * kGroupRec;
        const int d1 = ent_dim[g_e1[g2]];
        v += grad ? r[kGroupDim * kGroupDim + d1 + i] : RecJtJ(r, d1 + i, d1 + j);
      }
    }
    const int gs = single[a];
    if (gs >= 0) { const double* r = rec + (size_t)gs * kGroupRec; v += grad ? r[kGroupDim * kGroupDim + i] : RecJtJ(r, i, j); }
    if (grad) g[oa + i] = v; else H[(size_t)(oa + i) * n + oa + j] = v;
  }
}

__global__ void __launch_bounds__(256) dense_cost_kernel(int n_groups, const double* __restrict__ rec, double* __restrict__ cost) {
  __shared__ double sm[8];
  double s = 0.0;
  for (int g = threadIdx.x; g < n_groups; g += 256) s += rec[(size_t)g * kGroupRec + kGroupDim * kGroupDim + kGroupDim];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) { double t = 0.0; for (int i = 0; i < 8; ++i) t += sm[i]; *cost = t; }
}

int AssembleDense(const DenseGroups& G, double* d_H, double* d_g, double* d_cost, cudaStream_t st, int64_t* launches) {
  if (G.n > 0) {
    SBA_CUDA(cudaMemsetAsync(d_g, 0, (size_t)G.n * sizeof(double), st));
    dense_assemble_kernel<<<dim3(G.n_ent, G.n_ent), 128, 0, st>>>(G.n_ent, G.n, G.d_ent_off.p, G.d_ent_dim.p, G.d_table.p,
                                                                   G.d_single.p, G.d_g_e1.p, G.d_rec.p, d_H, d_g);
  }
  dense_cost_kernel<<<1, 256, 0, st>>>(G.n_groups, G.d_rec.p, d_cost);
  if (launches) *launches += 2;
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

// ---- the LM step on the dense system --------------------------------------------------------------------------
__global__ void dense_scale_kernel(int n, const double* __restrict__ H, double* __restrict__ scale, int jacobi) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) scale[j] = jacobi ? 1.0 / (1.0 + sqrt(H[(size_t)j * n + j])) : 1.0;
}
// S_chol (lower, leading dimension ld) = S H S + D^2, rhs row = S g; diag refreshed from diag(S H S) unless reused
__global__ void __launch_bounds__(256) dense_system_kernel(int n, int ld, const double* __restrict__ H, const double* __restrict__ g,
                                                           const double* __restrict__ scale, double* __restrict__ diagv,
                                                           int refresh, double lo, double hi, double radius,
                                                           double* __restrict__ A, double* __restrict__ rhs) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= (int64_t)n * n) return;
  const int i = (int)(t / n), j = (int)(t - (int64_t)i * n);
  if (j > i) return;
  double v = H[(size_t)i * n + j] * scale[i] * scale[j];
  if (i == j) {
    double d = diagv[i];
    if (refresh) { d = fmin(fmax(v, lo), hi); diagv[i] = d; }
    v += d / radius;
    rhs[i] = g[i] * scale[i];
  }
  A[(size_t)i * ld + j] = v;
}
// u = S y, delta = -u, model cost change = g.u - u^T H u / 2, gradient max norm = max |g_j|  (one CTA; n is small)
__global__ void __launch_bounds__(256) dense_model_kernel(int n, const double* __restrict__ H, const double* __restrict__ g,
                                                          const double* __restrict__ scale, const double* __restrict__ y,
                                                          double* __restrict__ u, double* __restrict__ delta,
                                                          double* __restrict__ out /* [0] model cost change, [1] |g|_max */) {
  __shared__ double sm[8], sm2[8];
  for (int j = threadIdx.x; j < n; j += 256) { const double v = scale[j] * y[j]; u[j] = v; delta[j] = -v; }
  __syncthreads();
  double acc = 0.0, gm = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) {
    double t = 0.0;
    for (int j = 0; j < n; ++j) t += (j <= i ? H[(size_t)i * n + j] : H[(size_t)j * n + i]) * u[j];
    acc += u[i] * (g[i] - 0.5 * t);
    gm = fmax(gm, fabs(g[i]));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { acc += __shfl_down_sync(0xffffffffu, acc, o); gm = fmax(gm, __shfl_down_sync(0xffffffffu, gm, o)); }
  if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = acc; sm2[threadIdx.x >> 5] = gm; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, m = 0.0;
    for (int i = 0; i < 8; ++i) { a += sm[i]; m = fmax(m, sm2[i]); }
    out[0] = a; out[1] = m;
  }
}

int DenseLMSolve(DenseGroups& G, const sba_solver_options& o, const DenseLMCallbacks& F, cudaStream_t st, sba_summary* sum,
                 std::vector<sba_iteration_summary>* trace) {
  const auto t_start = std::chrono::steady_clock::now();
  const int n = G.n;
  trace->clear();
  // the dense factorisation of csrc/cholesky.cu / chol_small.cu works on a Problem: give it one that owns just the
  // linear system (its stream is a side stream ordered behind `st` by events)
  Problem lin;
  lin.n_cdim = n;
  SBA_CUDA(cudaGetDevice(&lin.device));
  if (cudaDeviceGetAttribute(&lin.n_sm, cudaDevAttrMultiProcessorCount, lin.device) != cudaSuccess) lin.n_sm = 148;
  SBA_CUDA(cudaStreamCreateWithFlags(&lin.stream, cudaStreamNonBlocking));
  const int ld = DenseLd(n);
  const size_t nn = (size_t)std::max(1, n);
  SBA_CUDA(lin.d_S.Alloc((nn + 1) * std::max(1, ld))); SBA_CUDA(lin.d_rhs.Alloc(nn)); SBA_CUDA(lin.d_yc.Alloc(nn));
  SBA_CUDA(lin.d_tmp_c.Alloc(nn)); SBA_CUDA(lin.d_Dc.Alloc(nn));
  SBA_CUDA(lin.d_scalars.Alloc(1));
  lin.chol_add_diag = false;
  DevBuf<double> H, g, scale, diagv, u, delta, scal;
  SBA_CUDA(H.Alloc(nn * nn)); SBA_CUDA(g.Alloc(nn)); SBA_CUDA(scale.Alloc(nn)); SBA_CUDA(diagv.Alloc(nn));
  SBA_CUDA(u.Alloc(nn)); SBA_CUDA(delta.Alloc(nn)); SBA_CUDA(scal.Alloc(8));
  cudaEvent_t ev;
  SBA_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  struct EvGuard { cudaEvent_t e; ~EvGuard() { cudaEventDestroy(e); } } ev_guard{ev};
  int64_t launches = 0;
  int rc;
  auto linearize = [&](double* cost, double* gmax) -> int {
    int r = F.linearize();
    if (r) return r;
    if (F.assemble) { if ((r = F.assemble(H.p, g.p, scal.p + 2))) return r; }
    else if ((r = AssembleDense(G, H.p, g.p, scal.p + 2, st, &launches))) return r;
    double h[1];
    SBA_CUDA(cudaMemcpyAsync(h, scal.p + 2, sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    *cost = h[0];
    (void)gmax;
    return SBA_OK;
  };
  double x_cost = 0.0, gmax = 0.0;
  if ((rc = linearize(&x_cost, &gmax))) return rc;
  if (n > 0) dense_scale_kernel<<<DivUp(n, 256), 256, 0, st>>>(n, H.p, scale.p, o.jacobi_scaling);
  {  // gradient max norm of the start point (unscaled J): through the model kernel with y = 0
    SBA_CUDA(cudaMemsetAsync(lin.d_yc.p, 0, nn * sizeof(double), st));
    dense_model_kernel<<<1, 256, 0, st>>>(n, H.p, g.p, scale.p, lin.d_yc.p, u.p, delta.p, scal.p);
    double h[2];
    SBA_CUDA(cudaMemcpyAsync(h, scal.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    gmax = h[1];
  }
  sum->fixed_cost = 0.0;
  sum->initial_cost = x_cost;
  sba_iteration_summary it0;
  std::memset(&it0, 0, sizeof(it0));
  it0.cost = x_cost; it0.gradient_max_norm = gmax; it0.trust_region_radius = o.initial_trust_region_radius; it0.eta = o.eta;
  trace->push_back(it0);
  int termination = SBA_NO_CONVERGENCE, iteration = 0, num_consecutive_invalid = 0;
  double radius = o.initial_trust_region_radius, decrease_factor = 2.0;
  bool reuse_diagonal = false;
  auto user_stop = [&](const sba_iteration_summary& row) -> bool {
    if (!F.cb) return false;
    if (F.before_callback) F.before_callback();
    const int r = F.cb(&row, F.user);
    if (r == 1) { termination = SBA_USER_SUCCESS; return true; }
    if (r == 2) { termination = SBA_USER_FAILURE; return true; }
    return false;
  };
  bool stop = user_stop(it0);
  while (!stop) {
    if (iteration >= o.max_num_iterations) { termination = SBA_NO_CONVERGENCE; break; }
    if (trace->back().gradient_max_norm <= o.gradient_tolerance) { termination = SBA_CONVERGENCE; break; }
    if (radius < o.min_trust_region_radius) { termination = SBA_CONVERGENCE; break; }
    ++iteration;
    sba_iteration_summary it;
    std::memset(&it, 0, sizeof(it));
    it.iteration = iteration; it.eta = o.eta;
    // ---- LevenbergMarquardtStrategy::ComputeStep ----
    double mcc = 0.0;
    bool solver_ok = true;
    if (n > 0) {
      dense_system_kernel<<<DivUp((int64_t)n * n, 256), 256, 0, st>>>(n, ld, H.p, g.p, scale.p, diagv.p, reuse_diagonal ? 0 : 1,
                                                                    o.min_lm_diagonal, o.max_lm_diagonal, radius, lin.d_S.p,
                                                                    lin.d_rhs.p);
      SBA_CUDA(cudaMemsetAsync(lin.d_scalars.p, 0, sizeof(DeviceScalars), st));
      SBA_CUDA(cudaEventRecord(ev, st));
      SBA_CUDA(cudaStreamWaitEvent(lin.stream, ev, 0));
      if ((rc = CholeskySolve(&lin))) return rc;
      SBA_CUDA(cudaEventRecord(ev, lin.stream));
      SBA_CUDA(cudaStreamWaitEvent(st, ev, 0));
      dense_model_kernel<<<1, 256, 0, st>>>(n, H.p, g.p, scale.p, lin.d_yc.p, u.p, delta.p, scal.p);
      launches += 2 + lin.launches; lin.launches = 0;
      double h[2];
      DeviceScalars hs;
      SBA_CUDA(cudaMemcpyAsync(h, scal.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
      SBA_CUDA(cudaMemcpyAsync(&hs, lin.d_scalars.p, sizeof(hs), cudaMemcpyDeviceToHost, st));
      SBA_CUDA(cudaStreamSynchronize(st));
      mcc = h[0];
      solver_ok = !hs.chol_fail && !hs.solver_failed && std::isfinite(mcc);
    }
    reuse_diagonal = true;
    it.linear_solver_iterations = 1;
    sum->num_linear_solver_iterations += 1;
    const bool step_valid = solver_ok && mcc > 0.0;
    it.step_is_valid = step_valid;
    if (!step_valid) {
      ++num_consecutive_invalid;
      radius = radius / decrease_factor; decrease_factor *= 2.0;
      it.cost = x_cost; it.gradient_max_norm = trace->back().gradient_max_norm; it.trust_region_radius = radius;
      trace->push_back(it);
      sum->num_unsuccessful_steps++;
      if (num_consecutive_invalid >= o.max_num_consecutive_invalid_steps) { termination = SBA_FAILURE; break; }
      if (user_stop(it)) break;
      continue;
    }
    num_consecutive_invalid = 0;
    double cand_cost = 0.0, step2 = 0.0, x2 = 0.0;
    if ((rc = F.candidate(delta.p, &cand_cost, &step2, &x2))) return rc;
    if (!std::isfinite(cand_cost)) cand_cost = 1.7976931348623157e308;
    it.step_norm = std::sqrt(step2);
    it.cost_change = x_cost - cand_cost;
    const double x_norm = std::sqrt(x2);
    // (Minimize() returns from both tests before the iteration's row is appended to Summary::iterations)
    if (it.step_norm <= o.parameter_tolerance * (x_norm + o.parameter_tolerance)) { termination = SBA_CONVERGENCE; break; }
    if (std::fabs(it.cost_change) <= o.function_tolerance * x_cost) { termination = SBA_CONVERGENCE; break; }
    const double rho = it.cost_change / mcc;
    it.relative_decrease = rho;
    if (rho > o.min_relative_decrease) {
      F.accept();
      double new_cost = 0.0, unused = 0.0;
      if ((rc = linearize(&new_cost, &unused))) return rc;
      x_cost = new_cost;
      SBA_CUDA(cudaMemsetAsync(lin.d_yc.p, 0, nn * sizeof(double), st));
      dense_model_kernel<<<1, 256, 0, st>>>(n, H.p, g.p, scale.p, lin.d_yc.p, u.p, delta.p, scal.p);
      double h[2];
      SBA_CUDA(cudaMemcpyAsync(h, scal.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
      SBA_CUDA(cudaStreamSynchronize(st));
      it.gradient_max_norm = h[1];
      it.step_is_successful = 1;
      const double t = 2.0 * rho - 1.0;
      radius = std::min(o.max_trust_region_radius, radius / std::max(1.0 / 3.0, 1.0 - t * t * t));
      decrease_factor = 2.0; reuse_diagonal = false;
      sum->num_successful_steps++;
    } else {
      radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;
      it.gradient_max_norm = trace->back().gradient_max_norm;
      sum->num_unsuccessful_steps++;
    }
    it.cost = x_cost; it.trust_region_radius = radius;
    trace->push_back(it);
    if (user_stop(trace->back())) break;
  }
  SBA_CUDA(cudaStreamSynchronize(st));
  SBA_CUDA(cudaStreamSynchronize(lin.stream));
  sum->termination_type = termination;
  sum->final_cost = x_cost;
  sum->num_iterations = (int)trace->size();
  sum->total_time_in_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
  return SBA_OK;
}

}  // namespace sba
