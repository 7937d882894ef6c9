// ORACLE (test infrastructure only -- never linked into the product library).
//
// CPU restatement of the two semantic adjusters as solves:
//   SemanticBundleAdjuster            /root/reference/src/optim/semantic_bundle_adjustment.cc:430-544, 646-906
//   GeometricSemanticBundleAdjuster   /root/reference/src/optim/geometric_semantic_bundle_adjustment.cc:481-604, 714-996,
//                                     1126-1213; CylinderBy2Points::ToCylinder /root/reference/src/util/cylinder_by_2_points.h:84-107
// Residual blocks, their order, the functor chosen per block and the manifolds follow those files; the numeric
// differentiation, the loss corrector and the trust-region loop follow Ceres 2.1.0 (un-vendored) --
// the same restatement as ba_oracle.cpp, on dense normal equations because these problems have no point blocks
// (App. B.1).  Written residual block by residual block (Jacobian rows in compact tangent columns), not pair by pair
// as the device does.  PARITY: the residual blocks are UNPINNED (the reference has no test for any of this); the
// trust-region loop (SolveDense) is pinned by the iteration table Ceres' documentation publishes for Powell's function
// (oracle_powell_solve below, tests/test_oracle_golden.py).
#include <omp.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <functional>
#include <limits>
#include <vector>

#include "ba_oracle.h"
#include "projection.h"
#include "rotation.h"
#include "semantic_oracle.h"

namespace oracle {

namespace {

// ceres loss functions, App. B.2 (same as ba_oracle.cpp)
void LossEvalD(int type, double a, double s, double rho[3]) {
  if (type == SBA_LOSS_SOFT_L1) {
    const double b = a * a, c = 1.0 / b, sum = 1.0 + s * c, tmp = std::sqrt(sum);
    rho[0] = 2.0 * b * (tmp - 1.0); rho[1] = std::max(std::numeric_limits<double>::min(), 1.0 / tmp); rho[2] = -(c * rho[1]) / (2.0 * sum);
  } else if (type == SBA_LOSS_CAUCHY) {
    const double b = a * a, c = 1.0 / b, sum = 1.0 + s * c, inv = 1.0 / sum;
    rho[0] = b * std::log(sum); rho[1] = std::max(std::numeric_limits<double>::min(), inv); rho[2] = -c * (inv * inv);
  } else {
    rho[0] = s; rho[1] = 1.0; rho[2] = 0.0;
  }
}

// in-place lower Cholesky of row-major n x n; false if not SPD
bool CholLower(std::vector<double>& A, int n) {
  for (int j = 0; j < n; ++j) {
    double d = A[(size_t)j * n + j];
    for (int k = 0; k < j; ++k) d -= A[(size_t)j * n + k] * A[(size_t)j * n + k];
    if (!(d > 0.0) || !std::isfinite(d)) return false;
    d = std::sqrt(d);
    A[(size_t)j * n + j] = d;
    for (int i = j + 1; i < n; ++i) {
      double s = A[(size_t)i * n + j];
      for (int k = 0; k < j; ++k) s -= A[(size_t)i * n + k] * A[(size_t)j * n + k];
      A[(size_t)i * n + j] = s / d;
    }
  }
  return true;
}
void CholSolve(const std::vector<double>& L, int n, std::vector<double>& b) {
  for (int i = 0; i < n; ++i) { double s = b[i]; for (int k = 0; k < i; ++k) s -= L[(size_t)i * n + k] * b[k]; b[i] = s / L[(size_t)i * n + i]; }
  for (int i = n - 1; i >= 0; --i) { double s = b[i]; for (int k = i + 1; k < n; ++k) s -= L[(size_t)k * n + i] * b[k]; b[i] = s / L[(size_t)i * n + i]; }
}

// One residual block after manifold projection: r, columns and their Jacobian entries (before the loss corrector)
struct Row { double r; int n; int col[16]; double j[16]; double w = 0.0; };  // w > 0: this row's own ScaledLoss weight

struct DenseProblem {
  int n = 0;  // tangent dimension
  // all residual blocks at the current point (with Jacobians) / the cost of a candidate point
  std::function<void(std::vector<Row>*)> rows;
  std::function<double(const std::vector<double>& delta, double* step2, double* x2)> candidate;  // also stores the candidate
  std::function<void()> accept;
  int loss_type = SBA_LOSS_TRIVIAL; double loss_scale = 1.0; double loss_weight = 1.0;  // rho(s) = weight * loss(s)
  // optional: H, g, cost of the current point for problems whose blocks are not single rows (the rig solve)
  std::function<void(std::vector<double>*, std::vector<double>*, double*)> linearize;
};

void Linearize(const DenseProblem& P, std::vector<double>* H, std::vector<double>* g, double* cost) {
  if (P.linearize) { P.linearize(H, g, cost); return; }
  std::vector<Row> rows;
  P.rows(&rows);
  const int n = P.n;
  H->assign((size_t)n * n, 0.0); g->assign(n, 0.0);
  double c = 0.0;
  for (const Row& b : rows) {
    const double s = b.r * b.r;
    double rho[3];
    LossEvalD(P.loss_type, P.loss_scale, s, rho);
    for (double& v : rho) v *= (b.w > 0.0 ? b.w : P.loss_weight);  // ScaledLoss
    c += 0.5 * rho[0];
    const double sq = std::sqrt(rho[1]);
    double jscale = sq, rscale = sq;
    if (!(s == 0.0 || rho[2] <= 0.0)) {
      const double alpha = 1.0 - std::sqrt(1.0 + 2.0 * s * rho[2] / rho[1]);
      jscale = sq * (1.0 - alpha); rscale = sq / (1.0 - alpha);
    }
    const double rc = b.r * rscale;
    for (int a = 0; a < b.n; ++a) {
      const double ja = b.j[a] * jscale;
      if (ja == 0.0) continue;
      (*g)[b.col[a]] += ja * rc;
      for (int e = 0; e < b.n; ++e) (*H)[(size_t)b.col[a] * n + b.col[e]] += ja * (b.j[e] * jscale);
    }
  }
  *cost = c;
}

int SolveDense(DenseProblem& P, const sba_solver_options& o, sba_summary* sum, std::vector<sba_iteration_summary>* trace) {
  const auto t_start = std::chrono::steady_clock::now();
  const int n = P.n;
  std::vector<double> H, g;
  double x_cost = 0.0;
  Linearize(P, &H, &g, &x_cost);
  sum->initial_cost = x_cost; sum->fixed_cost = 0.0;
  std::vector<double> scale(n, 1.0);
  if (o.jacobi_scaling) for (int j = 0; j < n; ++j) scale[j] = 1.0 / (1.0 + std::sqrt(H[(size_t)j * n + j]));
  auto gmax_of = [&]() { double m = 0.0; for (int j = 0; j < n; ++j) m = std::max(m, std::fabs(g[j])); return m; };
  trace->clear();
  sba_iteration_summary it0;
  std::memset(&it0, 0, sizeof(it0));
  it0.cost = x_cost; it0.gradient_max_norm = gmax_of(); it0.trust_region_radius = o.initial_trust_region_radius; it0.eta = o.eta;
  trace->push_back(it0);
  double radius = o.initial_trust_region_radius, decrease_factor = 2.0;
  bool reuse_diagonal = false;
  std::vector<double> diagv(n, 0.0);
  int termination = SBA_NO_CONVERGENCE, iteration = 0, num_consecutive_invalid = 0;
  while (true) {
    if (iteration >= o.max_num_iterations) { termination = SBA_NO_CONVERGENCE; break; }
    if (trace->back().gradient_max_norm <= o.gradient_tolerance) { termination = SBA_CONVERGENCE; break; }
    if (radius < o.min_trust_region_radius) { termination = SBA_CONVERGENCE; break; }
    ++iteration;
    sba_iteration_summary it;
    std::memset(&it, 0, sizeof(it));
    it.iteration = iteration; it.eta = o.eta;
    std::vector<double> A((size_t)n * n, 0.0), y(n, 0.0);
    for (int i = 0; i < n; ++i)
      for (int j = 0; j <= i; ++j) A[(size_t)i * n + j] = H[(size_t)i * n + j] * scale[i] * scale[j];
    if (!reuse_diagonal) for (int j = 0; j < n; ++j) diagv[j] = std::min(std::max(A[(size_t)j * n + j], o.min_lm_diagonal), o.max_lm_diagonal);
    for (int j = 0; j < n; ++j) { A[(size_t)j * n + j] += diagv[j] / radius; y[j] = g[j] * scale[j]; }
    bool solver_ok = CholLower(A, n);
    if (solver_ok) CholSolve(A, n, y);
    reuse_diagonal = true;
    it.linear_solver_iterations = 1;
    sum->num_linear_solver_iterations += 1;
    std::vector<double> u(n), delta(n);
    double mcc = 0.0;
    if (solver_ok) {
      for (int j = 0; j < n; ++j) { u[j] = scale[j] * y[j]; delta[j] = -u[j]; }
      for (int i = 0; i < n; ++i) {
        double t = 0.0;
        for (int j = 0; j < n; ++j) t += H[(size_t)i * n + j] * u[j];
        mcc += u[i] * (g[i] - 0.5 * t);
      }
      solver_ok = std::isfinite(mcc);
    }
    const bool step_valid = solver_ok && mcc > 0.0;
    it.step_is_valid = step_valid;
    if (!step_valid) {
      ++num_consecutive_invalid;
      radius = radius / decrease_factor; decrease_factor *= 2.0;
      it.cost = x_cost; it.gradient_max_norm = trace->back().gradient_max_norm; it.trust_region_radius = radius;
      trace->push_back(it);
      sum->num_unsuccessful_steps++;
      if (num_consecutive_invalid >= o.max_num_consecutive_invalid_steps) { termination = SBA_FAILURE; break; }
      continue;
    }
    num_consecutive_invalid = 0;
    double step2 = 0.0, x2 = 0.0;
    double cand_cost = P.candidate(delta, &step2, &x2);
    if (!std::isfinite(cand_cost)) cand_cost = std::numeric_limits<double>::max();
    it.step_norm = std::sqrt(step2);
    it.cost_change = x_cost - cand_cost;
    const double x_norm = std::sqrt(x2);
    if (it.step_norm <= o.parameter_tolerance * (x_norm + o.parameter_tolerance)) { termination = SBA_CONVERGENCE; break; }
    if (std::fabs(it.cost_change) <= o.function_tolerance * x_cost) { termination = SBA_CONVERGENCE; break; }
    const double rho = it.cost_change / mcc;
    it.relative_decrease = rho;
    if (rho > o.min_relative_decrease) {
      P.accept();
      Linearize(P, &H, &g, &x_cost);
      it.gradient_max_norm = gmax_of();
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
  }
  sum->termination_type = termination;
  sum->final_cost = x_cost;
  sum->num_iterations = (int)trace->size();
  sum->total_time_in_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
  return SBA_OK;
}

void NormalizeQ(double* q) {
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n == 0.0) { q[0] = 1.0; q[1] = q[2] = q[3] = 0.0; return; }
  for (int a = 0; a < 4; ++a) q[a] /= n;
}
// J_local(3) = J_ambient(4) * PlusJacobian(q)
void QuatTangentRow(const double* q, const double* ja, double* jl) {
  double PJ[12];
  QuaternionPlusJacobian(q, PJ);
  for (int j = 0; j < 3; ++j) { double a = 0.0; for (int k = 0; k < 4; ++k) a += ja[k] * PJ[k * 3 + j]; jl[j] = a; }
}

struct PoseVars {  // tangent columns of the camera poses
  int n_img = 0;
  std::vector<uint8_t> pc, tm;
  std::vector<int> off, dim;
  int Build(int n, const uint8_t* pose_const, const uint8_t* tvec_mask, int start, sba_summary* s, bool any) {
    n_img = n; pc.assign(n, 0); tm.assign(n, 0); off.assign(n, 0); dim.assign(n, 0);
    int o = start;
    for (int i = 0; i < n; ++i) {
      pc[i] = pose_const ? pose_const[i] : 0; tm[i] = tvec_mask ? tvec_mask[i] : 0;
      int d = 0;
      if (!pc[i] && any) { d = 3; for (int a = 0; a < 3; ++a) d += !((tm[i] >> a) & 1); }
      off[i] = o; dim[i] = d; o += d;
      if (d) { s->num_parameter_blocks_reduced += 1 + (d > 3); s->num_parameters_reduced += 4 + (d > 3 ? 3 : 0); s->num_effective_parameters_reduced += d; }
    }
    return o;
  }
  // appends the pose columns of image i for the ambient row (jq[4], jt[3]) evaluated at q
  void Append(int i, const double* q, const double* jq, const double* jt, Row* row) const {
    if (!dim[i]) return;
    double jl[3];
    QuatTangentRow(q, jq, jl);
    for (int a = 0; a < 3; ++a) { row->col[row->n] = off[i] + a; row->j[row->n++] = jl[a]; }
    int c = 3;
    for (int a = 0; a < 3; ++a) if (!((tm[i] >> a) & 1)) { row->col[row->n] = off[i] + c++; row->j[row->n++] = jt[a]; }
  }
  void Plus(int i, const double* q, const double* t, const std::vector<double>& delta, double* qn, double* tn, double* step2, double* x2) const {
    for (int a = 0; a < 4; ++a) qn[a] = q[a];
    for (int a = 0; a < 3; ++a) tn[a] = t[a];
    if (!dim[i]) return;
    QuaternionPlus(q, &delta[off[i]], qn);
    int c = 3, free_t = 0;
    for (int a = 0; a < 3; ++a) if (!((tm[i] >> a) & 1)) { tn[a] += delta[off[i] + c++]; free_t++; }
    for (int a = 0; a < 4; ++a) { *step2 += (qn[a] - q[a]) * (qn[a] - q[a]); *x2 += q[a] * q[a]; }
    if (free_t > 0) for (int a = 0; a < 3; ++a) { *step2 += (tn[a] - t[a]) * (tn[a] - t[a]); *x2 += t[a] * t[a]; }
  }
};

int Policy(const sba_solver_options& o, int n_img) {
  if (o.linear_solver_type != SBA_SOLVER_AUTO) return o.linear_solver_type;
  const int nimg = o.num_config_images > 0 ? o.num_config_images : n_img;
  return nimg <= 50 ? SBA_DENSE_SCHUR : (nimg <= 1000 ? SBA_SPARSE_SCHUR : SBA_ITERATIVE_SCHUR);
}

// CylinderBy2Points(t1, t2, r).ToCylinder() -> q(4) t(3) r h
void TwoPointsToCylinder(const double* p7, double* c9) {
  double radius = p7[6];
  if (radius <= 0) radius = 1e-4;
  double d[3] = {p7[3] - p7[0], p7[4] - p7[1], p7[5] - p7[2]};
  const double dn = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  d[0] /= dn; d[1] /= dn; d[2] /= dn;
  double axis[3] = {-d[1], d[0], 0.0};
  const double an = std::sqrt(axis[0] * axis[0] + axis[1] * axis[1] + axis[2] * axis[2]);
  if (std::fabs(an) < 1e-10) { axis[0] = 1.0; axis[1] = 0.0; axis[2] = 0.0; }
  else { axis[0] /= an; axis[1] /= an; axis[2] /= an; }
  const double angle = std::acos(d[2]);
  const double aa[3] = {angle * axis[0], angle * axis[1], angle * axis[2]};
  const double th2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
  if (th2 > 0.0) {
    const double th = std::sqrt(th2), half = th * 0.5, k = std::sin(half) / th;
    c9[0] = std::cos(half); c9[1] = aa[0] * k; c9[2] = aa[1] * k; c9[3] = aa[2] * k;
  } else {
    c9[0] = 1.0; c9[1] = aa[0] * 0.5; c9[2] = aa[1] * 0.5; c9[3] = aa[2] * 0.5;
  }
  c9[4] = p7[0]; c9[5] = p7[1]; c9[6] = p7[2];
  c9[7] = radius;
  const double h[3] = {p7[0] - p7[3], p7[1] - p7[4], p7[2] - p7[5]};
  c9[8] = std::sqrt(h[0] * h[0] + h[1] * h[1] + h[2] * h[2]);
}

}  // namespace
}  // namespace oracle

using namespace oracle;

extern "C" {

int oracle_semantic_solve(const sba_semantic_problem* d, const sba_solver_options* o, int threads, sba_summary* summary,
                          sba_iteration_summary* trace_out, int cap, int* n_trace) {
  if (threads <= 0) threads = omp_get_max_threads();
  std::memset(summary, 0, sizeof(*summary));
  const int n_img = d->num_images, H = d->height, W = d->width;
  const size_t HW = (size_t)H * W;
  std::vector<double> q(d->image_qvec, d->image_qvec + 4 * (size_t)n_img), t(d->image_tvec, d->image_tvec + 3 * (size_t)n_img), qc, tc;
  if (n_img >= 2) for (int i = 0; i < n_img; ++i) NormalizeQ(&q[4 * i]);
  PoseVars V;
  DenseProblem P;
  P.n = V.Build(n_img, d->image_pose_constant, d->image_tvec_constant_mask, 0, summary, true);
  P.loss_type = o->loss_function_type; P.loss_scale = o->loss_function_scale;
  // residual blocks in the reference order: id1 outer, id2 inner, y outer, x inner (:656-661, :796-799)
  struct Blk { int i1, i2, x, y; };
  std::vector<Blk> blocks;
  for (int i1 = 0; i1 < n_img; ++i1)
    for (int i2 = 0; i2 < n_img; ++i2) {
      if (i1 == i2 || (V.pc[i1] && V.pc[i2])) continue;
      for (int y = 0; y < H; y += d->pixel_step)
        for (int x = 0; x < W; x += d->pixel_step) {
          const float depth = d->depth_maps[i1 * HW + (size_t)y * W + x];
          if (depth < 1e-4) continue;
          blocks.push_back(Blk{i1, i2, x, y});
        }
    }
  summary->num_residual_blocks = summary->num_residuals = summary->num_residuals_reduced = (int)blocks.size();
  summary->linear_solver_type_used = Policy(*o, n_img);
  if (blocks.empty()) return SBA_ERR_EMPTY_PROBLEM;
  auto residual = [&](const Blk& b, const double* x14) {
    SemanticBlock sb;
    sb.x1 = b.x; sb.y1 = b.y;
    sb.depth1 = d->depth_maps[b.i1 * HW + (size_t)b.y * W + b.x];
    sb.sem1 = d->semantic_maps[b.i1 * HW + (size_t)b.y * W + b.x];
    sb.depth2 = d->depth_maps + b.i2 * HW; sb.sem2 = d->semantic_maps + b.i2 * HW;
    sb.H = H; sb.W = W; sb.thr = d->depth_error_threshold;
    int st;
    return SemanticPairResidual(sb, x14, x14 + 4, x14 + 7, x14 + 11, d->camera_params + 3 * b.i1, d->camera_params + 3 * b.i2, &st);
  };
  auto load = [&](const std::vector<double>& Q, const std::vector<double>& T, const Blk& b, double* x) {
    for (int j = 0; j < 4; ++j) { x[j] = Q[4 * b.i1 + j]; x[7 + j] = Q[4 * b.i2 + j]; }
    for (int j = 0; j < 3; ++j) { x[4 + j] = T[3 * b.i1 + j]; x[11 + j] = T[3 * b.i2 + j]; }
  };
  P.rows = [&](std::vector<Row>* rows) {
    rows->assign(blocks.size(), Row());
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int64_t k = 0; k < (int64_t)blocks.size(); ++k) {
      const Blk& b = blocks[k];
      double x[14], ja[14];
      load(q, t, b, x);
      Row& row = (*rows)[k];
      row.n = 0;
      row.r = residual(b, x);
      for (int j = 0; j < 14; ++j) {
        ja[j] = 0.0;
        if ((j < 7 ? V.pc[b.i1] : V.pc[b.i2])) continue;  // a constant pose is not a parameter of the functor (:829-877)
        const double xj = x[j], delta = NumDiffDelta(xj, d->relative_step_size);
        x[j] = xj + delta; const double fp = residual(b, x);
        x[j] = xj - delta; const double fm = residual(b, x);
        x[j] = xj;
        ja[j] = (fp - fm) * ((1.0 / delta) / 2.0);
      }
      V.Append(b.i1, &q[4 * b.i1], ja, ja + 4, &row);
      V.Append(b.i2, &q[4 * b.i2], ja + 7, ja + 11, &row);
    }
  };
  P.candidate = [&](const std::vector<double>& delta, double* step2, double* x2) {
    qc = q; tc = t;
    for (int i = 0; i < n_img; ++i) V.Plus(i, &q[4 * i], &t[3 * i], delta, &qc[4 * i], &tc[3 * i], step2, x2);
    const int nt = std::max(1, threads);
    std::vector<double> part(nt * 8, 0.0);
#pragma omp parallel for num_threads(nt) schedule(static)
    for (int64_t k = 0; k < (int64_t)blocks.size(); ++k) {
      double x[14];
      load(qc, tc, blocks[k], x);
      const double r = residual(blocks[k], x);
      double rho[3];
      LossEvalD(P.loss_type, P.loss_scale, r * r, rho);
      part[8 * omp_get_thread_num()] += 0.5 * rho[0];
    }
    double c = 0.0;
    for (int i = 0; i < nt; ++i) c += part[8 * i];
    return c;
  };
  P.accept = [&]() { q.swap(qc); t.swap(tc); };
  std::vector<sba_iteration_summary> trace;
  const int rc = SolveDense(P, *o, summary, &trace);
  if (trace_out && n_trace) {
    const int m = std::min<int>(cap, (int)trace.size());
    for (int i = 0; i < m; ++i) trace_out[i] = trace[i];
    *n_trace = m;
  }
  std::memcpy(d->image_qvec, q.data(), q.size() * sizeof(double));
  std::memcpy(d->image_tvec, t.data(), t.size() * sizeof(double));
  return rc;
}

int oracle_cylinder_solve(const sba_cylinder_problem* d, const sba_solver_options* o, int threads, sba_summary* summary,
                          sba_iteration_summary* trace_out, int cap, int* n_trace) {
  if (threads <= 0) threads = omp_get_max_threads();
  std::memset(summary, 0, sizeof(*summary));
  const int n_img = d->num_images, n_cyl = d->num_cylinders, H = d->height, W = d->width, kind = d->cylinder_kind;
  const size_t HW = (size_t)H * W;
  const bool refine_geometry = d->refine_geometry != 0;
  std::vector<double> q(d->image_qvec, d->image_qvec + 4 * (size_t)n_img), t(d->image_tvec, d->image_tvec + 3 * (size_t)n_img), qc, tc;
  std::vector<double> cyl(d->cylinder_params, d->cylinder_params + 9 * (size_t)n_cyl), cylc;
  PoseVars V;
  DenseProblem P;
  // optional landmark blocks (:729-794, 1361-1469): ordinary SIMPLE_PINHOLE reprojection blocks, constant intrinsics,
  // ScaledLoss(nullptr, landmark_weight); the points join the parameter blocks
  const int64_t n_lm = d->num_landmarks;
  const int n_pts = n_lm > 0 ? d->num_points : 0;
  std::vector<uint8_t> has_lm(n_img, 0), pc2(n_img, 0);
  for (int64_t k = 0; k < n_lm; ++k) has_lm[d->landmark_image[k]] = 1;
  for (int i = 0; i < n_img; ++i) pc2[i] = (d->image_pose_constant && d->image_pose_constant[i]) || !(n_cyl > 0 || has_lm[i]);
  int off = V.Build(n_img, pc2.data(), d->image_tvec_constant_mask, 0, summary, true);
  for (int i = 0; i < n_img; ++i) if (has_lm[i]) NormalizeQ(&q[4 * i]);  // image.NormalizeQvec(), :1371
  std::vector<double> X, Xc;
  if (n_lm > 0) X.assign(d->points, d->points + 3 * (size_t)n_pts);
  struct Blk { int im, cy; };
  std::vector<Blk> blocks;
  for (int i = 0; i < n_img; ++i) {
    if (V.pc[i] && !refine_geometry) continue;
    NormalizeQ(&q[4 * i]);
    for (int c = 0; c < n_cyl; ++c) blocks.push_back(Blk{i, c});
  }
  const int cdim = (refine_geometry && !blocks.empty()) ? (kind == 0 ? 8 : 7) : 0;
  std::vector<int> coff(n_cyl, 0);
  for (int c = 0; c < n_cyl; ++c) {
    coff[c] = off; off += cdim;
    if (cdim) {
      summary->num_parameter_blocks_reduced += kind == 0 ? 4 : 3;
      summary->num_parameters_reduced += kind == 0 ? 9 : 7;
      summary->num_effective_parameters_reduced += cdim;
    }
  }
  std::vector<int> xcol(std::max(1, n_pts), -1);
  {
    std::vector<uint8_t> used(std::max(1, n_pts), 0);
    for (int64_t k = 0; k < n_lm; ++k) used[d->landmark_point[k]] = 1;
    for (int p = 0; p < n_pts; ++p) {
      if (!used[p] || (d->point_constant && d->point_constant[p])) continue;
      xcol[p] = off; off += 3;
      summary->num_parameter_blocks_reduced += 1; summary->num_parameters_reduced += 3; summary->num_effective_parameters_reduced += 3;
    }
  }
  P.n = off;
  P.loss_weight = 1.0 / n_img;  // ScaledLoss(nullptr, 1 / N_images)
  summary->num_residual_blocks = (int)blocks.size() + (int)n_lm;
  summary->num_residuals = summary->num_residuals_reduced = (int)blocks.size() + 2 * (int)n_lm;
  summary->linear_solver_type_used = n_img <= 50 ? SBA_DENSE_SCHUR : (n_img <= 1000 ? SBA_SPARSE_SCHUR : SBA_ITERATIVE_SCHUR);
  if (blocks.empty() && n_lm == 0) return SBA_ERR_EMPTY_PROBLEM;
  std::vector<int> pos(n_img, 0);
  for (int i = 0; i < n_img; ++i) for (size_t j = 0; j < HW; ++j) pos[i] += d->masks[i * HW + j] != 0;
  const int nc = kind == 0 ? 16 : 14;
  auto residual = [&](const Blk& b, const double* x16, std::vector<uint8_t>* scratch) {
    double x[16];
    for (int j = 0; j < 7; ++j) x[j] = x16[j];
    if (kind == 1) TwoPointsToCylinder(x16 + 7, x + 7); else for (int j = 0; j < 9; ++j) x[7 + j] = x16[7 + j];
    return CylinderResidual(x, d->camera_params + 3 * b.im, d->masks + b.im * HW, pos[b.im], H, W, scratch);
  };
  auto load = [&](const std::vector<double>& Q, const std::vector<double>& T, const std::vector<double>& C, const Blk& b, double* x) {
    for (int j = 0; j < 4; ++j) x[j] = Q[4 * b.im + j];
    for (int j = 0; j < 3; ++j) x[4 + j] = T[3 * b.im + j];
    for (int j = 0; j < 9; ++j) x[7 + j] = C[9 * b.cy + j];
  };
  const double ident_q[4] = {1.0, 0.0, 0.0, 0.0}, ident_t[3] = {0.0, 0.0, 0.0};
  auto landmark = [&](const std::vector<double>& Q, const std::vector<double>& T, const std::vector<double>& PX, int64_t k, double* r, double* J) {
    const int i = d->landmark_image[k], p = d->landmark_point[k];
    double prm[SBA_MAX_CAMERA_PARAMS] = {0};
    for (int a = 0; a < 3; ++a) prm[a] = d->camera_params[3 * (size_t)i + a];
    // the ordinary functor = the rig functor on a constant identity rig pose (exactly: q (x) 1 = q, R(q) 0 + t = t)
    EvalRigObservation(SBA_SIMPLE_PINHOLE, ident_q, ident_t, &Q[4 * (size_t)i], &T[3 * (size_t)i], &PX[3 * (size_t)p], prm,
                       d->landmark_xy + 2 * k, r, J);
  };
  P.rows = [&](std::vector<Row>* rows) {
    rows->assign(blocks.size() + 2 * (size_t)n_lm, Row());
    for (int64_t k = 0; k < n_lm; ++k) {
      double r[2], J[2 * 29];
      landmark(q, t, X, k, r, J);
      const int i = d->landmark_image[k], p = d->landmark_point[k];
      for (int rr = 0; rr < 2; ++rr) {
        Row& row = (*rows)[blocks.size() + 2 * (size_t)k + rr];
        row.n = 0; row.r = r[rr]; row.w = d->landmark_weight;
        const double* A = J + 29 * rr;
        V.Append(i, &q[4 * (size_t)i], A + 7, A + 11, &row);
        if (xcol[p] >= 0) for (int a = 0; a < 3; ++a) { row.col[row.n] = xcol[p] + a; row.j[row.n++] = A[14 + a]; }
      }
    }
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1)
    for (int64_t k = 0; k < (int64_t)blocks.size(); ++k) {
      std::vector<uint8_t> scratch;
      const Blk& b = blocks[k];
      double x[16], ja[16];
      load(q, t, cyl, b, x);
      Row& row = (*rows)[k];
      row.n = 0;
      row.r = residual(b, x, &scratch);
      for (int j = 0; j < 16; ++j) {
        ja[j] = 0.0;
        if (j >= nc || (j < 7 ? V.pc[b.im] != 0 : !refine_geometry)) continue;
        const double xj = x[j], delta = NumDiffDelta(xj, d->relative_step_size);
        x[j] = xj + delta; const double fp = residual(b, x, &scratch);
        x[j] = xj - delta; const double fm = residual(b, x, &scratch);
        x[j] = xj;
        ja[j] = (fp - fm) * ((1.0 / delta) / 2.0);
      }
      V.Append(b.im, &q[4 * b.im], ja, ja + 4, &row);
      if (cdim) {
        if (kind == 0) {
          double jl[3];
          QuatTangentRow(&cyl[9 * b.cy], ja + 7, jl);
          for (int a = 0; a < 3; ++a) { row.col[row.n] = coff[b.cy] + a; row.j[row.n++] = jl[a]; }
          for (int a = 0; a < 5; ++a) { row.col[row.n] = coff[b.cy] + 3 + a; row.j[row.n++] = ja[11 + a]; }
        } else {
          for (int a = 0; a < 7; ++a) { row.col[row.n] = coff[b.cy] + a; row.j[row.n++] = ja[7 + a]; }
        }
      }
    }
  };
  P.candidate = [&](const std::vector<double>& delta, double* step2, double* x2) {
    qc = q; tc = t; cylc = cyl; Xc = X;
    for (int p = 0; p < n_pts; ++p) if (xcol[p] >= 0) for (int a = 0; a < 3; ++a) {
      const double v = X[3 * (size_t)p + a];
      Xc[3 * (size_t)p + a] = v + delta[xcol[p] + a];
      const double e = Xc[3 * (size_t)p + a] - v; *step2 += e * e; *x2 += v * v;
    }
    for (int i = 0; i < n_img; ++i) V.Plus(i, &q[4 * i], &t[3 * i], delta, &qc[4 * i], &tc[3 * i], step2, x2);
    if (cdim)
      for (int c = 0; c < n_cyl; ++c) {
        const double* x = &cyl[9 * c];
        double* xn = &cylc[9 * c];
        const double* dl = &delta[coff[c]];
        if (kind == 0) {
          QuaternionPlus(x, dl, xn);
          for (int a = 0; a < 5; ++a) xn[4 + a] = x[4 + a] + dl[3 + a];
          xn[7] = std::max(xn[7], 0.0);  // radius lower bound 0, enforced in Plus (App. B.4)
          for (int a = 0; a < 9; ++a) { *step2 += (xn[a] - x[a]) * (xn[a] - x[a]); *x2 += x[a] * x[a]; }
        } else {
          for (int a = 0; a < 7; ++a) xn[a] = x[a] + dl[a];
          xn[6] = std::max(xn[6], 0.0);
          for (int a = 0; a < 7; ++a) { *step2 += (xn[a] - x[a]) * (xn[a] - x[a]); *x2 += x[a] * x[a]; }
        }
      }
    double c = 0.0;
    std::vector<uint8_t> scratch;
    for (const Blk& b : blocks) {
      double x[16];
      load(qc, tc, cylc, b, x);
      const double r = residual(b, x, &scratch);
      c += 0.5 * P.loss_weight * r * r;
    }
    for (int64_t k = 0; k < n_lm; ++k) {
      double r[2];
      landmark(qc, tc, Xc, k, r, nullptr);
      c += 0.5 * d->landmark_weight * (r[0] * r[0] + r[1] * r[1]);
    }
    return c;
  };
  P.accept = [&]() { q.swap(qc); t.swap(tc); cyl.swap(cylc); X.swap(Xc); };
  std::vector<sba_iteration_summary> trace;
  const int rc = SolveDense(P, *o, summary, &trace);
  if (trace_out && n_trace) {
    const int m = std::min<int>(cap, (int)trace.size());
    for (int i = 0; i < m; ++i) trace_out[i] = trace[i];
    *n_trace = m;
  }
  std::memcpy(d->image_qvec, q.data(), q.size() * sizeof(double));
  std::memcpy(d->image_tvec, t.data(), t.size() * sizeof(double));
  if (n_cyl > 0) std::memcpy(d->cylinder_params, cyl.data(), cyl.size() * sizeof(double));
  if (n_lm > 0) std::memcpy(d->points, X.data(), X.size() * sizeof(double));
  return rc;
}

// ---- RigBundleAdjuster as a solve (/root/reference/src/optim/bundle_adjustment.cc:789-1140) ---------------------------
// The flat problem of sba_rig_solve (include/sba_b200.h): two-row blocks over (rig pose, relative pose, point,
// intrinsics) with the Jet Jacobian of the rig functor (EvalRigObservation, ba_oracle.cpp), QuaternionManifold /
// SubsetManifold columns, the loss corrector on the 2-vector, dense normal equations over all blocks.
int oracle_rig_solve(sba_rig_problem* d, const sba_solver_options* o, int threads, sba_summary* summary,
                     sba_iteration_summary* trace, int32_t trace_cap, int32_t* trace_n) {
  (void)threads;
  std::memset(summary, 0, sizeof(*summary));
  if (trace_n) *trace_n = 0;
  const int n_rig = d->num_rig_poses, n_rel = d->num_rel_poses, n_pose = n_rig + n_rel, n_cam = d->num_cameras, n_pt = d->num_points;
  const int64_t n_obs = d->num_blocks;
  if (n_obs == 0) return SBA_ERR_EMPTY_PROBLEM;
  std::vector<uint8_t> use_pose(n_pose, 0), use_cam(n_cam, 0), use_pt(n_pt, 0);
  for (int64_t k = 0; k < n_obs; ++k) {
    use_pose[d->block_rig[k]] = use_pose[n_rig + d->block_rel[k]] = 1;
    use_cam[d->block_camera[k]] = 1; use_pt[d->block_point[k]] = 1;
  }
  std::vector<int> qcol(n_pose, -1), tcol(3 * (size_t)n_pose, -1), pcol(12 * (size_t)std::max(1, n_cam), -1), xcol(std::max(1, n_pt), -1);
  int n = 0;
  for (int i = 0; i < n_pose; ++i) {
    const bool rig = i < n_rig;
    const int j = rig ? i : i - n_rig;
    const bool konst = rig ? (d->rig_pose_constant && d->rig_pose_constant[j]) : (d->rel_pose_constant && d->rel_pose_constant[j]);
    if (!use_pose[i] || konst) continue;
    qcol[i] = n; n += 3;
    summary->num_parameter_blocks_reduced++; summary->num_parameters_reduced += 4;
    const unsigned mask = (!rig && d->rel_tvec_constant_mask) ? d->rel_tvec_constant_mask[j] & 7u : 0u;
    if (mask != 7u) {
      for (int c = 0; c < 3; ++c) if (!(mask >> c & 1u)) tcol[3 * (size_t)i + c] = n++;
      summary->num_parameter_blocks_reduced++; summary->num_parameters_reduced += 3;
    }
  }
  for (int c = 0; c < n_cam; ++c) {
    if (!use_cam[c]) continue;
    const int np = NumParams(d->camera_model[c]);
    const unsigned mask = d->camera_param_constant_mask ? d->camera_param_constant_mask[c] : 0u;
    bool any = false;
    for (int j = 0; j < np; ++j) if (!(mask >> j & 1u)) { pcol[12 * (size_t)c + j] = n++; any = true; }
    if (any) { summary->num_parameter_blocks_reduced++; summary->num_parameters_reduced += np; }
  }
  for (int p = 0; p < n_pt; ++p) {
    if (!use_pt[p] || (d->point_constant && d->point_constant[p])) continue;
    xcol[p] = n; n += 3;
    summary->num_parameter_blocks_reduced++; summary->num_parameters_reduced += 3;
  }
  summary->num_residual_blocks = (int)n_obs; summary->num_residuals = summary->num_residuals_reduced = 2 * (int)n_obs;
  summary->num_effective_parameters_reduced = n;
  summary->linear_solver_type_used = SBA_DENSE_SCHUR;
  // current and candidate state
  struct State { std::vector<double> q, t, cam, X; };
  State cur, cand;
  cur.q.resize(4 * (size_t)n_pose); cur.t.resize(3 * (size_t)n_pose);
  for (int i = 0; i < n_rig; ++i) { std::memcpy(&cur.q[4 * (size_t)i], d->rig_qvec + 4 * (size_t)i, 32); std::memcpy(&cur.t[3 * (size_t)i], d->rig_tvec + 3 * (size_t)i, 24); }
  for (int i = 0; i < n_rel; ++i) { std::memcpy(&cur.q[4 * (size_t)(n_rig + i)], d->rel_qvec + 4 * (size_t)i, 32); std::memcpy(&cur.t[3 * (size_t)(n_rig + i)], d->rel_tvec + 3 * (size_t)i, 24); }
  cur.cam.assign(d->camera_params, d->camera_params + 12 * (size_t)n_cam);
  cur.X.assign(d->points, d->points + 3 * (size_t)n_pt);
  cand = cur;
  auto block = [&](const State& S, int64_t k, double* r, double* J) {
    const int a = d->block_rig[k], b = n_rig + d->block_rel[k], c = d->block_camera[k], p = d->block_point[k];
    EvalRigObservation(d->camera_model[c], &S.q[4 * (size_t)a], &S.t[3 * (size_t)a], &S.q[4 * (size_t)b], &S.t[3 * (size_t)b],
                       &S.X[3 * (size_t)p], &S.cam[12 * (size_t)c], d->block_xy + 2 * k, r, J);
  };
  auto cost_of = [&](const State& S) {
    double c = 0.0;
    for (int64_t k = 0; k < n_obs; ++k) {
      double r[2], rho[3];
      block(S, k, r, nullptr);
      LossEvalD(o->loss_function_type, o->loss_function_scale, r[0] * r[0] + r[1] * r[1], rho);
      c += 0.5 * rho[0];
    }
    return c;
  };
  DenseProblem P;
  P.n = n;
  P.linearize = [&](std::vector<double>* H, std::vector<double>* g, double* cost) {
    H->assign((size_t)n * n, 0.0); g->assign(n, 0.0);
    double c = 0.0;
    constexpr int W = 4 + 3 + 4 + 3 + 3 + SBA_MAX_CAMERA_PARAMS;
    for (int64_t k = 0; k < n_obs; ++k) {
      const int a = d->block_rig[k], b = n_rig + d->block_rel[k], cm = d->block_camera[k], p = d->block_point[k];
      double r[2], J[2 * W], rho[3];
      block(cur, k, r, J);
      const double s = r[0] * r[0] + r[1] * r[1];
      LossEvalD(o->loss_function_type, o->loss_function_scale, s, rho);
      c += 0.5 * rho[0];
      const double sq = std::sqrt(rho[1]);
      double rscale = sq, asn = 0.0;
      if (!(s == 0.0 || rho[2] <= 0.0)) {
        const double alpha = 1.0 - std::sqrt(1.0 + 2.0 * s * rho[2] / rho[1]);
        rscale = sq / (1.0 - alpha); asn = alpha / s;
      }
      int col[27]; double j0[27], j1[27]; int m = 0;
      auto put = [&](int cidx, double x0, double x1) {
        if (cidx < 0) return;
        if (asn != 0.0) { const double rtj = r[0] * x0 + r[1] * x1; x0 -= asn * r[0] * rtj; x1 -= asn * r[1] * rtj; }
        col[m] = cidx; j0[m] = sq * x0; j1[m] = sq * x1; ++m;
      };
      const double* A = J; const double* B = J + W;
      double l0[3], l1[3];
      if (qcol[a] >= 0) { QuatTangentRow(&cur.q[4 * (size_t)a], A, l0); QuatTangentRow(&cur.q[4 * (size_t)a], B, l1); for (int j = 0; j < 3; ++j) put(qcol[a] + j, l0[j], l1[j]); }
      for (int j = 0; j < 3; ++j) put(tcol[3 * (size_t)a + j], A[4 + j], B[4 + j]);
      if (qcol[b] >= 0) { QuatTangentRow(&cur.q[4 * (size_t)b], A + 7, l0); QuatTangentRow(&cur.q[4 * (size_t)b], B + 7, l1); for (int j = 0; j < 3; ++j) put(qcol[b] + j, l0[j], l1[j]); }
      for (int j = 0; j < 3; ++j) put(tcol[3 * (size_t)b + j], A[11 + j], B[11 + j]);
      if (xcol[p] >= 0) for (int j = 0; j < 3; ++j) put(xcol[p] + j, A[14 + j], B[14 + j]);
      for (int j = 0; j < 12; ++j) put(pcol[12 * (size_t)cm + j], A[17 + j], B[17 + j]);
      const double r0 = r[0] * rscale, r1 = r[1] * rscale;
      for (int x = 0; x < m; ++x) {
        (*g)[col[x]] += j0[x] * r0 + j1[x] * r1;
        for (int y = 0; y < m; ++y) (*H)[(size_t)col[x] * n + col[y]] += j0[x] * j0[y] + j1[x] * j1[y];
      }
    }
    *cost = c;
  };
  P.candidate = [&](const std::vector<double>& delta, double* step2, double* x2) {
    *step2 = 0.0; *x2 = 0.0;
    cand = cur;
    for (int i = 0; i < n_pose; ++i) {
      if (qcol[i] >= 0) {
        const double* q = &cur.q[4 * (size_t)i];
        QuaternionPlus(q, &delta[qcol[i]], &cand.q[4 * (size_t)i]);
        for (int a = 0; a < 4; ++a) { const double e = cand.q[4 * (size_t)i + a] - q[a]; *step2 += e * e; *x2 += q[a] * q[a]; }
      }
      bool var = false;
      for (int a = 0; a < 3; ++a) var = var || tcol[3 * (size_t)i + a] >= 0;
      if (var) for (int a = 0; a < 3; ++a) {
        const int c = tcol[3 * (size_t)i + a];
        const double t = cur.t[3 * (size_t)i + a];
        if (c >= 0) cand.t[3 * (size_t)i + a] = t + delta[c];
        const double e = cand.t[3 * (size_t)i + a] - t; *step2 += e * e; *x2 += t * t;
      }
    }
    for (int c = 0; c < n_cam; ++c) {
      bool var = false;
      for (int j = 0; j < 12; ++j) var = var || pcol[12 * (size_t)c + j] >= 0;
      if (!var) continue;
      const int np = NumParams(d->camera_model[c]);
      for (int j = 0; j < np; ++j) {
        const int cc = pcol[12 * (size_t)c + j];
        const double v = cur.cam[12 * (size_t)c + j];
        if (cc >= 0) cand.cam[12 * (size_t)c + j] = v + delta[cc];
        const double e = cand.cam[12 * (size_t)c + j] - v; *step2 += e * e; *x2 += v * v;
      }
    }
    for (int p = 0; p < n_pt; ++p) if (xcol[p] >= 0) for (int a = 0; a < 3; ++a) {
      const double v = cur.X[3 * (size_t)p + a];
      cand.X[3 * (size_t)p + a] = v + delta[xcol[p] + a];
      const double e = cand.X[3 * (size_t)p + a] - v; *step2 += e * e; *x2 += v * v;
    }
    return cost_of(cand);
  };
  P.accept = [&]() { cur = cand; };
  std::vector<sba_iteration_summary> rows;
  const int rc = SolveDense(P, *o, summary, &rows);
  if (trace && trace_n) { const int k = std::min<int>((int)rows.size(), trace_cap); for (int i = 0; i < k; ++i) trace[i] = rows[i]; *trace_n = k; }
  for (int i = 0; i < n_rig; ++i) { std::memcpy(d->rig_qvec + 4 * (size_t)i, &cur.q[4 * (size_t)i], 32); std::memcpy(d->rig_tvec + 3 * (size_t)i, &cur.t[3 * (size_t)i], 24); }
  for (int i = 0; i < n_rel; ++i) { std::memcpy(d->rel_qvec + 4 * (size_t)i, &cur.q[4 * (size_t)(n_rig + i)], 32); std::memcpy(d->rel_tvec + 3 * (size_t)i, &cur.t[3 * (size_t)(n_rig + i)], 24); }
  if (n_cam) std::memcpy(d->camera_params, cur.cam.data(), cur.cam.size() * sizeof(double));
  if (n_pt) std::memcpy(d->points, cur.X.data(), cur.X.size() * sizeof(double));
  return rc;
}

// Powell's singular function through the same LM loop -- the one solve in Ceres' own documentation whose iteration
// table is published (docs/source/nnls_tutorial.rst, "Powell's Function": f1 = x1 + 10 x2, f2 = sqrt(5) (x3 - x4),
// f3 = (x2 - 2 x3)^2, f4 = sqrt(10) (x1 - x4)^2 from x = (3, -1, 0, 1), Solver::Options defaults, DENSE_QR).  A dense QR
// of the augmented Jacobian and the Cholesky of the normal equations solve the same regularised system, so the table
// pins the trust-region loop restated here (and, through the shared SolveDense, the loops of the semantic, cylinder
// and rig solves): radius update, step acceptance, LM diagonal with its clamp, Jacobi scaling, the termination tests.
int oracle_powell_solve(const sba_solver_options* o, double* x, sba_summary* summary, sba_iteration_summary* trace, int32_t trace_cap,
                        int32_t* trace_n) {
  if (!o || !x || !summary) return SBA_ERR_INVALID_ARGUMENT;
  std::memset(summary, 0, sizeof(*summary));
  std::vector<double> cur(x, x + 4), cand(4);
  auto residuals = [](const std::vector<double>& v, double f[4]) {
    f[0] = v[0] + 10.0 * v[1];
    f[1] = std::sqrt(5.0) * (v[2] - v[3]);
    f[2] = (v[1] - 2.0 * v[2]) * (v[1] - 2.0 * v[2]);
    f[3] = std::sqrt(10.0) * (v[0] - v[3]) * (v[0] - v[3]);
  };
  DenseProblem P;
  P.n = 4;
  P.rows = [&](std::vector<Row>* rows) {
    double f[4];
    residuals(cur, f);
    rows->clear();
    Row r;
    r.n = 4;
    for (int a = 0; a < 4; ++a) r.col[a] = a;
    auto push = [&](double res, double j0, double j1, double j2, double j3) { r.r = res; r.j[0] = j0; r.j[1] = j1; r.j[2] = j2; r.j[3] = j3; rows->push_back(r); };
    push(f[0], 1.0, 10.0, 0.0, 0.0);
    push(f[1], 0.0, 0.0, std::sqrt(5.0), -std::sqrt(5.0));
    const double a = 2.0 * (cur[1] - 2.0 * cur[2]);
    push(f[2], 0.0, a, -2.0 * a, 0.0);
    const double b = 2.0 * std::sqrt(10.0) * (cur[0] - cur[3]);
    push(f[3], b, 0.0, 0.0, -b);
  };
  P.candidate = [&](const std::vector<double>& delta, double* step2, double* x2) {
    *step2 = 0.0; *x2 = 0.0;
    for (int a = 0; a < 4; ++a) { cand[a] = cur[a] + delta[a]; *step2 += delta[a] * delta[a]; *x2 += cur[a] * cur[a]; }
    double f[4];
    residuals(cand, f);
    return 0.5 * (f[0] * f[0] + f[1] * f[1] + f[2] * f[2] + f[3] * f[3]);
  };
  P.accept = [&]() { cur = cand; };
  summary->num_residual_blocks = 4; summary->num_residuals = 4; summary->num_residuals_reduced = 4;
  summary->num_parameter_blocks_reduced = 4; summary->num_parameters_reduced = 4; summary->num_effective_parameters_reduced = 4;
  std::vector<sba_iteration_summary> rows;
  const int rc = SolveDense(P, *o, summary, &rows);
  if (trace && trace_n) { const int k = std::min<int>((int)rows.size(), trace_cap); for (int i = 0; i < k; ++i) trace[i] = rows[i]; *trace_n = k; }
  for (int a = 0; a < 4; ++a) x[a] = cur[a];
  return rc;
}

// The other table of the same tutorial page ("Hello World!"): r = 10 - x from x = 0.5, one parameter.
int oracle_helloworld_solve(const sba_solver_options* o, double* x, sba_summary* summary, sba_iteration_summary* trace,
                            int32_t trace_cap, int32_t* trace_n) {
  if (!o || !x || !summary) return SBA_ERR_INVALID_ARGUMENT;
  std::memset(summary, 0, sizeof(*summary));
  double cur = *x, cand = 0.0;
  DenseProblem P;
  P.n = 1;
  P.rows = [&](std::vector<Row>* rows) {
    rows->clear();
    Row r;
    r.n = 1; r.col[0] = 0; r.r = 10.0 - cur; r.j[0] = -1.0;
    rows->push_back(r);
  };
  P.candidate = [&](const std::vector<double>& delta, double* step2, double* x2) {
    cand = cur + delta[0]; *step2 = delta[0] * delta[0]; *x2 = cur * cur;
    return 0.5 * (10.0 - cand) * (10.0 - cand);
  };
  P.accept = [&]() { cur = cand; };
  std::vector<sba_iteration_summary> rows;
  const int rc = SolveDense(P, *o, summary, &rows);
  if (trace && trace_n) { const int k = std::min<int>((int)rows.size(), trace_cap); for (int i = 0; i < k; ++i) trace[i] = rows[i]; *trace_n = k; }
  *x = cur;
  return rc;
}

}  // extern "C"
