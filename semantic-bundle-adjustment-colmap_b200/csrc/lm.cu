// Levenberg-Marquardt driver: the device-resident replacement for ceres::Solve as the
// reference calls it at /root/reference/src/optim/bundle_adjustment.cc:306 with the
// options of src/optim/bundle_adjustment.h:78-93 and the solver policy of
// bundle_adjustment.cc:271-286.  Control flow follows Ceres 2.1.0's
// TrustRegionMinimizer + LevenbergMarquardtStrategy.
//
// The host only sequences kernels.  All trust-region state (radius, decrease factor,
// rho, accept / reject, diagonal reuse) lives in DeviceScalars and is updated by
// tr_update_kernel; the host reads that one small struct twice per iteration (after the
// linear solve, after the candidate cost) to decide which kernels to enqueue next and to
// fill the iteration table handed to the caller's callbacks.
#include <chrono>
#include <utility>

#include "solver.h"

namespace sba {

namespace {
struct StageTimer {
  cudaEvent_t ev[8];
  bool ok = false;
  StageTimer() {
    ok = true;
    for (auto& e : ev) if (cudaEventCreate(&e) != cudaSuccess) ok = false;
  }
  ~StageTimer() { for (auto& e : ev) cudaEventDestroy(e); }
  void Mark(int i, cudaStream_t s) { cudaEventRecord(ev[i], s); }
  double Ms(int a, int b) { float ms = 0.f; if (cudaEventElapsedTime(&ms, ev[a], ev[b]) != cudaSuccess) { cudaGetLastError(); return 0.0; } return ms; }
};
}  // namespace

static int ResolveSolver(const Problem* p, const sba_solver_options& o) {
  int solver = o.linear_solver_type;
  if (solver == SBA_SOLVER_AUTO) {
    const int nimg = o.num_config_images > 0 ? o.num_config_images : p->n_img;
    solver = nimg <= 50 ? SBA_DENSE_SCHUR : (nimg <= 1000 ? SBA_SPARSE_SCHUR : SBA_ITERATIVE_SCHUR);
  }
  return solver;
}

int SolveLM(Problem* p, const sba_solver_options* op, sba_iteration_callback cb, void* user, sba_summary* sum) {
  const sba_solver_options& o = *op;
  const auto t_start = std::chrono::steady_clock::now();
  HostTimer tmh("solve");
  *sum = p->counts;
  p->trace.clear();
  if (p->n_obs_global == 0) {
    SetError(SBA_ERR_EMPTY_PROBLEM, "no residual blocks");
    return SBA_ERR_EMPTY_PROBLEM;
  }
  if (p->solved) {
    SetError(SBA_ERR_ALREADY_SOLVED, "sba_solve called twice on one handle");
    return SBA_ERR_ALREADY_SOLVED;
  }
  p->solved = true;
  const int solver = ResolveSolver(p, o);
  sum->linear_solver_type_used = solver;
  const bool iterative = solver == SBA_ITERATIVE_SCHUR;
  cudaStream_t st = p->stream;
  DeviceScalars* sc = p->d_scalars.p;
  DeviceScalars* hs = p->h_scalars;
  const LossView loss = {o.loss_function_type, o.loss_function_scale, 1};
  StageTimer tm;
  cudaEvent_t ev_begin, ev_end;
  SBA_CUDA(cudaEventCreate(&ev_begin)); SBA_CUDA(cudaEventCreate(&ev_end));
  SBA_CUDA(cudaEventRecord(ev_begin, st));
  int rc;
  p->t_eval = p->t_elim = p->t_solve = p->t_update = 0.0;

  // ---- iteration 0: evaluate, gradient, Jacobi scaling ----
  std::memset(hs, 0, sizeof(*hs));
  hs->radius = o.initial_trust_region_radius;
  hs->decrease_factor = 2.0;
  SBA_CUDA(cudaMemcpyAsync(sc, hs, sizeof(*hs), cudaMemcpyHostToDevice, st));
  if ((rc = FillOnes(p))) return rc;
  tm.Mark(0, st);
  if ((rc = LaunchEvaluate(p, loss, false, false))) return rc;
  if ((rc = PointBlocks(p, true))) return rc;
  if ((rc = Gradient(p))) return rc;
  if (o.jacobi_scaling && (rc = InitJacobiScale(p))) return rc;
  tm.Mark(1, st);
  SBA_CUDA(cudaMemcpyAsync(hs, sc, sizeof(*hs), cudaMemcpyDeviceToHost, st));
  SBA_CUDA(cudaStreamSynchronize(st));
  p->t_eval += tm.Ms(0, 1) * 1e-3;
  double x_cost = hs->cost;
  const double fixed_cost = hs->fixed_cost;
  sum->fixed_cost = fixed_cost;
  sum->initial_cost = x_cost + fixed_cost;

  sba_iteration_summary it0;
  std::memset(&it0, 0, sizeof(it0));
  it0.cost = x_cost + fixed_cost;
  it0.gradient_max_norm = hs->gradient_max_norm;
  it0.trust_region_radius = hs->radius;
  it0.eta = o.eta;
  p->trace.push_back(it0);

  int termination = SBA_NO_CONVERGENCE;
  int iteration = 0, num_consecutive_invalid = 0;
  double radius = hs->radius;
  auto user_stop = [&](const sba_iteration_summary& row) -> bool {
    if (!cb) return false;
    const int r = cb(&row, user);
    if (r == 1) { termination = SBA_USER_SUCCESS; return true; }
    if (r == 2) { termination = SBA_USER_FAILURE; return true; }
    return false;
  };
  bool stop = user_stop(it0);
  tmh.Lap("iteration 0");

  // One host synchronisation per iteration.  The step-validity test (phase 0) and the acceptance test
  // (phase 1) both run on the device; the candidate point and its cost are enqueued unconditionally (an
  // invalid step -- rare -- makes phase 1 a no-op and wastes one cost pass).  After an accepted step the new
  // linearisation is enqueued and, unless a callback or a gradient tolerance needs its numbers right away, the
  // host does not wait for it: cost and gradient norm of that row are filled in at the next synchronisation.
  const bool eager = cb != nullptr || o.gradient_tolerance > 0.0;
  int pending_row = -1;       // trace row whose cost / gradient are still on the device
  bool pending_eval_time = false;
  auto settle = [&]() {       // hs is current: complete the deferred row
    if (pending_eval_time) { p->t_eval += tm.Ms(6, 7) * 1e-3; pending_eval_time = false; }
    if (pending_row >= 0) {
      x_cost = hs->cost;
      p->trace[pending_row].cost = x_cost + fixed_cost;
      p->trace[pending_row].gradient_max_norm = hs->gradient_max_norm;
      pending_row = -1;
    }
  };
  while (!stop) {
    if (iteration >= o.max_num_iterations) { termination = SBA_NO_CONVERGENCE; break; }
    if (pending_row < 0 && p->trace.back().gradient_max_norm <= o.gradient_tolerance) { termination = SBA_CONVERGENCE; break; }
    if (radius < o.min_trust_region_radius) { termination = SBA_CONVERGENCE; break; }
    ++iteration;
    sba_iteration_summary it;
    std::memset(&it, 0, sizeof(it));
    it.iteration = iteration;
    it.eta = o.eta;

    // ---- LevenbergMarquardtStrategy::ComputeStep ----
    tm.Mark(0, st);
    if ((rc = PrepareLMDiagonal(p, o))) return rc;
    if ((rc = PointBlocks(p, false))) return rc;
    int lin_iters = 0;
    if (p->n_cdim > 0) {
      if ((rc = ReducedRhs(p))) return rc;
      if (iterative) {
        if ((rc = BuildSchurJacobi(p))) return rc;
        tm.Mark(1, st);
        if ((rc = SolvePCG(p, o, &lin_iters))) return rc;
      } else {
        if ((rc = BuildDenseSchur(p))) return rc;
        tm.Mark(1, st);
        if ((rc = CholeskySolve(p))) return rc;
        lin_iters = 1;
      }
    } else {
      tm.Mark(1, st);
    }
    tm.Mark(2, st);
    if ((rc = BackSubstitute(p))) return rc;
    if ((rc = TrustRegionUpdate(p, o, 0, false))) return rc;
    // ---- candidate point and its cost ----
    tm.Mark(3, st);
    if ((rc = ApplyStep(p))) return rc;
    tm.Mark(4, st);
    if ((rc = LaunchCostOnly(p, loss, true, &sc->cand_cost))) return rc;
    tm.Mark(5, st);
    if ((rc = TrustRegionUpdate(p, o, 1, true))) return rc;  // syncs; hs is current
    settle();
    p->t_elim += tm.Ms(0, 1) * 1e-3; p->t_solve += tm.Ms(1, 2) * 1e-3;
    p->t_update += (tm.Ms(2, 3) + tm.Ms(3, 4)) * 1e-3; p->t_eval += tm.Ms(4, 5) * 1e-3;
    it.linear_solver_iterations = lin_iters;
    sum->num_linear_solver_iterations += lin_iters;
    it.step_is_valid = hs->step_valid;
    if (!hs->step_valid) {
      ++num_consecutive_invalid;
      radius = hs->radius;
      it.cost = x_cost + fixed_cost;
      it.gradient_max_norm = p->trace.back().gradient_max_norm;
      it.trust_region_radius = radius;
      p->trace.push_back(it);
      sum->num_unsuccessful_steps++;
      if (num_consecutive_invalid >= o.max_num_consecutive_invalid_steps) { termination = SBA_FAILURE; break; }
      if (user_stop(it)) break;
      continue;
    }
    num_consecutive_invalid = 0;
    it.step_norm = hs->step_norm;
    it.cost_change = hs->cost_change;
    if (hs->converged) {  // parameter / function tolerance
      it.cost = x_cost + fixed_cost;
      it.trust_region_radius = radius;
      it.gradient_max_norm = p->trace.back().gradient_max_norm;
      p->trace.push_back(it);
      termination = SBA_CONVERGENCE;
      break;
    }
    it.relative_decrease = hs->relative_decrease;
    radius = hs->radius;
    it.trust_region_radius = radius;
    if (hs->step_accepted) {
      std::swap(p->d_pose.p, p->d_pose2.p);
      std::swap(p->d_cam.p, p->d_cam2.p);
      std::swap(p->d_X.p, p->d_X2.p);
      tm.Mark(6, st);
      if ((rc = LaunchEvaluate(p, loss, o.jacobi_scaling != 0, false))) return rc;
      if ((rc = PointBlocks(p, true))) return rc;
      if ((rc = Gradient(p))) return rc;
      tm.Mark(7, st);
      it.step_is_successful = 1;
      sum->num_successful_steps++;
      it.cost = x_cost + fixed_cost;  // provisional; settled below or at the next synchronisation
      it.gradient_max_norm = p->trace.back().gradient_max_norm;
      p->trace.push_back(it);
      pending_row = (int)p->trace.size() - 1;
      pending_eval_time = true;
      if (eager || iteration >= o.max_num_iterations) {
        SBA_CUDA(cudaMemcpyAsync(hs, sc, sizeof(*hs), cudaMemcpyDeviceToHost, st));
        SBA_CUDA(cudaStreamSynchronize(st));
        settle();
      }
    } else {
      it.gradient_max_norm = p->trace.back().gradient_max_norm;
      sum->num_unsuccessful_steps++;
      it.cost = x_cost + fixed_cost;
      p->trace.push_back(it);
    }
    tmh.Lap("iteration");
    if (eager && user_stop(p->trace.back())) break;
  }
  if (pending_row >= 0 || pending_eval_time) {
    SBA_CUDA(cudaMemcpyAsync(hs, sc, sizeof(*hs), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    settle();
  }
  {
    float ms = 0.f;
    cudaEventRecord(ev_end, st);
    cudaEventSynchronize(ev_end);
    cudaEventElapsedTime(&ms, ev_begin, ev_end);
    sum->device_time_s = ms * 1e-3;
    cudaEventDestroy(ev_begin); cudaEventDestroy(ev_end);
  }
  sum->termination_type = termination;
  sum->final_cost = x_cost + fixed_cost;
  sum->num_iterations = (int)p->trace.size();
  sum->time_evaluate_s = p->t_eval;
  sum->time_eliminate_s = p->t_elim;
  sum->time_linear_solve_s = p->t_solve;
  sum->time_update_s = p->t_update;
  sum->total_time_in_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
  return SBA_OK;
}

}  // namespace sba
