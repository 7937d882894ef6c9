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
#include <cstdlib>
#include <cstring>
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

// Replays `body` (a fixed launch sequence on p->stream, no host synchronisation inside) as a CUDA graph: captured
// the first time a key is seen, launched as one graph afterwards.  An LM iteration at C2 is ~45 launches of
// 3-150 us kernels; as a graph the launch gaps and the host's per-launch cost leave the critical path.
template <class F>
static int RunSegment(Problem* p, int key, bool use_graph, F&& body) {
  if (!use_graph || p->graphs_unusable) return body();
  cudaStream_t st = p->stream;
  if (!p->graph_exec[key]) {
    const int64_t l0 = p->launches;
    SBA_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    const int rc = body();
    cudaGraph_t g = nullptr;
    const cudaError_t e = cudaStreamEndCapture(st, &g);
    if (rc || e != cudaSuccess || !g) {
      if (g) cudaGraphDestroy(g);
      cudaGetLastError();
      if (rc && e == cudaSuccess) return rc;
      // a launch that cannot be captured on this driver: nothing ran, so run the segment as plain launches from now on
      p->graphs_unusable = true;
      p->launches = l0;
      return body();
    }
    const cudaError_t ei = cudaGraphInstantiate(&p->graph_exec[key], g, 0);
    cudaGraphDestroy(g);
    if (ei != cudaSuccess) {
      p->graph_exec[key] = nullptr;
      cudaGetLastError();
      p->graphs_unusable = true;
      p->launches = l0;
      return body();
    }
    p->graph_launches[key] = p->launches - l0;
  } else {
    p->launches += p->graph_launches[key];
  }
  SBA_CUDA(cudaGraphLaunch(p->graph_exec[key], st));
  return SBA_OK;
}

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
  // Graph replay: direct solvers on one GPU (the PCG path polls its state from the host; NCCL stays outside).
  // SBA_NO_GRAPHS=1 keeps plain launches with the per-stage event timers (time_*_s of the summary).
  static const bool graphs_off = getenv("SBA_NO_GRAPHS") != nullptr || getenv("SBA_CHOL_TRACE") != nullptr;
  // Capturing and instantiating the (up to six) graphs costs about as much as they save over ~10 iterations, so a
  // fresh handle asked for a short solve keeps plain launches; a handle that is solved again (sba_reset) or a long
  // solve (the reference's default is max_num_iterations = 100) replays graphs.
  const bool worth_it = p->solve_count > 0 || o.max_num_iterations > 25;
  p->solve_count++;
  // The iterative solver qualifies once it runs on the block-sparse S: its PCG is then one persistent launch with no
  // host poll inside (decided in the first iteration, which always runs as plain launches).
  static const bool pcg_host_driven = getenv("SBA_PCG_LAUNCHES") != nullptr || getenv("SBA_PCG_TRACE") != nullptr;
  const bool multi = p->comm != nullptr;  // sharded by point: partial sums meet in as few collectives as the data flow allows
  static const bool comm_graphs = getenv("SBA_NO_COMM_GRAPHS") == nullptr;  // the collectives (NCCL or peer-memory) are captured with the rest
  const bool graphs_ok = !graphs_off && worth_it && (!multi || comm_graphs) && p->n_cdim > 0 && !(iterative && pcg_host_driven);
  if (graphs_ok && (!p->graph_opts_valid || std::memcmp(&p->graph_opts, &o, sizeof(o)) != 0)) {
    p->DropGraphs();
    p->graph_opts = o; p->graph_opts_valid = true;
  }
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

    // ---- LevenbergMarquardtStrategy::ComputeStep, candidate point and its cost: one launch sequence ----
    // (the first iteration runs as plain launches: it allocates, builds the pair lists and sets kernel attributes)
    const bool use_graph = graphs_ok && iteration >= 2 && (!iterative || p->bsr.enabled) && !p->graphs_unusable;
    const bool marks = !use_graph;
    int lin_iters = 0;
    auto step_segment = [&]() -> int {
      int r;
      if (marks) tm.Mark(0, st);
      if ((r = PrepareLMDiagonal(p, o))) return r;
      if ((r = PointBlocks(p, false))) return r;
      if (p->n_cdim > 0) {
        if (iterative && (r = DecideBlockSchur(p))) return r;  // once per problem: explicit block-sparse S or implicit products
        // the reduced right-hand side: formed by the explicit Schur build's own pass over the observations where that pass
        // reads everything it needs (dc <= 8), else by a pass of its own.  Multi-GPU: summed across ranks by the Schur
        // build's all-reduce either way.
        const bool fused_rhs = SchurBuildsRhs(p) && (!iterative || p->bsr.enabled);
        if (!fused_rhs && (r = ReducedRhs(p, multi))) return r;
        if (iterative) {
          if ((r = p->bsr.enabled ? BuildExplicitSchur(p, fused_rhs) : BuildSchurJacobi(p))) return r;
          if (marks) tm.Mark(1, st);
          if ((r = SolvePCG(p, o, &lin_iters))) return r;
        } else {
          if ((r = BuildDenseSchur(p, fused_rhs))) return r;
          if (marks) tm.Mark(1, st);
          if ((r = CholeskySolve(p))) return r;
          lin_iters = 1;
        }
        if (p->rhs_unreduced) {  // no Schur build carried it (no preconditioner blocks)
          if ((r = AllReduceSum(p, p->d_rhs.p, p->n_cdim))) return r;
          p->rhs_unreduced = false;
        }
      } else if (marks) {
        tm.Mark(1, st);
      }
      if (marks) tm.Mark(2, st);
      // multi-GPU: model_cost_change and the point-side norms stay per-rank partial sums until the candidate cost's
      // all-reduce, which carries all four; the step's validity (phase 0) is then decided after it.  Nothing between reads it.
      if ((r = BackSubstitute(p, multi))) return r;
      if (!multi && (r = TrustRegionUpdate(p, o, 0, false))) return r;
      if (marks) tm.Mark(3, st);
      if ((r = ApplyStep(p, multi))) return r;
      if (marks) tm.Mark(4, st);
      if ((r = LaunchCostOnly(p, loss, true, &sc->cand_cost, multi))) return r;
      if (marks) tm.Mark(5, st);
      if (multi && (r = TrustRegionUpdate(p, o, 0, false))) return r;
      if ((r = TrustRegionUpdate(p, o, 1, false))) return r;
      SBA_CUDA(cudaMemcpyAsync(hs, sc, sizeof(*hs), cudaMemcpyDeviceToHost, st));
      return SBA_OK;
    };
    const int refresh = hs->reuse_diagonal ? 0 : 1;  // PrepareLMDiagonal branches on the same host flag
    if ((rc = RunSegment(p, p->parity * 2 + refresh, use_graph, step_segment))) return rc;
    SBA_CUDA(cudaStreamSynchronize(st));  // hs is current
    if (use_graph) lin_iters = iterative ? hs->cg_iter : 1;
    settle();
    if (marks) {
      p->t_elim += tm.Ms(0, 1) * 1e-3; p->t_solve += tm.Ms(1, 2) * 1e-3;
      p->t_update += (tm.Ms(2, 3) + tm.Ms(3, 4)) * 1e-3; p->t_eval += tm.Ms(4, 5) * 1e-3;
    }
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
    if (hs->converged) {
      // parameter / function tolerance: TrustRegionMinimizer::Minimize returns right here, BEFORE the next
      // FinalizeIterationAndCheckIfMinimizerCanContinue would append this iteration's row -- Summary::iterations (and
      // num_iterations) do not contain it
      termination = SBA_CONVERGENCE;
      break;
    }
    it.relative_decrease = hs->relative_decrease;
    radius = hs->radius;
    it.trust_region_radius = radius;
    if (hs->step_accepted) {
      // (the true block size travels with the pointer: the block cache may have handed one of the two a larger block)
      std::swap(p->d_pose.p, p->d_pose2.p); std::swap(p->d_pose.cap_bytes, p->d_pose2.cap_bytes);
      std::swap(p->d_cam.p, p->d_cam2.p);   std::swap(p->d_cam.cap_bytes, p->d_cam2.cap_bytes);
      std::swap(p->d_X.p, p->d_X2.p);       std::swap(p->d_X.cap_bytes, p->d_X2.cap_bytes);
      p->parity ^= 1;
      auto linearise_segment = [&]() -> int {
        int r;
        if (marks) tm.Mark(6, st);
        if ((r = LaunchEvaluate(p, loss, o.jacobi_scaling != 0, false, multi))) return r;  // cost sums ride on the gradient's all-reduce
        if ((r = PointBlocks(p, true))) return r;
        if ((r = Gradient(p, true))) return r;  // + the camera column norms the next LM diagonal needs, same pass over F
        if (marks) tm.Mark(7, st);
        return SBA_OK;
      };
      if ((rc = RunSegment(p, 4 + p->parity, use_graph, linearise_segment))) return rc;
      p->colnorm_current = true;  // (also when the segment was replayed as a graph: host code inside it did not run)
      it.step_is_successful = 1;
      sum->num_successful_steps++;
      it.cost = x_cost + fixed_cost;  // provisional; settled below or at the next synchronisation
      it.gradient_max_norm = p->trace.back().gradient_max_norm;
      p->trace.push_back(it);
      pending_row = (int)p->trace.size() - 1;
      pending_eval_time = marks;
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
