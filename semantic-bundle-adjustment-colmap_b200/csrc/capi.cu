// extern "C" surface declared in include/sba_b200.h: the drop-in boundary behind
//   BundleAdjuster::Solve   /root/reference/src/optim/bundle_adjustment.cc:258-320
// No Eigen / Ceres / torch types cross it.  There is no CPU fallback: every compute entry
// point needs a CUDA device and says so when there is none.
#include <algorithm>
#include <cstdio>
#include <functional>

#include "device_math.cuh"
#include "solver.h"

namespace sba {
const char* LastErrorMessage();
void ShardPointRange(const std::vector<int64_t>& pt_obs_prefix, int nranks, int rank, int* p0, int* p1);
}  // namespace sba

using namespace sba;

struct sba_handle {
  Problem* p;
};

extern "C" {

void sba_default_solver_options(sba_solver_options* o) {
  // src/optim/bundle_adjustment.h:78-93 on top of ceres::Solver::Options 2.1.0 defaults
  std::memset(o, 0, sizeof(*o));
  o->linear_solver_type = SBA_SOLVER_AUTO;
  o->num_config_images = 0;
  o->max_num_iterations = 100;
  o->max_linear_solver_iterations = 200;
  o->min_linear_solver_iterations = 0;
  o->max_num_consecutive_invalid_steps = 10;
  o->function_tolerance = 0.0;
  o->gradient_tolerance = 0.0;
  o->parameter_tolerance = 0.0;
  o->loss_function_type = SBA_LOSS_TRIVIAL;
  o->loss_function_scale = 1.0;
  o->initial_trust_region_radius = 1e4;
  o->max_trust_region_radius = 1e16;
  o->min_trust_region_radius = 1e-32;
  o->min_relative_decrease = 1e-3;
  o->min_lm_diagonal = 1e-6;
  o->max_lm_diagonal = 1e32;
  o->eta = 1e-1;
  o->jacobi_scaling = 1;
  o->update_state_every_iteration = 0;
}

const char* sba_last_error(void) { return LastErrorMessage(); }
const char* sba_version(void) { return "sba_b200 0.1 (sm_100a, fp64)"; }

int sba_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int sba_problem_counts(const sba_problem_desc* desc, sba_summary* counts) {
  if (!counts) { SetError(SBA_ERR_INVALID_ARGUMENT, "null output"); return SBA_ERR_INVALID_ARGUMENT; }
  return ProblemCounts(desc, counts);
}

int sba_create(const sba_problem_desc* desc, sba_handle** out) {
  if (!out) { SetError(SBA_ERR_INVALID_ARGUMENT, "null output handle"); return SBA_ERR_INVALID_ARGUMENT; }
  *out = nullptr;
  Problem* p = nullptr;
  const int rc = BuildProblem(desc, &p, false);
  if (rc) return rc;
  *out = new sba_handle{p};
  return SBA_OK;
}

struct sba_maps { MapSet* m; };

int sba_maps_create(int32_t num_images, int32_t height, int32_t width, const float* maps, sba_maps** out) {
  if (!out) { SetError(SBA_ERR_INVALID_ARGUMENT, "null output map set"); return SBA_ERR_INVALID_ARGUMENT; }
  *out = nullptr;
  MapSet* m = nullptr;
  const int rc = BuildMapSet(num_images, height, width, maps, &m);
  if (rc) return rc;
  *out = new sba_maps{m};
  return SBA_OK;
}

void sba_maps_destroy(sba_maps* maps) {
  if (!maps) return;
  MapSetRelease(maps->m);  // the device block goes once the last handle created over it is destroyed
  delete maps;
}

int sba_create_with_maps(const sba_problem_desc* desc, const sba_maps* maps, sba_handle** out) {
  if (!out) { SetError(SBA_ERR_INVALID_ARGUMENT, "null output handle"); return SBA_ERR_INVALID_ARGUMENT; }
  *out = nullptr;
  if (!maps || !maps->m) { SetError(SBA_ERR_INVALID_ARGUMENT, "null map set"); return SBA_ERR_INVALID_ARGUMENT; }
  Problem* p = nullptr;
  const int rc = BuildProblem(desc, &p, false, maps->m);
  if (rc) return rc;
  *out = new sba_handle{p};
  return SBA_OK;
}

void* sba_host_alloc(uint64_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    SetError(SBA_ERR_CUDA, "cudaHostAlloc failed (no device or out of page-locked memory)");
    return nullptr;
  }
  return p;
}
void sba_host_free(void* ptr) { if (ptr) cudaFreeHost(ptr); }

void sba_destroy(sba_handle* h) {
  if (!h) return;
  delete h->p;
  delete h;
}

int sba_set_device(int32_t device) {
  SBA_CUDA(cudaSetDevice(device));
  return SBA_OK;
}

static int CheckHandle(sba_handle* h) {
  if (!h || !h->p) { SetError(SBA_ERR_INVALID_ARGUMENT, "null handle"); return SBA_ERR_INVALID_ARGUMENT; }
  return SBA_OK;
}

int sba_cost_no_alloc(sba_handle* h, int32_t loss_type, double loss_scale, double* cost_host, double* cost_device) {
  int rc = CheckHandle(h);
  if (rc) return rc;
  if (!cost_host && !cost_device) { SetError(SBA_ERR_INVALID_ARGUMENT, "sba_cost_no_alloc: no output pointer"); return SBA_ERR_INVALID_ARGUMENT; }
  const LossView loss = {loss_type, loss_scale, 1};
  return CostNoAlloc(h->p, loss, cost_host, cost_device);
}

int sba_evaluate(sba_handle* h, int32_t loss_type, double loss_scale, int32_t apply_loss, double* cost,
                 double* residuals, double* jac_pose, double* jac_intr, double* jac_point,
                 double* jac_qvec_ambient, double* sem_residuals) {
  int rc = CheckHandle(h);
  if (rc) return rc;
  Problem* p = h->p;
  CommScope scope(p);
  if (!scope.ok) return SBA_ERR_COMM;
  cudaStream_t st = p->stream;
  const int n = p->n_obs, dc = p->dc;
  DevBuf<double> d_jq;
  if (jac_qvec_ambient) { SBA_CUDA(d_jq.Alloc((size_t)std::max(1, n) * 8)); p->jq_out = d_jq.p; }
  const LossView loss = {loss_type, loss_scale, apply_loss ? 1 : 0};
  rc = LaunchEvaluate(p, loss, false, false);
  p->jq_out = nullptr;
  if (rc) return rc;
  if (sem_residuals && p->nr != 3) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "sem_residuals requested but the problem has no semantic maps");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  std::vector<D4> ER(TilePad(n) * 2), F(TilePad(n) * (dc / 2));  // tile-blocked, common.cuh
  std::vector<D4> ES(sem_residuals ? (size_t)n : 0);
  std::vector<double> JQ;
  if (n > 0) {
    if (sem_residuals) SBA_CUDA(cudaMemcpyAsync(ES.data(), p->d_ES.p, ES.size() * sizeof(D4), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaMemcpyAsync(ER.data(), p->d_ER.p, ER.size() * sizeof(D4), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaMemcpyAsync(F.data(), p->d_F.p, F.size() * sizeof(D4), cudaMemcpyDeviceToHost, st));
    if (jac_qvec_ambient) {
      JQ.resize((size_t)n * 8);
      SBA_CUDA(cudaMemcpyAsync(JQ.data(), d_jq.p, JQ.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    }
  }
  SBA_CUDA(cudaMemcpyAsync(p->h_scalars, p->d_scalars.p, sizeof(DeviceScalars), cudaMemcpyDeviceToHost, st));
  SBA_CUDA(cudaStreamSynchronize(st));
  if (cost) *cost = p->h_scalars->cost + p->h_scalars->fixed_cost;
  if ((rc = FetchHostOrder(p))) return rc;
  auto slot = [&](int k, int j, double* d0, double* d1) {  // (d r0 / d s_j, d r1 / d s_j)
    const D4& u = F[FUnit(k, j >> 1, dc / 2)];
    if (j & 1) { *d0 = u.z; *d1 = u.w; } else { *d0 = u.x; *d1 = u.y; }
  };
  for (int k = 0; k < n; ++k) {
    const int64_t r = p->h_local2ref[k];
    const int i = p->h_obs_img[k], c = p->h_img_cam[i];
    const D4& ea = ER[EUnit(k, 0)];
    const D4& eb = ER[EUnit(k, 1)];
    if (residuals) { residuals[2 * r] = eb.z; residuals[2 * r + 1] = eb.w; }
    if (jac_point) {
      const double e[6] = {ea.x, ea.y, ea.z, ea.w, eb.x, eb.y};
      for (int j = 0; j < 6; ++j) jac_point[6 * r + j] = e[j];
    }
    if (jac_pose) {
      for (int j = 0; j < 6; ++j) slot(k, j, &jac_pose[12 * r + j], &jac_pose[12 * r + 6 + j]);
    }
    if (jac_intr) {
      double* out = &jac_intr[(size_t)r * 2 * SBA_MAX_CAMERA_PARAMS];
      for (int j = 0; j < 2 * SBA_MAX_CAMERA_PARAMS; ++j) out[j] = 0.0;
      for (int s2 = 0; s2 < p->h_k_dim[c] && 6 + s2 < dc; ++s2) {
        const int a = p->h_k_cols[12 * (size_t)c + s2];
        slot(k, 6 + s2, &out[a], &out[SBA_MAX_CAMERA_PARAMS + a]);
      }
    }
    if (jac_qvec_ambient) for (int j = 0; j < 8; ++j) jac_qvec_ambient[8 * r + j] = JQ[(size_t)k * 8 + j];
    if (sem_residuals) sem_residuals[r] = ES[k].w;
  }
  return SBA_OK;
}

int sba_get_semantic_jacobian(sba_handle* h, double* jac_pose, double* jac_intr, double* jac_point) {
  int rc = CheckHandle(h);
  if (rc) return rc;
  Problem* p = h->p;
  if (p->nr != 3 || !p->linearized) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "no semantic row: the problem has no semantic maps, or sba_evaluate has not run");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  cudaStream_t st = p->stream;
  const int n = p->n_obs, dc = p->dc, nq = FSUnits(dc);
  std::vector<D4> ES((size_t)n), FS(TilePad(n) * nq);
  if (n > 0) {
    SBA_CUDA(cudaMemcpyAsync(ES.data(), p->d_ES.p, ES.size() * sizeof(D4), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaMemcpyAsync(FS.data(), p->d_FS.p, FS.size() * sizeof(D4), cudaMemcpyDeviceToHost, st));
  }
  SBA_CUDA(cudaStreamSynchronize(st));
  if ((rc = FetchHostOrder(p))) return rc;
  auto slot = [&](int k, int j) { return reinterpret_cast<const double*>(&FS[FSUnit(k, j >> 2, nq)])[j & 3]; };
  for (int k = 0; k < n; ++k) {
    const int64_t r = p->h_local2ref[k];
    const int c = p->h_img_cam[p->h_obs_img[k]];
    if (jac_point) { jac_point[3 * r] = ES[k].x; jac_point[3 * r + 1] = ES[k].y; jac_point[3 * r + 2] = ES[k].z; }
    if (jac_pose) for (int j = 0; j < 6; ++j) jac_pose[6 * r + j] = slot(k, j);
    if (jac_intr) {
      double* out = &jac_intr[(size_t)r * SBA_MAX_CAMERA_PARAMS];
      for (int j = 0; j < SBA_MAX_CAMERA_PARAMS; ++j) out[j] = 0.0;
      for (int s2 = 0; s2 < p->h_k_dim[c] && 6 + s2 < dc; ++s2) out[p->h_k_cols[12 * (size_t)c + s2]] = slot(k, 6 + s2);
    }
  }
  return SBA_OK;
}

int sba_solve(sba_handle* h, const sba_solver_options* opts, sba_iteration_callback cb, void* user,
              sba_summary* summary) {
  int rc = CheckHandle(h);
  if (rc) return rc;
  if (!summary) { SetError(SBA_ERR_INVALID_ARGUMENT, "null summary"); return SBA_ERR_INVALID_ARGUMENT; }
  sba_solver_options o;
  if (opts) o = *opts; else sba_default_solver_options(&o);
  if (o.max_num_iterations < 0 || o.max_linear_solver_iterations < 0 || !(o.initial_trust_region_radius > 0.0) ||
      !(o.loss_function_scale >= 0.0) || (o.loss_function_scale == 0.0 && o.loss_function_type != SBA_LOSS_TRIVIAL) ||
      o.loss_function_type < 0 || o.loss_function_type > SBA_LOSS_CAUCHY ||
      o.linear_solver_type < SBA_SOLVER_AUTO || o.linear_solver_type > SBA_ITERATIVE_SCHUR) {
    // Solver::Options::IsValid / BundleAdjustmentOptions::Check (bundle_adjustment.cc:70-73, 303-304)
    SetError(SBA_ERR_INVALID_ARGUMENT, "invalid solver options");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  CommScope scope(h->p);
  if (!scope.ok) return SBA_ERR_COMM;
  return SolveLM(h->p, &o, cb, user, summary);
}

int sba_reset(sba_handle* h) {
  int rc = CheckHandle(h);
  if (rc) return rc;
  Problem* p = h->p;
  cudaStream_t st = p->stream;
  auto cp = [&](double* dst, const double* src, size_t n) {
    return n ? cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyDeviceToDevice, st) : cudaSuccess;
  };
  SBA_CUDA(cp(p->d_pose.p, p->d_pose0.p, (size_t)p->n_img * 8)); SBA_CUDA(cp(p->d_pose2.p, p->d_pose0.p, (size_t)p->n_img * 8));
  SBA_CUDA(cp(p->d_cam.p, p->d_cam0.p, (size_t)p->n_cam * 12)); SBA_CUDA(cp(p->d_cam2.p, p->d_cam0.p, (size_t)p->n_cam * 12));
  SBA_CUDA(cp(p->d_X.p, p->d_X0.p, (size_t)p->n_pt * 3)); SBA_CUDA(cp(p->d_X2.p, p->d_X0.p, (size_t)p->n_pt * 3));
  SBA_CUDA(cudaStreamSynchronize(st));
  p->solved = false;
  p->trace.clear();
  return SBA_OK;
}

int sba_get_trace(sba_handle* h, sba_iteration_summary* rows, int32_t cap, int32_t* n) {
  int rc = CheckHandle(h);
  if (rc) return rc;
  const int m = std::min<int>(cap, (int)h->p->trace.size());
  for (int i = 0; i < m; ++i) rows[i] = h->p->trace[i];
  if (n) *n = m;
  return SBA_OK;
}

int sba_get_parameters(sba_handle* h, double* image_qvec, double* image_tvec, double* camera_params, double* points) {
  int rc = CheckHandle(h);
  if (rc) return rc;
  Problem* p = h->p;
  cudaStream_t st = p->stream;
  std::vector<double> pose((size_t)p->n_img * 8);
  if ((image_qvec || image_tvec) && p->n_img > 0)
    SBA_CUDA(cudaMemcpyAsync(pose.data(), p->d_pose.p, pose.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (camera_params && p->n_cam > 0)
    SBA_CUDA(cudaMemcpyAsync(camera_params, p->d_cam.p, (size_t)p->n_cam * 12 * sizeof(double), cudaMemcpyDeviceToHost, st));
  SBA_CUDA(cudaStreamSynchronize(st));
  // the shard's points are one contiguous range of the caller's array (points [p0, p0 + n_pt)): they land in place
  if (points && p->n_pt > 0)
    SBA_CUDA(DownloadLarge(points + 3 * (size_t)p->pt_base, p->d_X.p, (size_t)p->n_pt * 3 * sizeof(double), st));
  for (int i = 0; i < p->n_img; ++i) {
    if (image_qvec) for (int j = 0; j < 4; ++j) image_qvec[4 * (size_t)i + j] = pose[8 * (size_t)i + j];
    if (image_tvec) for (int j = 0; j < 3; ++j) image_tvec[3 * (size_t)i + j] = pose[8 * (size_t)i + 4 + j];
  }
  return SBA_OK;
}

int64_t sba_kernel_launch_count(sba_handle* h) { return (h && h->p) ? h->p->launches : 0; }

// ---- stage benchmarks --------------------------------------------------------------------
static int PrepareLinearSystem(Problem* p, const sba_solver_options& o) {
  DeviceScalars* hs = p->h_scalars;
  std::memset(hs, 0, sizeof(*hs));
  hs->radius = o.initial_trust_region_radius;
  hs->decrease_factor = 2.0;
  SBA_CUDA(cudaMemcpyAsync(p->d_scalars.p, hs, sizeof(*hs), cudaMemcpyHostToDevice, p->stream));
  const LossView loss = {o.loss_function_type, o.loss_function_scale, 1};
  int rc;
  if ((rc = FillOnes(p))) return rc;
  if ((rc = LaunchEvaluate(p, loss, false, false))) return rc;
  if ((rc = PointBlocks(p, true))) return rc;
  if ((rc = Gradient(p))) return rc;
  if ((rc = InitJacobiScale(p))) return rc;
  if ((rc = PrepareLMDiagonal(p, o))) return rc;
  if ((rc = PointBlocks(p, false))) return rc;
  if ((rc = ReducedRhs(p))) return rc;
  SBA_CUDA(cudaStreamSynchronize(p->stream));
  return SBA_OK;
}

int sba_bench_stage(sba_handle* h, const char* stage, int32_t repeat, double* ms_per_launch, int64_t* algorithmic_bytes) {
  int rc = CheckHandle(h);
  if (rc) return rc;
  Problem* p = h->p;
  if (!stage || repeat <= 0) { SetError(SBA_ERR_INVALID_ARGUMENT, "bad stage / repeat"); return SBA_ERR_INVALID_ARGUMENT; }
  CommScope scope(p);
  if (!scope.ok) return SBA_ERR_COMM;
  sba_solver_options o;
  sba_default_solver_options(&o);
  const std::string s(stage);
  const LossView loss = {SBA_LOSS_TRIVIAL, 1.0, 1};
  const int64_t n_obs = p->n_obs, dc = p->dc;
  const int64_t gathers = (int64_t)p->n_img * 64 + (int64_t)p->n_cam * 96 + (int64_t)p->n_pt * 24;
  std::function<int()> prep = []() { return SBA_OK; }, run;
  int64_t bytes = 0;
  bool blocks_built = false;
  if (s == "eval") {
    run = [&]() { return LaunchEvaluate(p, loss, true, false); };
    // idx 8 + xy 16 in; r 16 + J 16*(dc+3) out; + parameter gathers once
    bytes = n_obs * (8 + 16 + 16 + 16 * (dc + 3)) + gathers;
    if (p->nr == 3) bytes += n_obs * (32 + 32 * FSUnits((int)dc) + 4 * 4);  // third row out (ES + FS units) + four texels in
  } else if (s == "cost") {
    run = [&]() { return LaunchCostOnly(p, loss, false, &p->d_scalars.p->cand_cost); };
    bytes = n_obs * 24 + gathers;
  } else if (s == "point_blocks") {
    run = [&]() { p->ete_valid = false; return PointBlocks(p, false); };  // the pass over ER (after an accepted step)
    bytes = n_obs * (64 + 4) + (int64_t)p->n_pvar * (24 + 48 + 24 + 48);
  } else if (s == "point_refresh") {
    run = [&]() { return PointBlocks(p, false); };  // the per-iteration form: cached E^T E + the new LM diagonal
    bytes = (int64_t)p->n_pvar * (48 + 24 + 48 + 48);
  } else if (s == "rhs") {
    run = [&]() { return ReducedRhs(p); };
    bytes = n_obs * (64 + 16 * dc + 16 + 16 + 8);
  } else if (s == "matvec") {
    // the IMPLICIT product (what Ceres' ITERATIVE_SCHUR does), also when the problem solves with the block-sparse S
    run = [&]() {
      const bool keep = p->bsr.enabled;
      p->bsr.enabled = false;
      const int r = SchurMatVec(p, p->d_rhs.p, p->d_tmp_c.p);
      p->bsr.enabled = keep;
      return r;
    };
    // two passes over the stored F (16*dc) + E (48) per observation + 48 B / point
    bytes = n_obs * (2 * 16 * (dc + 3)) + (int64_t)p->n_pvar * 48;
  } else if (s == "backsub") {
    run = [&]() { return BackSubstitute(p); };
    bytes = n_obs * (64 + 16 * dc + 8) + (int64_t)p->n_pvar * (48 + 24 + 24);
  } else if (s == "schur_dense") {
    run = [&]() { return BuildDenseSchur(p, true); };  // as the LM step runs it: the reduced rhs rides on the same pass
    bytes = n_obs * (16 * (dc + 3) + 16) + (int64_t)p->n_pvar * 96 + 8 * (int64_t)p->n_cdim * p->n_cdim;
  } else if (s == "schur_jacobi") {
    run = [&]() { return BuildSchurJacobi(p); };
    bytes = n_obs * (16 * (dc + 3) + 16) + (int64_t)p->n_pvar * 48;
  } else if (s == "schur_blocks" || s == "matvec_blocks") {
    // explicit block-sparse S (bsr.cu): build = one pass over the stored Jacobian + the blocks written once;
    // product = off-diagonal blocks twice (straight + transposed), diagonal once, three camera vectors
    if ((rc = DecideBlockSchur(p))) return rc;
    if (!p->bsr.enabled) { SetError(SBA_ERR_INVALID_ARGUMENT, "the block-sparse Schur complement is not in use for this problem"); return SBA_ERR_INVALID_ARGUMENT; }
    const int64_t ne = dc * dc;
    if (s == "schur_blocks") {
      run = [&]() { return BuildExplicitSchur(p, true); };  // as the LM step runs it: the reduced rhs rides on the same pass
      bytes = n_obs * (16 * (dc + 3) + 16) + (int64_t)p->n_pvar * 48 + 8 * p->bsr.n_blocks * ne;
    } else {
      prep = [&]() { if (blocks_built) return (int)SBA_OK; blocks_built = true; return BuildExplicitSchur(p); };
      run = [&]() { return BlockSchurMatVec(p, p->d_rhs.p, p->d_tmp_c.p, nullptr); };
      bytes = 8 * ((int64_t)p->n_img + 2 * (int64_t)p->bsr.n_keys) * ne + 24 * (int64_t)p->n_cdim;
    }
  } else if (s == "cholesky") {
    prep = [&]() { return BuildDenseSchur(p); };
    run = [&]() { return CholeskySolve(p); };
    bytes = 8 * (int64_t)p->n_cdim * p->n_cdim;  // flops are n^3/3; bytes reported for completeness
  } else {
    SetError(SBA_ERR_INVALID_ARGUMENT, "unknown stage: " + s);
    return SBA_ERR_INVALID_ARGUMENT;
  }
  if ((rc = PrepareLinearSystem(p, o))) return rc;
  const size_t flush_bytes = 256u << 20;  // > 126 MB L2
  SBA_CUDA(p->d_flush.Alloc(flush_bytes));
  cudaEvent_t e0, e1;
  SBA_CUDA(cudaEventCreate(&e0)); SBA_CUDA(cudaEventCreate(&e1));
  double total = 0.0;
  for (int it = -1; it < repeat; ++it) {  // it == -1: untimed warm-up
    if ((rc = prep())) return rc;
    SBA_CUDA(cudaMemsetAsync(p->d_flush.p, it & 0xff, flush_bytes, p->stream));
    SBA_CUDA(cudaEventRecord(e0, p->stream));
    if ((rc = run())) return rc;
    SBA_CUDA(cudaEventRecord(e1, p->stream));
    SBA_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    SBA_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    if (it >= 0) total += ms;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (ms_per_launch) *ms_per_launch = total / repeat;
  if (algorithmic_bytes) *algorithmic_bytes = bytes;
  return SBA_OK;
}

void sba_release_cached_memory(void) { ReleaseCachedBlocks(); }

int sba_schur_product(sba_handle* h, int32_t use_blocks, const double* x, double* y, int32_t* n) {
  int rc = CheckHandle(h);
  if (rc) return rc;
  Problem* p = h->p;
  if (n) *n = p->n_cdim;
  if (!x || !y || p->n_cdim == 0) return SBA_OK;
  CommScope scope(p);
  if (!scope.ok) return SBA_ERR_COMM;
  sba_solver_options o;
  sba_default_solver_options(&o);
  if ((rc = PrepareLinearSystem(p, o))) return rc;
  SBA_CUDA(cudaMemcpyAsync(p->d_cg_p.p, x, (size_t)p->n_cdim * sizeof(double), cudaMemcpyHostToDevice, p->stream));
  if (use_blocks) {
    if ((rc = DecideBlockSchur(p))) return rc;
    if (!p->bsr.enabled) { SetError(SBA_ERR_INVALID_ARGUMENT, "the block-sparse Schur complement is not in use for this problem"); return SBA_ERR_INVALID_ARGUMENT; }
    if ((rc = BuildExplicitSchur(p))) return rc;
    if ((rc = BlockSchurMatVec(p, p->d_cg_p.p, p->d_cg_q.p, nullptr))) return rc;
  } else {
    const bool keep = p->bsr.enabled;
    p->bsr.enabled = false;
    rc = SchurMatVec(p, p->d_cg_p.p, p->d_cg_q.p);
    p->bsr.enabled = keep;
    if (rc) return rc;
  }
  SBA_CUDA(cudaMemcpyAsync(y, p->d_cg_q.p, (size_t)p->n_cdim * sizeof(double), cudaMemcpyDeviceToHost, p->stream));
  SBA_CUDA(cudaStreamSynchronize(p->stream));
  return SBA_OK;
}

// ---- multi-GPU ------------------------------------------------------------------------------
int sba_comm_get_unique_id(void* id128) { return CommGetUniqueId(id128); }
int sba_comm_init(sba_handle* h, int32_t nranks, int32_t rank, const void* id128) {
  int rc = CheckHandle(h);
  if (rc) return rc;
  return CommInit(h->p, nranks, rank, id128);
}

int sba_shard_points(const sba_problem_desc* d, int32_t nranks, int32_t rank, uint8_t* point_keep, uint8_t* obs_keep) {
  if (!d || nranks <= 0 || rank < 0 || rank >= nranks) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "bad shard arguments");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  const int n_pt = d->num_points;
  std::vector<int64_t> prefix(n_pt + 1, 0);
  for (int64_t k = 0; k < d->num_observations; ++k) {
    const int q = d->obs_point[k];
    if (q < 0 || q >= n_pt) { SetError(SBA_ERR_INVALID_ARGUMENT, "observation index out of range"); return SBA_ERR_INVALID_ARGUMENT; }
    prefix[q + 1]++;
  }
  for (int q = 0; q < n_pt; ++q) prefix[q + 1] += prefix[q];
  int p0, p1;
  ShardPointRange(prefix, nranks, rank, &p0, &p1);
  if (point_keep) for (int q = 0; q < n_pt; ++q) point_keep[q] = (q >= p0 && q < p1) ? 1 : 0;
  if (obs_keep)
    for (int64_t k = 0; k < d->num_observations; ++k) obs_keep[k] = (d->obs_point[k] >= p0 && d->obs_point[k] < p1) ? 1 : 0;
  return SBA_OK;
}

}  // extern "C"
