// K-PCG: preconditioned conjugate gradients on the implicit reduced camera system
// with the SCHUR_JACOBI block preconditioner -- the linear solver the reference selects
// for > 1000 images (ITERATIVE_SCHUR + SCHUR_JACOBI,
//   /root/reference/src/optim/bundle_adjustment.cc:283-286) and Ceres runs inside
// ceres::Solve (:306).  The iteration, its failure rules and the quadratic-model stopping
// test (zeta = k (Q_k - Q_{k-1}) / Q_k < eta) follow Ceres 2.1.0's
// ConjugateGradientsSolver.
//
// All CG scalars (rho, alpha, beta, Q, zeta, state) live in DeviceScalars; the BLAS-1
// work of one iteration is two single-CTA kernels (the camera-space vectors have at
// most a few 10^4 entries and are replicated on every GPU, so no collective is needed
// for the dot products); the product with S is SchurMatVec (linalg.cu), whose
// camera-side reduction is the one allreduce per iteration.
#include "reduce.cuh"
#include "solver.h"

namespace sba {

constexpr int kCgThreads = 1024;

// deterministic block-wide sum, result broadcast to every thread
__device__ inline double BlockAllSum(double v) {
  __shared__ double sm[32];
  __shared__ double total;
  v = WarpSum(v);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    double a = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : 0.0;
    a = WarpSum(a);
    if (threadIdx.x == 0) total = a;
  }
  __syncthreads();
  const double r = total;
  __syncthreads();
  return r;
}

__device__ inline bool IsZeroOrInf(double x) { return x == 0.0 || isinf(x); }

// z = M^{-1} r (block diagonal), one thread per column
__device__ inline double PrecondRow(int col, const int* __restrict__ col_blk, const int* __restrict__ blk_off,
                                    const int* __restrict__ blk_dim, const int* __restrict__ blk_moff,
                                    const double* __restrict__ Minv, const double* __restrict__ r) {
  const int b = col_blk[col], o = blk_off[b], d = blk_dim[b];
  const double* m = Minv + blk_moff[b] + (size_t)(col - o) * d;
  double s = 0.0;
  for (int j = 0; j < d; ++j) s += m[j] * r[o + j];
  return s;
}

__global__ void __launch_bounds__(kCgThreads) cg_init_kernel(int n, const double* __restrict__ b, double* __restrict__ x,
                                                             double* __restrict__ r, DeviceScalars* sc) {
  double s = 0.0;
  for (int j = threadIdx.x; j < n; j += kCgThreads) { const double v = b[j]; x[j] = 0.0; r[j] = v; s += v * v; }
  s = BlockAllSum(s);
  if (threadIdx.x == 0) {
    sc->cg_norm_b = sqrt(s);
    sc->cg_rho = 1.0; sc->cg_last_rho = 1.0; sc->cg_Q0 = 0.0; sc->cg_Q1 = 0.0; sc->cg_zeta = 0.0;
    sc->cg_iter = 0;
    sc->cg_state = (s == 0.0) ? 1 : 0;  // b == 0: x = 0 is the solution
  }
}

// z = M^{-1} r ; rho = r.z ; p = z + (rho / last_rho) p
__global__ void __launch_bounds__(kCgThreads) cg_direction_kernel(int n, int it, const int* __restrict__ col_blk,
                                                                  const int* __restrict__ blk_off, const int* __restrict__ blk_dim,
                                                                  const int* __restrict__ blk_moff, const double* __restrict__ Minv,
                                                                  const double* __restrict__ r, double* __restrict__ z,
                                                                  double* __restrict__ p, DeviceScalars* sc) {
  if (sc->cg_state != 0) return;
  double s = 0.0;
  for (int j = threadIdx.x; j < n; j += kCgThreads) {
    const double zj = PrecondRow(j, col_blk, blk_off, blk_dim, blk_moff, Minv, r);
    z[j] = zj;
    s += r[j] * zj;
  }
  const double rho = BlockAllSum(s);
  const double last_rho = sc->cg_rho;
  __syncthreads();
  if (IsZeroOrInf(rho) || isnan(rho)) {
    if (threadIdx.x == 0) { sc->cg_state = 3; sc->cg_iter = it; }
    return;
  }
  double beta = 0.0;
  if (it > 1) {
    beta = rho / last_rho;
    if (IsZeroOrInf(beta) || isnan(beta)) {
      if (threadIdx.x == 0) { sc->cg_state = 3; sc->cg_iter = it; }
      return;
    }
  }
  for (int j = threadIdx.x; j < n; j += kCgThreads) p[j] = (it == 1) ? z[j] : z[j] + beta * p[j];
  if (threadIdx.x == 0) { sc->cg_last_rho = last_rho; sc->cg_rho = rho; sc->cg_beta = beta; }
}

// finishing part shared by the plain and the residual-reset iteration
__device__ inline void CgFinish(int n, int it, int min_it, int max_it, double q_tol, const double* __restrict__ b,
                                const double* __restrict__ x, const double* __restrict__ r, DeviceScalars* sc) {
  double s = 0.0;
  for (int j = threadIdx.x; j < n; j += kCgThreads) s += x[j] * (b[j] + r[j]);
  const double Q1 = -1.0 * BlockAllSum(s);
  if (threadIdx.x == 0) {
    const double zeta = it * (Q1 - sc->cg_Q0) / Q1;
    sc->cg_zeta = zeta; sc->cg_Q1 = Q1; sc->cg_Q0 = Q1; sc->cg_iter = it;
    if (zeta < q_tol && it >= min_it) sc->cg_state = 1;
    else if (it >= max_it) sc->cg_state = 2;
  }
}

// alpha = rho / p.q ; x += alpha p ; r -= alpha q (unless this is a reset iteration)
__global__ void __launch_bounds__(kCgThreads) cg_update_kernel(int n, int it, int reset, int min_it, int max_it, double q_tol,
                                                               const double* __restrict__ b, const double* __restrict__ p,
                                                               const double* __restrict__ q, double* __restrict__ x,
                                                               double* __restrict__ r, DeviceScalars* sc) {
  if (sc->cg_state != 0) return;
  double s = 0.0;
  for (int j = threadIdx.x; j < n; j += kCgThreads) s += p[j] * q[j];
  const double pq = BlockAllSum(s);
  if (pq <= 0.0 || isinf(pq)) {
    if (threadIdx.x == 0) { sc->cg_state = 2; sc->cg_iter = it; }
    return;
  }
  const double alpha = sc->cg_rho / pq;
  if (isinf(alpha)) {
    if (threadIdx.x == 0) { sc->cg_state = 3; sc->cg_iter = it; }
    return;
  }
  for (int j = threadIdx.x; j < n; j += kCgThreads) {
    x[j] += alpha * p[j];
    if (!reset) r[j] -= alpha * q[j];
  }
  if (threadIdx.x == 0) { sc->cg_pq = pq; sc->cg_alpha = alpha; }
  if (reset) return;  // r is recomputed from scratch by cg_reset_kernel after S x
  __syncthreads();
  CgFinish(n, it, min_it, max_it, q_tol, b, x, r, sc);
}

// r = b - S x (tmp = S x), then the stopping test
__global__ void __launch_bounds__(kCgThreads) cg_reset_kernel(int n, int it, int min_it, int max_it, double q_tol,
                                                              const double* __restrict__ b, const double* __restrict__ tmp,
                                                              const double* __restrict__ x, double* __restrict__ r,
                                                              DeviceScalars* sc) {
  if (sc->cg_state != 0 || sc->cg_iter == it) return;  // cg_iter == it: this iteration already stopped
  for (int j = threadIdx.x; j < n; j += kCgThreads) r[j] = b[j] - tmp[j];
  __syncthreads();
  CgFinish(n, it, min_it, max_it, q_tol, b, x, r, sc);
}

__global__ void __launch_bounds__(256) cg_zero_kernel(int n, double* __restrict__ x) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j < n) x[j] = 0.0;
}

// Solves S y = rhs into d_yc.  *iters = CG iterations used.  Sets scalars.solver_failed on
// a CG breakdown (LinearSolver FAILURE); "no convergence" leaves a usable iterate.
int SolvePCG(Problem* p, const sba_solver_options& o, int* iters) {
  const int n = p->n_cdim;
  *iters = 0;
  if (n == 0) return SBA_OK;
  cudaStream_t st = p->stream;
  DeviceScalars* sc = p->d_scalars.p;
  const double* b = p->d_rhs.p;
  double* x = p->d_yc.p;
  const int min_it = o.min_linear_solver_iterations, max_it = std::max(1, o.max_linear_solver_iterations);
  const int poll_every = p->cg_poll_every > 0 ? p->cg_poll_every : 1;
  const bool persistent = p->bsr.enabled && !std::getenv("SBA_PCG_LAUNCHES");
  if (persistent) {  // explicit block-sparse S: the whole solve is one cooperative kernel (pcg_blocks.cu)
    const int rc = SolvePCGBlocks(p, o);
    if (rc) return rc;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    SBA_CUDA(cudaStreamIsCapturing(st, &cap));
    if (cap != cudaStreamCaptureStatusNone) return SBA_OK;  // graph capture: the kernel itself flags a breakdown; the caller reads cg_iter later
  } else {
    cg_init_kernel<<<1, kCgThreads, 0, st>>>(n, b, x, p->d_cg_r.p, sc);
    p->launches++;
  }
  for (int it = 1; it <= max_it && !persistent; ++it) {
    cg_direction_kernel<<<1, kCgThreads, 0, st>>>(n, it, p->d_col_blk.p, p->d_blk_off.p, p->d_blk_dim.p, p->d_blk_moff.p,
                                                  p->d_Minv.p, p->d_cg_r.p, p->d_cg_z.p, p->d_cg_p.p, sc);
    p->launches++;
    int rc = SchurMatVec(p, p->d_cg_p.p, p->d_cg_q.p, &sc->cg_state);
    if (rc) return rc;
    const int reset = (it % 10 == 0) ? 1 : 0;
    cg_update_kernel<<<1, kCgThreads, 0, st>>>(n, it, reset, min_it, max_it, o.eta, b, p->d_cg_p.p, p->d_cg_q.p, x,
                                               p->d_cg_r.p, sc);
    p->launches++;
    if (reset) {
      rc = SchurMatVec(p, x, p->d_tmp_c.p, &sc->cg_state);
      if (rc) return rc;
      cg_reset_kernel<<<1, kCgThreads, 0, st>>>(n, it, min_it, max_it, o.eta, b, p->d_tmp_c.p, x, p->d_cg_r.p, sc);
      p->launches++;
    }
    if (it % poll_every == 0 || it == max_it) {
      SBA_CUDA(cudaMemcpyAsync(p->h_scalars, sc, sizeof(DeviceScalars), cudaMemcpyDeviceToHost, st));
      SBA_CUDA(cudaStreamSynchronize(st));
      if (p->h_scalars->cg_state != 0) break;
    }
  }
  SBA_CUDA(cudaMemcpyAsync(p->h_scalars, sc, sizeof(DeviceScalars), cudaMemcpyDeviceToHost, st));
  SBA_CUDA(cudaStreamSynchronize(st));
  *iters = p->h_scalars->cg_iter;
  if (p->h_scalars->cg_state == 3) {
    // LinearSolver FAILURE: the step is unusable (TrustRegionMinimizer treats it as invalid)
    p->h_scalars->solver_failed = 1;
    SBA_CUDA(cudaMemcpyAsync(&sc->solver_failed, &p->h_scalars->solver_failed, sizeof(int), cudaMemcpyHostToDevice, st));
  }
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

}  // namespace sba
