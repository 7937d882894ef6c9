// K-CHOL: blocked right-looking fp64 Cholesky of the reduced camera matrix and the
// two triangular solves, entirely on the device.
//
// This replaces the dense / sparse Cholesky Ceres runs on the Schur complement for
// DENSE_SCHUR and SPARSE_SCHUR (selected at
//   /root/reference/src/optim/bundle_adjustment.cc:276-282, solved inside ceres::Solve :306;
// SURVEY.md App. B.5).  SPARSE_SCHUR sizes (<= 1000 images, n <= ~12k) run the same
// dense factorisation: the reduced camera matrix of a 500-image scene is 4000^2 fp64
// = 128 MB, a fraction of HBM, and 21 GFLOP.
//
// Layout: S is row-major, leading dimension n, LOWER triangle valid, with one extra
// row (row n) that holds the right-hand side.  Factoring the augmented matrix
//   [ S  . ]        [ L   . ]
//   [ b' . ]   ->   [ y'  . ]   with  L y = b
// performs the forward substitution as part of the panel updates; only the backward
// substitution L' x = y runs as a separate sweep.
//
// Per block column k (NB = 64):
//   potrf_diag   1 CTA      : L_kk = chol(A_kk), Linv_kk = L_kk^{-1}
//   trsm_panel   1 CTA/tile : L_ik = A_ik Linv_kk'                (i > k, incl. rhs row)
//   syrk_update  1 CTA/tile : A_ij -= L_ik L_jk'                  (i >= j > k) -- fp64 DMMA
#include "reduce.cuh"
#include "solver.h"

namespace sba {

constexpr int NB = 64;
constexpr int LDS = NB + 1;

__global__ void __launch_bounds__(256) chol_load_rhs_kernel(int n, const double* __restrict__ rhs, double* __restrict__ S) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j < n) S[(size_t)n * n + j] = rhs[j];
}

// ---- 64x64 diagonal block: warp-register Cholesky -----------------------------------------
// The block is split 2x2 into 32x32 sub-blocks.  A 32x32 factorisation runs in ONE warp with
// lane r holding row r in registers (fully unrolled, shuffles broadcast the pivot column), so
// the 32-step dependency chain costs a shuffle + FMA per step instead of block-wide barriers.
constexpr unsigned kFull = 0xffffffffu;

// in: a[c] = A[lane][c] (lower part valid).  out: a[c] = L[lane][c] (zero above the diagonal).
__device__ inline bool Chol32(double (&a)[32], int lane) {
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    double d = __shfl_sync(kFull, a[j], j);
    if (!(d > 0.0) || !isfinite(d)) { ok = false; d = 1.0; }
    const double s = sqrt(d), inv = 1.0 / s;
    const double l = (lane == j) ? s : (lane > j ? a[j] * inv : 0.0);
    a[j] = l;
#pragma unroll
    for (int c = j + 1; c < 32; ++c) {
      const double lc = __shfl_sync(kFull, l, c);
      a[c] -= l * lc;
    }
  }
  return ok;
}

// lane c: column c of the inverse of the lower-triangular 32x32 L stored in shared memory
// (row stride LDS, at L[0][0]); x[i] = Linv[i][c].
__device__ inline void TriInv32(const double* __restrict__ L, int lane, double (&x)[32]) {
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    double s = (i == lane) ? 1.0 : 0.0;
#pragma unroll
    for (int k = 0; k < i; ++k) s -= L[i * LDS + k] * x[k];   // x[k] == 0 for k < lane
    x[i] = (i >= lane) ? s / L[i * LDS + i] : 0.0;
  }
}

__global__ void __launch_bounds__(128) potrf_diag_kernel(double* __restrict__ S, int n, int k0,
                                                         double* __restrict__ Linv, DeviceScalars* sc) {
  extern __shared__ double smem[];
  double(*A)[LDS] = reinterpret_cast<double(*)[LDS]>(smem);             // block, then its factor
  double(*X)[LDS] = reinterpret_cast<double(*)[LDS]>(smem + NB * LDS);  // inverse of the factor
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  const int nb = min(NB, n - k0);
  for (int idx = tid; idx < NB * NB; idx += 128) {
    const int r = idx / NB, c = idx - r * NB;
    double v = (r == c) ? 1.0 : 0.0;   // identity padding keeps a partial block well posed
    if (r < nb && c <= r) v = S[(size_t)(k0 + r) * n + k0 + c];
    A[r][c] = v;
    X[r][c] = 0.0;
  }
  __syncthreads();
  bool ok = true;
  // (1) L11 = chol(A11)
  if (w == 0) {
    double a[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) a[c] = A[lane][c];
    ok = Chol32(a, lane);
#pragma unroll
    for (int c = 0; c < 32; ++c) A[lane][c] = a[c];
  }
  __syncthreads();
  // (2) warp 0: L21 = A21 L11^{-T} (one row per lane, forward substitution along the row);
  //     warp 1: X11 = L11^{-1}
  if (w == 0) {
    double x[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      double s = A[32 + lane][c];
#pragma unroll
      for (int m = 0; m < c; ++m) s -= x[m] * A[c][m];
      x[c] = s / A[c][c];
    }
#pragma unroll
    for (int c = 0; c < 32; ++c) A[32 + lane][c] = x[c];
  } else if (w == 1) {
    double x[32];
    TriInv32(&A[0][0], lane, x);
#pragma unroll
    for (int i = 0; i < 32; ++i) X[i][lane] = x[i];
  }
  __syncthreads();
  // (3) A22 -= L21 L21^T   (4 warps: row = lane, 8 columns each)
  {
    double acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.0;
    for (int m = 0; m < 32; ++m) {
      const double l = A[32 + lane][m];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += l * A[32 + w * 8 + e][m];
    }
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 8; ++e) A[32 + lane][32 + w * 8 + e] -= acc[e];
  }
  __syncthreads();
  // (4) L22 = chol(A22)
  if (w == 0) {
    double a[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) a[c] = A[32 + lane][32 + c];
    ok = Chol32(a, lane) && ok;
#pragma unroll
    for (int c = 0; c < 32; ++c) A[32 + lane][32 + c] = a[c];
    if (!ok && lane == 0) { sc->chol_fail = 1; sc->solver_failed = 1; }
  }
  __syncthreads();
  // (5) warp 0: X22 = L22^{-1};  warps 1..3 + 0 afterwards: T = L21 X11 (into the unused upper-right of A)
  if (w == 0) {
    double x[32];
    TriInv32(&A[32][32], lane, x);
#pragma unroll
    for (int i = 0; i < 32; ++i) X[32 + i][32 + lane] = x[i];
  }
  {
    double acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.0;
    for (int m = 0; m < 32; ++m) {
      const double l = A[32 + lane][m];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += l * X[m][w * 8 + e];
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) A[lane][32 + w * 8 + e] = acc[e];   // T[lane][.] parked above the diagonal
  }
  __syncthreads();
  // (6) X21 = -X22 T
  {
    double acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.0;
    for (int m = 0; m < 32; ++m) {
      const double l = X[32 + lane][32 + m];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += l * A[m][32 + w * 8 + e];
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) X[32 + lane][w * 8 + e] = -acc[e];
  }
  __syncthreads();
  for (int idx = tid; idx < NB * NB; idx += 128) {
    const int r = idx / NB, c = idx - r * NB;
    if (r < nb && c <= r) S[(size_t)(k0 + r) * n + k0 + c] = A[r][c];
    Linv[idx] = X[r][c];
  }
}

// L_ik = A_ik * Linv_kk^T for the row tile starting at r0 = k0 + NB*(1 + blockIdx.x)
__global__ void __launch_bounds__(256) trsm_panel_kernel(double* __restrict__ S, int n, int n_rows, int k0,
                                                         const double* __restrict__ Linv) {
  extern __shared__ double smem[];
  double(*A)[LDS] = reinterpret_cast<double(*)[LDS]>(smem);
  double(*Li)[LDS] = reinterpret_cast<double(*)[LDS]>(smem + NB * LDS);
  const int tid = threadIdx.x;
  const int nb = min(NB, n - k0);
  const int r0 = k0 + nb + NB * blockIdx.x;
  const int nr = min(NB, n_rows - r0);
  for (int idx = tid; idx < NB * NB; idx += 256) {
    const int r = idx / NB, c = idx - r * NB;
    A[r][c] = (r < nr && c < nb) ? S[(size_t)(r0 + r) * n + k0 + c] : 0.0;
    Li[r][c] = Linv[idx];
  }
  __syncthreads();
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, 4 x 4 outputs each (strided by 16)
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
  for (int m = 0; m < nb; ++m) {
    double av[4], bv[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) av[a] = A[ty + 16 * a][m];
#pragma unroll
    for (int b = 0; b < 4; ++b) bv[b] = Li[tx + 16 * b][m];  // Linv[c][m], zero for m > c
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) acc[a][b] += av[a] * bv[b];
  }
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int r = ty + 16 * a, c = tx + 16 * b;
      if (r < nr && c < nb) S[(size_t)(r0 + r) * n + k0 + c] = acc[a][b];
    }
}

// fp64 tensor-core MMA (DMMA): D(8x8) += A(8x4, row) * B(4x8, col)
__device__ inline void Dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// A_ij -= L_ik L_jk^T on 64x64 tiles of the trailing lower triangle (rows up to n_rows).
// 8 warps; warp w owns the 8-row strip w of the tile and sweeps the eight 8x8 column
// blocks with DMMA: A fragment = L_ik[8 rows][4 k], B fragment = L_jk[8 cols][4 k].
__global__ void __launch_bounds__(256) syrk_update_kernel(double* __restrict__ S, int n, int n_rows, int k0, int nb,
                                                          int t0 /* first trailing row */, int m_tiles) {
  constexpr int KC = 32, LDK = KC + 4;  // k-chunk staged in shared memory, padded to spread banks
  __shared__ double As[NB][LDK];        // [row][k]
  __shared__ double Bs[NB][LDK];
  // linear tile id -> (bi >= bj) in the lower triangle of an m_tiles x m_tiles grid
  const int t = blockIdx.x;
  int bi = (int)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
  while ((bi + 1) * (bi + 2) / 2 <= t) ++bi;
  while (bi * (bi + 1) / 2 > t) --bi;
  const int bj = t - bi * (bi + 1) / 2;
  (void)m_tiles;
  const int i0 = t0 + NB * bi, j0 = t0 + NB * bj;
  const int tid = threadIdx.x;
  const int w = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;  // fragment coordinates: row/col g, k index q
  double acc[8][2];
#pragma unroll
  for (int cb = 0; cb < 8; ++cb) { acc[cb][0] = 0.0; acc[cb][1] = 0.0; }
  for (int kc = 0; kc < nb; kc += KC) {
    for (int idx = tid; idx < NB * KC; idx += 256) {
      const int r = idx / KC, c = idx - r * KC;
      As[r][c] = (i0 + r < n_rows && kc + c < nb) ? S[(size_t)(i0 + r) * n + k0 + kc + c] : 0.0;
      Bs[r][c] = (j0 + r < n && kc + c < nb) ? S[(size_t)(j0 + r) * n + k0 + kc + c] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int m = 0; m < KC; m += 4) {
      const double a = As[8 * w + g][m + q];
#pragma unroll
      for (int cb = 0; cb < 8; ++cb) {
        const double b = Bs[8 * cb + g][m + q];
        Dmma884(acc[cb][0], acc[cb][1], a, b);
      }
    }
    __syncthreads();
  }
  // D fragment: row g, cols 2q, 2q+1
  const int r = i0 + 8 * w + g;
  if (r < n_rows) {
#pragma unroll
    for (int cb = 0; cb < 8; ++cb) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int c = j0 + 8 * cb + 2 * q + e;
        if (c < n && c <= r) S[(size_t)r * n + c] -= acc[cb][e];
      }
    }
  }
}

// backward substitution sweep for block column k: every CTA recomputes
// x_k = Linv_kk^T y_k (64x64, cheap) and then updates its slice of y_j -= L_kj^T x_k, j < k.
__global__ void __launch_bounds__(256) back_sweep_kernel(const double* __restrict__ S, int n, int k0, int nb,
                                                         const double* __restrict__ Linv, double* __restrict__ y,
                                                         double* __restrict__ x) {
  __shared__ double yk[NB], xk[NB];
  const int tid = threadIdx.x;
  if (tid < NB) yk[tid] = tid < nb ? y[k0 + tid] : 0.0;
  __syncthreads();
  if (tid < NB) {
    double s = 0.0;
    for (int m = tid; m < nb; ++m) s += Linv[m * NB + tid] * yk[m];  // (Linv^T)[tid][m]
    xk[tid] = s;
    if (blockIdx.x == 0 && tid < nb) x[k0 + tid] = s;
  }
  __syncthreads();
  const int c = blockIdx.x * 256 + tid;
  if (c < k0) {
    double s = 0.0;
    for (int r = 0; r < nb; ++r) s += S[(size_t)(k0 + r) * n + c] * xk[r];
    y[c] -= s;
  }
}

__global__ void __launch_bounds__(256) chol_fetch_y_kernel(int n, const double* __restrict__ S, double* __restrict__ y) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j < n) y[j] = S[(size_t)n * n + j];
}

int CholeskySolve(Problem* p) {
  const int n = p->n_cdim;
  if (n == 0) return SBA_OK;
  if (n <= kCholSmallMax) return CholeskySolveSmall(p);
  static bool attr_set = false;
  const int dyn = 2 * NB * LDS * (int)sizeof(double);
  if (!attr_set) {
    SBA_CUDA(cudaFuncSetAttribute(potrf_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
    SBA_CUDA(cudaFuncSetAttribute(trsm_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
    attr_set = true;
  }
  cudaStream_t st = p->stream;
  double* S = p->d_S.p;
  const int n_rows = n + 1;
  const int nblk = DivUp(n, NB);
  SBA_CUDA(p->d_Linv.Alloc((size_t)nblk * NB * NB));
  chol_load_rhs_kernel<<<DivUp(n, 256), 256, 0, st>>>(n, p->d_rhs.p, S);
  p->launches++;
  for (int kb = 0; kb < nblk; ++kb) {
    const int k0 = kb * NB, nb = std::min(NB, n - k0);
    double* Linv = p->d_Linv.p + (size_t)kb * NB * NB;
    potrf_diag_kernel<<<1, 128, dyn, st>>>(S, n, k0, Linv, p->d_scalars.p);
    const int rem_rows = n_rows - (k0 + nb);
    const int row_tiles = DivUp(rem_rows, NB);
    trsm_panel_kernel<<<row_tiles, 256, dyn, st>>>(S, n, n_rows, k0, Linv);
    p->launches += 2;
    if (k0 + nb < n) {
      const int m_tiles = row_tiles;  // tiles cover rows [k0+nb, n_rows); columns [k0+nb, n)
      const int n_lower = m_tiles * (m_tiles + 1) / 2;
      syrk_update_kernel<<<n_lower, 256, 0, st>>>(S, n, n_rows, k0, nb, k0 + nb, m_tiles);
      p->launches++;
    }
  }
  // y = row n of the factor; back substitution into d_yc
  chol_fetch_y_kernel<<<DivUp(n, 256), 256, 0, st>>>(n, S, p->d_tmp_c.p);
  p->launches++;
  for (int kb = nblk - 1; kb >= 0; --kb) {
    const int k0 = kb * NB, nb = std::min(NB, n - k0);
    back_sweep_kernel<<<std::max(1, DivUp(k0, 256)), 256, 0, st>>>(S, n, k0, nb, p->d_Linv.p + (size_t)kb * NB * NB,
                                                                  p->d_tmp_c.p, p->d_yc.p);
    p->launches++;
  }
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

}  // namespace sba
