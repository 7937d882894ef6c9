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
#include <cstdio>
#include <cstdlib>

#include "reduce.cuh"
#include "solver.h"

namespace sba {

constexpr int NB = 64;

__global__ void __launch_bounds__(256) chol_load_rhs_kernel(int n, int ld, const double* __restrict__ rhs, double* __restrict__ S) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j < n) S[(size_t)n * ld + j] = rhs[j];
}

__device__ __forceinline__ void CpAsync16(void* smem_dst, const void* gsrc, bool pred) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int bytes = pred ? 16 : 0;  // src-size 0: the 16 bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(bytes) : "memory");
}

__device__ __forceinline__ void CpAsync8(void* smem_dst, const void* gsrc, bool pred) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int bytes = pred ? 8 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(d), "l"(gsrc), "r"(bytes) : "memory");
}

// ---- panel diagonal block (PB = 128): L_pp = chol(A_pp) and Linv_pp = L_pp^{-1}, one CTA ---------------------------
// This kernel is the critical path of the factorisation (one launch per panel, nothing else can run until it ends), so
// it is built for latency:
//   * the block lives in ONE shared 128 x 129 array P: the factor L in the lower part (P[r][c], c <= r), the inverse
//     X = L^{-1} transposed into the upper part (X[r][c] at P[c][r + 1], c <= r);
//   * a 32 x 32 diagonal sub-block is factored by ONE warp in registers (lane r holds row r; the pivot column travels by
//     shuffle; 1 / sqrt(d) is one rsqrt -- no fp64 division anywhere on the chain) and inverted by the same warp, a
//     column per lane;
//   * everything around it (rows below, trailing update, off-diagonal blocks of X) is 4 x 4 register-tiled GEMM work
//     spread over all 256 threads.
//   for each 32-column block kb:  warp 0: L_kk = chol(A_kk), X_kk = L_kk^{-1}
//                                 all   : L_ik = A_ik X_kk^T (rows below), A_ij -= L_ik L_jk^T (rest of the block)
//   then the off-diagonal blocks of X by block rows:  X_ij = -X_ii sum_{k=j..i-1} L_ik X_kj
// A partial last panel (nb < 128) is padded with the identity, which keeps it well posed.
constexpr int PB = 128, LDP = PB + 1, SB = 32;
constexpr unsigned kFull = 0xffffffffu;
struct DiagSmem { double P[PB][LDP]; double T[64 * 65]; };  // T: scratch, [96][33] or [64][65]

// in: a[c] = A[lane][c] (lower part valid).  out: a[c] = L[lane][c] (zero above the diagonal); lane j returns 1 / L[j][j].
__device__ inline bool Chol32(double (&a)[32], int lane, double* my_inv) {
  bool ok = true;
  double mine = 1.0;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    double d = __shfl_sync(kFull, a[j], j);
    if (!(d > 0.0) || !isfinite(d)) { ok = false; d = 1.0; }
    const double inv = rsqrt(d);
    const double l = (lane == j) ? d * inv : (lane > j ? a[j] * inv : 0.0);
    if (lane == j) mine = inv;
    a[j] = l;
#pragma unroll
    for (int c = j + 1; c < 32; ++c) {
      const double lc = __shfl_sync(kFull, l, c);
      a[c] -= l * lc;
    }
  }
  *my_inv = mine;
  return ok;
}

// acc[i][j] += sum_{m in [m0, m1)} A(i, m) * B(j, m)
template <class FA, class FB>
__device__ __forceinline__ void Acc4x4(FA A, FB B, int m0, int m1, double (&acc)[4][4]) {
#pragma unroll 4
  for (int m = m0; m < m1; ++m) {
    double a[4], b[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { a[i] = A(i, m); b[i] = B(i, m); }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] += a[i] * b[j];
  }
}

__global__ void __launch_bounds__(256) chol_diag128_kernel(double* __restrict__ S, int n, int ld, int k0, double* __restrict__ Linv,
                                                           DeviceScalars* sc, long long* __restrict__ trace) {
  extern __shared__ double smem[];
  DiagSmem& M = *reinterpret_cast<DiagSmem*>(smem);
  int n_stamp = 0;
  auto stamp = [&]() {  // SBA_CHOL_TRACE=1: globaltimer at the phase boundaries of the first panel
    if (trace && threadIdx.x == 0) { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); trace[n_stamp++] = t; }
  };
  stamp();
  double(*P)[LDP] = M.P;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  const int nb = min(PB, n - k0);
  auto X = [&](int r, int c) -> double& { return P[c][r + 1]; };  // entry (r, c), c <= r, of the inverse
  // the lower part of the block: 8-byte cp.async, all of a thread's pieces in flight together (a plain load loop paid a
  // global-memory round trip per element: 18 us); the upper part receives the inverse, every entry of which is written
  // before it is read
  for (int idx = tid; idx < PB * PB; idx += 256) {
    const int r = idx / PB, c = idx - r * PB;
    if (c > r) continue;
    if (r < nb) CpAsync8(&P[r][c], S + (size_t)(k0 + r) * ld + k0 + c, true);
    else P[r][c] = (r == c ? 1.0 : 0.0);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  stamp();
  bool ok = true;
  for (int kb = 0; kb < PB / SB; ++kb) {
    const int o = kb * SB;
    if (w == 0) {
      double a[32], my_inv;
#pragma unroll
      for (int c = 0; c < 32; ++c) a[c] = P[o + lane][o + c];   // (entries above the diagonal: ignored by Chol32)
      ok = Chol32(a, lane, &my_inv) && ok;
#pragma unroll
      for (int c = 0; c < 32; ++c) if (c <= lane) P[o + lane][o + c] = a[c];
      __syncwarp();
      // X_kk = L_kk^{-1}: lane c owns column c; x[i] = (delta_ic - sum_{k<i} L[i][k] x[k]) / L[i][i]
      double x[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        double s0 = (i == lane) ? 1.0 : 0.0, s1 = 0.0;
#pragma unroll
        for (int k = 0; k + 1 < i; k += 2) { s0 -= P[o + i][o + k] * x[k]; s1 -= P[o + i][o + k + 1] * x[k + 1]; }
        if (i & 1) s0 -= P[o + i][o + i - 1] * x[i - 1];
        const double di = __shfl_sync(kFull, my_inv, i);
        x[i] = (i >= lane) ? (s0 + s1) * di : 0.0;   // x[k] == 0 for k < lane
      }
#pragma unroll
      for (int i = 0; i < 32; ++i) if (i >= lane) X(o + i, o + lane) = x[i];
    }
    __syncthreads();
    stamp();
    const int R = PB - o - SB;  // rows below the diagonal block
    if (R > 0) {
      // L_ik = A_ik X_kk^T into T (the outputs of a row overwrite its inputs): 4 x 4 tiles, (R / 4) x 8 of them
      for (int tile = tid; tile < (R / 4) * 8; tile += 256) {
        const int ti = tile >> 3, tj = tile & 7;
        double acc[4][4] = {};
        auto A = [&](int i, int m) { return P[o + SB + 4 * ti + i][o + m]; };
        Acc4x4(A, [&](int j, int m) { return X(o + 4 * tj + j, o + m); }, 0, 4 * tj, acc);
        Acc4x4(A, [&](int j, int m) { return m <= 4 * tj + j ? X(o + (m <= 4 * tj + j ? 4 * tj + j : m), o + m) : 0.0; }, 4 * tj, 4 * tj + 4, acc);
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) M.T[(4 * ti + i) * (SB + 1) + 4 * tj + j] = acc[i][j];
      }
      __syncthreads();
      for (int idx = tid; idx < R * SB; idx += 256) { const int r = idx / SB, c = idx - r * SB; P[o + SB + r][o + c] = M.T[r * (SB + 1) + c]; }
      __syncthreads();
      stamp();
      // A_ij -= L_ik L_jk^T on the lower part of the remaining R x R block: 4 x 4 tiles (ti >= tj)
      const int TT = R / 4;
      for (int tile = tid; tile < TT * (TT + 1) / 2; tile += 256) {
        int ti = (int)((sqrtf(8.0f * (float)tile + 1.0f) - 1.0f) * 0.5f);
        while ((ti + 1) * (ti + 2) / 2 <= tile) ++ti;
        while (ti * (ti + 1) / 2 > tile) --ti;
        const int tj = tile - ti * (ti + 1) / 2;
        double acc[4][4] = {};
        Acc4x4([&](int i, int m) { return P[o + SB + 4 * ti + i][o + m]; },
               [&](int j, int m) { return P[o + SB + 4 * tj + j][o + m]; }, 0, SB, acc);
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (4 * tj + j <= 4 * ti + i) P[o + SB + 4 * ti + i][o + SB + 4 * tj + j] -= acc[i][j];
      }
      __syncthreads();
      stamp();
    }
  }
  if (tid == 0 && !ok) { sc->chol_fail = 1; sc->solver_failed = 1; }
  // the factor goes back while the inverse is completed
  for (int idx = tid; idx < PB * PB; idx += 256) {
    const int r = idx / PB, c = idx - r * PB;
    if (r < nb && c <= r) S[(size_t)(k0 + r) * ld + k0 + c] = P[r][c];
  }
  // off-diagonal blocks of X in two levels of the 2 x 2 block inverse  X_lo,hi = -X_hi L_hi,lo X_lo  (4 x 4 tiles):
  //   level 1: the 32-blocks (1,0) and (3,2), side by side;  level 2: the 64-block (rows 64.., columns ..63)
  for (int level = 1; level <= 2; ++level) {
    const int bs = level == 1 ? SB : 2 * SB, tpb = bs / 4;     // block size, tiles per block side
    const int n_pairs = level == 1 ? 2 : 1, ldt = bs + 1;
    for (int tile = tid; tile < n_pairs * tpb * tpb; tile += 256) {   // T = L_hi,lo X_lo
      const int pr = tile / (tpb * tpb), rem = tile - pr * tpb * tpb, ti = rem / tpb, tj = rem - ti * tpb;
      const int lo = pr * 2 * bs, hi = lo + bs;              // first row / column of the two diagonal blocks of the pair
      const int col0 = lo + 4 * tj;                          // X[m][col] is zero for m < col
      double acc[4][4] = {};
      auto A = [&](int i, int m) { return P[hi + 4 * ti + i][m]; };
      Acc4x4(A, [&](int j, int m) { return m >= col0 + j ? X(m >= col0 + j ? m : col0 + j, col0 + j) : 0.0; }, col0, col0 + 4, acc);
      Acc4x4(A, [&](int j, int m) { return X(m, col0 + j); }, col0 + 4, hi, acc);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) M.T[(pr * bs + 4 * ti + i) * ldt + 4 * tj + j] = acc[i][j];
    }
    __syncthreads();
    for (int tile = tid; tile < n_pairs * tpb * tpb; tile += 256) {   // X_hi,lo = -X_hi T, X_hi lower triangular
      const int pr = tile / (tpb * tpb), rem = tile - pr * tpb * tpb, ti = rem / tpb, tj = rem - ti * tpb;
      const int lo = pr * 2 * bs, hi = lo + bs;
      double acc[4][4] = {};
      auto B = [&](int j, int m) { return M.T[(pr * bs + m) * ldt + 4 * tj + j]; };
      Acc4x4([&](int i, int m) { return X(hi + 4 * ti + i, hi + m); }, B, 0, 4 * ti, acc);
      Acc4x4([&](int i, int m) { return m <= 4 * ti + i ? X(hi + (m <= 4 * ti + i ? 4 * ti + i : m), hi + m) : 0.0; }, B,
             4 * ti, 4 * ti + 4, acc);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) X(hi + 4 * ti + i, lo + 4 * tj + j) = -acc[i][j];
    }
    __syncthreads();
    stamp();
  }
  for (int idx = tid; idx < PB * PB; idx += 256) {
    const int r = idx / PB, c = idx - r * PB;
    Linv[idx] = c <= r ? X(r, c) : 0.0;
  }
  stamp();
}

// ---- panel rows below the diagonal block: L_ik = A_ik Linv_pp^T for a tile of 16 * RPT rows (incl. the rhs row) ------
// RPT = 4 (64-row tiles) for the bulk of a panel; RPT = 1 (16-row tiles, eight CTAs for 128 rows) for the rows of the
// NEXT diagonal tile, which sit on the factorisation's critical chain: there the latency of one CTA is what counts.
template <int RPT>
__global__ void __launch_bounds__(256) chol_trsm128_kernel(double* __restrict__ S, int n, int ld, int n_rows, int k0, int nb,
                                                           const double* __restrict__ Linv, int row_off) {
  constexpr int NBT = 16 * RPT;
  extern __shared__ double smem[];
  double(*A)[LDP] = reinterpret_cast<double(*)[LDP]>(smem);                // [NBT][129]
  double(*Li)[LDP] = reinterpret_cast<double(*)[LDP]>(smem + NBT * LDP);   // [128][129]
  const int tid = threadIdx.x;
  const int r0 = k0 + nb + row_off + NBT * blockIdx.x;
  const int nr = min(NBT, n_rows - r0);
  // all loads of a batch are issued before the first shared-memory store (these kernels sit on the factorisation's
  // critical chain: with the loads trickling out four at a time the 16-row variant took 24 us, most of it waiting)
  {
    double va[(NBT * PB) / 256];
#pragma unroll
    for (int it = 0; it < (NBT * PB) / 256; ++it) {
      const int idx = tid + 256 * it, r = idx / PB, c = idx - r * PB;
      va[it] = (r < nr && c < nb) ? S[(size_t)(r0 + r) * ld + k0 + c] : 0.0;
    }
#pragma unroll
    for (int it = 0; it < (NBT * PB) / 256; ++it) { const int idx = tid + 256 * it; A[idx / PB][idx % PB] = va[it]; }
  }
#pragma unroll 1
  for (int b0 = 0; b0 < PB * PB; b0 += 256 * 16) {
    double vl[16];
#pragma unroll
    for (int it = 0; it < 16; ++it) vl[it] = Linv[b0 + tid + 256 * it];
#pragma unroll
    for (int it = 0; it < 16; ++it) { const int idx = b0 + tid + 256 * it; Li[idx / PB][idx % PB] = vl[it]; }
  }
  __syncthreads();
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, RPT rows x 8 columns each (strided by 16)
  double acc[RPT][8];
#pragma unroll
  for (int a = 0; a < RPT; ++a)
#pragma unroll
    for (int b = 0; b < 8; ++b) acc[a][b] = 0.0;
  const int mend = min(nb, tx + 113);  // Linv[c][m] = 0 for m > c; the thread's largest column is tx + 112
  for (int m = 0; m < mend; ++m) {
    double av[RPT], bv[8];
#pragma unroll
    for (int a = 0; a < RPT; ++a) av[a] = A[ty + 16 * a][m];
#pragma unroll
    for (int b = 0; b < 8; ++b) bv[b] = Li[tx + 16 * b][m];
#pragma unroll
    for (int a = 0; a < RPT; ++a)
#pragma unroll
      for (int b = 0; b < 8; ++b) acc[a][b] += av[a] * bv[b];
  }
#pragma unroll
  for (int a = 0; a < RPT; ++a)
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int r = ty + 16 * a, c = tx + 16 * b;
      if (r < nr && c < nb) S[(size_t)(r0 + r) * ld + k0 + c] = acc[a][b];
    }
}

// ---- update of the NEXT diagonal tile alone: A_tt -= L_tk L_tk^T (rows / columns [t0, t0 + 128)) in 32 x 32 pieces,
// one CTA per piece of the lower triangle (grid 4 x 4, ten CTAs work).  On the critical chain: ten small CTAs finish in
// a few microseconds where the 128 x 128 DMMA tile of chol_syrk128_kernel takes ~40 in one.
__global__ void __launch_bounds__(256) chol_syrk_next_diag_kernel(double* __restrict__ S, int n, int ld, int n_rows, int k0, int nb,
                                                                  int t0) {
  if (blockIdx.x > blockIdx.y) return;
  extern __shared__ double smem[];
  double(*As)[LDP] = reinterpret_cast<double(*)[LDP]>(smem);             // [32][129]
  double(*Bs)[LDP] = reinterpret_cast<double(*)[LDP]>(smem + 32 * LDP);  // [32][129]
  const int tid = threadIdx.x;
  const int i0 = t0 + 32 * blockIdx.y, j0 = t0 + 32 * blockIdx.x;
  {
    double va[16], vb[16];  // 32 x 128 of each operand: 16 loads per thread and operand, all in flight together
#pragma unroll
    for (int it = 0; it < 16; ++it) {
      const int idx = tid + 256 * it, r = idx / PB, c = idx - r * PB;
      va[it] = (i0 + r < n_rows && c < nb) ? S[(size_t)(i0 + r) * ld + k0 + c] : 0.0;
      vb[it] = (j0 + r < n_rows && c < nb) ? S[(size_t)(j0 + r) * ld + k0 + c] : 0.0;
    }
#pragma unroll
    for (int it = 0; it < 16; ++it) { const int idx = tid + 256 * it; As[idx / PB][idx % PB] = va[it]; Bs[idx / PB][idx % PB] = vb[it]; }
  }
  __syncthreads();
  const int tx = tid & 15, ty = tid >> 4;
  double acc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
  for (int m = 0; m < nb; ++m) {
    const double a0 = As[ty][m], a1 = As[ty + 16][m], b0 = Bs[tx][m], b1 = Bs[tx + 16][m];
    acc[0][0] += a0 * b0; acc[0][1] += a0 * b1; acc[1][0] += a1 * b0; acc[1][1] += a1 * b1;
  }
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int r = i0 + ty + 16 * a, c = j0 + tx + 16 * b;
      if (r < n_rows && c < n && c <= r && r < t0 + PB && c < t0 + PB) S[(size_t)r * ld + c] -= acc[a][b];
    }
}

// fp64 tensor-core MMA (DMMA): D(8x8) += A(8x4, row) * B(4x8, col)
__device__ __forceinline__ void Dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}
// ---- trailing update: A_ij -= L_ik L_jk^T on 128 x 128 tiles of the lower triangle (rows up to n_rows) -------------
// The one dense contraction of the path, on the fp64 tensor cores.  8 warps as 4 (rows) x 2 (columns): a warp owns a
// 32 x 64 part of the tile = 4 x 8 DMMA accumulators (64 registers); per k-step of 4 it loads 4 A and 8 B fragments
// from shared memory for 32 DMMAs.  The K = 128 panel is streamed in chunks of 32 columns through a double-buffered
// cp.async pipeline (rows of S are contiguous in k: 256-byte runs).
constexpr int TB = 128, KC = 32, LDK = KC + 4;
struct SyrkSmem { double As[2][TB][LDK], Bs[2][TB][LDK]; };

__global__ void __launch_bounds__(256) chol_syrk128_kernel(double* __restrict__ S, int n, int ld, int n_rows, int k0, int nb,
                                                           int t0 /* first trailing row */, int first_column, int bi0) {
  extern __shared__ double smem[];
  SyrkSmem& M = *reinterpret_cast<SyrkSmem*>(smem);
  // first_column: the tiles of the first tile column (they hold the next panel: updated first, so that the next panel's
  // factorisation can start while the rest of the update runs); else the rest of the lower triangle: linear tile id ->
  // (bi >= bj >= 1)
  int bi, bj;
  if (first_column) { bi = blockIdx.x + bi0; bj = 0; }
  else {
    const int t = blockIdx.x;
    bi = (int)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
    while ((bi + 1) * (bi + 2) / 2 <= t) ++bi;
    while (bi * (bi + 1) / 2 > t) --bi;
    bj = t - bi * (bi + 1) / 2 + 1;
    bi += 1;
  }
  const int i0 = t0 + TB * bi, j0 = t0 + TB * bj;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  const int wr = w >> 1, wc = w & 1;        // warp's 32-row strip, 64-column half
  const int g = lane >> 2, q = lane & 3;    // fragment coordinates: row/col g, k index q
  // The accumulators start as -C: all of a thread's 32 loads of the tile are issued back to back here, in flight under
  // the first operand chunk, and the result leaves as C_new = -(L L^T - C).  (A read-modify-write epilogue measured
  // 64 % of the kernel's stall samples: 32 dependent load -> DADD -> store round trips per thread.)
  double acc[4][8][2];
  const bool pair_ok = (ld & 1) == 0;  // rows of S are 16-byte aligned: (c, c + 1) moves as one double2
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int r = i0 + 32 * wr + 8 * a + g;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int c = j0 + 64 * wc + 8 * b + 2 * q;
      double v0 = 0.0, v1 = 0.0;
      if (r < n_rows) {
        const double* src = S + (size_t)r * ld + c;  // (dereferenced only where the guards below hold)
        if (pair_ok && c + 1 < n && c + 1 <= r) { const double2 v = *reinterpret_cast<const double2*>(src); v0 = v.x; v1 = v.y; }
        else {
          if (c < n && c <= r) v0 = src[0];
          if (c + 1 < n && c + 1 <= r) v1 = src[1];
        }
      }
      acc[a][b][0] = -v0; acc[a][b][1] = -v1;
    }
  }
  const int n_chunks = (nb + KC - 1) / KC;
  // 128 rows x 32 doubles of each operand per chunk.  Rows of S start 16-byte aligned only when n is even: then 16-byte
  // pieces (8 per thread and operand), else 8-byte ones (16 per thread and operand); a piece outside the matrix or the
  // panel is zero-filled (src-size 0).
  const bool wide = (ld & 1) == 0;
  auto stage = [&](int buf, int kc) {
    if (wide) {
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int piece = tid + 256 * it;
        const int r = piece >> 4, c2 = (piece & 15) * 2;
        const bool kin = kc + c2 < nb;  // (nb is even whenever n is)
        const int kk = k0 + (kin ? kc + c2 : 0);  // (an address inside the matrix also when nothing is read)
        CpAsync16(&M.As[buf][r][c2], S + (size_t)min(i0 + r, n_rows - 1) * ld + kk, kin && (i0 + r < n_rows));
        CpAsync16(&M.Bs[buf][r][c2], S + (size_t)min(j0 + r, n - 1) * ld + kk, kin && (j0 + r < n));
      }
    } else {
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int piece = tid + 256 * it;
        const int r = piece >> 5, c = piece & 31;
        const bool kin = kc + c < nb;
        CpAsync8(&M.As[buf][r][c], S + (size_t)min(i0 + r, n_rows - 1) * ld + k0 + min(kc + c, nb - 1), kin && (i0 + r < n_rows));
        CpAsync8(&M.Bs[buf][r][c], S + (size_t)min(j0 + r, n - 1) * ld + k0 + min(kc + c, nb - 1), kin && (j0 + r < n));
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  stage(0, 0);
  for (int ch = 0; ch < n_chunks; ++ch) {
    const int buf = ch & 1;
    if (ch + 1 < n_chunks) { stage(buf ^ 1, (ch + 1) * KC); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
    else asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
#pragma unroll
    for (int m = 0; m < KC; m += 4) {
      double af[4], bf[8];
#pragma unroll
      for (int a = 0; a < 4; ++a) af[a] = M.As[buf][32 * wr + 8 * a + g][m + q];
#pragma unroll
      for (int b = 0; b < 8; ++b) bf[b] = M.Bs[buf][64 * wc + 8 * b + g][m + q];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) Dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    }
    __syncthreads();  // the buffer is refilled two chunks later
  }
  // D fragment: row g, cols 2q, 2q+1
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int r = i0 + 32 * wr + 8 * a + g;
    if (r >= n_rows) continue;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int c = j0 + 64 * wc + 8 * b + 2 * q;
      double* dst = S + (size_t)r * ld + c;
      if (pair_ok && c + 1 < n && c + 1 <= r) {
        *reinterpret_cast<double2*>(dst) = make_double2(-acc[a][b][0], -acc[a][b][1]);
      } else {
        if (c < n && c <= r) dst[0] = -acc[a][b][0];
        if (c + 1 < n && c + 1 <= r) dst[1] = -acc[a][b][1];
      }
    }
  }
}

// backward substitution sweep for panel k: every CTA recomputes x_k = Linv_pp^T y_k (128 x 128, cheap) and then
// updates its slice of y_j -= L_kj^T x_k, j < k.
__global__ void __launch_bounds__(256) chol_back128_kernel(const double* __restrict__ S, int n, int ld, int k0, int nb,
                                                           const double* __restrict__ Linv, double* __restrict__ y,
                                                           double* __restrict__ x) {
  __shared__ double yk[PB], xk[PB];
  const int tid = threadIdx.x;
  if (tid < PB) yk[tid] = tid < nb ? y[k0 + tid] : 0.0;
  __syncthreads();
  if (tid < PB) {
    double s = 0.0;
    for (int m = tid; m < nb; ++m) s += Linv[m * PB + tid] * yk[m];  // (Linv^T)[tid][m]
    xk[tid] = s;
    if (blockIdx.x == 0 && tid < nb) x[k0 + tid] = s;
  }
  __syncthreads();
  const int c = blockIdx.x * 256 + tid;
  if (c < k0) {
    double s = 0.0;
    for (int r = 0; r < nb; ++r) s += S[(size_t)(k0 + r) * ld + c] * xk[r];
    y[c] -= s;
  }
}

__global__ void __launch_bounds__(256) chol_fetch_y_kernel(int n, int ld, const double* __restrict__ S, double* __restrict__ y) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j < n) y[j] = S[(size_t)n * ld + j];
}

int CholeskySolve(Problem* p) {
  const int n = p->n_cdim;
  if (n == 0) return SBA_OK;
  if (n <= kCholSmallMax) return CholeskySolveSmall(p);
  static bool attr_set[64] = {};
  const int dyn_diag = (int)sizeof(DiagSmem), dyn_trsm = (NB + PB) * LDP * (int)sizeof(double), dyn_trsm1 = (16 + PB) * LDP * (int)sizeof(double), dyn_syrk = (int)sizeof(SyrkSmem), dyn_next = 64 * LDP * (int)sizeof(double);
  if (!attr_set[p->device & 63]) {
    SBA_CUDA(cudaFuncSetAttribute(chol_diag128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_diag));
    SBA_CUDA(cudaFuncSetAttribute(chol_trsm128_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_trsm));
    SBA_CUDA(cudaFuncSetAttribute(chol_trsm128_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_trsm1));
    SBA_CUDA(cudaFuncSetAttribute(chol_syrk_next_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_next));
    SBA_CUDA(cudaFuncSetAttribute(chol_syrk128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_syrk));
    attr_set[p->device & 63] = true;
  }
  cudaStream_t st = p->stream;
  if (!p->stream2) {  // (first factorisation of the handle: plain launches, never inside a capture)
    // The chain kernels are a handful of CTAs that must not queue behind the hundreds of pending CTAs of a trailing
    // update (a syrk CTA owns a whole SM: 254 registers x 256 threads): the chain stream gets the highest priority, the
    // panel stream the next, the main stream keeps the default (lowest).
    int prio_lo = 0, prio_hi = 0;
    SBA_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    SBA_CUDA(cudaStreamCreateWithPriority(&p->stream2, cudaStreamNonBlocking, prio_hi));
    SBA_CUDA(cudaStreamCreateWithPriority(&p->stream3, cudaStreamNonBlocking, std::min(prio_lo, prio_hi + 1)));
    for (auto& e : p->ev_chol) SBA_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  double* S = p->d_S.p;
  const int ld = DenseLd(n);  // even: every row of S starts 16-byte aligned
  const int n_rows = n + 1;
  const int nblk = DivUp(n, PB);
  SBA_CUDA(p->d_Linv.Alloc((size_t)nblk * PB * PB));
  chol_load_rhs_kernel<<<DivUp(n, 256), 256, 0, st>>>(n, ld, p->d_rhs.p, S);
  p->launches++;
  static const bool want_trace = getenv("SBA_CHOL_TRACE") != nullptr;
  DevBuf<long long> trace;
  if (want_trace) { SBA_CUDA(trace.Alloc(32)); SBA_CUDA(trace.Zero(st)); }
  // Three streams.  What is serial in a right-looking factorisation is the chain
  //   diag(k) -> L of the NEXT diagonal tile's rows -> update of the next diagonal tile -> diag(k + 1),
  // so only those small pieces (one CTA, two CTAs, one CTA) ride on the chain stream sD; the rest of panel k's rows and
  // the first tile column of its update go to sT, the bulk of the update (tiles bi >= bj >= 1) to the main stream.  Per
  // panel the chain is diag + two small kernels instead of diag + whole panel solve + first tile column (round 2's first
  // version: 199 us per panel on average at n = 4000, profiles/r02_launches_c3_chol.csv; this schedule: 154 us, of which
  // the chain alone -- measured with the other kernels left out -- is ~140: the diagonal factor 82-95 us, the two chain
  // kernels and three dependent-launch gaps the rest).
  // Who writes what, per panel k with t0 = k0 + nb (tile coordinates relative to t0, 128 rows each):
  //   D2   [sD] rows [t0, t0 + 128) of column k            needs diag(k), first(k-1)
  //   D3   [sD] tile (0, 0)                                needs D2, rest(k-1)            -> diag(k + 1)
  //   trsm [sT] rows >= t0 + 128 of column k               needs diag(k), first(k-1)
  //   first[sT] tiles (bi >= 1, 0)                         needs trsm, D2, rest(k-1)
  //   rest [st] tiles (bi >= bj >= 1)                      needs trsm, rest(k-1)
  // The same dependencies are captured when the LM iteration is recorded as a CUDA graph.
  cudaStream_t sD = p->stream2, sT = p->stream3;
  cudaEvent_t eFork = p->ev_chol[0], eDiag = p->ev_chol[1], eD2 = p->ev_chol[2], eTrsm = p->ev_chol[3], eFirst = p->ev_chol[4],
              eRest = p->ev_chol[5], eJoinD = p->ev_chol[6], eJoinT = p->ev_chol[7];
  SBA_CUDA(cudaEventRecord(eFork, st));
  SBA_CUDA(cudaStreamWaitEvent(sD, eFork, 0));
  SBA_CUDA(cudaStreamWaitEvent(sT, eFork, 0));
  bool have_first = false, have_rest = false;
  for (int kb = 0; kb < nblk; ++kb) {
    const int k0 = kb * PB, nb = std::min(PB, n - k0);
    double* Linv = p->d_Linv.p + (size_t)kb * PB * PB;
    const int rem_rows = n_rows - (k0 + nb);      // >= 1: the rhs row
    const int d2_rows = std::min(rem_rows, TB);   // rows of the next diagonal tile: on the chain, in 16-row tiles
    const int row_tiles = DivUp(rem_rows - d2_rows, NB);  // 64-row tiles of the rest of the panel
    const bool trailing = k0 + nb < n;
    const int m_tiles = DivUp(rem_rows, TB);      // tiles cover rows [k0+nb, n_rows); columns [k0+nb, n)
    chol_diag128_kernel<<<1, 256, dyn_diag, sD>>>(S, n, ld, k0, Linv, p->d_scalars.p, kb == 1 ? trace.p : nullptr);
    SBA_CUDA(cudaEventRecord(eDiag, sD));
    if (have_first) SBA_CUDA(cudaStreamWaitEvent(sD, eFirst, 0));
    chol_trsm128_kernel<1><<<DivUp(d2_rows, 16), 256, dyn_trsm1, sD>>>(S, n, ld, n_rows, k0, nb, Linv, 0);
    SBA_CUDA(cudaEventRecord(eD2, sD));
    p->launches += 2;
    SBA_CUDA(cudaStreamWaitEvent(sT, eDiag, 0));
    if (row_tiles > 0) {
      chol_trsm128_kernel<4><<<row_tiles, 256, dyn_trsm, sT>>>(S, n, ld, n_rows, k0, nb, Linv, d2_rows);
      p->launches++;
    }
    SBA_CUDA(cudaEventRecord(eTrsm, sT));
    if (trailing) {
      if (have_rest) SBA_CUDA(cudaStreamWaitEvent(sD, eRest, 0));
      chol_syrk_next_diag_kernel<<<dim3(4, 4), 256, dyn_next, sD>>>(S, n, ld, n_rows, k0, nb, k0 + nb);
      p->launches++;
      if (m_tiles > 1) {
        SBA_CUDA(cudaStreamWaitEvent(sT, eD2, 0));
        if (have_rest) SBA_CUDA(cudaStreamWaitEvent(sT, eRest, 0));
        chol_syrk128_kernel<<<m_tiles - 1, 256, dyn_syrk, sT>>>(S, n, ld, n_rows, k0, nb, k0 + nb, 1, 1);
        SBA_CUDA(cudaEventRecord(eFirst, sT));
        have_first = true;
        SBA_CUDA(cudaStreamWaitEvent(st, eTrsm, 0));
        chol_syrk128_kernel<<<m_tiles * (m_tiles - 1) / 2, 256, dyn_syrk, st>>>(S, n, ld, n_rows, k0, nb, k0 + nb, 0, 0);
        SBA_CUDA(cudaEventRecord(eRest, st));
        have_rest = true;
        p->launches += 2;
      }
    }
  }
  SBA_CUDA(cudaEventRecord(eJoinD, sD));
  SBA_CUDA(cudaEventRecord(eJoinT, sT));
  SBA_CUDA(cudaStreamWaitEvent(st, eJoinD, 0));
  SBA_CUDA(cudaStreamWaitEvent(st, eJoinT, 0));
  // y = row n of the factor; back substitution into d_yc
  chol_fetch_y_kernel<<<DivUp(n, 256), 256, 0, st>>>(n, ld, S, p->d_tmp_c.p);
  p->launches++;
  for (int kb = nblk - 1; kb >= 0; --kb) {
    const int k0 = kb * PB, nb = std::min(PB, n - k0);
    chol_back128_kernel<<<std::max(1, DivUp(k0, 256)), 256, 0, st>>>(S, n, ld, k0, nb, p->d_Linv.p + (size_t)kb * PB * PB,
                                                                   p->d_tmp_c.p, p->d_yc.p);
    p->launches++;
  }
  SBA_CUDA(cudaGetLastError());
  if (want_trace) {
    long long h[32];
    SBA_CUDA(cudaMemcpyAsync(h, trace.p, sizeof(h), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    std::fprintf(stderr, "[chol trace] diag128 panel 1, us between phase stamps (load | per 32-block: chol+inv, rows, update | X rows | store):");
    for (int i = 1; i < 32 && h[i]; ++i) std::fprintf(stderr, " %.1f", (h[i] - h[i - 1]) * 1e-3);
    std::fprintf(stderr, "\n");
  }
  return SBA_OK;
}

}  // namespace sba
