// K-CHOL (small systems): single-launch dataflow Cholesky + both triangular solves for the
// reduced camera matrix of DENSE_SCHUR-sized problems (n <= kCholSmallMax).
//
// Replaces the dense LLT Ceres runs on the Schur complement inside ceres::Solve
// (/root/reference/src/optim/bundle_adjustment.cc:276-282, :306).
//
// At n ~ 300 (C2: 50 images x 6) a factorisation is 9 MFLOP: nothing but the dependency chain of
// n pivots matters.  A launch per panel step (cholesky.cu) pays a launch gap and a cold L2 read
// per step; here ONE kernel runs the whole solve:
//   * one CTA per 32-row block row i (+ one CTA for the right-hand side, carried as an extra row);
//   * CTA i walks its tiles (i,0..i) left to right; tile (i,j) waits -- by polling row j's progress
//     counter in global memory -- only for the tiles of row j it multiplies with, so the trailing
//     update of later rows overlaps the pivot chain of earlier ones (look-ahead for free);
//   * the 32x32 diagonal block is factored by ONE warp holding a row per lane in registers
//     (LDL^T elimination: the per-pivot critical path is shuffle -> reciprocal -> multiply -> fma,
//     the square roots are taken once per block, in parallel, afterwards);
//   * an off-diagonal solve X L_jj^T = T runs in one warp, a row per lane, as a right-looking
//     substitution (per step: multiply by 1/L_cc, then independent FMAs), so nothing but the factor
//     itself sits between two pivots blocks; the inverse of each diagonal factor is computed after
//     the factor has been published and is used only by the backward substitution;
//   * a row keeps its own finished tiles in shared memory (<= kOwnCap of them), the tiles of other
//     rows stream from L2 through a register prefetch that overlaps the previous product;
//   * the right-hand-side CTA finishes with the backward substitution.
// All CTAs (<= 33) are co-resident on the 148 SMs, which the flag protocol requires; every wait has
// a bounded spin budget and reports through DeviceScalars::chol_fail instead of hanging.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "solver.h"

namespace sba {

constexpr int SB = 32;            // tile edge
constexpr int SLD = SB + 1;       // padded shared-memory leading dimension
constexpr int kCholThreads = 256;
constexpr unsigned kAll = 0xffffffffu;

__device__ inline int LoadAcquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ inline void StoreRelease(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

struct CholSmall {
  double* S;            // (n+1) x n row-major, lower triangle valid; row n receives L^{-1} rhs
  const double* Dc;     // LM diagonal, added squared to the diagonal while loading (may be null)
  const double* rhs;    // n
  double* y;            // n   solution
  double* Linv;         // nbr x 32 x 32 inverses of the diagonal factors, then nbr x 32 reciprocal diagonals
  int* flags;           // [nbr+1] per-row progress counters (see WaitProgress)
  DeviceScalars* sc;
  int n, nbr, epoch;
  long long* trace;     // optional [nbr+1][8] globaltimer stamps (SBA_CHOL_TRACE=1), else null
  int warm_icache;      // rows >= 1 pre-run the diagonal factorisation while they wait (SBA_CHOL_NO_WARM=1 turns it off)
};

__device__ inline void Stamp(const CholSmall& a, int slot) {
  if (a.trace && threadIdx.x == 0) {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    a.trace[blockIdx.x * 8 + slot] = t;
  }
}


constexpr int kTileD = SB * SLD;   // doubles per shared-memory tile
constexpr int kOwnCap = 20;       // own-row tiles kept in shared memory (rows beyond re-read theirs from L2)

// Progress of block row i is ONE monotone counter, epoch * 256 + (tiles finished): tiles (i,0..i-1) in order,
// then the diagonal factor (count i+1), then its inverse (count i+2).  A consumer re-polls only when the
// count it last saw does not cover what it needs, so a row that is far ahead costs one poll per tile, not
// one per product.  Every thread polls for itself: its own acquire orders its own subsequent loads, so no
// barrier and no single-thread relay sits between "published" and "loaded".  Bounded: a lost producer
// becomes chol_fail, never a hang.
__device__ inline int WaitProgress(const CholSmall& a, int row, int need, int have) {
  if (have >= need) return have;
  const int* f = a.flags + row;
  const int base = a.epoch << 8;
  int spins = 0, v;
  while ((v = LoadAcquire(f) - base) < need) {
    if (++spins > (1 << 22)) { a.sc->chol_fail = 1; a.sc->solver_failed = 1; return need; }
  }
  return v;
}
__device__ inline void Publish(const CholSmall& a, int row, int count) {
  __syncthreads();  // every thread's stores precede thread 0's release (cumulativity through the barrier)
  if (threadIdx.x == 0) StoreRelease(a.flags + row, (a.epoch << 8) + count);
}

// 1/d for a positive normal d: hardware seed + two Newton steps (2^-20 -> 2^-40 -> 2^-80 relative).
// The IEEE division sequence costs ~250 cycles of dependent issue on this part; this is ~60, and it
// sits on the per-pivot critical path of the factorisation.
__device__ inline double FastRcp(double d) {
  double x;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(d));
  double e = fma(-d, x, 1.0);
  x = fma(x, e, x);
  e = fma(-d, x, 1.0);
  x = fma(x, e, x);
  return x;
}

// fp64 tensor-core MMA (DMMA): D(8x8) += A(8x4, row) * B(4x8, col)
__device__ inline void Dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// Which four outputs of a 32x32 tile a thread owns (the DMMA accumulator layout): warp w works on row
// block w & 3 and on the column half w >> 2; with g = lane >> 2, q = lane & 3 the thread holds
//   acc[2 b + e]  <->  out[8 (w & 3) + g][16 (w >> 2) + 8 b + 2 q + e],   b, e in {0, 1}.
struct OutMap {
  int row, col0;  // col(b, e) = col0 + 8 b + e
  int g, q, rb, ch;
  __device__ OutMap() {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    g = lane >> 2; q = lane & 3; rb = w & 3; ch = w >> 2;
    row = 8 * rb + g; col0 = 16 * ch + 2 * q;
  }
  __device__ int col(int x) const { return col0 + 8 * (x >> 1) + (x & 1); }
};

// acc -= A * B^T for 32x32 tiles in shared memory (ld SLD): out[r][c] -= sum_m A[r][m] B[c][m].
// 16 DMMAs and 24 shared loads per warp, against 128 FMAs and 160 loads per THREAD for the scalar form.
__device__ inline void TileMulSubNT(const double* __restrict__ A, const double* __restrict__ B, const OutMap& o,
                                    double (&acc)[4]) {
  const double* ap = A + (8 * o.rb + o.g) * SLD + o.q;
  const double* b0 = B + (16 * o.ch + o.g) * SLD + o.q;
  const double* b1 = b0 + 8 * SLD;
#pragma unroll
  for (int m = 0; m < SB; m += 4) {
    const double na = -ap[m];
    Dmma884(acc[0], acc[1], na, b0[m]);
    Dmma884(acc[2], acc[3], na, b1[m]);
  }
}

// this thread's four elements (idx = tid + 256 q) of the 32x32 tile at (row0, col0) of S, read from L2
// (ld.global.cg: a tile written by another CTA must never be served from a stale L1 line)
__device__ inline void FetchTile(const double* __restrict__ S, int n, int row0, int col0, int row_lim, bool lower_only,
                                 double (&v)[4]) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int idx = threadIdx.x + kCholThreads * q;
    const int rr = idx >> 5, cc = idx & 31;
    const int gr = row0 + rr, gc = col0 + cc;
    v[q] = (gr < row_lim && gc < n && (!lower_only || cc <= rr)) ? __ldcg(S + (size_t)gr * n + gc) : 0.0;
  }
}
__device__ inline void StashTile(const double (&v)[4], double* __restrict__ dst) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int idx = threadIdx.x + kCholThreads * q;
    dst[(idx >> 5) * SLD + (idx & 31)] = v[q];
  }
}

// One warp, a row per lane: t <- t L^{-T}, with the factor given in UNIT-lower form
// Ms[r][c] = L[r][c] / L[c][c] (shared, ld SLD) and rd[c] = 1 / L[c][c].  With y_c = t_c - sum_{m<c} x_m L[c][m]
// and x_c = rd_c y_c, the right-looking update is y_c' -= y_c M[c'][c]: ONE dependent FMA per column (an fp64
// FMA has ~40 cycles of latency here, and this chain sits between two pivot blocks); the scaling by rd is
// done once at the end.  M[c'][c] is a warp-wide broadcast read.
__device__ inline void WarpTrsm32(double (&t)[SB], const double* __restrict__ Ms, const double* __restrict__ rd) {
#pragma unroll
  for (int c = 0; c < SB; ++c) {
    const double y = t[c];
#pragma unroll
    for (int c2 = c + 1; c2 < SB; ++c2) t[c2] = fma(-y, Ms[c2 * SLD + c], t[c2]);
  }
#pragma unroll
  for (int c = 0; c < SB; ++c) t[c] *= rd[c];
}

// Inverse of the lower-triangular 32x32 factor in Ls (shared, ld SLD) into Xs, block-recursively:
//   [A 0; B C]^{-1} = [A^{-1} 0; -C^{-1} B A^{-1}  C^{-1}]   with 16x16 blocks.
// rd[i] = 1 / L[i][i].  Called by all kCholThreads threads.  Off the critical path (backward sweep only).
__device__ inline void TriInv32Blocked(const double* __restrict__ Ls, const double* __restrict__ rd,
                                       double* __restrict__ Xs, double* __restrict__ Ts) {
  const int tid = threadIdx.x;
  if (tid < 32) {  // lanes 0..15: columns of A^{-1}; lanes 16..31: columns of C^{-1}
    const int h = tid >> 4, c = tid & 15;
    const double* Lb = Ls + (16 * h) * SLD + 16 * h;
    double x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      double s = (i == c) ? 1.0 : 0.0;
#pragma unroll
      for (int k = 0; k < i; ++k) s -= Lb[i * SLD + k] * x[k];  // x[k] == 0 for k < c
      x[i] = (i >= c) ? s * rd[16 * h + i] : 0.0;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      Xs[(16 * h + i) * SLD + 16 * h + c] = x[i];
      if (h == 1) Xs[i * SLD + 16 + c] = 0.0;  // upper-right block
    }
  }
  __syncthreads();
  const int r = tid >> 4, c = tid & 15;
  {  // T = B A^{-1}
    double s = 0.0;
#pragma unroll
    for (int m = 0; m < 16; ++m) s += Ls[(16 + r) * SLD + m] * Xs[m * SLD + c];
    Ts[r * 17 + c] = s;
  }
  __syncthreads();
  {  // X21 = -C^{-1} T
    double s = 0.0;
#pragma unroll
    for (int m = 0; m < 16; ++m) s += Xs[(16 + r) * SLD + 16 + m] * Ts[m * 17 + c];
    Xs[(16 + r) * SLD + c] = -s;
  }
  __syncthreads();
}

// Warp-level tile drivers, kept out of line so that their 32-double register rows do not add to the
// register pressure of the tile loop.
__device__ __noinline__ void SolveTileWarp(const double* __restrict__ sC, const double* __restrict__ sB,
                                           const double* __restrict__ sRd, double* __restrict__ sA,
                                           double* __restrict__ own) {
  const int lane = threadIdx.x & 31;
  double t[SB];
#pragma unroll
  for (int c = 0; c < SB; ++c) t[c] = sC[lane * SLD + c];
  WarpTrsm32(t, sB, sRd);
#pragma unroll
  for (int c = 0; c < SB; ++c) {
    sA[lane * SLD + c] = t[c];
    if (own) own[lane * SLD + c] = t[c];
  }
}
__device__ __noinline__ bool FactorTileWarp(double* __restrict__ sC, double* __restrict__ sRd, double* __restrict__ gRd) {
  // One warp, lane = row, the row held in registers (fully unrolled; a shared-memory version with rolled
  // loops measured 3x slower: every store -> load round trip lands on the pivot chain).
  // LDL^T elimination with the scaling deferred: a[r][c] -= (a[r][j] a[c][j]) / d_j -- the product is formed
  // while the reciprocal is still in flight, so the chain per pivot is shuffle d -> reciprocal -> one FMA.
  // Columns stay unscaled until the end: L[r][c] = a[r][c] / sqrt(d_c).
  const int lane = threadIdx.x & 31;
  bool ok = true;
  double a[SB];
#pragma unroll
  for (int c = 0; c < SB; ++c) a[c] = (c <= lane) ? sC[lane * SLD + c] : 0.0;
#pragma unroll
  for (int j = 0; j < SB; ++j) {
    double d = __shfl_sync(kAll, a[j], j);
    if (!(d > 0.0) || !isfinite(d)) { ok = false; d = 1.0; }
    const double inv = FastRcp(d);
    const double col = a[j];  // unscaled A[lane][j]; zero for lane < j
#pragma unroll
    for (int c = j + 1; c < SB; ++c) {
      const double ac = __shfl_sync(kAll, col, c);  // unscaled A[c][j]
      a[c] = fma(-(col * ac), inv, a[c]);
    }
    if (lane == j) a[j] = d;
  }
  double dd = 1.0;  // own pivot, picked with selects (a dynamic index would spill the row to local memory)
#pragma unroll
  for (int c = 0; c < SB; ++c) if (c == lane) dd = a[c];
  const double sq = sqrt(dd), rs = 1.0 / sq;
  sRd[lane] = rs;
  gRd[lane] = rs;
#pragma unroll
  for (int c = 0; c < SB; ++c) {
    const double rc = __shfl_sync(kAll, rs, c);
    sC[lane * SLD + c] = (c < lane) ? a[c] * rc : (c == lane ? sq : 0.0);
  }
  return __all_sync(kAll, ok);
}

__global__ void __launch_bounds__(kCholThreads, 1) chol_small_kernel(CholSmall a, int own_cap) {
  extern __shared__ double smem[];
  double* sA = smem;                 // X staging (solve result of the current tile)
  double* sB = sA + kTileD;           // remote tile L(j,k) / diagonal factor L_jj
  double* sC = sB + kTileD;           // T tile / diagonal tile
  double* sT = sC + kTileD;           // 16 x 17 scratch of the inverse
  double* sRd = sT + 16 * 17;        // 32 reciprocal diagonals
  double* sX = sRd + SB;             // 32
  double* sV = sX + SB;              // 1024 + 32 (backward sweep)
  double* sOwn = sV + 1024 + SB;     // own_cap tiles L(i,k)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const OutMap om;
  const int n = a.n, nbr = a.nbr, i = blockIdx.x;
  const bool is_rhs = i == nbr;
  const int row0 = is_rhs ? n : SB * i;
  const int row_lim = is_rhs ? n + 1 : n;
  double* S = a.S;
  double* Rd = a.Linv + (size_t)nbr * SB * SB;  // reciprocal diagonals of every factor
  Stamp(a, 0);
  if (is_rhs) {  // the right-hand side rides along as row n of the matrix
    for (int c = tid; c < n; c += kCholThreads) S[(size_t)n * n + c] = a.rhs[c];
    __syncthreads();
  }
  // Rows >= 1 cannot start before row 0's factor is published (~28 us): one otherwise idle warp spends that wait running
  // the (fully unrolled, ~100 KB of SASS) diagonal factorisation on an identity tile, so the instruction cache of this
  // SM is warm when the real factor -- which sits on the critical path of every later row -- arrives.
  if (!is_rhs && i > 0 && warp == (kCholThreads >> 5) - 1 && a.warm_icache) {
    for (int idx = lane; idx < SB * SLD; idx += 32) sV[idx] = (idx / SLD == idx % SLD) ? 1.0 : 0.0;
    __syncwarp();
    FactorTileWarp(sV, sX, a.Linv + (size_t)i * SB * SB);  // scratch outputs: overwritten by this row's inverse later
  }
  double accD[4] = {0.0, 0.0, 0.0, 0.0};
  if (!is_rhs) {  // diagonal tile, lower part, + D^2; identity padding keeps a partial block well posed
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int gr = row0 + om.row, gc = row0 + om.col(e);
      double v = 0.0;
      if (gr < n && gc < n) {
        if (om.col(e) <= om.row) v = __ldcg(S + (size_t)gr * n + gc);
        if (om.col(e) == om.row && a.Dc) v += a.Dc[gr] * a.Dc[gr];
      } else if (om.col(e) == om.row) {
        v = 1.0;
      }
      accD[e] = v;
    }
  }
  const int jmax = is_rhs ? nbr : i;
  for (int j = 0; j < jmax; ++j) {
    double acc[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int gr = row0 + om.row, gc = SB * j + om.col(e);
      acc[e] = (gr < row_lim && gc < n) ? __ldcg(S + (size_t)gr * n + gc) : 0.0;
    }
    // T = A(i,j) - sum_{k<j} L(i,k) L(j,k)^T, remote tiles prefetched one step ahead
    double pre[4], pre2[4];
    int have = 0;  // tiles of row j known to be published
    if (j > 0) { have = WaitProgress(a, j, 1, have); FetchTile(S, n, SB * j, 0, n, false, pre); }
    if (j > 1) { have = WaitProgress(a, j, 2, have); FetchTile(S, n, SB * j, SB, n, false, pre2); }
    for (int k = 0; k < j; ++k) {
      __syncthreads();  // the previous product has finished reading sB (and sA)
      StashTile(pre, sB);
      const double* own = sOwn + (size_t)k * kTileD;
      if (k >= own_cap) {  // row too long for the shared-memory cache: re-read the own tile from L2
        double o[4];
        FetchTile(S, n, row0, SB * k, row_lim, false, o);
        StashTile(o, sA);
        own = sA;
      }
      __syncthreads();
#pragma unroll
      for (int q = 0; q < 4; ++q) pre[q] = pre2[q];
      if (k + 2 < j) { have = WaitProgress(a, j, k + 3, have); FetchTile(S, n, SB * j, SB * (k + 2), n, false, pre2); }
      TileMulSubNT(own, sB, om, acc);
    }
    // solve against the diagonal factor j
    have = WaitProgress(a, j, j + 1, have);
    if (j == jmax - 1) Stamp(a, 5);
    FetchTile(S, n, SB * j, SB * j, n, true, pre);
    const double rdv = (tid < SB) ? __ldcg(Rd + j * SB + tid) : 0.0;
    __syncthreads();  // last product done: sB / sC free
#pragma unroll
    for (int e = 0; e < 4; ++e) sC[om.row * SLD + om.col(e)] = acc[e];
    if (tid < SB) sRd[tid] = rdv;
    {  // unit-lower form of the factor for the substitution: M[r][c] = L[r][c] * rd[c]  (c = idx & 31 = lane)
      const double rc = __ldcg(Rd + j * SB + lane);
#pragma unroll
      for (int q = 0; q < 4; ++q) pre[q] *= rc;
    }
    StashTile(pre, sB);
    __syncthreads();
    if (warp == 0) SolveTileWarp(sC, sB, sRd, sA, j < own_cap ? sOwn + (size_t)j * kTileD : nullptr);
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int gr = row0 + om.row, gc = SB * j + om.col(e);
      if (gr < row_lim && gc < n) S[(size_t)gr * n + gc] = sA[om.row * SLD + om.col(e)];
    }
    if (!is_rhs) TileMulSubNT(sA, sA, om, accD);  // running update of the diagonal tile
    Publish(a, i, j + 1);
  }
  Stamp(a, 1);  // all off-diagonal tiles of this row done
  if (!is_rhs) {
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 4; ++e) sC[om.row * SLD + om.col(e)] = accD[e];
    __syncthreads();
    if (warp == 0) {
      const bool ok = FactorTileWarp(sC, sRd, Rd + i * SB);
      if (!ok && lane == 0) { a.sc->chol_fail = 1; a.sc->solver_failed = 1; }
    }
    __syncthreads();
    Stamp(a, 2);  // factor done
    for (int idx = tid; idx < SB * SB; idx += kCholThreads) {
      const int rr = idx >> 5, cc = idx & 31;
      const int gr = row0 + rr, gc = row0 + cc;
      if (gr < n && gc < n && cc <= rr) S[(size_t)gr * n + gc] = sC[rr * SLD + cc];
    }
    Publish(a, i, i + 1);
    Stamp(a, 4);
    // inverse of the factor: used by the backward sweep only
    TriInv32Blocked(sC, sRd, sB, sT);
    for (int idx = tid; idx < SB * SB; idx += kCholThreads)
      a.Linv[(size_t)i * SB * SB + idx] = sB[(idx >> 5) * SLD + (idx & 31)];
    Publish(a, i, i + 2);  // "inverse of factor i stored"
    Stamp(a, 3);
    return;
  }
  // ---- right-hand-side CTA: z = row n (still in sA row 0 per tile, re-read from L2); L^T y = z ----
  __syncthreads();
  for (int c = tid; c < n; c += kCholThreads) sV[c] = __ldcg(S + (size_t)n * n + c);
  __syncthreads();
  // Every fp64 FMA costs ~40 cycles of latency, so no sum below is longer than 4-8 dependent terms:
  // the 32-term dot products are split 8 ways across lanes and folded with shuffles.
  double li[4];
  WaitProgress(a, nbr - 1, nbr + 1, 0);
#pragma unroll
  for (int q = 0; q < 4; ++q) li[q] = __ldcg(a.Linv + (size_t)(nbr - 1) * SB * SB + tid + kCholThreads * q);
  for (int kb = nbr - 1; kb >= 0; --kb) {
    const int k0 = SB * kb, nb = min(SB, n - k0);
    StashTile(li, sB);
    __syncthreads();
    if (kb > 0) {  // next inverse block: in flight while this step computes
      WaitProgress(a, kb - 1, kb + 1, 0);
#pragma unroll
      for (int q = 0; q < 4; ++q) li[q] = __ldcg(a.Linv + (size_t)(kb - 1) * SB * SB + tid + kCholThreads * q);
    }
    {  // x_k[c] = sum_m Linv[m][c] v[k0+m]: thread (c = tid / 8, part = tid % 8) takes m = part, part+8, ...
      const int c = tid >> 3, part = tid & 7;
      double s = 0.0;
#pragma unroll
      for (int m = part; m < SB; m += 8) if (m < nb) s += sB[m * SLD + c] * sV[k0 + m];
      s += __shfl_xor_sync(kAll, s, 1);
      s += __shfl_xor_sync(kAll, s, 2);
      s += __shfl_xor_sync(kAll, s, 4);
      if (part == 0) { sX[c] = s; if (c < nb) a.y[k0 + c] = s; }
    }
    __syncthreads();
    for (int c = tid; c < k0; c += kCholThreads) {
      double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
      const double* col = S + (size_t)k0 * n + c;
#pragma unroll
      for (int rr = 0; rr < SB; rr += 4) {
        if (rr < nb) s0 += __ldcg(col + (size_t)rr * n) * sX[rr];
        if (rr + 1 < nb) s1 += __ldcg(col + (size_t)(rr + 1) * n) * sX[rr + 1];
        if (rr + 2 < nb) s2 += __ldcg(col + (size_t)(rr + 2) * n) * sX[rr + 2];
        if (rr + 3 < nb) s3 += __ldcg(col + (size_t)(rr + 3) * n) * sX[rr + 3];
      }
      sV[c] -= (s0 + s1) + (s2 + s3);
    }
    __syncthreads();
  }
  Stamp(a, 6);
}

int CholeskySolveSmall(Problem* p) {
  const int n = p->n_cdim;
  const int nbr = DivUp(n, SB);
  cudaStream_t st = p->stream;
  if (p->chol_flags.n < (size_t)(nbr + 1)) SBA_CUDA(p->chol_flags.Alloc((size_t)(kCholSmallMax / SB + 2)));
  // progress counters start from zero on every solve (a memset node: the launch sequence of an LM iteration has no
  // per-call kernel argument, so that it can be replayed as a CUDA graph -- lm.cu)
  SBA_CUDA(p->chol_flags.Zero(st));
  SBA_CUDA(p->d_Linv.Alloc((size_t)std::max(nbr, 1) * 64 * 64));
  CholSmall a;
  a.S = p->d_S.p; a.Dc = p->chol_add_diag ? p->d_Dc.p : nullptr; a.rhs = p->d_rhs.p; a.y = p->d_yc.p;
  a.Linv = p->d_Linv.p; a.flags = p->chol_flags.p; a.sc = p->d_scalars.p;
  a.n = n; a.nbr = nbr; a.epoch = 1;
  const int own_cap = std::min(nbr, kOwnCap);
  const size_t smem = ((size_t)(3 + own_cap) * kTileD + 16 * 17 + 2 * SB + 1024 + SB) * sizeof(double);
  static size_t smem_set = 0;
  if (smem > smem_set) {
    SBA_CUDA(cudaFuncSetAttribute(chol_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set = smem;
  }
  static const bool trace = getenv("SBA_CHOL_TRACE") != nullptr;
  static const bool no_warm = getenv("SBA_CHOL_NO_WARM") != nullptr;
  a.warm_icache = no_warm ? 0 : 1;
  DevBuf<long long> d_trace;
  a.trace = nullptr;
  if (trace) { SBA_CUDA(d_trace.Alloc((size_t)(nbr + 1) * 8)); SBA_CUDA(d_trace.Zero(st)); a.trace = d_trace.p; }
  chol_small_kernel<<<nbr + 1, kCholThreads, smem, st>>>(a, own_cap);
  p->launches++;
  if (trace) {
    std::vector<long long> h((size_t)(nbr + 1) * 8);
    SBA_CUDA(cudaMemcpyAsync(h.data(), d_trace.p, h.size() * sizeof(long long), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    long long t0 = h[0];
    for (int i = 0; i <= nbr; ++i) t0 = std::min(t0, h[8 * i]);
    for (int i = 0; i <= nbr; ++i) {
      fprintf(stderr, "chol row %2d:", i);
      for (int s = 0; s < 7; ++s) fprintf(stderr, " %8.2f", h[8 * i + s] ? (h[8 * i + s] - t0) * 1e-3 : -1.0);
      fprintf(stderr, "  us [start, offdiag, factor, inverse, publish, lastwait, backsub]\n");
    }
  }
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

}  // namespace sba
