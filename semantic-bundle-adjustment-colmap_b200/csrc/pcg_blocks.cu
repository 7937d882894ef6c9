// K-PCG (block-sparse S): the whole preconditioned conjugate-gradient solve as ONE persistent cooperative kernel.
//
// Same iteration, failure rules and quadratic-model stopping test as pcg.cu (Ceres 2.1.0 ConjugateGradientsSolver,
// SURVEY.md App. B.5; the solver /root/reference/src/optim/bundle_adjustment.cc:283-286 selects), but with the
// explicit block-sparse S of bsr.cu a product costs microseconds, so three launches and a host poll per iteration
// would be most of the time.  Here a CTA owns block rows (images) i = blockIdx.x, blockIdx.x + gridDim.x, ...:
//   * every vector operation of a row (x, r, z, p, q and the SCHUR_JACOBI blocks, which never cross an image) is
//     local to its CTA; only the product reads other rows' p (or x);
//   * the three scalars of an iteration (p.q, r.z, x.(b + r)) are per-CTA partials summed by EVERY CTA in the same
//     fixed order after a grid barrier, so all CTAs take the same branch without a broadcast;
//   * two grid barriers per iteration (p.q | r.z and the stopping test), none of them on the host: the product forms
//     the neighbours' direction p_j = z_j + beta p_j(previous) on the fly from two slot-space vectors every CTA
//     published before the last barrier, instead of waiting for a third barrier after the direction update.
// Deterministic: fixed ownership, fixed summation order, no atomics.
//
// N > 1 GPUs (DIST): the block rows are split into contiguous ranges, one per rank, and the SAME kernel runs on every
// rank over its range -- one solve, N cooperative grids.  The slot-space vectors (p, z, x) and the solution live in an
// area that every rank maps through CUDA IPC (comm.cu CommPcgArea): a CTA stores the entries of its rows into every
// rank's copy over NVLink, so the product reads neighbours' rows from local L2.  A scalar of the iteration is reduced in
// two levels: grid barrier + the fixed-order sum of this rank's partials (as on one GPU), then CTA 0 stores the rank's
// total and a release flag into every peer, every CTA spins on its LOCAL flags and adds the N totals in rank order --
// all ranks take the same branches with bit-identical scalars.  The replicated solve of round 1 cost the same at any N
// (C5: 63 us per CG iteration, 7.8 of the 23 ms of an 8-GPU solve).
#include <cooperative_groups.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "reduce.cuh"
#include "solver.h"

namespace cg = cooperative_groups;

namespace sba {

#ifndef SBA_PCG_UNROLL
#define SBA_PCG_UNROLL 4
#endif
constexpr int kPcgUnroll = SBA_PCG_UNROLL;  // blocks of a row a warp has in flight in the product

struct PcgBlocks {
  const int* row_ptr; const int2* entries; const double* B; const int* slot_col;
  const double* Minv_s;  // [n_img][DC x DC] SCHUR_JACOBI inverse blocks laid out in slot space (zero outside a block)
  const double* Dc; const double* b;
  double* y;             // compact solution (n_cdim)
  double* p_s;           // slot-space vectors read by other CTAs (layout at the kernel): p [2][n_img * DC] (ping-pong) | z | x
                         // (x only on residual-reset iterations); DIST: this rank's copy inside the peer-mapped area
  double* partials;      // [3][gridDim.x]
  DeviceScalars* sc;
  int n_img, rows_per_cta, min_it, max_it;
  double q_tol;
  long long* trace;      // optional [64][8] globaltimer stamps of CTA 0 (SBA_PCG_TRACE=1), else null
  // DIST only
  int i_lo, i_hi;        // this rank's block rows
  int nranks, rank, n_cdim;
  double* peer[8];       // every rank's area: p | z | x | y | total words   (peer[rank] is the local one)
  unsigned* peer_flags[8];  // the first 128 B of every area; [8] of the local one holds the sync counter
  unsigned* epoch;       // local: syncs completed by this rank (persists across launches)
  double* y_out;         // compact solution of this rank (copied from the area after the last sync)
};

__device__ inline void PcgStamp(const PcgBlocks& a, int it, int slot) {
  if (a.trace && blockIdx.x == 0 && threadIdx.x == 0 && it < 64) {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    a.trace[it * 8 + slot] = t;
  }
}

__device__ inline double BlockAllSum128(double v) {
  __shared__ double sm[4];
  v = WarpSum(v);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  const double t = (sm[0] + sm[1]) + (sm[2] + sm[3]);
  __syncthreads();
  return t;
}

// identical in every CTA: strided per-thread sums, then the same tree
__device__ inline double GridTotal(const double* partials, int G) {
  double s = 0.0;
  for (int k = threadIdx.x; k < G; k += 128) s += partials[k];
  return BlockAllSum128(s);
}

__device__ inline bool BadScalar(double x) { return x == 0.0 || isinf(x) || isnan(x); }

// (S u) restricted to the CTA's rows i = i_first + lr * G < i_hi, u = v + beta vold in slot space (vold may be null);
// the results land in qs[lr * DC + slot].
// A warp takes a contiguous quarter of every row's entries: one coalesced load brings 32 of them, the loop below then
// has no load that depends on another load, so several blocks are in flight per warp.  All rows of the CTA are
// accumulated before the ONE barrier and reduction at the end (a barrier pair per row serialised the rows: C5, five rows
// per CTA, spent 45 of its 63 us per CG iteration here).
// v is written by other CTAs during the kernel: plain loads only.
template <int DC, bool DIST>
__device__ inline void RowsProduct(const PcgBlocks& a, int i_first, int i_hi, int R, const double* v, const double* vold,
                                   double beta, double* red /* [R][4][2][DC * DC] */, double* qs) {
  constexpr int NE = DC * DC, N = (NE + 31) / 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, G = gridDim.x;
  int er[N], ec[N];
#pragma unroll
  for (int q = 0; q < N; ++q) {
    const int e = min(lane + 32 * q, NE - 1);
    er[q] = e / DC; ec[q] = e - er[q] * DC;
  }
  int rows = 0;
  for (int lr = 0; lr < R; ++lr) {
    const int i = i_first + lr * G;
    if (i >= i_hi) break;
    rows = lr + 1;
    double as[N], at[N];
#pragma unroll
    for (int q = 0; q < N; ++q) { as[q] = 0.0; at[q] = 0.0; }
    const int e0 = a.row_ptr[i], cnt = a.row_ptr[i + 1] - e0;
    const int per = (cnt + 3) >> 2;
    const int w0 = e0 + warp * per, w1 = min(e0 + cnt, w0 + per);
    for (int base = w0; base < w1; base += 32) {
      const int m = min(32, w1 - base);
      const int2 mine = lane < m ? a.entries[base + lane] : make_int2(0, 0);
#pragma unroll kPcgUnroll
      for (int k = 0; k < m; ++k) {
        const int blk_id = __shfl_sync(0xffffffffu, mine.x, k), jy = __shfl_sync(0xffffffffu, mine.y, k);
        const int j = jy & 0x7fffffff;
        double xv = 0.0;
        if (lane < DC) {
          // (DIST: rows of other ranks arrive over NVLink in L2 -- L2 loads, never a stale L1 line)
          xv = DIST ? __ldcg(v + (size_t)j * DC + lane) : v[(size_t)j * DC + lane];
          if (vold) xv += beta * (DIST ? __ldcg(vold + (size_t)j * DC + lane) : vold[(size_t)j * DC + lane]);  // the direction of row j, as its owner forms it
        }
        const double* blk = a.B + (size_t)blk_id * NE;
        double w[N];
#pragma unroll
        for (int q = 0; q < N; ++q) w[q] = (lane + 32 * q < NE) ? blk[lane + 32 * q] : 0.0;
        if (jy < 0) {
#pragma unroll
          for (int q = 0; q < N; ++q) at[q] += w[q] * __shfl_sync(0xffffffffu, xv, er[q]);
        } else {
#pragma unroll
          for (int q = 0; q < N; ++q) as[q] += w[q] * __shfl_sync(0xffffffffu, xv, ec[q]);
        }
      }
    }
    double* rd = red + ((size_t)lr * 4 + warp) * 2 * NE;
#pragma unroll
    for (int q = 0; q < N; ++q) {
      const int e = lane + 32 * q;
      if (e < NE) { rd[e] = as[q]; rd[NE + e] = at[q]; }
    }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < rows * DC; k += 128) {
    const int lr = k / DC, r = k - lr * DC;
    double out = 0.0;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const double* rd = red + ((size_t)lr * 4 + w) * 2 * NE;
#pragma unroll
      for (int c = 0; c < DC; ++c) out += rd[r * DC + c];
#pragma unroll
      for (int c = 0; c < DC; ++c) out += rd[NE + c * DC + r];
    }
    qs[k] = out;
  }
  __syncthreads();  // qs complete; red is reused by the next product
}

// Grid-wide (DIST: all ranks') totals of up to two per-CTA partial arrays, identical in every thread of every CTA of
// every rank.  e: this rank's sync counter (DIST).
// DIST, after the local barrier and the local totals: CTA 0 sends this rank's totals to every rank as words that carry
// their own flag -- {epoch : high half of the double} and {epoch : low half}, two 8-byte stores, each delivered whole
// (the NCCL "LL" idea) -- so a total needs no release store behind it, i.e. no second NVLink round trip.  Every CTA
// polls the words in its LOCAL area.  `fence`: the CTAs have pushed vector entries to the peers that the phase after this sync
// reads; every thread then makes its own stores visible system-wide before the barrier, the sender's fence is
// cumulative over them, and the poller's fence after the last word closes the acquire side.
// Consecutive syncs never use the same slot, so a rank that runs ahead cannot overwrite a word a slower rank still needs.
constexpr int kPcgSlotNone = 3;  // syncs that carry no scalar
template <bool DIST>
__device__ inline void SyncTotals(cg::grid_group& grid, const PcgBlocks& a, unsigned& e, size_t tot_off, bool fence, int nt,
                                  const double* P0, int slot0, const double* P1, int slot1, double* out, double (*s_tot)[8]) {
  const int G = gridDim.x, t = threadIdx.x;
  if (DIST && fence) {  // one fence per CTA, cumulative over the CTA's stores through the barrier (a fence in every
    __syncthreads();    // thread measured ~15 us: 10^5 system-scope fences queue up behind each other)
    if (t == 0) __threadfence_system();
  }
  grid.sync();
  double loc[2] = {0.0, 0.0};
  if (nt >= 1) loc[0] = GridTotal(P0, G);
  if (nt == 2) loc[1] = GridTotal(P1, G);
  if (!DIST) { out[0] = loc[0]; out[1] = loc[1]; return; }
  e += 1;
  const int ns = nt == 0 ? 1 : nt;
  const int slots[2] = {nt == 0 ? kPcgSlotNone : slot0, slot1};
  if (t < a.nranks) {
    if (blockIdx.x == 0) {
      __threadfence_system();
      unsigned long long* dst = reinterpret_cast<unsigned long long*>(a.peer[t] + tot_off);
      for (int j = 0; j < ns; ++j) {
        const unsigned long long bits = (unsigned long long)__double_as_longlong(loc[j]);
        unsigned long long* w = dst + (size_t)(slots[j] * 8 + a.rank) * 2;
        const unsigned long long w0 = ((unsigned long long)e << 32) | (bits >> 32), w1 = ((unsigned long long)e << 32) | (bits & 0xffffffffull);
        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(w), "l"(w0) : "memory");
        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(w + 1), "l"(w1) : "memory");
      }
    }
    const unsigned long long* src = reinterpret_cast<const unsigned long long*>(a.peer[a.rank] + tot_off);
    unsigned long long t0 = 0;
    for (int j = 0; j < ns; ++j) {
      const unsigned long long* w = src + (size_t)(slots[j] * 8 + t) * 2;
      unsigned long long w0, w1;
      for (unsigned spins = 0;; ++spins) {
        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w0) : "l"(w) : "memory");
        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w1) : "l"(w + 1) : "memory");
        if ((unsigned)(w0 >> 32) == e && (unsigned)(w1 >> 32) == e) break;
        if ((spins & 0xfff) != 0xfff) continue;
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (t0 == 0) t0 = now;
        else if (now - t0 > 30000000000ull) __trap();  // a peer left the sequence: fail loudly instead of spinning for ever
      }
      s_tot[j][t] = __longlong_as_double((long long)(((w0 & 0xffffffffull) << 32) | (w1 & 0xffffffffull)));
    }
    __threadfence_system();
  }
  __syncthreads();
  out[0] = 0.0; out[1] = 0.0;
  for (int r = 0; r < a.nranks; ++r) {
    if (nt >= 1) out[0] += s_tot[0][r];
    if (nt == 2) out[1] += s_tot[1][r];
  }
  __syncthreads();  // s_tot is reused by the next sync
}

// The CTA's own rows live in shared memory for the whole solve: x, r, z, q, b, Dc^2 and the compact column of each
// (row, slot) item k = local_row * DC + slot.  Items of constant slots (col < 0) and of rows past the range stay zero.
// Shared slot-space vectors, offsets in doubles from a.p_s (DIST: from every a.peer[r]): p [2][NV] | z [NV] | x [NV] |
// y [n_cdim] | total words [4 slots][8 ranks][2], NV = n_img * DC.
template <int DC, bool DIST>
__global__ void __launch_bounds__(128) pcg_blocks_kernel(PcgBlocks a) {
  cg::grid_group grid = cg::this_grid();
  __shared__ double s_tot_mem[2][8];
  extern __shared__ double dyn[];
  const int G = gridDim.x, t = threadIdx.x, R = a.rows_per_cta, M = R * DC;
  double *xs = dyn, *rs = xs + M, *zs = rs + M, *qs = zs + M, *bs = qs + M, *ds = bs + M, *ps = ds + M;
  double* red = ps + M;             // [R][4][2][DC * DC] product partials
  int* cols = reinterpret_cast<int*>(red + (size_t)R * 8 * DC * DC);
  double* PA = a.partials;          // p.q   (and b.b at the start)
  double* PB = a.partials + G;      // r.z
  double* PC = a.partials + 2 * G;  // x.(b + r)
  const int i_lo = DIST ? a.i_lo : 0, i_hi = DIST ? a.i_hi : a.n_img;
  const size_t NV = (size_t)a.n_img * DC, z_off = 2 * NV, x_off = 3 * NV, y_off = 4 * NV;
  const size_t tot_off = (y_off + a.n_cdim + 1) & ~(size_t)1;
  const double* z_s = a.p_s + z_off;
  const double* x_s = a.p_s + x_off;
  unsigned epoch = DIST ? *a.epoch : 0u;
  double (*s_tot)[8] = s_tot_mem;
  auto image_of = [&](int k) { return i_lo + blockIdx.x + (k / DC) * G; };
  // an entry of a shared vector: into this GPU's copy, or (DIST) into every rank's
  auto publish = [&](size_t off, double v) {
    if (!DIST) {
      a.p_s[off] = v;
    } else {
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < a.nranks) a.peer[r][off] = v;
    }
  };
  // z = M^-1 r on the CTA's rows (rs complete), returns this thread's part of r.z
  auto precondition = [&]() {
    double rz = 0.0;
    for (int k = t; k < M; k += 128) {
      if (cols[k] < 0) continue;
      const int slot = k % DC, r0 = k - slot;
      const double* m = a.Minv_s + ((size_t)image_of(k) * DC + slot) * DC;
      double zj = 0.0;
#pragma unroll
      for (int c = 0; c < DC; ++c) zj += m[c] * rs[r0 + c];
      zs[k] = zj;
      publish(z_off + (size_t)image_of(k) * DC + slot, zj);
      rz += rs[k] * zj;
    }
    return rz;
  };
  // ---- x = 0, r = b, |b|^2 ----
  double s = 0.0;
  for (int k = t; k < M; k += 128) {
    const int i = image_of(k);
    const int col = i < i_hi ? a.slot_col[(size_t)i * DC + k % DC] : -1;
    cols[k] = col;
    const double v = col >= 0 ? a.b[col] : 0.0, d = col >= 0 ? a.Dc[col] : 0.0;
    xs[k] = 0.0; rs[k] = v; bs[k] = v; zs[k] = 0.0; qs[k] = 0.0; ps[k] = 0.0; ds[k] = d * d;
    s += v * v;
    if (i < i_hi) {
      const size_t g = (size_t)i * DC + k % DC;
      publish(g, 0.0); publish(NV + g, 0.0); publish(z_off + g, 0.0); publish(x_off + g, 0.0);
    }
  }
  s = BlockAllSum128(s);  // (also orders rs for the preconditioner below)
  if (t == 0) PA[blockIdx.x] = s;
  double tot[2];
  SyncTotals<DIST>(grid, a, epoch, tot_off, false, 1, PA, 0, nullptr, 0, tot, s_tot);
  const double bb = tot[0];
  int state = 0, iter = 0;
  double rho = 1.0, last_rho = 1.0, Q0 = 0.0, zeta = 0.0, alpha = 0.0, beta = 0.0, pq = 0.0;
  if (bb == 0.0) {
    state = 1;  // b == 0: x = 0 is the solution
  } else {
    double rz = BlockAllSum128(precondition());
    if (t == 0) PB[blockIdx.x] = rz;
    SyncTotals<DIST>(grid, a, epoch, tot_off, true, 1, PB, 1, nullptr, 0, tot, s_tot);
    double rho_next = tot[0];
    for (int it = 1;; ++it) {
      PcgStamp(a, it, 0);
      const double rho_new = rho_next;
      if (BadScalar(rho_new)) { state = 3; iter = it; break; }
      beta = 0.0;
      if (it > 1) {
        beta = rho_new / rho;
        if (BadScalar(beta)) { state = 3; iter = it; break; }
      }
      last_rho = rho; rho = rho_new;
      // p = z + beta p: own rows here (published for the NEXT iteration's products), neighbours inside the product
      const double* p_prev = a.p_s + (size_t)((it - 1) & 1) * NV;
      const size_t p_cur = (size_t)(it & 1) * NV;
      for (int k = t; k < M; k += 128) {
        if (cols[k] < 0) continue;
        const double pv = zs[k] + beta * ps[k];  // beta = 0 and p = 0 in the first iteration
        ps[k] = pv;
        publish(p_cur + (size_t)image_of(k) * DC + k % DC, pv);
      }
      PcgStamp(a, it, 1);
      PcgStamp(a, it, 2);
      RowsProduct<DC, DIST>(a, i_lo + blockIdx.x, i_hi, R, z_s, p_prev, beta, red, qs);
      double part = 0.0;
      for (int k = t; k < M; k += 128) {
        if (cols[k] < 0) continue;
        const double qv = qs[k] + ds[k] * ps[k];
        qs[k] = qv;
        part += ps[k] * qv;
      }
      part = BlockAllSum128(part);
      if (t == 0) PA[blockIdx.x] = part;
      PcgStamp(a, it, 3);
      SyncTotals<DIST>(grid, a, epoch, tot_off, false, 1, PA, 0, nullptr, 0, tot, s_tot);  // p.q (nothing pushed since the last sync is read before the next one)
      PcgStamp(a, it, 4);
      pq = tot[0];
      if (pq <= 0.0 || isinf(pq)) { state = 2; iter = it; break; }
      alpha = rho / pq;
      if (isinf(alpha)) { state = 3; iter = it; break; }
      const bool reset = (it % 10) == 0;  // residual recomputed from scratch every 10th iteration
      for (int k = t; k < M; k += 128) {
        if (cols[k] < 0) continue;
        xs[k] += alpha * ps[k];
        if (!reset) rs[k] -= alpha * qs[k];
        else publish(x_off + (size_t)image_of(k) * DC + k % DC, xs[k]);
      }
      if (reset) {
        SyncTotals<DIST>(grid, a, epoch, tot_off, true, 0, nullptr, 0, nullptr, 0, tot, s_tot);  // x complete
        RowsProduct<DC, DIST>(a, i_lo + blockIdx.x, i_hi, R, x_s, nullptr, 0.0, red, qs);
        for (int k = t; k < M; k += 128)
          if (cols[k] >= 0) rs[k] = bs[k] - (qs[k] + ds[k] * xs[k]);
      }
      __syncthreads();  // rs complete
      double qp = 0.0;
      for (int k = t; k < M; k += 128)
        if (cols[k] >= 0) qp += xs[k] * (bs[k] + rs[k]);
      rz = precondition();
      qp = BlockAllSum128(qp); rz = BlockAllSum128(rz);
      if (t == 0) { PC[blockIdx.x] = qp; PB[blockIdx.x] = rz; }
      PcgStamp(a, it, 5);
      SyncTotals<DIST>(grid, a, epoch, tot_off, true, 2, PC, 2, PB, 1, tot, s_tot);  // x.(b + r) and r.z
      PcgStamp(a, it, 6);
      const double Q1 = -1.0 * tot[0];
      rho_next = tot[1];
      zeta = it * (Q1 - Q0) / Q1;
      Q0 = Q1;
      iter = it;
      if (zeta < a.q_tol && it >= a.min_it) { state = 1; break; }
      if (it >= a.max_it) { state = 2; break; }
    }
  }
  if (!DIST) {
    for (int k = t; k < M; k += 128)
      if (cols[k] >= 0) a.y[cols[k]] = xs[k];
  } else {
    for (int k = t; k < M; k += 128)
      if (cols[k] >= 0) publish(y_off + cols[k], xs[k]);
    SyncTotals<DIST>(grid, a, epoch, tot_off, true, 0, nullptr, 0, nullptr, 0, tot, s_tot);  // every rank's part of y has arrived
    const double* y_all = a.p_s + y_off;
    for (int k = blockIdx.x * 128 + t; k < a.n_cdim; k += G * 128) a.y_out[k] = __ldcg(y_all + k);
    if (blockIdx.x == 0 && t == 0) *a.epoch = epoch;
  }
  if (blockIdx.x == 0 && t == 0) {
    DeviceScalars* sc = a.sc;
    sc->cg_norm_b = sqrt(bb);
    sc->cg_rho = rho; sc->cg_last_rho = last_rho; sc->cg_alpha = alpha; sc->cg_beta = beta; sc->cg_pq = pq;
    sc->cg_Q0 = Q0; sc->cg_Q1 = Q0; sc->cg_zeta = zeta;
    sc->cg_iter = iter; sc->cg_state = state;
    if (state == 3) sc->solver_failed = 1;  // LinearSolver FAILURE: the step is unusable (TrustRegionMinimizer treats it as invalid)
  }
}

// Launches the persistent solve of S y = rhs into d_yc; the caller reads DeviceScalars afterwards (pcg.cu).
int SolvePCGBlocks(Problem* p, const sba_solver_options& o) {
  BlockSchur& B = p->bsr;
  const int dc = p->dc;
  const size_t NV = (size_t)p->n_img * dc;
  const size_t tot_off = (4 * NV + p->n_cdim + 1) & ~(size_t)1;
  const size_t area_doubles = tot_off + 4 * 8 * 2, area_bytes = 128 + area_doubles * sizeof(double);  // flags [8] + sync counter in the first 128 B (fixed, whatever the problem)
  // N > 1 with peer memory: SBA_PCG_SHARDED=1 splits the rows across the ranks.  Not the default: measured on this pool a
  // cross-GPU sync costs 10 us (23 us when vector entries were pushed before it), twice per CG iteration, which is what the
  // replicated product costs -- C4 on 2 GPUs 17.9 vs 16.8 ms per solve, C5 on 8 GPUs 24.0 vs 23.3 ms.
  static const bool sharded = std::getenv("SBA_PCG_SHARDED") != nullptr;
  const bool dist = CommHasPeerMemory(p) && sharded;
  auto kernel_of = [&](bool d) -> const void* {
#define SBA_PCG_FN(DC) (d ? (const void*)pcg_blocks_kernel<DC, true> : (const void*)pcg_blocks_kernel<DC, false>)
    switch (dc) {
      case 6: return SBA_PCG_FN(6); case 8: return SBA_PCG_FN(8); case 10: return SBA_PCG_FN(10);
      case 14: return SBA_PCG_FN(14); default: return SBA_PCG_FN(18);
    }
#undef SBA_PCG_FN
  };
  const void* fn = kernel_of(dist);
  void* area[8] = {nullptr};
  int nranks = 1, rank = 0;
  if (dist) {  // (collective on the first call of a group; a cached lookup afterwards)
    const int rc = CommPcgArea(p, area_bytes, area, &nranks, &rank);
    if (rc) return rc;
  }
  const int i_lo = (int)((int64_t)p->n_img * rank / nranks), i_hi = (int)((int64_t)p->n_img * (rank + 1) / nranks);
  const int rows_max = DivUp(p->n_img, nranks);  // the same bound on every rank
  auto smem_for = [&](int rows) { return (size_t)rows * dc * (7 * sizeof(double) + sizeof(int)) + (size_t)rows * 8 * dc * dc * sizeof(double); };
  if (B.pcg_grid == 0 || B.pcg_dist != dist) {
    int coop = 0;
    SBA_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, p->device));
    if (!coop) { SetError(SBA_ERR_CUDA, "device lacks cooperative launch"); return SBA_ERR_CUDA; }
    const char* env = std::getenv("SBA_PCG_CTAS_PER_SM");
    const int cap = (env && std::atoi(env) > 0) ? std::atoi(env) : 1 << 20;
    // the grid must be co-resident: occupancy depends on the shared memory of the rows a CTA owns, which depends on
    // the grid -- two rounds settle it
    int grid = std::max(1, std::min(rows_max, p->n_sm));
    for (int round = 0; round < 3; ++round) {
      const int rows = DivUp(rows_max, grid);
      int per_sm = 0;
      SBA_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem_for(rows), 1024)));
      SBA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 128, smem_for(rows)));
      per_sm = std::min(per_sm, cap);
      if (per_sm < 1) { SetError(SBA_ERR_CUDA, "persistent PCG kernel does not fit"); return SBA_ERR_CUDA; }
      const int next = std::max(1, std::min(rows_max, per_sm * p->n_sm));
      if (next <= grid && round > 0) break;
      grid = next;
    }
    B.pcg_grid = grid;
    B.pcg_rows = DivUp(rows_max, grid);
    B.pcg_dist = dist;
    SBA_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem_for(B.pcg_rows), 1024)));
    SBA_CUDA(B.pcg_partials.Alloc((size_t)3 * B.pcg_grid));
    if (!dist) SBA_CUDA(B.p_s.Alloc(4 * NV));
  }
  PcgBlocks a;
  a.row_ptr = B.row_ptr.p; a.entries = B.entries.p; a.B = B.blocks.p; a.slot_col = B.slot_col.p;
  a.Minv_s = B.minv_s.p; a.Dc = p->d_Dc.p; a.b = p->d_rhs.p; a.y = p->d_yc.p;
  a.p_s = dist ? reinterpret_cast<double*>(static_cast<char*>(area[rank]) + 128) : B.p_s.p;
  a.partials = B.pcg_partials.p; a.sc = p->d_scalars.p;
  a.n_img = p->n_img; a.rows_per_cta = B.pcg_rows;
  a.min_it = o.min_linear_solver_iterations; a.max_it = std::max(1, o.max_linear_solver_iterations);
  a.q_tol = o.eta;
  a.i_lo = i_lo; a.i_hi = i_hi; a.nranks = nranks; a.rank = rank; a.n_cdim = p->n_cdim;
  for (int r = 0; r < 8; ++r) {
    a.peer[r] = r < nranks && dist ? reinterpret_cast<double*>(static_cast<char*>(area[r]) + 128) : nullptr;
    a.peer_flags[r] = r < nranks && dist ? static_cast<unsigned*>(area[r]) : nullptr;
  }
  // the sync counter lives with the flags (per communicator group: every handle of the group continues one sequence)
  a.epoch = dist ? a.peer_flags[rank] + 8 : nullptr; a.y_out = p->d_yc.p;
  a.trace = nullptr;
  static const bool want_trace = std::getenv("SBA_PCG_TRACE") != nullptr;
  DevBuf<long long> trace;
  if (want_trace) { SBA_CUDA(trace.Alloc(64 * 8)); SBA_CUDA(trace.Zero(p->stream)); a.trace = trace.p; }
  void* args[] = {&a};
  SBA_CUDA(cudaLaunchCooperativeKernel(fn, dim3(B.pcg_grid), dim3(128), args, smem_for(B.pcg_rows), p->stream));
  p->launches++;
  if (want_trace) {
    std::vector<long long> h(64 * 8);
    SBA_CUDA(cudaMemcpyAsync(h.data(), trace.p, h.size() * sizeof(long long), cudaMemcpyDeviceToHost, p->stream));
    SBA_CUDA(cudaStreamSynchronize(p->stream));
    std::fprintf(stderr, "[pcg trace] grid %d rows/cta %d : it | totals+p | - | product | sync(p.q) | update+precond | sync(r.z) (us)\n", B.pcg_grid, B.pcg_rows);
    for (int it = 1; it < 16 && h[it * 8 + 6]; ++it) {
      const long long* r = &h[it * 8];
      std::fprintf(stderr, "[pcg trace] %2d | %6.2f | %6.2f | %6.2f | %6.2f | %6.2f | %6.2f\n", it, (r[1] - r[0]) * 1e-3,
                   (r[2] - r[1]) * 1e-3, (r[3] - r[2]) * 1e-3, (r[4] - r[3]) * 1e-3, (r[5] - r[4]) * 1e-3, (r[6] - r[5]) * 1e-3);
    }
  }
  return SBA_OK;
}

}  // namespace sba
