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
#include <cooperative_groups.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "reduce.cuh"
#include "solver.h"

namespace cg = cooperative_groups;

namespace sba {

struct PcgBlocks {
  const int* row_ptr; const int2* entries; const double* B; const int* slot_col;
  const double* Minv_s;  // [n_img][DC x DC] SCHUR_JACOBI inverse blocks laid out in slot space (zero outside a block)
  const double* Dc; const double* b;
  double* y;             // compact solution (n_cdim)
  double *p_s, *z_s, *x_s;  // slot-space vectors read by other CTAs: p_s [2][n_img * DC] (ping-pong), z_s, x_s [n_img * DC]
                            // (x_s only on residual-reset iterations)
  double* partials;      // [3][gridDim.x]
  DeviceScalars* sc;
  int n_img, rows_per_cta, min_it, max_it;
  double q_tol;
  long long* trace;      // optional [64][8] globaltimer stamps of CTA 0 (SBA_PCG_TRACE=1), else null
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

// (S u) restricted to the rows of image i, u = v + beta vold in slot space (vold may be null); thread r < DC returns
// the entry of slot r.
// A warp takes a contiguous quarter of the row's entries: one coalesced load brings 32 of them, the loop below then
// has no load that depends on another load, so several blocks are in flight per warp.
// v is written by other CTAs during the kernel: plain loads only.
template <int DC>
__device__ inline double RowProduct(const PcgBlocks& a, int i, const double* v, const double* vold, double beta,
                                    double (*red)[2][DC * DC]) {
  constexpr int NE = DC * DC, N = (NE + 31) / 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int er[N], ec[N];
  double as[N], at[N];
#pragma unroll
  for (int q = 0; q < N; ++q) {
    const int e = min(lane + 32 * q, NE - 1);
    er[q] = e / DC; ec[q] = e - er[q] * DC;
    as[q] = 0.0; at[q] = 0.0;
  }
  const int e0 = a.row_ptr[i], cnt = a.row_ptr[i + 1] - e0;
  const int per = (cnt + 3) >> 2;
  const int w0 = e0 + warp * per, w1 = min(e0 + cnt, w0 + per);
  for (int base = w0; base < w1; base += 32) {
    const int m = min(32, w1 - base);
    const int2 mine = lane < m ? a.entries[base + lane] : make_int2(0, 0);
#pragma unroll 4
    for (int k = 0; k < m; ++k) {
      const int blk_id = __shfl_sync(0xffffffffu, mine.x, k), jy = __shfl_sync(0xffffffffu, mine.y, k);
      const int j = jy & 0x7fffffff;
      double xv = 0.0;
      if (lane < DC) {
        xv = v[(size_t)j * DC + lane];
        if (vold) xv += beta * vold[(size_t)j * DC + lane];  // the direction of row j, as its owner forms it
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
#pragma unroll
  for (int q = 0; q < N; ++q) {
    const int e = lane + 32 * q;
    if (e < NE) { red[warp][0][e] = as[q]; red[warp][1][e] = at[q]; }
  }
  __syncthreads();
  double out = 0.0;
  if (threadIdx.x < DC) {
    const int r = threadIdx.x;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
#pragma unroll
      for (int c = 0; c < DC; ++c) out += red[w][0][r * DC + c];
#pragma unroll
      for (int c = 0; c < DC; ++c) out += red[w][1][c * DC + r];
    }
  }
  __syncthreads();  // red is reused by the next row
  return out;
}

// The CTA's own rows live in shared memory for the whole solve: x, r, z, q, b, Dc^2 and the compact column of each
// (row, slot) item k = local_row * DC + slot.  Items of constant slots (col < 0) and of rows past n_img stay zero.
template <int DC>
__global__ void __launch_bounds__(128) pcg_blocks_kernel(PcgBlocks a) {
  cg::grid_group grid = cg::this_grid();
  __shared__ double red[4][2][DC * DC];
  extern __shared__ double dyn[];
  const int G = gridDim.x, t = threadIdx.x, R = a.rows_per_cta, M = R * DC;
  double *xs = dyn, *rs = xs + M, *zs = rs + M, *qs = zs + M, *bs = qs + M, *ds = bs + M, *ps = ds + M;
  int* cols = reinterpret_cast<int*>(ps + M);
  double* PA = a.partials;          // p.q   (and b.b at the start)
  double* PB = a.partials + G;      // r.z
  double* PC = a.partials + 2 * G;  // x.(b + r)
  auto image_of = [&](int k) { return blockIdx.x + (k / DC) * G; };
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
      a.z_s[(size_t)image_of(k) * DC + slot] = zj;
      rz += rs[k] * zj;
    }
    return rz;
  };
  // ---- x = 0, r = b, |b|^2 ----
  double s = 0.0;
  for (int k = t; k < M; k += 128) {
    const int i = image_of(k);
    const int col = i < a.n_img ? a.slot_col[(size_t)i * DC + k % DC] : -1;
    cols[k] = col;
    const double v = col >= 0 ? a.b[col] : 0.0, d = col >= 0 ? a.Dc[col] : 0.0;
    xs[k] = 0.0; rs[k] = v; bs[k] = v; zs[k] = 0.0; qs[k] = 0.0; ps[k] = 0.0; ds[k] = d * d;
    s += v * v;
    if (i < a.n_img) {
      const size_t g = (size_t)i * DC + k % DC;
      a.p_s[g] = 0.0; a.p_s[(size_t)a.n_img * DC + g] = 0.0; a.z_s[g] = 0.0; a.x_s[g] = 0.0;
    }
  }
  s = BlockAllSum128(s);  // (also orders rs for the preconditioner below)
  if (t == 0) PA[blockIdx.x] = s;
  grid.sync();
  const double bb = GridTotal(PA, G);
  int state = 0, iter = 0;
  double rho = 1.0, last_rho = 1.0, Q0 = 0.0, zeta = 0.0, alpha = 0.0, beta = 0.0, pq = 0.0;
  if (bb == 0.0) {
    state = 1;  // b == 0: x = 0 is the solution
  } else {
    double rz = BlockAllSum128(precondition());
    if (t == 0) PB[blockIdx.x] = rz;
    grid.sync();
    for (int it = 1;; ++it) {
      PcgStamp(a, it, 0);
      const double rho_new = GridTotal(PB, G);
      if (BadScalar(rho_new)) { state = 3; iter = it; break; }
      beta = 0.0;
      if (it > 1) {
        beta = rho_new / rho;
        if (BadScalar(beta)) { state = 3; iter = it; break; }
      }
      last_rho = rho; rho = rho_new;
      // p = z + beta p: own rows here (published for the NEXT iteration's products), neighbours inside the product
      const double* p_prev = a.p_s + (size_t)((it - 1) & 1) * a.n_img * DC;
      double* p_cur = a.p_s + (size_t)(it & 1) * a.n_img * DC;
      for (int k = t; k < M; k += 128) {
        if (cols[k] < 0) continue;
        const double pv = zs[k] + beta * ps[k];  // beta = 0 and p = 0 in the first iteration
        ps[k] = pv;
        p_cur[(size_t)image_of(k) * DC + k % DC] = pv;
      }
      PcgStamp(a, it, 1);
      PcgStamp(a, it, 2);
      for (int lr = 0; lr < R; ++lr) {
        const int i = blockIdx.x + lr * G;
        if (i >= a.n_img) break;
        const double v = RowProduct<DC>(a, i, a.z_s, p_prev, beta, red);
        if (t < DC) qs[lr * DC + t] = v;
      }
      __syncthreads();
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
      grid.sync();  // p.q
      PcgStamp(a, it, 4);
      pq = GridTotal(PA, G);
      if (pq <= 0.0 || isinf(pq)) { state = 2; iter = it; break; }
      alpha = rho / pq;
      if (isinf(alpha)) { state = 3; iter = it; break; }
      const bool reset = (it % 10) == 0;  // residual recomputed from scratch every 10th iteration
      for (int k = t; k < M; k += 128) {
        if (cols[k] < 0) continue;
        xs[k] += alpha * ps[k];
        if (!reset) rs[k] -= alpha * qs[k];
        else a.x_s[(size_t)image_of(k) * DC + k % DC] = xs[k];
      }
      if (reset) {
        grid.sync();  // x complete
        for (int lr = 0; lr < R; ++lr) {
          const int i = blockIdx.x + lr * G;
          if (i >= a.n_img) break;
          const double v = RowProduct<DC>(a, i, a.x_s, nullptr, 0.0, red);
          if (t < DC) qs[lr * DC + t] = v;
        }
        __syncthreads();
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
      grid.sync();  // x.(b + r) and r.z
      PcgStamp(a, it, 6);
      const double Q1 = -1.0 * GridTotal(PC, G);
      zeta = it * (Q1 - Q0) / Q1;
      Q0 = Q1;
      iter = it;
      if (zeta < a.q_tol && it >= a.min_it) { state = 1; break; }
      if (it >= a.max_it) { state = 2; break; }
    }
  }
  for (int k = t; k < M; k += 128)
    if (cols[k] >= 0) a.y[cols[k]] = xs[k];
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
  const void* fn = nullptr;
  switch (dc) {
    case 6: fn = (const void*)pcg_blocks_kernel<6>; break;   case 8: fn = (const void*)pcg_blocks_kernel<8>; break;
    case 10: fn = (const void*)pcg_blocks_kernel<10>; break; case 14: fn = (const void*)pcg_blocks_kernel<14>; break;
    default: fn = (const void*)pcg_blocks_kernel<18>; break;
  }
  auto smem_for = [&](int rows) { return (size_t)rows * dc * (7 * sizeof(double) + sizeof(int)); };
  if (B.pcg_grid == 0) {
    int coop = 0;
    SBA_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, p->device));
    if (!coop) { SetError(SBA_ERR_CUDA, "device lacks cooperative launch"); return SBA_ERR_CUDA; }
    const char* env = std::getenv("SBA_PCG_CTAS_PER_SM");
    const int cap = (env && std::atoi(env) > 0) ? std::atoi(env) : 1 << 20;
    // the grid must be co-resident: occupancy depends on the shared memory of the rows a CTA owns, which depends on
    // the grid -- two rounds settle it
    int grid = std::max(1, std::min(p->n_img, p->n_sm));
    for (int round = 0; round < 3; ++round) {
      const int rows = DivUp(p->n_img, grid);
      int per_sm = 0;
      SBA_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem_for(rows), 1024)));
      SBA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 128, smem_for(rows)));
      per_sm = std::min(per_sm, cap);
      if (per_sm < 1) { SetError(SBA_ERR_CUDA, "persistent PCG kernel does not fit"); return SBA_ERR_CUDA; }
      const int next = std::max(1, std::min(p->n_img, per_sm * p->n_sm));
      if (next <= grid && round > 0) break;
      grid = next;
    }
    B.pcg_grid = grid;
    B.pcg_rows = DivUp(p->n_img, grid);
    SBA_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem_for(B.pcg_rows), 1024)));
    SBA_CUDA(B.pcg_partials.Alloc((size_t)3 * B.pcg_grid));
    SBA_CUDA(B.p_s.Alloc((size_t)2 * p->n_img * dc)); SBA_CUDA(B.x_s.Alloc((size_t)p->n_img * dc));
    SBA_CUDA(B.z_s.Alloc((size_t)p->n_img * dc));
  }
  PcgBlocks a;
  a.row_ptr = B.row_ptr.p; a.entries = B.entries.p; a.B = B.blocks.p; a.slot_col = B.slot_col.p;
  a.Minv_s = B.minv_s.p; a.Dc = p->d_Dc.p; a.b = p->d_rhs.p; a.y = p->d_yc.p;
  a.p_s = B.p_s.p; a.x_s = B.x_s.p; a.z_s = B.z_s.p;
  a.partials = B.pcg_partials.p; a.sc = p->d_scalars.p;
  a.n_img = p->n_img; a.rows_per_cta = B.pcg_rows;
  a.min_it = o.min_linear_solver_iterations; a.max_it = std::max(1, o.max_linear_solver_iterations);
  a.q_tol = o.eta;
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
