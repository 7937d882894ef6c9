// K-BSR: explicit block-sparse reduced camera matrix for the iterative solver.
//
// The reference selects ITERATIVE_SCHUR + SCHUR_JACOBI above 1000 images
// (/root/reference/src/optim/bundle_adjustment.cc:283-286); Ceres then multiplies by the Schur complement
// IMPLICITLY -- two passes over the whole Jacobian per CG iteration, ~350 B per observation.
// In structure-from-motion a camera shares points with a bounded set of neighbours, so S itself is block sparse
// and far smaller than the Jacobian: C4 (1778 cameras, dc = 8, 5 M observations) has ~57 k lower blocks = 29 MB
// against 1.8 GB of Jacobian traffic per implicit product.  S is the same matrix either way (Ceres'
// use_explicit_schur_complement computes exactly this); only the rounding order of a product differs.
//
//   structure : the camera-pair keys of the dense pair list (pairs.cu), already sorted by (gi, gj), ARE the block
//               list; a row's neighbour list = its diagonal block, its (gi = i) keys, its (gj = i) keys transposed.
//               With several GPUs the key sets of the shards are united through a presence table (one int
//               allreduce), so every rank holds the same block layout and the blocks are summed by ONE fp64
//               allreduce per LM iteration instead of one camera vector per CG product.
//   values    : schur.cu accumulates into the blocks (Sink kind 2), single writer per entry, fixed order.
//   product   : one CTA per block row; warp w streams entries w, w+4, ...; lanes own block elements; the only
//               reductions are one shared-memory pass per row.  No atomics: bit-reproducible.
// Blocks are stored in SLOT space (pose 6 + intrinsics slots), so constant slots are simply columns nobody reads.
// Requires every free intrinsics block to belong to one image (otherwise the implicit product stays in use).
#include <cub/cub.cuh>

#include "solver.h"

namespace sba {

__global__ void bsr_slot_col_kernel(StructView S, int dc, int* __restrict__ slot_col) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= S.n_img * dc) return;
  const int i = t / dc, j = t - i * dc;
  slot_col[t] = j < 6 ? S.col_img[6 * (size_t)i + j] : S.col_cam[12 * (size_t)S.img_cam[i] + j - 6];
}

__global__ void bsr_mark_kernel(int n_keys, int n_img, const int* __restrict__ gi, const int* __restrict__ gj,
                                int* __restrict__ present) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n_keys) present[(size_t)gi[k] * n_img + gj[k]] = 1;
}

__global__ void bsr_compact_kernel(int64_t cells, int n_img, const int* __restrict__ present, const int* __restrict__ pos,
                                   int* __restrict__ ggi, int* __restrict__ ggj) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cells || !present[c]) return;
  ggi[pos[c]] = (int)(c / n_img); ggj[pos[c]] = (int)(c % n_img);
}

__global__ void bsr_key_block_kernel(int n_keys, int n_img, const int* __restrict__ gi, const int* __restrict__ gj,
                                     const int* __restrict__ pos, int* __restrict__ key_block) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n_keys) key_block[k] = n_img + pos[(size_t)gi[k] * n_img + gj[k]];
}

// y = (S + Dc^2) x for the rows of image blockIdx.x
template <int DC>
__global__ void __launch_bounds__(128) bsr_matvec_kernel(const int* __restrict__ row_ptr, const int2* __restrict__ entries,
                                                         const double* __restrict__ B, const int* __restrict__ slot_col,
                                                         const double* __restrict__ x, const double* __restrict__ Dc,
                                                         double* __restrict__ y, const int* __restrict__ skip) {
  if (skip && *skip != 0) return;
  constexpr int NE = DC * DC, N = (NE + 31) / 32;
  __shared__ double red[4][2][NE];
  const int i = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int er[N], ec[N];
  double as[N], at[N];
#pragma unroll
  for (int q = 0; q < N; ++q) {
    const int e = min(lane + 32 * q, NE - 1);
    er[q] = e / DC; ec[q] = e - er[q] * DC;
    as[q] = 0.0; at[q] = 0.0;
  }
  const int e0 = row_ptr[i], e1 = row_ptr[i + 1];
  int2 nxt = (e0 + warp < e1) ? entries[e0 + warp] : make_int2(0, 0);
  for (int t = e0 + warp; t < e1; t += 4) {
    const int2 en = nxt;
    if (t + 4 < e1) nxt = entries[t + 4];
    const int j = en.y & 0x7fffffff;
    const bool tr = en.y < 0;
    double xv = 0.0;
    if (lane < DC) { const int c = slot_col[(size_t)j * DC + lane]; if (c >= 0) xv = x[c]; }
    const double* blk = B + (size_t)en.x * NE;
    double v[N];
#pragma unroll
    for (int q = 0; q < N; ++q) v[q] = (lane + 32 * q < NE) ? blk[lane + 32 * q] : 0.0;
    if (tr) {
#pragma unroll
      for (int q = 0; q < N; ++q) at[q] += v[q] * __shfl_sync(0xffffffffu, xv, er[q]);
    } else {
#pragma unroll
      for (int q = 0; q < N; ++q) as[q] += v[q] * __shfl_sync(0xffffffffu, xv, ec[q]);
    }
  }
#pragma unroll
  for (int q = 0; q < N; ++q) {
    const int e = lane + 32 * q;
    if (e < NE) { red[warp][0][e] = as[q]; red[warp][1][e] = at[q]; }
  }
  __syncthreads();
  if (threadIdx.x < DC) {
    const int r = threadIdx.x;
    const int col = slot_col[(size_t)i * DC + r];
    if (col >= 0) {
      double s = 0.0;
#pragma unroll
      for (int w = 0; w < 4; ++w) {
#pragma unroll
        for (int c = 0; c < DC; ++c) s += red[w][0][r * DC + c];
#pragma unroll
        for (int c = 0; c < DC; ++c) s += red[w][1][c * DC + r];
      }
      const double d = Dc[col];
      y[col] = s + d * d * x[col];
    }
  }
}

// Builds the dense pair list, unites the key sets across ranks, lays out the blocks and the row lists, and decides
// whether the explicit product is the cheaper one.  Runs once per problem (first iterative LM iteration).
int DecideBlockSchur(Problem* p) {
  BlockSchur& B = p->bsr;
  if (B.decided) return SBA_OK;
  B.decided = true;
  B.enabled = false;
  const char* env = std::getenv("SBA_IMPLICIT_SCHUR");
  if (env && env[0] == '1') return SBA_OK;
  if (p->shared_intrinsics || p->n_cdim == 0) return SBA_OK;
  const int n_img = p->n_img, dc = p->dc;
  const int64_t cells = (int64_t)n_img * n_img;
  if (cells > 0x7fffffffLL || (p->comm && cells > (1ll << 27))) return SBA_OK;
  {  // the pair list grows with the square of the track length: bound it before building it
    const double pairs = p->pair_estimate;  // (from the point pass of sba_create)
    size_t free_b = 0, total_b = 0;
    SBA_CUDA(cudaMemGetInfo(&free_b, &total_b));
    double too_many = (pairs > 1.5e9 || pairs * 40.0 > 0.25 * (double)free_b) ? 1.0 : 0.0;  // 32-bit pair offsets; 40 B per pair while sorting
    if (p->comm) {  // every rank must take the same path
      double* flag = p->d_red.p;
      SBA_CUDA(cudaMemcpyAsync(flag, &too_many, sizeof(double), cudaMemcpyHostToDevice, p->stream));
      int rcf = AllReduceMax(p, flag, 1);
      if (rcf) return rcf;
      SBA_CUDA(cudaMemcpyAsync(&too_many, flag, sizeof(double), cudaMemcpyDeviceToHost, p->stream));
      SBA_CUDA(cudaStreamSynchronize(p->stream));
    }
    if (too_many != 0.0) return SBA_OK;
  }
  HostTimer tm("block-sparse S structure");
  int rc = BuildPairLists(p, true);
  if (rc) return rc;
  cudaStream_t st = p->stream;
  const PairList& L = p->pairs_dense;
  std::vector<int> gi, gj;
  if (!p->comm) {
    B.n_keys = L.n_keys;
    gi.resize(L.n_keys); gj.resize(L.n_keys);
    if (L.n_keys > 0) {
      SBA_CUDA(cudaMemcpyAsync(gi.data(), L.key_gi.p, (size_t)L.n_keys * sizeof(int), cudaMemcpyDeviceToHost, st));
      SBA_CUDA(cudaMemcpyAsync(gj.data(), L.key_gj.p, (size_t)L.n_keys * sizeof(int), cudaMemcpyDeviceToHost, st));
    }
    SBA_CUDA(cudaStreamSynchronize(st));
  } else {
    // union of the shards' key sets: presence table, int max-allreduce, prefix sum = global block numbering
    DevBuf<int> present, pos, ggi, ggj;
    DevBuf<char> tmp;
    SBA_CUDA(present.Alloc(cells)); SBA_CUDA(pos.Alloc(cells));
    SBA_CUDA(present.Zero(st));
    if (L.n_keys > 0) bsr_mark_kernel<<<DivUp(L.n_keys, 256), 256, 0, st>>>(L.n_keys, n_img, L.key_gi.p, L.key_gj.p, present.p);
    rc = AllReduceMaxInt(p, present.p, cells);
    if (rc) return rc;
    size_t bytes = 0;
    SBA_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, bytes, present.p, pos.p, (int)cells, st));
    SBA_CUDA(tmp.Alloc(bytes));
    SBA_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, bytes, present.p, pos.p, (int)cells, st));
    int last[2] = {0, 0};
    SBA_CUDA(cudaMemcpyAsync(&last[0], pos.p + cells - 1, sizeof(int), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaMemcpyAsync(&last[1], present.p + cells - 1, sizeof(int), cudaMemcpyDeviceToHost, st));
    SBA_CUDA(cudaStreamSynchronize(st));
    B.n_keys = last[0] + last[1];
    SBA_CUDA(ggi.Alloc(std::max(1, B.n_keys))); SBA_CUDA(ggj.Alloc(std::max(1, B.n_keys)));
    bsr_compact_kernel<<<DivUp(cells, 256), 256, 0, st>>>(cells, n_img, present.p, pos.p, ggi.p, ggj.p);
    SBA_CUDA(B.key_block.Alloc(std::max(1, L.n_keys)));
    if (L.n_keys > 0) bsr_key_block_kernel<<<DivUp(L.n_keys, 256), 256, 0, st>>>(L.n_keys, n_img, L.key_gi.p, L.key_gj.p, pos.p, B.key_block.p);
    p->launches += 3;
    gi.resize(B.n_keys); gj.resize(B.n_keys);
    if (B.n_keys > 0) {
      SBA_CUDA(cudaMemcpyAsync(gi.data(), ggi.p, (size_t)B.n_keys * sizeof(int), cudaMemcpyDeviceToHost, st));
      SBA_CUDA(cudaMemcpyAsync(gj.data(), ggj.p, (size_t)B.n_keys * sizeof(int), cudaMemcpyDeviceToHost, st));
    }
    SBA_CUDA(cudaStreamSynchronize(st));
  }
  tm.Lap("keys");
  // cost model: bytes one product streams.  Explicit: every off-diagonal block twice (straight + transposed) and the
  // diagonal once; implicit: two passes over F (2 x dc) and E (2 x 3) of every observation of the whole problem.
  const double ne = (double)dc * dc;
  const double explicit_bytes = ((double)n_img + 2.0 * B.n_keys) * ne * 8.0;
  const double implicit_bytes = 2.0 * (double)p->n_obs_global * (2.0 * dc + 6.0) * 8.0;
  size_t free_b = 0, total_b = 0;
  SBA_CUDA(cudaMemGetInfo(&free_b, &total_b));
  const bool force = env && env[0] == '0';
  // (total, not free, memory: every rank must reach the same verdict)
  if (!force && (explicit_bytes > implicit_bytes || explicit_bytes > 0.125 * (double)total_b)) return SBA_OK;

  // row lists (host; n_keys entries): diagonal, lower neighbours in key order, upper neighbours in key order
  std::vector<int> cnt((size_t)n_img + 1, 0);
  for (int i = 0; i < n_img; ++i) cnt[i + 1] = 1;
  for (int k = 0; k < B.n_keys; ++k)
    if (gi[k] != gj[k]) { cnt[gi[k] + 1]++; cnt[gj[k] + 1]++; }
  for (int i = 0; i < n_img; ++i) cnt[i + 1] += cnt[i];
  std::vector<int2> ent((size_t)cnt[n_img]);
  std::vector<int> at(cnt.begin(), cnt.end() - 1);
  for (int i = 0; i < n_img; ++i) ent[at[i]++] = make_int2(i, i);
  for (int k = 0; k < B.n_keys; ++k)
    if (gi[k] != gj[k]) ent[at[gi[k]]++] = make_int2(n_img + k, gj[k]);
  for (int k = 0; k < B.n_keys; ++k)
    if (gi[k] != gj[k]) ent[at[gj[k]]++] = make_int2(n_img + k, (int)((unsigned)gi[k] | 0x80000000u));
  B.n_entries = (int64_t)ent.size();
  B.n_blocks = (int64_t)n_img + B.n_keys;
  SBA_CUDA(B.row_ptr.Upload(cnt, st));
  SBA_CUDA(B.entries.Upload(ent, st));
  SBA_CUDA(B.slot_col.Alloc((size_t)n_img * dc));
  bsr_slot_col_kernel<<<DivUp((int64_t)n_img * dc, 256), 256, 0, st>>>(p->Struct(), dc, B.slot_col.p);
  p->launches++;
  SBA_CUDA(B.blocks.Alloc((size_t)B.n_blocks * dc * dc + p->n_cdim));  // + room for the rhs riding on the all-reduce
  SBA_CUDA(B.blocks.Zero(st));  // blocks of keys other ranks own stay zero here until the allreduce
  SBA_CUDA(cudaGetLastError());
  SBA_CUDA(cudaStreamSynchronize(st));  // the host vectors above go out of scope
  B.enabled = true;
  tm.Lap("rows + blocks");
  return SBA_OK;
}

int BlockSchurMatVec(Problem* p, const double* d_x, double* d_y, const int* skip) {
  const BlockSchur& B = p->bsr;
#define CALL(DC) bsr_matvec_kernel<DC><<<p->n_img, 128, 0, p->stream>>>(B.row_ptr.p, B.entries.p, B.blocks.p, B.slot_col.p, d_x, p->d_Dc.p, d_y, skip)
  switch (p->dc) {
    case 6: CALL(6); break; case 8: CALL(8); break; case 10: CALL(10); break;
    case 14: CALL(14); break; default: CALL(18); break;
  }
#undef CALL
  p->launches++;
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

}  // namespace sba
