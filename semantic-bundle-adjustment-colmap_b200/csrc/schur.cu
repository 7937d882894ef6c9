// K-SCHUR: explicit reduced camera matrix (DENSE_SCHUR / SPARSE_SCHUR sizes) and the
// SCHUR_JACOBI block diagonal (ITERATIVE_SCHUR), built without atomics wherever the
// integer indexing gives every matrix entry a single owner.
//
//   S = F^T F + D_c^2 - sum_p (F^T E)_p C_p^{-1} (E^T F)_p
// is what ceres::SchurEliminator<2,3,Dynamic>::Eliminate produces for the problem of
//   /root/reference/src/optim/bundle_adjustment.cc:326-530;
// block-diag(S) at parameter-block granularity is Ceres' SCHUR_JACOBI preconditioner
// selected at bundle_adjustment.cc:283-286.
//
// Work decomposition
//   * "diagonal" contributions (observation a with itself): image-major chunks of <=256
//     observations; a warp walks 32 observations, lanes own the DCxDC output elements;
//     chunk partials are summed per image in a fixed order.
//   * "off-diagonal" contributions: host-built list of observation pairs (a,b) of one
//     point with image(a) > image(b), sorted by (image(a), image(b)) key; one warp per
//     key accumulates its block in registers and is its only writer.
//   * intrinsics shared by several images (and duplicate observations of one point in
//     one image) have several writers per entry; only those entries use fp64 atomics.
#include <algorithm>

#include "solver.h"
#include "tiles.cuh"

namespace sba {

#define SBA_DC_SWITCH(dc, CALL)                                                       \
  switch (dc) {                                                                       \
    case 6: CALL(6); break; case 8: CALL(8); break; case 10: CALL(10); break;         \
    case 14: CALL(14); break; default: CALL(18); break;                               \
  }

template <int DC> struct Elems { static constexpr int N = (DC * DC + 31) / 32; };

struct PairSmem {            // per warp
  double Ha[kMaxDC][3];
  double Wb[kMaxDC][3];
  double2 Fa[kMaxDC];
};

// accumulate the contribution of the ordered observation pair (a,b) of one point:
//   acc[e] -= (W_a C^{-1}) [ja] . W_b[jb]     (+ F_a[ja].F_a[jb] when a == b)
template <int DC>
__device__ inline void AccumulatePair(const StructView& S, const double* __restrict__ ER, const double2* __restrict__ F,
                                      const double* __restrict__ Cinv, int a, int b, PairSmem* sm,
                                      double (&acc)[Elems<DC>::N]) {
  const int lane = threadIdx.x & 31;
  const int po = S.p_off[S.obs_pt[a]];
  if (lane < DC) {
    const double2 fa = F[(size_t)a * DC + lane];
    const double2* ea = reinterpret_cast<const double2*>(ER + (size_t)a * 8);
    const double2 e0 = ea[0], e1 = ea[1], e2 = ea[2];
    double wa[3] = {fa.x * e0.x + fa.y * e1.y, fa.x * e0.y + fa.y * e2.x, fa.x * e1.x + fa.y * e2.y};
    double h[3] = {0.0, 0.0, 0.0};
    if (po >= 0) Sym3Mul(Cinv + 6 * (size_t)po, wa, h);
    sm->Ha[lane][0] = h[0]; sm->Ha[lane][1] = h[1]; sm->Ha[lane][2] = h[2];
    if (a == b) {
      sm->Wb[lane][0] = wa[0]; sm->Wb[lane][1] = wa[1]; sm->Wb[lane][2] = wa[2];
      sm->Fa[lane] = fa;
    } else {
      const double2 fb = F[(size_t)b * DC + lane];
      const double2* eb = reinterpret_cast<const double2*>(ER + (size_t)b * 8);
      const double2 g0 = eb[0], g1 = eb[1], g2 = eb[2];
      sm->Wb[lane][0] = fb.x * g0.x + fb.y * g1.y;
      sm->Wb[lane][1] = fb.x * g0.y + fb.y * g2.x;
      sm->Wb[lane][2] = fb.x * g1.x + fb.y * g2.y;
    }
  }
  __syncwarp();
#pragma unroll
  for (int q = 0; q < Elems<DC>::N; ++q) {
    const int e = lane + 32 * q;
    if (e < DC * DC) {
      const int ja = e / DC, jb = e - ja * DC;
      double v = -(sm->Ha[ja][0] * sm->Wb[jb][0] + sm->Ha[ja][1] * sm->Wb[jb][1] + sm->Ha[ja][2] * sm->Wb[jb][2]);
      if (a == b) v += sm->Fa[ja].x * sm->Fa[jb].x + sm->Fa[ja].y * sm->Fa[jb].y;
      acc[q] += v;
    }
  }
  __syncwarp();
}

// compact column of slot j of an observation of image i (camera c)
__device__ inline int SlotCol(const StructView& S, int i, int c, int j) {
  return j < 6 ? S.col_img[6 * (size_t)i + j] : S.col_cam[12 * (size_t)c + j - 6];
}

// Destination of a block element.  kind 0: dense lower-triangular S (row-major n x n);
// kind 1: SCHUR_JACOBI block-diagonal storage.
struct Sink {
  int kind;
  double* S; int n;                 // dense
  double* M; const int* col_blk; const int* blk_off; const int* blk_dim; const int* blk_moff;  // block diagonal
  int shared_intr;                  // intrinsics columns have several writers -> atomics
};

// same_image: the block is a diagonal (i,i) block made of ordered pairs -> keep only row >= col;
// otherwise (i > j) mirror entries that fall above the diagonal and count a shared column twice.
__device__ inline void Emit(const Sink& k, int row, int col, double v, bool same_image, bool atomic) {
  if (row < 0 || col < 0) return;
  if (same_image) { if (row < col) return; }
  else {
    if (row < col) { const int t = row; row = col; col = t; }
    else if (row == col) v *= 2.0;
  }
  double* dst;
  if (k.kind == 0) {
    dst = k.S + (size_t)row * k.n + col;
  } else {
    const int b = k.col_blk[row];
    if (b != k.col_blk[col]) return;
    const int o = k.blk_off[b], d = k.blk_dim[b];
    dst = k.M + k.blk_moff[b] + (size_t)(row - o) * d + (col - o);
  }
  if (atomic) atomicAdd(dst, v); else *dst = v;
}

// ---- diagonal contributions: chunk partial blocks -------------------------------------
template <int DC>
__global__ void __launch_bounds__(256) schur_diag_chunk_kernel(StructView S, ChunkView C, const double* __restrict__ ER,
                                                               const double2* __restrict__ F, const double* __restrict__ Cinv,
                                                               double* __restrict__ partial) {
  __shared__ PairSmem sm[8];
  __shared__ double red[8][DC * DC];
  const int ch = blockIdx.x, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int s0 = C.chunk_start[ch], n = C.chunk_start[ch + 1] - s0;
  double acc[Elems<DC>::N];
#pragma unroll
  for (int q = 0; q < Elems<DC>::N; ++q) acc[q] = 0.0;
  const int lo = w * 32, hi = min(n, lo + 32);
  for (int t = lo; t < hi; ++t) {
    const int a = C.cam_order[s0 + t];
    AccumulatePair<DC>(S, ER, F, Cinv, a, a, &sm[w], acc);
  }
#pragma unroll
  for (int q = 0; q < Elems<DC>::N; ++q) { const int e = lane + 32 * q; if (e < DC * DC) red[w][e] = acc[q]; }
  __syncthreads();
  for (int e = threadIdx.x; e < DC * DC; e += 256) {
    double a = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) a += red[i][e];
    partial[(size_t)ch * DC * DC + e] = a;
  }
}

template <int DC>
__global__ void __launch_bounds__(128) schur_diag_finalize_kernel(StructView S, ChunkView C, const double* __restrict__ partial,
                                                                  Sink sink) {
  const int i = blockIdx.x;
  const int c = S.img_cam[i];
  const int c0 = C.img_chunk_ptr[i], c1 = C.img_chunk_ptr[i + 1];
  if (c0 == c1) return;
  for (int e = threadIdx.x; e < DC * DC; e += 128) {
    double a = 0.0;
    for (int ch = c0; ch < c1; ++ch) a += partial[(size_t)ch * DC * DC + e];
    const int ja = e / DC, jb = e - ja * DC;
    const bool atomic = sink.shared_intr && (ja >= 6 || jb >= 6);
    Emit(sink, SlotCol(S, i, c, ja), SlotCol(S, i, c, jb), a, true, atomic);
  }
}

// ---- off-diagonal contributions: one warp per (image a, image b) key --------------------
template <int DC>
__global__ void __launch_bounds__(128) schur_pair_kernel(StructView S, int n_keys, const int* __restrict__ key_ptr,
                                                         const int* __restrict__ key_gi, const int* __restrict__ key_gj,
                                                         const int* __restrict__ pa, const int* __restrict__ pb,
                                                         const double* __restrict__ ER, const double2* __restrict__ F,
                                                         const double* __restrict__ Cinv, Sink sink, int cam_only) {
  __shared__ PairSmem sm[4];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int key = blockIdx.x * 4 + w;
  if (key >= n_keys) return;
  double acc[Elems<DC>::N];
#pragma unroll
  for (int q = 0; q < Elems<DC>::N; ++q) acc[q] = 0.0;
  for (int t = key_ptr[key]; t < key_ptr[key + 1]; ++t) AccumulatePair<DC>(S, ER, F, Cinv, pa[t], pb[t], &sm[w], acc);
  const int gi = key_gi[key], gj = key_gj[key];
  const int ci = S.img_cam[gi], cj = S.img_cam[gj];
  const bool same = gi == gj;
#pragma unroll
  for (int q = 0; q < Elems<DC>::N; ++q) {
    const int e = lane + 32 * q;
    if (e < DC * DC) {
      const int ja = e / DC, jb = e - ja * DC;
      const bool intr = (ja >= 6 || jb >= 6);
      if (cam_only && !(ja >= 6 && jb >= 6)) continue;
      // duplicate-observation keys (same image) add on top of the chunk result -> atomics
      const bool atomic = same || (sink.shared_intr && intr);
      Emit(sink, SlotCol(S, gi, ci, ja), SlotCol(S, gj, cj, jb), acc[q], same, atomic);
    }
  }
}

__global__ void add_diag_dense_kernel(int n, double* __restrict__ S, const double* __restrict__ Dc) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) S[(size_t)j * n + j] += Dc[j] * Dc[j];
}

// block-Jacobi: add D^2, invert every block (dim <= 12) through its Cholesky factor
__global__ void __launch_bounds__(64) invert_blocks_kernel(int n_blk, const int* __restrict__ blk_off, const int* __restrict__ blk_dim,
                                                           const int* __restrict__ blk_moff, const double* __restrict__ M,
                                                           const double* __restrict__ Dc, double* __restrict__ Minv,
                                                           DeviceScalars* sc) {
  const int b = blockIdx.x * 64 + threadIdx.x;
  if (b >= n_blk) return;
  const int d = blk_dim[b], o = blk_off[b];
  const double* m = M + blk_moff[b];
  double L[12 * 12];
  for (int r = 0; r < d; ++r)
    for (int c = 0; c <= r; ++c) L[r * 12 + c] = m[r * d + c] + (r == c ? Dc[o + r] * Dc[o + r] : 0.0);
  bool ok = true;
  for (int j = 0; j < d && ok; ++j) {
    double s = L[j * 12 + j];
    for (int k = 0; k < j; ++k) s -= L[j * 12 + k] * L[j * 12 + k];
    if (!(s > 0.0) || !isfinite(s)) { ok = false; break; }
    s = sqrt(s);
    L[j * 12 + j] = s;
    const double inv = 1.0 / s;
    for (int i = j + 1; i < d; ++i) {
      double t = L[i * 12 + j];
      for (int k = 0; k < j; ++k) t -= L[i * 12 + k] * L[j * 12 + k];
      L[i * 12 + j] = t * inv;
    }
  }
  double* out = Minv + blk_moff[b];
  if (!ok) { sc->solver_failed = 1; for (int i = 0; i < d * d; ++i) out[i] = 0.0; return; }
  for (int c = 0; c < d; ++c) {
    double e[12];
    for (int i = 0; i < d; ++i) {
      double s = (i == c) ? 1.0 : 0.0;
      for (int k = 0; k < i; ++k) s -= L[i * 12 + k] * e[k];
      e[i] = s / L[i * 12 + i];
    }
    for (int i = d - 1; i >= 0; --i) {
      double s = e[i];
      for (int k = i + 1; k < d; ++k) s -= L[k * 12 + i] * e[k];
      e[i] = s / L[i * 12 + i];
    }
    for (int i = 0; i < d; ++i) out[i * d + c] = e[i];
  }
}

// ---- host: pair lists --------------------------------------------------------------------
// dense == true : every pair of observations of one point in different images (key = image pair,
//                 larger image first) plus duplicate observations inside one image.
// dense == false: only what SCHUR_JACOBI needs beyond the per-observation diagonal terms:
//                 duplicates inside one image, and -- when intrinsics are shared -- all pairs
//                 whose two images use the same camera (camera x camera block only).
int BuildPairLists(Problem* p, bool dense) {
  if (dense ? p->pairs_built : p->diag_built) return SBA_OK;
  const int n_pt = p->n_pt;
  std::vector<int> pt_ptr(n_pt + 1, 0);
  for (int k = 0; k < p->n_obs; ++k) pt_ptr[p->h_obs_pt[k] + 1]++;
  for (int q = 0; q < n_pt; ++q) pt_ptr[q + 1] += pt_ptr[q];
  const std::vector<int>& img = p->h_obs_img;
  auto wanted = [&](int a, int b) -> bool {  // a != b, ordered so that img[a] >= img[b]
    if (p->h_p_off[p->h_obs_pt[a]] < 0) return false;   // constant point: E = 0
    if (img[a] == img[b]) return true;
    if (dense) return true;
    if (!p->shared_intrinsics) return false;
    const int ca = p->h_img_cam[img[a]], cb = p->h_img_cam[img[b]];
    return ca == cb && p->h_k_dim[ca] > 0;
  };
  struct Ent { int64_t key; int a, b; };
  std::vector<Ent> ents;
  for (int q = 0; q < n_pt; ++q) {
    for (int s = pt_ptr[q]; s < pt_ptr[q + 1]; ++s)
      for (int t = pt_ptr[q]; t < pt_ptr[q + 1]; ++t) {
        if (s == t) continue;
        int a = s, b = t;
        if (img[a] < img[b]) continue;                 // the (b,a) visit handles it
        if (img[a] == img[b]) { /* ordered both ways */ }
        if (!wanted(a, b)) continue;
        ents.push_back({(int64_t)img[a] * p->n_img + img[b], a, b});
      }
  }
  std::stable_sort(ents.begin(), ents.end(), [](const Ent& x, const Ent& y) { return x.key < y.key; });
  std::vector<int> key_ptr, key_gi, key_gj, pa(ents.size()), pb(ents.size());
  for (size_t e = 0; e < ents.size(); ++e) {
    if (e == 0 || ents[e].key != ents[e - 1].key) {
      key_ptr.push_back((int)e);
      key_gi.push_back((int)(ents[e].key / p->n_img)); key_gj.push_back((int)(ents[e].key % p->n_img));
    }
    pa[e] = ents[e].a; pb[e] = ents[e].b;
  }
  key_ptr.push_back((int)ents.size());
  cudaStream_t st = p->stream;
  if (dense) {
    SBA_CUDA(p->d_pair_key_ptr.Upload(key_ptr, st)); SBA_CUDA(p->d_pair_key_gi.Upload(key_gi, st));
    SBA_CUDA(p->d_pair_key_gj.Upload(key_gj, st)); SBA_CUDA(p->d_pair_a.Upload(pa, st)); SBA_CUDA(p->d_pair_b.Upload(pb, st));
    p->n_pair_keys = (int)key_gi.size(); p->n_pairs = (int64_t)ents.size(); p->pairs_built = true;
  } else {
    SBA_CUDA(p->d_diag_key_ptr.Upload(key_ptr, st)); SBA_CUDA(p->d_diag_key_gi.Upload(key_gi, st));
    SBA_CUDA(p->d_diag_key_gj.Upload(key_gj, st)); SBA_CUDA(p->d_diag_a.Upload(pa, st)); SBA_CUDA(p->d_diag_b.Upload(pb, st));
    p->n_diag_keys = (int)key_gi.size(); p->n_diag_pairs = (int64_t)ents.size(); p->diag_built = true;
  }
  SBA_CUDA(cudaStreamSynchronize(st));
  return SBA_OK;
}

static int EnsureBlockStorage(Problem* p) {
  if (p->d_blk_moff.p) return SBA_OK;
  std::vector<int> moff(p->n_blk + 1, 0);
  for (int b = 0; b < p->n_blk; ++b) moff[b + 1] = moff[b] + p->h_blk_dim[b] * p->h_blk_dim[b];
  p->m_size = moff[p->n_blk];
  SBA_CUDA(p->d_blk_moff.Upload(moff, p->stream));
  SBA_CUDA(p->d_M.Alloc(std::max(1, p->m_size))); SBA_CUDA(p->d_Minv.Alloc(std::max(1, p->m_size)));
  SBA_CUDA(cudaStreamSynchronize(p->stream));
  return SBA_OK;
}

static int DiagChunks(Problem* p, const Sink& sink) {
  if (p->n_chunks == 0) return SBA_OK;
  SBA_CUDA(p->d_chunk_block.Alloc((size_t)p->n_chunks * p->dc * p->dc));
#define CALL(DC)                                                                                                     \
  do {                                                                                                               \
    schur_diag_chunk_kernel<DC><<<p->n_chunks, 256, 0, p->stream>>>(p->Struct(), p->Chunks(), p->d_ER.p, p->d_F.p,    \
                                                                    p->d_Cinv.p, p->d_chunk_block.p);                 \
    schur_diag_finalize_kernel<DC><<<p->n_img, 128, 0, p->stream>>>(p->Struct(), p->Chunks(), p->d_chunk_block.p, sink); \
  } while (0)
  SBA_DC_SWITCH(p->dc, CALL)
#undef CALL
  p->launches += 2;
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

int BuildDenseSchur(Problem* p) {
  int rc = BuildPairLists(p, true);
  if (rc) return rc;
  const size_t n = p->n_cdim;
  SBA_CUDA(p->d_S.Alloc(std::max<size_t>(1, (n + 1) * n)));  // + one row: the right-hand side (cholesky.cu)
  SBA_CUDA(cudaMemsetAsync(p->d_S.p, 0, n * n * sizeof(double), p->stream));
  Sink sink;
  std::memset(&sink, 0, sizeof(sink));
  sink.kind = 0; sink.S = p->d_S.p; sink.n = (int)n; sink.shared_intr = p->shared_intrinsics ? 1 : 0;
  rc = DiagChunks(p, sink);
  if (rc) return rc;
  if (p->n_pair_keys > 0) {
#define CALL(DC) schur_pair_kernel<DC><<<DivUp(p->n_pair_keys, 4), 128, 0, p->stream>>>(p->Struct(), p->n_pair_keys, p->d_pair_key_ptr.p, p->d_pair_key_gi.p, p->d_pair_key_gj.p, p->d_pair_a.p, p->d_pair_b.p, p->d_ER.p, p->d_F.p, p->d_Cinv.p, sink, 0)
    SBA_DC_SWITCH(p->dc, CALL)
#undef CALL
    p->launches++;
  }
  rc = AllReduceSum(p, p->d_S.p, n * n);
  if (rc) return rc;
  if (n > 0) { add_diag_dense_kernel<<<DivUp(n, 256), 256, 0, p->stream>>>((int)n, p->d_S.p, p->d_Dc.p); p->launches++; }
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

int BuildSchurJacobi(Problem* p) {
  int rc = BuildPairLists(p, false);
  if (rc) return rc;
  rc = EnsureBlockStorage(p);
  if (rc) return rc;
  if (p->n_blk == 0) return SBA_OK;
  SBA_CUDA(cudaMemsetAsync(p->d_M.p, 0, (size_t)std::max(1, p->m_size) * sizeof(double), p->stream));
  Sink sink;
  std::memset(&sink, 0, sizeof(sink));
  sink.kind = 1; sink.M = p->d_M.p; sink.col_blk = p->d_col_blk.p; sink.blk_off = p->d_blk_off.p;
  sink.blk_dim = p->d_blk_dim.p; sink.blk_moff = p->d_blk_moff.p; sink.shared_intr = p->shared_intrinsics ? 1 : 0;
  rc = DiagChunks(p, sink);
  if (rc) return rc;
  if (p->n_diag_keys > 0) {
    // keys of one image: all elements; keys of two images sharing a camera: camera x camera elements only
#define CALL(DC) schur_pair_kernel<DC><<<DivUp(p->n_diag_keys, 4), 128, 0, p->stream>>>(p->Struct(), p->n_diag_keys, p->d_diag_key_ptr.p, p->d_diag_key_gi.p, p->d_diag_key_gj.p, p->d_diag_a.p, p->d_diag_b.p, p->d_ER.p, p->d_F.p, p->d_Cinv.p, sink, 0)
    SBA_DC_SWITCH(p->dc, CALL)
#undef CALL
    p->launches++;
  }
  rc = AllReduceSum(p, p->d_M.p, p->m_size);
  if (rc) return rc;
  invert_blocks_kernel<<<DivUp(p->n_blk, 64), 64, 0, p->stream>>>(p->n_blk, p->d_blk_off.p, p->d_blk_dim.p, p->d_blk_moff.p,
                                                                 p->d_M.p, p->d_Dc.p, p->d_Minv.p, p->d_scalars.p);
  p->launches++;
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

}  // namespace sba
