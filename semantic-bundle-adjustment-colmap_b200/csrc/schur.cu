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
//   * per observation a (one pass): W_a = F_a^T E_a (DC x 3) and H_a = W_a C_p^{-1}.
//   * "diagonal" contributions F_a^T F_a - H_a W_a^T: image-major chunks of <= 256
//     observations; lanes own the DC x DC output elements; chunk partials are summed per
//     image in a fixed order.
//   * "off-diagonal" contributions -H_a W_b^T: host-built list of observation pairs (a,b)
//     of one point with image(a) > image(b), counting-sorted by the (image(a), image(b)) key
//     and cut into segments of <= kSegPairs pairs.  One warp reduces one segment in registers
//     (a warp-level segmented reduction keyed by camera pair); the segment partials of a key
//     are then summed in a fixed order by the key's single writer.  Bit-reproducible.
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

constexpr int kSegPairs = 32;  // one warp-load of pair ids per segment (schur_pair_segment_kernel)
template <int DC> struct Elems { static constexpr int N = (DC * DC + 31) / 32; };

// W_a = F_a^T E_a, H_a = W_a C^{-1}; one thread per (observation, slot)
template <int DC>
__global__ void __launch_bounds__(256) schur_wh_kernel(StructView S, const D4* __restrict__ ER,
                                                       const D4* __restrict__ F, const double* __restrict__ Cinv,
                                                       double* __restrict__ W, double* __restrict__ H) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= (int64_t)S.n_obs * DC) return;
  const int a = (int)(t / DC);
  const double2 f = LoadSlot(F, a, (int)(t - (int64_t)a * DC), DC / 2);
  const D4 ea = Ld256(ER + EUnit(a, 0));                                          // E0 E1 E2 E3
  const double2 eb = reinterpret_cast<const double2*>(ER + EUnit(a, 1))[0];       // E4 E5
  const double w[3] = {f.x * ea.x + f.y * ea.w, f.x * ea.y + f.y * eb.x, f.x * ea.z + f.y * eb.y};
  double h[3] = {0.0, 0.0, 0.0};
  const int po = S.p_off[S.obs_pt[a]];
  if (po >= 0) Sym3Mul(Cinv + 6 * (size_t)po, w, h);
  W[3 * t] = w[0]; W[3 * t + 1] = w[1]; W[3 * t + 2] = w[2];
  H[3 * t] = h[0]; H[3 * t + 1] = h[1]; H[3 * t + 2] = h[2];
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
__global__ void __launch_bounds__(256) schur_diag_chunk_kernel(ChunkView C, const D4* __restrict__ F,
                                                               const double* __restrict__ W, const double* __restrict__ H,
                                                               double* __restrict__ partial) {
  __shared__ double red[8][DC * DC];
  const int ch = blockIdx.x, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int s0 = C.chunk_start[ch], n = C.chunk_start[ch + 1] - s0;
  double acc[Elems<DC>::N];
  int ja[Elems<DC>::N], jb[Elems<DC>::N];
#pragma unroll
  for (int q = 0; q < Elems<DC>::N; ++q) {
    acc[q] = 0.0;
    const int e = min(lane + 32 * q, DC * DC - 1);
    ja[q] = e / DC; jb[q] = e - ja[q] * DC;
  }
  const int lo = w * 32, hi = min(n, lo + 32);
  for (int t = lo; t < hi; ++t) {
    const int obs = C.cam_order[s0 + t];
    const size_t a = (size_t)obs * DC;
#pragma unroll
    for (int q = 0; q < Elems<DC>::N; ++q) {
      const double2 fa = LoadSlot(F, obs, ja[q], DC / 2), fb = LoadSlot(F, obs, jb[q], DC / 2);
      const double* h = H + 3 * (a + ja[q]);
      const double* wv = W + 3 * (a + jb[q]);
      acc[q] += (fa.x * fb.x + fa.y * fb.y) - (h[0] * wv[0] + h[1] * wv[1] + h[2] * wv[2]);
    }
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

// ---- off-diagonal contributions ---------------------------------------------------------
// one warp per segment (<= kSegPairs pairs of one key): partial[seg] = - sum H_a W_b^T
template <int DC>
__global__ void __launch_bounds__(128) schur_pair_segment_kernel(int n_seg, const int* __restrict__ seg_ptr,
                                                                 const int* __restrict__ pa, const int* __restrict__ pb,
                                                                 const double* __restrict__ W, const double* __restrict__ H,
                                                                 double* __restrict__ partial) {
  const int seg = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (seg >= n_seg) return;
  double acc[Elems<DC>::N];
  int ja[Elems<DC>::N], jb[Elems<DC>::N];
#pragma unroll
  for (int q = 0; q < Elems<DC>::N; ++q) {
    acc[q] = 0.0;
    const int e = min(lane + 32 * q, DC * DC - 1);
    ja[q] = e / DC; jb[q] = e - ja[q] * DC;
  }
  const int t0 = seg_ptr[seg], t1 = seg_ptr[seg + 1];
  // a segment holds <= kSegPairs (= 32) pairs: one coalesced index load per lane, then the pair ids are
  // broadcast by shuffle, so the operand loads of several pairs are in flight together instead of
  // each waiting for its own index load.
  const int np = t1 - t0;
  const int my_a = lane < np ? pa[t0 + lane] : 0, my_b = lane < np ? pb[t0 + lane] : 0;
#pragma unroll 4
  for (int t = 0; t < np; ++t) {
    const size_t a = (size_t)__shfl_sync(0xffffffffu, my_a, t) * DC, b = (size_t)__shfl_sync(0xffffffffu, my_b, t) * DC;
#pragma unroll
    for (int q = 0; q < Elems<DC>::N; ++q) {
      const double* h = H + 3 * (a + ja[q]);
      const double* wv = W + 3 * (b + jb[q]);
      acc[q] -= h[0] * wv[0] + h[1] * wv[1] + h[2] * wv[2];
    }
  }
#pragma unroll
  for (int q = 0; q < Elems<DC>::N; ++q) {
    const int e = lane + 32 * q;
    if (e < DC * DC) partial[(size_t)seg * DC * DC + e] = acc[q];
  }
}

// one warp per key: fixed-order sum of its segment partials, then the single write per entry
template <int DC>
__global__ void __launch_bounds__(128) schur_pair_emit_kernel(StructView S, int n_keys, const int* __restrict__ key_seg_ptr,
                                                              const int* __restrict__ key_gi, const int* __restrict__ key_gj,
                                                              const double* __restrict__ partial, Sink sink) {
  const int key = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (key >= n_keys) return;
  const int s0 = key_seg_ptr[key], s1 = key_seg_ptr[key + 1];
  const int gi = key_gi[key], gj = key_gj[key];
  const int ci = S.img_cam[gi], cj = S.img_cam[gj];
  const bool same = gi == gj;
#pragma unroll
  for (int q = 0; q < Elems<DC>::N; ++q) {
    const int e = lane + 32 * q;
    if (e >= DC * DC) continue;
    double a = 0.0;
    for (int sg = s0; sg < s1; ++sg) a += partial[(size_t)sg * DC * DC + e];
    const int ja = e / DC, jb = e - ja * DC;
    // duplicate-observation keys (same image) add on top of the chunk result -> atomics
    const bool atomic = same || (sink.shared_intr && (ja >= 6 || jb >= 6));
    Emit(sink, SlotCol(S, gi, ci, ja), SlotCol(S, gj, cj, jb), a, same, atomic);
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
// Pairs are counting-sorted by key (stable in point order) and cut into <= kSegPairs segments.
int BuildPairLists(Problem* p, bool dense) {
  PairList& L = dense ? p->pairs_dense : p->pairs_diag;
  if (L.built) return SBA_OK;
  HostTimer tm("pair lists");
  const int n_pt = p->n_pt, n_img = p->n_img;
  const std::vector<int>& img = p->h_obs_img;
  const std::vector<int>& pt_ptr = p->h_pt_ptr;
  // which points can contribute at all
  bool any = dense;
  std::vector<uint8_t> has_dup(n_pt, 0);
  {
    std::vector<int> stamp(std::max(1, n_img), -1);
    for (int q = 0; q < n_pt; ++q) {
      if (p->h_p_off[q] < 0) continue;
      for (int s = pt_ptr[q]; s < pt_ptr[q + 1]; ++s) {
        if (stamp[img[s]] == q) { has_dup[q] = 1; any = true; }
        stamp[img[s]] = q;
      }
    }
  }
  const bool shared = !dense && p->shared_intrinsics;
  if (shared) any = true;
  tm.Lap("duplicates");
  // The key (image(a), image(b)) is owned by image(a): every image enumerates, over its own observations in
  // image-major order, the partner observations of the same point and groups them by image(b) with a local
  // counting sort.  Images are independent, so the enumeration runs on all host threads; concatenating the
  // per-image results in image order IS the key-sorted list, and the order inside a key (point order, then
  // partner order) does not depend on the thread count -- the reductions stay bit-reproducible.
  struct Part { std::vector<int> a, b, key_j, key_n; };
  std::vector<Part> parts(std::max(1, n_img));
  if (any && p->n_obs > 0) {
    ParallelFor(n_img, [&](int i) {
      Part& P = parts[i];
      const int s0 = p->h_img_ptr[i], s1 = p->h_img_ptr[i + 1];
      if (s0 == s1) return;
      const int ci = p->h_img_cam[i];
      std::vector<int> ta, tb, tj;
      for (int pos = s0; pos < s1; ++pos) {
        const int s = p->h_cam_order[pos];
        const int q = p->h_obs_pt[s];
        if (p->h_p_off[q] < 0) continue;               // constant point: E = 0
        if (!dense && !shared && !has_dup[q]) continue;
        for (int t = pt_ptr[q]; t < pt_ptr[q + 1]; ++t) {
          const int j = img[t];
          if (t == s || j > i) continue;                // the partner's own visit handles image(b) > image(a)
          if (j != i && !dense) {
            if (!shared) continue;
            if (p->h_img_cam[j] != ci || p->h_k_dim[ci] <= 0) continue;
          }
          ta.push_back(s); tb.push_back(t); tj.push_back(j);
        }
      }
      if (ta.empty()) return;
      std::vector<int> cnt(i + 2, 0);                   // partner images are <= i
      for (int j : tj) cnt[j + 1]++;
      for (int j = 0; j <= i; ++j) {
        if (cnt[j + 1] > 0) { P.key_j.push_back(j); P.key_n.push_back(cnt[j + 1]); }
        cnt[j + 1] += cnt[j];
      }
      P.a.resize(ta.size()); P.b.resize(ta.size());
      for (size_t e = 0; e < ta.size(); ++e) { const int pos = cnt[tj[e]]++; P.a[pos] = ta[e]; P.b[pos] = tb[e]; }
    });
  }
  tm.Lap("enumerate + sort");
  size_t n_pairs = 0, n_keys = 0;
  std::vector<size_t> pair_off(n_img + 1, 0), key_off(n_img + 1, 0);
  for (int i = 0; i < n_img; ++i) {
    pair_off[i] = n_pairs; key_off[i] = n_keys;
    n_pairs += parts[i].a.size(); n_keys += parts[i].key_j.size();
  }
  pair_off[n_img] = n_pairs; key_off[n_img] = n_keys;
  std::vector<int> pa(n_pairs), pb(n_pairs), key_ptr(n_keys + 1), key_gi(n_keys), key_gj(n_keys);
  ParallelFor(n_img, [&](int i) {
    const Part& P = parts[i];
    if (P.a.empty()) return;
    std::copy(P.a.begin(), P.a.end(), pa.begin() + pair_off[i]);
    std::copy(P.b.begin(), P.b.end(), pb.begin() + pair_off[i]);
    size_t at = pair_off[i];
    for (size_t k = 0; k < P.key_j.size(); ++k) {
      key_ptr[key_off[i] + k] = (int)at; key_gi[key_off[i] + k] = i; key_gj[key_off[i] + k] = P.key_j[k];
      at += P.key_n[k];
    }
  });
  key_ptr[n_keys] = (int)n_pairs;
  // segments
  std::vector<int> seg_ptr, key_seg_ptr;
  seg_ptr.reserve(n_pairs / kSegPairs + n_keys + 1); key_seg_ptr.reserve(n_keys + 1);
  for (size_t k = 0; k < n_keys; ++k) {
    key_seg_ptr.push_back((int)seg_ptr.size());
    for (int s = key_ptr[k]; s < key_ptr[k + 1]; s += kSegPairs) seg_ptr.push_back(s);
  }
  key_seg_ptr.push_back((int)seg_ptr.size());
  L.n_seg = (int)seg_ptr.size();
  seg_ptr.push_back((int)n_pairs);
  L.n_keys = (int)n_keys;
  L.n_pairs = (int64_t)n_pairs;
  tm.Lap("merge + segments");
  cudaStream_t st = p->stream;
  SBA_CUDA(L.seg_ptr.Upload(seg_ptr, st)); SBA_CUDA(L.key_seg_ptr.Upload(key_seg_ptr, st));
  SBA_CUDA(L.key_gi.Upload(key_gi, st)); SBA_CUDA(L.key_gj.Upload(key_gj, st));
  SBA_CUDA(L.a.Upload(pa, st)); SBA_CUDA(L.b.Upload(pb, st));
  SBA_CUDA(L.partial.Alloc((size_t)std::max(1, L.n_seg) * p->dc * p->dc));
  SBA_CUDA(cudaStreamSynchronize(st));
  L.built = true;
  tm.Lap("upload");
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

// W/H pass + diagonal chunks + pair segments into `sink`
static int Accumulate(Problem* p, const PairList& L, const Sink& sink) {
  const int dc = p->dc;
  cudaStream_t st = p->stream;
  if (p->n_obs == 0) return SBA_OK;
  SBA_CUDA(p->d_W.Alloc((size_t)p->n_obs * dc * 3)); SBA_CUDA(p->d_H.Alloc((size_t)p->n_obs * dc * 3));
  SBA_CUDA(p->d_chunk_block.Alloc((size_t)std::max(1, p->n_chunks) * dc * dc));
#define CALL(DC)                                                                                                          \
  do {                                                                                                                    \
    schur_wh_kernel<DC><<<DivUp((int64_t)p->n_obs * DC, 256), 256, 0, st>>>(p->Struct(), p->d_ER.p, p->d_F.p, p->d_Cinv.p, \
                                                                          p->d_W.p, p->d_H.p);                            \
    schur_diag_chunk_kernel<DC><<<p->n_chunks, 256, 0, st>>>(p->Chunks(), p->d_F.p, p->d_W.p, p->d_H.p, p->d_chunk_block.p); \
    schur_diag_finalize_kernel<DC><<<p->n_img, 128, 0, st>>>(p->Struct(), p->Chunks(), p->d_chunk_block.p, sink);          \
    p->launches += 3;                                                                                                     \
    if (L.n_seg > 0) {                                                                                                    \
      schur_pair_segment_kernel<DC><<<DivUp(L.n_seg, 4), 128, 0, st>>>(L.n_seg, L.seg_ptr.p, L.a.p, L.b.p, p->d_W.p,       \
                                                                      p->d_H.p, L.partial.p);                             \
      schur_pair_emit_kernel<DC><<<DivUp(L.n_keys, 4), 128, 0, st>>>(p->Struct(), L.n_keys, L.key_seg_ptr.p, L.key_gi.p,   \
                                                                    L.key_gj.p, L.partial.p, sink);                       \
      p->launches += 2;                                                                                                   \
    }                                                                                                                     \
  } while (0)
  SBA_DC_SWITCH(dc, CALL)
#undef CALL
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
  rc = Accumulate(p, p->pairs_dense, sink);
  if (rc) return rc;
  rc = AllReduceSum(p, p->d_S.p, n * n);
  if (rc) return rc;
  p->chol_add_diag = (int)n <= kCholSmallMax;  // the small-system Cholesky adds D^2 while loading its diagonal tiles
  if (n > 0 && !p->chol_add_diag) { add_diag_dense_kernel<<<DivUp(n, 256), 256, 0, p->stream>>>((int)n, p->d_S.p, p->d_Dc.p); p->launches++; }
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
  rc = Accumulate(p, p->pairs_diag, sink);
  if (rc) return rc;
  rc = AllReduceSum(p, p->d_M.p, p->m_size);
  if (rc) return rc;
  invert_blocks_kernel<<<DivUp(p->n_blk, 64), 64, 0, p->stream>>>(p->n_blk, p->d_blk_off.p, p->d_blk_dim.p, p->d_blk_moff.p,
                                                                 p->d_M.p, p->d_Dc.p, p->d_Minv.p, p->d_scalars.p);
  p->launches++;
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

}  // namespace sba
