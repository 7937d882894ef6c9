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
//   * "off-diagonal" contributions -H_a W_b^T: device-built list (pairs.cu) of observation pairs (a,b)
//     of one point with image(a) > image(b), counting-sorted by the (image(a), image(b)) key
//     and cut into segments of <= PairSegmentSize(dc) pairs (pairs.cu).  One warp reduces one segment in registers
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

template <int DC> struct Elems { static constexpr int N = (DC * DC + 31) / 32; };

// W_a = F_a^T E_a, H_a = W_a C^{-1}; one thread per (image-major position, slot) -- the rows of W / H are stored in
// image-major order so that a chunk's rows are contiguous
template <int DC, int NR>
__global__ void __launch_bounds__(256) schur_wh_kernel(StructView S, const int* __restrict__ cam_order, const D4* __restrict__ ER,
                                                       const D4* __restrict__ F, const D4* __restrict__ ES,
                                                       const D4* __restrict__ FS, const double* __restrict__ Cinv,
                                                       double* __restrict__ W, double* __restrict__ H) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= (int64_t)S.n_obs * DC) return;
  const int pos = (int)(t / DC);
  const int a = cam_order[pos];
  const int slot = (int)(t - (int64_t)pos * DC);
  const double2 f = LoadSlot(F, a, slot, DC / 2);
  const D4 ea = Ld256(ER + EUnit(a, 0));                                          // E0 E1 E2 E3
  const double2 eb = reinterpret_cast<const double2*>(ER + EUnit(a, 1))[0];       // E4 E5
  double w[3] = {f.x * ea.x + f.y * ea.w, f.x * ea.y + f.y * eb.x, f.x * ea.z + f.y * eb.y};
  if constexpr (NR == 3) {
    const double fs = LoadSlotS(FS, a, slot, FSUnits(DC));
    const D4 es = Ld256(ES + a);
    w[0] += fs * es.x; w[1] += fs * es.y; w[2] += fs * es.z;
  }
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
// kind 1: SCHUR_JACOBI block-diagonal storage; kind 2: block-sparse S in SLOT space (bsr.cu): block i < n_img is
// the full diagonal block of image i, block n_img + key the (gi > gj) block of a camera-pair key.
struct Sink {
  int kind;
  double* S; int n;                 // dense
  double* M; const int* col_blk; const int* blk_off; const int* blk_dim; const int* blk_moff;  // block diagonal
  int shared_intr;                  // intrinsics columns have several writers -> atomics
  double* B; int n_img; const int* key_block;  // block-sparse: key -> block index (null: n_img + key)
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
template <int DC, int NR>
__global__ void __launch_bounds__(256) schur_diag_chunk_kernel(ChunkView C, const D4* __restrict__ F, const D4* __restrict__ FS,
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
    const size_t a = (size_t)(s0 + t) * DC;  // row of W / H
#pragma unroll
    for (int q = 0; q < Elems<DC>::N; ++q) {
      const double2 fa = LoadSlot(F, obs, ja[q], DC / 2), fb = LoadSlot(F, obs, jb[q], DC / 2);
      const double* h = H + 3 * (a + ja[q]);
      const double* wv = W + 3 * (a + jb[q]);
      acc[q] += (fa.x * fb.x + fa.y * fb.y) - (h[0] * wv[0] + h[1] * wv[1] + h[2] * wv[2]);
      if constexpr (NR == 3) acc[q] += LoadSlotS(FS, obs, ja[q], FSUnits(DC)) * LoadSlotS(FS, obs, jb[q], FSUnits(DC));
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
    if (sink.kind == 2) { sink.B[(size_t)i * DC * DC + e] = a; continue; }
    const int ja = e / DC, jb = e - ja * DC;
    const bool atomic = sink.shared_intr && (ja >= 6 || jb >= 6);
    Emit(sink, SlotCol(S, i, c, ja), SlotCol(S, i, c, jb), a, true, atomic);
  }
}

// ---- off-diagonal contributions ---------------------------------------------------------
// one warp per segment (<= 32 pairs of one key): partial[seg] = - sum H_a W_b^T
template <int DC>
__global__ void __launch_bounds__(128) schur_pair_segment_kernel(int n_seg, const int* __restrict__ seg_ptr,
                                                                 const unsigned long long* __restrict__ ab,
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
  // a segment holds <= 32 pairs (PairSegmentSize, dc > 8): one coalesced index load per lane, then the pair ids are
  // broadcast by shuffle, so the operand loads of several pairs are in flight together instead of
  // each waiting for its own index load.
  const int np = t1 - t0;
  const unsigned long long my_ab = lane < np ? ab[t0 + lane] : 0ull;
  const int my_a = (int)(my_ab >> 32), my_b = (int)(my_ab & 0xffffffffull);
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
    if (sink.kind == 2) {  // single writer per entry: the diagonal blocks were written by the finalize kernel before
      if (same) sink.B[(size_t)gi * DC * DC + e] += a;
      else sink.B[(size_t)(sink.key_block ? sink.key_block[key] : sink.n_img + key) * DC * DC + e] = a;
      continue;
    }
    const int ja = e / DC, jb = e - ja * DC;
    // duplicate-observation keys (same image) add on top of the chunk result -> atomics
    const bool atomic = same || (sink.shared_intr && (ja >= 6 || jb >= 6));
    Emit(sink, SlotCol(S, gi, ci, ja), SlotCol(S, gj, cj, jb), a, same, atomic);
  }
}

// ---- lane-per-unit kernels (DC <= 8: const intrinsics / f,k free -- the C1..C3 regimes) --------------------------
// The kernels above spread the DC x DC outputs of ONE observation / pair over the lanes of a warp; with DC = 6 that
// keeps 36 of 64 lane-slots busy and spends most instructions on addressing.  Here a lane owns a whole observation
// (or pair): it loads its records with wide accesses, forms the full block in registers, and the warp pays ONE
// transposing reduction through shared memory per 32 observations (per <= 128 pairs).  Fixed summation order.

// doubles per W / H row: the lane kernels (DC <= 8) pad a row to a multiple of 32 B, so that a row starts on a sector
// boundary (DC = 6: 144 -> 160 B = exactly five sectors per gather instead of five or six) and is written with
// 256-bit stores; the generic kernels keep DC * 3.
__host__ __device__ constexpr int WRow(int dc) { return dc <= 8 ? (dc * 3 + 3) / 4 * 4 : dc * 3; }

template <int DC> struct ObsLane {
  static constexpr int NT = DC * (DC + 1) / 2;
  static constexpr size_t kSmem = (size_t)(NT * 256 + NT * 8) * 8;
};

// DC <= 8.  ONE row per observation instead of two: with C_p^{-1} = M^T M (M = L^{-1}, stored by point_blocks_kernel)
//   U_a = (F_a^T E_a) M^T       (DC x 3)
// gives H_a W_b^T = U_a U_b^T, so the pair pass gathers two rows of ONE array (half the bytes written here, half the
// footprint the pair gathers have to find in L2) and the diagonal contribution is F_a^T F_a - U_a U_a^T.
// partial[ch][e], e over the lower triangle (ja >= jb, e = ja (ja + 1) / 2 + jb).
// A lane owns an observation and writes its own row (image-major, sector-aligned, 256-bit stores).
// rhs_partial != null: the same pass also forms this chunk's part of the reduced right-hand side
//   F_a^T (r_a - E_a C_p^{-1} g_p)
// -- r rides in the ER record the lane loads anyway, g_p is one more 24 B gather -- which replaces rhs_e_kernel and a
// cam_reduce pass over F (0.27 ms per LM iteration at C4).
template <int DC, int NR>
__global__ void __launch_bounds__(256, 2) schur_obs_u_kernel(StructView S, ChunkView C, const D4* __restrict__ ER,
                                                          const D4* __restrict__ F, const D4* __restrict__ ES,
                                                          const D4* __restrict__ FS, const double* __restrict__ Cfac,
                                                          const double* __restrict__ gp, double* __restrict__ U,
                                                          double* __restrict__ partial, double* __restrict__ rhs_partial) {
  constexpr int NT = DC * (DC + 1) / 2, DH = DC / 2, RW = WRow(DC);
  extern __shared__ __align__(16) double sm[];  // [NT][256] blocks + [NT][8] partials
  double* part = sm + NT * 256;
  const int ch = blockIdx.x, t = threadIdx.x;
  const int s0 = C.chunk_start[ch], n = C.chunk_start[ch + 1] - s0;
  double rhs[DC];
#pragma unroll
  for (int j = 0; j < DC; ++j) rhs[j] = 0.0;
  if (t < n) {
    const int po = C.cam_po[s0 + t];
    const int a = C.cam_order[s0 + t];
    double2 f[DC];
#pragma unroll
    for (int jj = 0; jj < DH; ++jj) {
      const D4 u = Ld256(F + FUnit(a, jj, DH));
      f[2 * jj] = make_double2(u.x, u.y); f[2 * jj + 1] = make_double2(u.z, u.w);
    }
    const D4 ea = Ld256(ER + EUnit(a, 0)), eb = Ld256(ER + EUnit(a, 1));
    const double E[6] = {ea.x, ea.y, ea.z, ea.w, eb.x, eb.y};
    double fs[NR == 3 ? DC : 1], Es[3] = {0.0, 0.0, 0.0}, rs = 0.0;  // third row (X1): F_s (DC), E_s (3), r_s
    if constexpr (NR == 3) {
      constexpr int NQ = FSUnits(DC);
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        const D4 u = Ld256(FS + FSUnit(a, q, NQ));
        fs[4 * q] = u.x;
        if (4 * q + 1 < DC) fs[4 * q + 1 < DC ? 4 * q + 1 : 0] = u.y;
        if (4 * q + 2 < DC) fs[4 * q + 2 < DC ? 4 * q + 2 : 0] = u.z;
        if (4 * q + 3 < DC) fs[4 * q + 3 < DC ? 4 * q + 3 : 0] = u.w;
      }
      const D4 es = Ld256(ES + a);
      Es[0] = es.x; Es[1] = es.y; Es[2] = es.z; rs = es.w;
    }
    double m[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};  // m00 m10 m11 m20 m21 m22
    if (po >= 0) {
      const double2* cp = reinterpret_cast<const double2*>(Cfac + 6 * (size_t)po);
      const double2 c0 = cp[0], c1 = cp[1], c2 = cp[2];
      m[0] = c0.x; m[1] = c0.y; m[2] = c1.x; m[3] = c1.y; m[4] = c2.x; m[5] = c2.y;
    }
    if (rhs_partial) {
      double e0 = eb.z, e1 = eb.w, e2 = rs;
      if (po >= 0) {
        const double g0 = gp[3 * (size_t)po], g1 = gp[3 * (size_t)po + 1], g2 = gp[3 * (size_t)po + 2];
        const double v0 = m[0] * g0, v1 = m[1] * g0 + m[2] * g1, v2 = m[3] * g0 + m[4] * g1 + m[5] * g2;  // M g
        const double w0 = m[0] * v0 + m[1] * v1 + m[3] * v2, w1 = m[2] * v1 + m[4] * v2, w2 = m[5] * v2;  // M^T M g
        e0 -= E[0] * w0 + E[1] * w1 + E[2] * w2;
        e1 -= E[3] * w0 + E[4] * w1 + E[5] * w2;
        if constexpr (NR == 3) e2 -= Es[0] * w0 + Es[1] * w1 + Es[2] * w2;
      }
#pragma unroll
      for (int j = 0; j < DC; ++j) {
        rhs[j] = f[j].x * e0 + f[j].y * e1;
        if constexpr (NR == 3) rhs[j] += fs[j] * e2;
      }
    }
    D4* urow = reinterpret_cast<D4*>(U + (size_t)(s0 + t) * RW);
    double u[RW];
#pragma unroll
    for (int q = DC * 3; q < RW; ++q) u[q] = 0.0;
#pragma unroll
    for (int ja = 0; ja < DC; ++ja) {
      double w[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        w[c] = f[ja].x * E[c] + f[ja].y * E[3 + c];
        if constexpr (NR == 3) w[c] += fs[ja] * Es[c];
      }
      u[3 * ja] = w[0] * m[0];
      u[3 * ja + 1] = w[0] * m[1] + w[1] * m[2];
      u[3 * ja + 2] = w[0] * m[3] + w[1] * m[4] + w[2] * m[5];
      // the entries of slot ja are final: they go straight to the reduction area, not through registers
#pragma unroll
      for (int jb = 0; jb <= ja; ++jb)
        sm[(ja * (ja + 1) / 2 + jb) * 256 + t] = (f[ja].x * f[jb].x + f[ja].y * f[jb].y + (NR == 3 ? fs[NR == 3 ? ja : 0] * fs[NR == 3 ? jb : 0] : 0.0)) -
                                                 (u[3 * ja] * u[3 * jb] + u[3 * ja + 1] * u[3 * jb + 1] + u[3 * ja + 2] * u[3 * jb + 2]);
      // 32-byte units of the row leave as soon as they are complete
#pragma unroll
      for (int q = 0; q < RW / 4; ++q)
        if (4 * q + 3 < 3 * (ja + 1) && 4 * q + 3 >= 3 * ja) St256(urow + q, D4{u[4 * q], u[4 * q + 1], u[4 * q + 2], u[4 * q + 3]});
    }
#pragma unroll
    for (int q = 0; q < RW / 4; ++q)
      if (4 * q + 3 >= 3 * DC) St256(urow + q, D4{u[4 * q], u[4 * q + 1], u[4 * q + 2], u[4 * q + 3]});  // the unit(s) that hold the padding
  } else {
#pragma unroll
    for (int e = 0; e < NT; ++e) sm[e * 256 + t] = 0.0;
  }
  __syncthreads();
  for (int k = t; k < NT * 8; k += 256) {  // task (e, warp): 32 values; the rotation keeps the 16 lanes of a phase on 16 banks
    const int e = k >> 3, wq = k & 7;
    const double* src = sm + e * 256 + wq * 32;
    const int rot = wq + 8 * (e & 1);
    double acc = 0.0;
#pragma unroll 8
    for (int j = 0; j < 32; ++j) acc += src[(j + rot) & 31];
    part[k] = acc;
  }
  __syncthreads();
  if (t < NT) {
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += part[t * 8 + i];
    partial[(size_t)ch * NT + t] = acc;
  }
  if (rhs_partial) {  // fixed tree: warp shuffles, then the eight warp sums in order (as cam_reduce_kernel)
    __syncthreads();
    double* rsm = sm;  // [8][DC]
    const int lane = t & 31, w = t >> 5;
#pragma unroll
    for (int j = 0; j < DC; ++j) {
      double a = rhs[j];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
      if (lane == 0) rsm[w * DC + j] = a;
    }
    __syncthreads();
    if (t < DC) {
      double a = 0.0;
#pragma unroll
      for (int i = 0; i < 8; ++i) a += rsm[i * DC + t];
      rhs_partial[(size_t)ch * DC + t] = a;
    }
  }
}

template <int DC>
__global__ void __launch_bounds__(64) schur_diag_finalize_tri_kernel(StructView S, ChunkView C, const double* __restrict__ partial,
                                                                     Sink sink) {
  constexpr int NT = DC * (DC + 1) / 2;
  const int i = blockIdx.x;
  const int c = S.img_cam[i];
  const int c0 = C.img_chunk_ptr[i], c1 = C.img_chunk_ptr[i + 1];
  if (c0 == c1) return;
  for (int e = threadIdx.x; e < NT; e += 64) {
    double a = 0.0;
    for (int ch = c0; ch < c1; ++ch) a += partial[(size_t)ch * NT + e];
    int ja = 0;
    while ((ja + 1) * (ja + 2) / 2 <= e) ++ja;
    const int jb = e - ja * (ja + 1) / 2;
    if (sink.kind == 2) {
      double* blk = sink.B + (size_t)i * DC * DC;
      blk[ja * DC + jb] = a; blk[jb * DC + ja] = a;
      continue;
    }
    int row = SlotCol(S, i, c, ja), col = SlotCol(S, i, c, jb);
    if (row < col) { const int x = row; row = col; col = x; }  // the block is symmetric: store its lower image
    const bool atomic = sink.shared_intr && (ja >= 6 || jb >= 6);
    Emit(sink, row, col, a, true, atomic);
  }
}

// fp64 tensor-core MMA (DMMA): D(8x8) += A(8x4, row) * B(4x8, col)
__device__ __forceinline__ void Dmma884(double& d0, double& d1, double a, double b) {
  // (not volatile: a pure function of its operands, so that the loads of the following pairs can be scheduled above it)
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
      : "+d"(d0), "+d"(d1)
      : "d"(a), "d"(b));
}

// One warp per segment (<= 128 pairs of one key): partial[seg] = - sum_pairs U_a U_b^T, one DMMA per pair.
// U_a (DC x 3, DC <= 8) is the A operand of an m8n8k4 product as it lies in memory: lane l holds A[l / 4][l % 4] =
// U_a[slot l / 4][coordinate l % 4] (the fourth coordinate and slots >= DC are zero), and the B operand B[k][n] =
// U_b[slot n][coordinate k] sits at the SAME offset of row b -- so a pair is two 192-byte row reads (consecutive lanes
// on consecutive doubles: whole sectors, no staging through shared memory) and one tensor-core instruction; the 8 x 8
// accumulator stays in two registers per lane for the whole segment.  The pair ids are read ahead in blocks of 32
// (one coalesced load, broadcast by shuffle), so the row loads of several pairs are in flight together.
template <int DC>
__global__ void __launch_bounds__(128) schur_pair_mma_kernel(int n_seg, const int* __restrict__ seg_ptr,
                                                             const unsigned long long* __restrict__ ab,
                                                             const double* __restrict__ U, double* __restrict__ partial) {
  constexpr int RW = WRow(DC);
  const int seg = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (seg >= n_seg) return;  // warps are independent
  const int slot = lane >> 2, k = lane & 3;
  const bool live = slot < DC && k < 3;
  const int off = live ? slot * 3 + k : 0;
  const int t0 = seg_ptr[seg], np = seg_ptr[seg + 1] - t0;
  double c0 = 0.0, c1 = 0.0;
  for (int base = 0; base < np; base += 32) {
    const int m = min(32, np - base);
    const unsigned long long mine = lane < m ? ab[t0 + base + lane] : 0ull;
    const int my_a = (int)(mine >> 32), my_b = (int)(mine & 0xffffffffull);
    // batches of 8 pairs: all 16 row loads are issued before the first product consumes one (a batch's tail repeats
    // the last pair with zero operands)
    for (int t = 0; t < m; t += 8) {
      double av[8], bv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int src = min(t + u, m - 1);
        const int a = __shfl_sync(0xffffffffu, my_a, src), b = __shfl_sync(0xffffffffu, my_b, src);
        const bool on = live && t + u < m;
        av[u] = on ? __ldg(U + (size_t)a * RW + off) : 0.0;
        bv[u] = on ? __ldg(U + (size_t)b * RW + off) : 0.0;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) Dmma884(c0, c1, av[u], bv[u]);
    }
  }
  // accumulator layout: lane holds C[l / 4][2 (l % 4)], C[l / 4][2 (l % 4) + 1]
  const int row = lane >> 2, col = 2 * (lane & 3);
  if (row < DC) {
    double* out = partial + (size_t)seg * DC * DC + row * DC;
    if (col < DC) out[col] = -c0;
    if (col + 1 < DC) out[col + 1] = -c1;
  }
}

__global__ void add_diag_dense_kernel(int n, int ld, double* __restrict__ S, const double* __restrict__ Dc) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) S[(size_t)j * ld + j] += Dc[j] * Dc[j];
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

static int EnsureBlockStorage(Problem* p) {
  if (p->d_blk_moff.p) return SBA_OK;
  std::vector<int> moff(p->n_blk + 1, 0);
  for (int b = 0; b < p->n_blk; ++b) moff[b + 1] = moff[b] + p->h_blk_dim[b] * p->h_blk_dim[b];
  p->m_size = moff[p->n_blk];
  SBA_CUDA(p->d_blk_moff.Upload(moff, p->stream));
  SBA_CUDA(p->d_M.Alloc(std::max(1, p->m_size) + p->n_cdim)); SBA_CUDA(p->d_Minv.Alloc(std::max(1, p->m_size)));
  SBA_CUDA(cudaStreamSynchronize(p->stream));
  return SBA_OK;
}

// U pass (+ rhs) + diagonal chunks + pair segments into `sink`
template <int DC, int NR>
static cudaError_t AccumulateLane(Problem* p, const PairList& L, const Sink& sink, bool with_rhs) {
  constexpr size_t smem = ObsLane<DC>::kSmem;
  static bool attr_set[64] = {};
  if (!attr_set[p->device & 63]) {
    cudaError_t e = cudaFuncSetAttribute(schur_obs_u_kernel<DC, NR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_set[p->device & 63] = true;
  }
  cudaStream_t st = p->stream;
  schur_obs_u_kernel<DC, NR><<<p->n_chunks, 256, smem, st>>>(p->Struct(), p->Chunks(), p->d_ER.p, p->d_F.p, p->d_ES.p, p->d_FS.p,
                                                             p->d_Cfac.p, p->d_gp.p, p->d_W.p, p->d_chunk_block.p,
                                                             with_rhs ? p->d_chunk_partial.p : nullptr);
  schur_diag_finalize_tri_kernel<DC><<<p->n_img, 64, 0, st>>>(p->Struct(), p->Chunks(), p->d_chunk_block.p, sink);
  p->launches += 2;
  if (L.n_seg > 0) {
    schur_pair_mma_kernel<DC><<<DivUp(L.n_seg, 4), 128, 0, st>>>(L.n_seg, L.seg_ptr.p, L.ab.p, p->d_W.p, L.partial.p);
    schur_pair_emit_kernel<DC><<<DivUp(L.n_keys, 4), 128, 0, st>>>(p->Struct(), L.n_keys, L.key_seg_ptr.p, L.key_gi.p,
                                                                  L.key_gj.p, L.partial.p, sink);
    p->launches += 2;
  }
  return cudaGetLastError();
}

bool SchurBuildsRhs(const Problem* p) { return p->dc <= 8 && p->n_obs_global > 0; }

static int Accumulate(Problem* p, const PairList& L, const Sink& sink, bool with_rhs = false) {
  const int dc = p->dc;
  cudaStream_t st = p->stream;
  with_rhs = with_rhs && SchurBuildsRhs(p);
  p->rhs_from_schur = false;
  if (p->n_obs == 0) {
    if (with_rhs && p->n_cdim > 0) {  // an empty shard contributes a zero right-hand side
      SBA_CUDA(cudaMemsetAsync(p->d_rhs.p, 0, (size_t)p->n_cdim * sizeof(double), st));
      p->rhs_from_schur = true;
    }
    return SBA_OK;
  }
  SBA_CUDA(p->d_W.Alloc((size_t)p->n_obs * WRow(dc)));
  if (dc > 8) SBA_CUDA(p->d_H.Alloc((size_t)p->n_obs * WRow(dc)));
  SBA_CUDA(p->d_chunk_block.Alloc((size_t)std::max(1, p->n_chunks) * dc * dc));
  const bool r3 = p->nr == 3;
  if (dc == 6 || dc == 8) {
    cudaError_t e;
    if (dc == 6) e = r3 ? AccumulateLane<6, 3>(p, L, sink, with_rhs) : AccumulateLane<6, 2>(p, L, sink, with_rhs);
    else e = r3 ? AccumulateLane<8, 3>(p, L, sink, with_rhs) : AccumulateLane<8, 2>(p, L, sink, with_rhs);
    SBA_CUDA(e);
    if (with_rhs) {
      const int rc = FinalizeCameraVector(p, p->d_chunk_partial.p, p->d_rhs.p);
      if (rc) return rc;
      p->rhs_from_schur = true;
    }
    return SBA_OK;
  }
#define CALL(DC)                                                                                                          \
  do {                                                                                                                    \
    if (r3) {                                                                                                             \
      schur_wh_kernel<DC, 3><<<DivUp((int64_t)p->n_obs * DC, 256), 256, 0, st>>>(p->Struct(), p->d_cam_order.p, p->d_ER.p, p->d_F.p, p->d_ES.p, p->d_FS.p, \
                                                                               p->d_Cinv.p, p->d_W.p, p->d_H.p);          \
      schur_diag_chunk_kernel<DC, 3><<<p->n_chunks, 256, 0, st>>>(p->Chunks(), p->d_F.p, p->d_FS.p, p->d_W.p, p->d_H.p, p->d_chunk_block.p); \
    } else {                                                                                                              \
      schur_wh_kernel<DC, 2><<<DivUp((int64_t)p->n_obs * DC, 256), 256, 0, st>>>(p->Struct(), p->d_cam_order.p, p->d_ER.p, p->d_F.p, p->d_ES.p, p->d_FS.p, \
                                                                               p->d_Cinv.p, p->d_W.p, p->d_H.p);          \
      schur_diag_chunk_kernel<DC, 2><<<p->n_chunks, 256, 0, st>>>(p->Chunks(), p->d_F.p, p->d_FS.p, p->d_W.p, p->d_H.p, p->d_chunk_block.p); \
    }                                                                                                                     \
    schur_diag_finalize_kernel<DC><<<p->n_img, 128, 0, st>>>(p->Struct(), p->Chunks(), p->d_chunk_block.p, sink);          \
    p->launches += 3;                                                                                                     \
    if (L.n_seg > 0) {                                                                                                    \
      schur_pair_segment_kernel<DC><<<DivUp(L.n_seg, 4), 128, 0, st>>>(L.n_seg, L.seg_ptr.p, L.ab.p, p->d_W.p,             \
                                                                      p->d_H.p, L.partial.p);                             \
      schur_pair_emit_kernel<DC><<<DivUp(L.n_keys, 4), 128, 0, st>>>(p->Struct(), L.n_keys, L.key_seg_ptr.p, L.key_gi.p,   \
                                                                    L.key_gj.p, L.partial.p, sink);                       \
      p->launches += 2;                                                                                                   \
    }                                                                                                                     \
  } while (0)
  switch (dc) {
    case 10: CALL(10); break; case 14: CALL(14); break; default: CALL(18); break;
  }
#undef CALL
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

int BuildDenseSchur(Problem* p, bool with_rhs) {
  int rc = BuildPairLists(p, true);
  if (rc) return rc;
  const size_t n = p->n_cdim, ld = DenseLd((int)n);
  SBA_CUDA(p->d_S.Alloc(std::max<size_t>(1, (n + 1) * ld)));  // + one row: the right-hand side (cholesky.cu)
  SBA_CUDA(cudaMemsetAsync(p->d_S.p, 0, n * ld * sizeof(double), p->stream));
  Sink sink;
  std::memset(&sink, 0, sizeof(sink));
  sink.kind = 0; sink.S = p->d_S.p; sink.n = (int)ld; sink.shared_intr = p->shared_intrinsics ? 1 : 0;
  rc = Accumulate(p, p->pairs_dense, sink, with_rhs);
  if (rc) return rc;
  if (p->rhs_from_schur) p->rhs_unreduced = p->comm != nullptr;
  rc = AllReduceSumWithRhs(p, p->d_S.p, n * ld);  // the row behind S carries the reduced right-hand side along
  if (rc) return rc;
  p->chol_add_diag = (int)n <= kCholSmallMax;  // the small-system Cholesky adds D^2 while loading its diagonal tiles
  if (n > 0 && !p->chol_add_diag) { add_diag_dense_kernel<<<DivUp(n, 256), 256, 0, p->stream>>>((int)n, (int)ld, p->d_S.p, p->d_Dc.p); p->launches++; }
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
  rc = AllReduceSumWithRhs(p, p->d_M.p, p->m_size);
  if (rc) return rc;
  invert_blocks_kernel<<<DivUp(p->n_blk, 64), 64, 0, p->stream>>>(p->n_blk, p->d_blk_off.p, p->d_blk_dim.p, p->d_blk_moff.p,
                                                                 p->d_M.p, p->d_Dc.p, p->d_Minv.p, p->d_scalars.p);
  p->launches++;
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

// SCHUR_JACOBI blocks out of the diagonal blocks of the block-sparse S (every parameter block lies inside one image)
template <int DC>
__global__ void __launch_bounds__(128) bsr_extract_jacobi_kernel(StructView S, const double* __restrict__ B, Sink sink) {
  const int i = blockIdx.x, c = S.img_cam[i];
  for (int e = threadIdx.x; e < DC * DC; e += 128) {
    const int ja = e / DC, jb = e - ja * DC;
    Emit(sink, SlotCol(S, i, c, ja), SlotCol(S, i, c, jb), B[(size_t)i * DC * DC + e], true, false);
  }
}

// the inverted SCHUR_JACOBI blocks re-laid per image in slot space, for the persistent PCG kernel (pcg_blocks.cu)
template <int DC>
__global__ void __launch_bounds__(128) bsr_precond_slots_kernel(StructView S, const int* __restrict__ col_blk,
                                                                const int* __restrict__ blk_off, const int* __restrict__ blk_dim,
                                                                const int* __restrict__ blk_moff, const double* __restrict__ Minv,
                                                                double* __restrict__ out) {
  const int i = blockIdx.x, c = S.img_cam[i];
  for (int e = threadIdx.x; e < DC * DC; e += 128) {
    const int ja = e / DC, jb = e - ja * DC;
    const int row = SlotCol(S, i, c, ja), col = SlotCol(S, i, c, jb);
    double v = 0.0;
    if (row >= 0 && col >= 0 && col_blk[row] == col_blk[col]) {
      const int b = col_blk[row], o = blk_off[b], d = blk_dim[b];
      v = Minv[blk_moff[b] + (size_t)(row - o) * d + (col - o)];
    }
    out[(size_t)i * DC * DC + e] = v;
  }
}

// Block-sparse S (bsr.cu) and the SCHUR_JACOBI preconditioner read off its diagonal blocks.
int BuildExplicitSchur(Problem* p, bool with_rhs) {
  BlockSchur& B = p->bsr;
  int rc = EnsureBlockStorage(p);
  if (rc) return rc;
  const size_t ne = (size_t)p->dc * p->dc;
  SBA_CUDA(cudaMemsetAsync(B.blocks.p, 0, (size_t)B.n_blocks * ne * sizeof(double), p->stream));
  Sink sink;
  std::memset(&sink, 0, sizeof(sink));
  sink.kind = 2; sink.B = B.blocks.p; sink.n_img = p->n_img; sink.key_block = p->comm ? B.key_block.p : nullptr;
  rc = Accumulate(p, p->pairs_dense, sink, with_rhs);
  if (rc) return rc;
  if (p->rhs_from_schur) p->rhs_unreduced = p->comm != nullptr;
  rc = AllReduceSumWithRhs(p, B.blocks.p, (size_t)B.n_blocks * ne);
  if (rc) return rc;
  if (p->n_blk == 0) return SBA_OK;
  Sink m;
  std::memset(&m, 0, sizeof(m));
  m.kind = 1; m.M = p->d_M.p; m.col_blk = p->d_col_blk.p; m.blk_off = p->d_blk_off.p;
  m.blk_dim = p->d_blk_dim.p; m.blk_moff = p->d_blk_moff.p;
#define CALL(DC) bsr_extract_jacobi_kernel<DC><<<p->n_img, 128, 0, p->stream>>>(p->Struct(), B.blocks.p, m)
  SBA_DC_SWITCH(p->dc, CALL)
#undef CALL
  invert_blocks_kernel<<<DivUp(p->n_blk, 64), 64, 0, p->stream>>>(p->n_blk, p->d_blk_off.p, p->d_blk_dim.p, p->d_blk_moff.p,
                                                                 p->d_M.p, p->d_Dc.p, p->d_Minv.p, p->d_scalars.p);
  SBA_CUDA(B.minv_s.Alloc((size_t)p->n_img * ne));
#define CALL(DC) bsr_precond_slots_kernel<DC><<<p->n_img, 128, 0, p->stream>>>(p->Struct(), p->d_col_blk.p, p->d_blk_off.p, p->d_blk_dim.p, p->d_blk_moff.p, p->d_Minv.p, B.minv_s.p)
  SBA_DC_SWITCH(p->dc, CALL)
#undef CALL
  p->launches += 3;
  SBA_CUDA(cudaGetLastError());
  return SBA_OK;
}

}  // namespace sba
