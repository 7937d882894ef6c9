// Observation-pair lists for the off-diagonal contributions of the explicit reduced camera matrix
// (K-SCHUR, schur.cu), built ON THE DEVICE once per problem.
//
// Ceres' SchurEliminator visits, for every point, all pairs of its f-blocks and adds -(F_a^T E_a) C^-1 (E_b^T F_b) into
// the S block of the two cameras under a per-block lock (problem of
// /root/reference/src/optim/bundle_adjustment.cc:326-530).  Here the same pairs are enumerated once, keyed by the
// camera pair (image(a), image(b)) with image(a) >= image(b), and STABLY sorted by key, so that every S block is
// the fixed-order sum of a contiguous run of pairs: single writer, no atomics, bit-reproducible.
//
//   dense == true : every pair of observations of one variable point in different images plus duplicate
//                   observations inside one image (both orders).
//   dense == false: only what SCHUR_JACOBI needs beyond the per-observation diagonal terms: duplicates inside
//                   one image and -- when intrinsics are shared -- pairs whose images use the same camera.
//
// Order inside a key: ascending (a, b) in the point-sorted observation numbering -- the enumeration emits pairs in
// that order and the radix sort is stable.  All of this is integer work; cub's scan / radix sort / run-length
// primitives do the data movement.
#include <cub/cub.cuh>

#include "solver.h"

namespace sba {

struct PairRule {
  int dense, shared;
  const int* img_cam;
  const int* k_dim;
};

__device__ inline bool PairEligible(const PairRule& r, int s, int t, int i, int j) {
  if (t == s || j > i) return false;  // the partner's own visit handles image(b) > image(a)
  if (j != i && !r.dense) {
    if (!r.shared) return false;
    const int ci = r.img_cam[i];
    if (r.img_cam[j] != ci || r.k_dim[ci] <= 0) return false;
  }
  return true;
}

// mode 0: count the pairs owned by observation s; mode 1: write them at off[s]
template <int MODE>
__global__ void __launch_bounds__(256) pair_enum_kernel(int n_obs, int n_img, const int* __restrict__ obs_img,
                                                        const int* __restrict__ obs_pt, const int* __restrict__ pt_ptr,
                                                        const int* __restrict__ p_off, const int* __restrict__ cam_pos,
                                                        PairRule rule,
                                                        int* __restrict__ cnt, const int* __restrict__ off,
                                                        unsigned* __restrict__ keys, unsigned long long* __restrict__ ab) {
  const int s = blockIdx.x * 256 + threadIdx.x;
  if (s >= n_obs) return;
  const int q = obs_pt[s];
  int c = 0;
  if (p_off[q] >= 0) {  // constant point: E = 0, nothing to eliminate
    const int i = obs_img[s];
    const int t0 = pt_ptr[q], t1 = pt_ptr[q + 1];
    const int base = MODE ? off[s] : 0;
    for (int t = t0; t < t1; ++t) {
      const int j = obs_img[t];
      if (!PairEligible(rule, s, t, i, j)) continue;
      if (MODE) {
        keys[base + c] = (unsigned)i * (unsigned)n_img + (unsigned)j;
        // the ids are ROWS of W / H (image-major positions); the order of the pairs is that of the enumeration
        ab[base + c] = ((unsigned long long)(unsigned)cam_pos[s] << 32) | (unsigned)cam_pos[t];
      }
      ++c;
    }
  }
  if (!MODE) cnt[s] = c;
}

__global__ void pair_total_kernel(int n_obs, const int* __restrict__ cnt, const int* __restrict__ off, int* __restrict__ total) {
  if (threadIdx.x == 0 && blockIdx.x == 0) total[0] = n_obs > 0 ? off[n_obs - 1] + cnt[n_obs - 1] : 0;
}

// per key: image pair and number of segments
__global__ void pair_key_kernel(const int* __restrict__ n_keys_p, int n_img, int seg_pairs, const unsigned* __restrict__ ukeys,
                                const int* __restrict__ kcnt, int* __restrict__ key_gi, int* __restrict__ key_gj,
                                int* __restrict__ key_nseg, int cap) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= cap) return;
  if (k >= n_keys_p[0]) { key_nseg[k] = 0; return; }
  const unsigned key = ukeys[k];
  key_gi[k] = (int)(key / (unsigned)n_img);
  key_gj[k] = (int)(key % (unsigned)n_img);
  key_nseg[k] = (kcnt[k] + seg_pairs - 1) / seg_pairs;
}

// seg_ptr of every segment of every key (+ the terminator), and the totals for the host
__global__ void pair_seg_kernel(const int* __restrict__ n_keys_p, int seg_pairs, const int* __restrict__ key_ptr,
                                const int* __restrict__ kcnt, int* __restrict__ key_seg_ptr, const int* __restrict__ key_nseg,
                                int* __restrict__ seg_ptr, int n_pairs, int* __restrict__ totals) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int n_keys = n_keys_p[0];
  if (k > n_keys) return;
  if (k == n_keys) {
    const int n_seg = n_keys > 0 ? key_seg_ptr[n_keys - 1] + key_nseg[n_keys - 1] : 0;
    key_seg_ptr[n_keys] = n_seg;
    seg_ptr[n_seg] = n_pairs;
    totals[0] = n_keys; totals[1] = n_seg;
    return;
  }
  const int s0 = key_seg_ptr[k], p0 = key_ptr[k];
  for (int x = 0; x < key_nseg[k]; ++x) seg_ptr[s0 + x] = p0 + x * seg_pairs;
}

int BuildPairLists(Problem* p, bool dense) {
  PairList& L = dense ? p->pairs_dense : p->pairs_diag;
  if (L.built) return SBA_OK;
  HostTimer tm("pair lists (device)");
  cudaStream_t st = p->stream;
  const int n_obs = p->n_obs, n_img = p->n_img;
  L.n_seg = 0; L.n_keys = 0; L.n_pairs = 0;
  L.seg_pairs = PairSegmentSize(p->dc);
  L.built = true;
  if (n_obs == 0 || (int64_t)n_img * n_img > 0x7fffffffLL) {
    if (n_obs == 0) return SBA_OK;
    SetError(SBA_ERR_INVALID_ARGUMENT, "too many images for 32-bit camera-pair keys");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  PairRule rule{dense ? 1 : 0, (!dense && p->shared_intrinsics) ? 1 : 0, p->d_img_cam.p, p->d_k_dim.p};
  DevBuf<int> cnt, off, totals;
  SBA_CUDA(cnt.Alloc(n_obs)); SBA_CUDA(off.Alloc(n_obs)); SBA_CUDA(totals.Alloc(4));
  int h_total_pairs = 0;
  pair_enum_kernel<0><<<DivUp(n_obs, 256), 256, 0, st>>>(n_obs, n_img, p->d_obs_img.p, p->d_obs_pt.p, p->d_pt_ptr.p,
                                                         p->d_p_off.p, p->d_cam_pos.p, rule, cnt.p, nullptr, nullptr, nullptr);
  size_t tmp_bytes = 0;
  SBA_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, cnt.p, off.p, n_obs, st));
  DevBuf<char> tmp;
  SBA_CUDA(tmp.Alloc(tmp_bytes));
  SBA_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, cnt.p, off.p, n_obs, st));
  pair_total_kernel<<<1, 32, 0, st>>>(n_obs, cnt.p, off.p, totals.p);
  p->launches += 3;
  SBA_CUDA(cudaMemcpyAsync(&h_total_pairs, totals.p, sizeof(int), cudaMemcpyDeviceToHost, st));
  SBA_CUDA(cudaStreamSynchronize(st));
  const int n_pairs = h_total_pairs;
  tm.Lap("count");
  if (n_pairs == 0) return SBA_OK;

  DevBuf<unsigned> keys, keys_sorted, ukeys;
  DevBuf<unsigned long long> ab;
  SBA_CUDA(keys.Alloc(n_pairs)); SBA_CUDA(keys_sorted.Alloc(n_pairs));
  SBA_CUDA(ab.Alloc(n_pairs)); SBA_CUDA(L.ab.Alloc(n_pairs));
  tm.Lap("pair buffers");
  pair_enum_kernel<1><<<DivUp(n_obs, 256), 256, 0, st>>>(n_obs, n_img, p->d_obs_img.p, p->d_obs_pt.p, p->d_pt_ptr.p,
                                                         p->d_p_off.p, p->d_cam_pos.p, rule, nullptr, off.p, keys.p, ab.p);
  int key_bits = 1;
  while (key_bits < 32 && (1ll << key_bits) < (int64_t)n_img * n_img) ++key_bits;
  size_t sort_bytes = 0, rle_bytes = 0, scan_bytes = 0;
  SBA_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys.p, keys_sorted.p, ab.p,
                                           L.ab.p, n_pairs, 0, key_bits, st));
  const int key_cap = (int)std::min<int64_t>(n_pairs, (int64_t)n_img * (n_img + 1) / 2);
  DevBuf<int> kcnt, key_ptr, key_nseg, n_keys_d;
  SBA_CUDA(ukeys.Alloc(key_cap)); SBA_CUDA(kcnt.Alloc(key_cap)); SBA_CUDA(key_ptr.Alloc(key_cap));
  SBA_CUDA(key_nseg.Alloc(key_cap)); SBA_CUDA(n_keys_d.Alloc(1));
  SBA_CUDA(cub::DeviceRunLengthEncode::Encode(nullptr, rle_bytes, keys_sorted.p, ukeys.p, kcnt.p, n_keys_d.p, n_pairs, st));
  SBA_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, kcnt.p, key_ptr.p, key_cap, st));
  SBA_CUDA(tmp.Alloc(std::max(sort_bytes, std::max(rle_bytes, scan_bytes))));
  SBA_CUDA(cudaStreamSynchronize(st));
  tm.Lap("enumerate + key buffers");
  size_t b = sort_bytes;
  SBA_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, b, keys.p, keys_sorted.p, ab.p,
                                           L.ab.p, n_pairs, 0, key_bits, st));
  b = rle_bytes;
  SBA_CUDA(cub::DeviceRunLengthEncode::Encode(tmp.p, b, keys_sorted.p, ukeys.p, kcnt.p, n_keys_d.p, n_pairs, st));
  b = scan_bytes;
  SBA_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, b, kcnt.p, key_ptr.p, key_cap, st));
  const int seg_cap = n_pairs / L.seg_pairs + key_cap + 1;
  SBA_CUDA(L.key_gi.Alloc(key_cap)); SBA_CUDA(L.key_gj.Alloc(key_cap));
  SBA_CUDA(L.key_seg_ptr.Alloc(key_cap + 1)); SBA_CUDA(L.seg_ptr.Alloc(seg_cap + 1));
  pair_key_kernel<<<DivUp(key_cap, 256), 256, 0, st>>>(n_keys_d.p, n_img, L.seg_pairs, ukeys.p, kcnt.p, L.key_gi.p,
                                                       L.key_gj.p, key_nseg.p, key_cap);
  b = scan_bytes;
  SBA_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, b, key_nseg.p, L.key_seg_ptr.p, key_cap, st));
  pair_seg_kernel<<<DivUp(key_cap + 1, 256), 256, 0, st>>>(n_keys_d.p, L.seg_pairs, key_ptr.p, kcnt.p, L.key_seg_ptr.p,
                                                           key_nseg.p, L.seg_ptr.p, n_pairs, totals.p);
  p->launches += 7;
  int h_totals[2] = {0, 0};
  SBA_CUDA(cudaMemcpyAsync(h_totals, totals.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
  SBA_CUDA(cudaStreamSynchronize(st));  // also: the temporaries above may go out of scope
  L.n_keys = h_totals[0]; L.n_seg = h_totals[1]; L.n_pairs = n_pairs;
  tm.Lap("sort + segments");
  SBA_CUDA(L.partial.Alloc((size_t)std::max(1, L.n_seg) * p->dc * p->dc));
  tm.Lap("segment partials buffer");
  return SBA_OK;
}

}  // namespace sba
