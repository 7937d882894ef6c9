// Device-side ordering of the observations (set-up, once per problem).
//
// The reference's residual order is the loop nest of BundleAdjuster::SetUp
//   /root/reference/src/optim/bundle_adjustment.cc:326-342 (images outer, Points2D inner, then extra track elements);
// the hot path wants the same observations POINT-sorted (Schur elimination is per point) and an IMAGE-major index
// list (camera-side reductions).  Both orders are stable radix sorts (cub) of the caller's own index arrays, made on
// the device from a raw copy of obs_image / obs_point / obs_xy:
//   local2ref : global stable sort by point, then the shard's contiguous slice [obs_base, obs_base + n_obs)
//   cam_order : stable sort of the point-sorted local observations by image
// Integer work only; bit-exact by construction (stable sorts of the reference order).
#include <cub/cub.cuh>

#include "solver.h"

namespace sba {

// cam_pos[cam_order[k]] = k: image-major position of a point-sorted observation;
// cam_po[k] = p_off of the point of the k-th observation in image-major order (-1: constant point), so that an
// image-major pass reaches its point's records with one dependent load instead of three
__global__ void invert_order_kernel(int n, const int* __restrict__ order, int* __restrict__ pos,
                                    const int* __restrict__ obs_pt, const int* __restrict__ p_off, int* __restrict__ cam_po) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) { const int a = order[k]; pos[a] = k; cam_po[k] = p_off[obs_pt[a]]; }
}

__global__ void iota_kernel(unsigned* __restrict__ v, int64_t n) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) v[k] = (unsigned)k;
}

// per local observation k: gather the caller's arrays through local2ref; obs_active as Analyse() defines it
__global__ void __launch_bounds__(256) order_gather_kernel(int n_obs, int p0, const unsigned* __restrict__ l2r,
                                                           const int* __restrict__ raw_img, const int* __restrict__ raw_pt,
                                                           const double2* __restrict__ raw_xy, const int* __restrict__ img_cam,
                                                           const int* __restrict__ col_img, const int* __restrict__ k_dim,
                                                           const int* __restrict__ p_off, int* __restrict__ obs_img,
                                                           int* __restrict__ obs_pt, double2* __restrict__ obs_xy,
                                                           uint8_t* __restrict__ obs_active, unsigned* __restrict__ img_key) {
  const int k = blockIdx.x * 256 + threadIdx.x;
  if (k >= n_obs) return;
  const unsigned r = l2r[k];
  const int i = raw_img[r], q = raw_pt[r] - p0;
  obs_img[k] = i; obs_pt[k] = q; obs_xy[k] = raw_xy[r];
  img_key[k] = (unsigned)i;
  obs_active[k] = (col_img[6 * (size_t)i] >= 0 || k_dim[img_cam[i]] > 0 || p_off[q] >= 0) ? 1 : 0;
}

static int Bits(int64_t n) {
  int b = 1;
  while (b < 32 && (1ll << b) < n) ++b;
  return b;
}

// d_raw_*: the caller's arrays already on the device (whole problem, reference order).
int BuildDeviceOrder(Problem* p, int p0, int64_t obs_base, int64_t n_obs_g, int n_pt_g, const int* d_raw_img,
                     const int* d_raw_pt, const double* d_raw_xy) {
  cudaStream_t st = p->stream;
  const int n_obs = p->n_obs;
  SBA_CUDA(p->d_local2ref.Alloc(std::max(1, n_obs)));
  SBA_CUDA(p->d_obs_img.Alloc(std::max(1, n_obs))); SBA_CUDA(p->d_obs_pt.Alloc(std::max(1, n_obs)));
  SBA_CUDA(p->d_obs_xy.Alloc((size_t)std::max(1, n_obs) * 2)); SBA_CUDA(p->d_obs_active.Alloc(std::max(1, n_obs)));
  SBA_CUDA(p->d_cam_order.Alloc(std::max(1, n_obs))); SBA_CUDA(p->d_cam_pos.Alloc(std::max(1, n_obs)));
  SBA_CUDA(p->d_cam_po.Alloc(std::max(1, n_obs)));
  if (n_obs == 0 || n_obs_g == 0) return SBA_OK;
  DevBuf<unsigned> idx, keys_out, vals_out, img_key;
  DevBuf<char> tmp;
  SBA_CUDA(idx.Alloc(n_obs_g)); SBA_CUDA(keys_out.Alloc(n_obs_g)); SBA_CUDA(vals_out.Alloc(n_obs_g));
  SBA_CUDA(img_key.Alloc(n_obs));
  const unsigned* raw_pt_u = reinterpret_cast<const unsigned*>(d_raw_pt);  // validated non-negative
  size_t b1 = 0, b2 = 0;
  SBA_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, b1, raw_pt_u, keys_out.p, idx.p, vals_out.p, (int)n_obs_g, 0, Bits(n_pt_g), st));
  SBA_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, b2, img_key.p, keys_out.p, idx.p,
                                           reinterpret_cast<unsigned*>(p->d_cam_order.p), n_obs, 0, Bits(p->n_img), st));
  SBA_CUDA(tmp.Alloc(std::max(b1, b2)));
  iota_kernel<<<DivUp(n_obs_g, 256), 256, 0, st>>>(idx.p, n_obs_g);
  SBA_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, b1, raw_pt_u, keys_out.p, idx.p, vals_out.p, (int)n_obs_g, 0, Bits(n_pt_g), st));
  SBA_CUDA(cudaMemcpyAsync(p->d_local2ref.p, vals_out.p + obs_base, (size_t)n_obs * sizeof(unsigned), cudaMemcpyDeviceToDevice, st));
  order_gather_kernel<<<DivUp(n_obs, 256), 256, 0, st>>>(
      n_obs, p0, p->d_local2ref.p, d_raw_img, d_raw_pt, reinterpret_cast<const double2*>(d_raw_xy), p->d_img_cam.p,
      p->d_col_img.p, p->d_k_dim.p, p->d_p_off.p, p->d_obs_img.p, p->d_obs_pt.p, reinterpret_cast<double2*>(p->d_obs_xy.p),
      p->d_obs_active.p, img_key.p);
  // idx[0..n_obs) is still 0,1,2,...: the values of the image sort are the local point-sorted positions
  SBA_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, b2, img_key.p, keys_out.p, idx.p,
                                           reinterpret_cast<unsigned*>(p->d_cam_order.p), n_obs, 0, Bits(p->n_img), st));
  invert_order_kernel<<<DivUp(n_obs, 256), 256, 0, st>>>(n_obs, p->d_cam_order.p, p->d_cam_pos.p, p->d_obs_pt.p, p->d_p_off.p,
                                                             p->d_cam_po.p);
  p->launches += 3;
  SBA_CUDA(cudaGetLastError());
  SBA_CUDA(cudaStreamSynchronize(st));  // the temporaries above are released on return
  return SBA_OK;
}

// Host copies of the two orders, needed only by the parity surface (sba_evaluate returns reference order).
int FetchHostOrder(Problem* p) {
  if (p->host_order_ready) return SBA_OK;
  const int n = p->n_obs;
  std::vector<unsigned> l2r(std::max(1, n));
  p->h_obs_img.resize(n); p->h_obs_pt.resize(n);
  if (n > 0) {
    SBA_CUDA(cudaMemcpyAsync(l2r.data(), p->d_local2ref.p, (size_t)n * sizeof(unsigned), cudaMemcpyDeviceToHost, p->stream));
    SBA_CUDA(cudaMemcpyAsync(p->h_obs_img.data(), p->d_obs_img.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, p->stream));
    SBA_CUDA(cudaMemcpyAsync(p->h_obs_pt.data(), p->d_obs_pt.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, p->stream));
    SBA_CUDA(cudaStreamSynchronize(p->stream));
  }
  p->h_local2ref.resize(n);
  for (int k = 0; k < n; ++k) p->h_local2ref[k] = (int64_t)l2r[k];
  p->host_order_ready = true;
  return SBA_OK;
}

}  // namespace sba
