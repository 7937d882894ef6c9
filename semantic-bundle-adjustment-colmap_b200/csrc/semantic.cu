// K-SEM-PAIR, K-CYL-IOU and K-SEM-BILIN: the fork's semantic residuals as CUDA kernels.
//
//  * semantic pair residual (0/1 label mismatch with depth gate)
//      /root/reference/src/base/semantic_cost_functions.h:88-208, pose algebra
//      /root/reference/src/util/rotation_extension.h:44-101
//  * cylinder IoU residual (1 - IoU of the rasterised tangent quadrilateral vs a mask)
//      /root/reference/src/util/cylinder.h:27-121 (drawQuadrilateral), :352-474, :497-540,
//      /root/reference/src/util/xywh.h:85-131, /root/reference/src/util/utils.h:23-55,
//      /root/reference/src/base/geometric_semantic_cost_functions.h:55-93
//  * both differentiated the way the reference asks Ceres to: NumericDiff CENTRAL,
//    step max(sqrt(eps), |x| * relative_step_size) (semantic_cost_functions.h:248-258,
//    SURVEY.md App. B.6)
//  * extension X1 (not in the reference): bilinear map sample at the projection with the
//    rule of Bitmap::InterpolateBilinear, /root/reference/src/util/bitmap.cc:238-269
//
// This file is compiled with -fmad=false: the pair residual is an integer-valued function
// of rounded pixel coordinates, and IEEE mul/add/div/sqrt without contraction reproduce the
// CPU evaluation order bit for bit, so rounding ties fall the same way.
#include <cfloat>

#include "device_math.cuh"
#include "reduce.cuh"
#include "semantic_device.cuh"
#include "solver.h"

namespace sba {
template <bool JAC>
__global__ void __launch_bounds__(128) semantic_pair_kernel(SemView V, int64_t n_blocks, double* __restrict__ res,
                                                            int* __restrict__ status, double* __restrict__ jac) {
  const int lane = threadIdx.x & 31;
  const int64_t k = JAC ? ((int64_t)blockIdx.x * 4 + (threadIdx.x >> 5)) : ((int64_t)blockIdx.x * 128 + threadIdx.x);
  if (k >= n_blocks) return;
  const int i1 = V.i1[k], i2 = V.i2[k], x1 = V.x1[k], y1 = V.y1[k];
  const size_t HW = (size_t)V.H * V.W;
  const float depth1 = __ldg(V.depth + i1 * HW + (size_t)y1 * V.W + x1);
  const float sem1 = __ldg(V.label + i1 * HW + (size_t)y1 * V.W + x1);
  double x[14];
#pragma unroll
  for (int j = 0; j < 4; ++j) { x[j] = V.q[4 * i1 + j]; x[7 + j] = V.q[4 * i2 + j]; }
#pragma unroll
  for (int j = 0; j < 3; ++j) { x[4 + j] = V.t[3 * i1 + j]; x[11 + j] = V.t[3 * i2 + j]; }
  // step of MY column (lane j < 14 owns Jacobian column j), taken before any perturbation
  const double my_delta = (JAC && lane < 14) ? NumDiffDelta(x[lane], V.rel) : 1.0;
  if (JAC && lane >= 1 && lane <= 28) {
    const int j = (lane - 1) >> 1;
    const double delta = NumDiffDelta(x[j], V.rel);
    x[j] = ((lane - 1) & 1) ? x[j] - delta : x[j] + delta;
  }
  int st = 0;
  double r = 0.0;
  if (!JAC || lane <= 28)
    r = PairResidual(V, x1, y1, depth1, sem1, V.depth + i2 * HW, V.label + i2 * HW, x, V.cam + 3 * i1, V.cam + 3 * i2, &st);
  if (!JAC) {
    if (res) res[k] = r;
    if (status) status[k] = st;
    return;
  }
  if (lane == 0) {
    if (res) res[k] = r;
    if (status) status[k] = st;
  }
  // lane j (0..13) gathers f(x + d e_j) from lane 1+2j and f(x - d e_j) from lane 2+2j
  const int src_p = (1 + 2 * lane) & 31, src_m = (2 + 2 * lane) & 31;
  const double fp = __shfl_sync(0xffffffffu, r, src_p);
  const double fm = __shfl_sync(0xffffffffu, r, src_m);
  if (lane < 14) jac[k * 14 + lane] = (fp - fm) * ((1.0 / my_delta) / 2.0);
}

}  // namespace sba

using namespace sba;

extern "C" int sba_semantic_evaluate(const sba_semantic_desc* d, double* residuals, int32_t* status, double* jac,
                                     double* ms_device) {
  if (!d || d->num_images <= 0 || d->height <= 0 || d->width <= 0 || d->num_blocks < 0 || !d->image_qvec ||
      !d->image_tvec || !d->camera_params || !d->depth_maps || !d->semantic_maps ||
      (d->num_blocks > 0 && (!d->block_image1 || !d->block_image2 || !d->block_x1 || !d->block_y1))) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "bad semantic description");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  const int64_t nb = d->num_blocks;
  for (int64_t k = 0; k < nb; ++k)
    if (d->block_image1[k] < 0 || d->block_image1[k] >= d->num_images || d->block_image2[k] < 0 ||
        d->block_image2[k] >= d->num_images || d->block_x1[k] < 0 || d->block_x1[k] >= d->width ||
        d->block_y1[k] < 0 || d->block_y1[k] >= d->height) {
      SetError(SBA_ERR_INVALID_ARGUMENT, "semantic block index out of range");
      return SBA_ERR_INVALID_ARGUMENT;
    }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    return SBA_ERR_NO_DEVICE;
  }
  cudaStream_t st;
  SBA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  const size_t HW = (size_t)d->height * d->width, n_img = d->num_images;
  DevBuf<double> dq, dt, dc, dres, djac;
  DevBuf<float> ddepth, dlabel;
  DevBuf<int> di1, di2, dx1, dy1, dstatus;
  auto up = [&](auto& buf, const auto* src, size_t n) -> cudaError_t {
    cudaError_t e = buf.Alloc(n);
    if (e != cudaSuccess) return e;
    return cudaMemcpyAsync(buf.p, src, n * sizeof(*src), cudaMemcpyHostToDevice, st);
  };
  SBA_CUDA(up(dq, d->image_qvec, n_img * 4)); SBA_CUDA(up(dt, d->image_tvec, n_img * 3));
  SBA_CUDA(up(dc, d->camera_params, n_img * 3));
  SBA_CUDA(up(ddepth, d->depth_maps, n_img * HW)); SBA_CUDA(up(dlabel, d->semantic_maps, n_img * HW));
  if (nb > 0) {
    SBA_CUDA(up(di1, d->block_image1, nb)); SBA_CUDA(up(di2, d->block_image2, nb));
    SBA_CUDA(up(dx1, d->block_x1, nb)); SBA_CUDA(up(dy1, d->block_y1, nb));
  }
  SBA_CUDA(dres.Alloc(std::max<int64_t>(1, nb))); SBA_CUDA(dstatus.Alloc(std::max<int64_t>(1, nb)));
  if (jac) SBA_CUDA(djac.Alloc(std::max<int64_t>(1, nb * 14)));
  SemView V;
  V.H = d->height; V.W = d->width; V.q = dq.p; V.t = dt.p; V.cam = dc.p; V.depth = ddepth.p; V.label = dlabel.p;
  V.i1 = di1.p; V.i2 = di2.p; V.x1 = dx1.p; V.y1 = dy1.p; V.thr = d->depth_error_threshold; V.rel = d->relative_step_size;
  cudaEvent_t e0, e1;
  SBA_CUDA(cudaEventCreate(&e0)); SBA_CUDA(cudaEventCreate(&e1));
  SBA_CUDA(cudaEventRecord(e0, st));
  if (nb > 0) {
    if (jac) semantic_pair_kernel<true><<<DivUp(nb, 4), 128, 0, st>>>(V, nb, dres.p, dstatus.p, djac.p);
    else semantic_pair_kernel<false><<<DivUp(nb, 128), 128, 0, st>>>(V, nb, dres.p, dstatus.p, nullptr);
  }
  SBA_CUDA(cudaEventRecord(e1, st));
  SBA_CUDA(cudaGetLastError());
  if (nb > 0) {
    if (residuals) SBA_CUDA(cudaMemcpyAsync(residuals, dres.p, nb * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (status) SBA_CUDA(cudaMemcpyAsync(status, dstatus.p, nb * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (jac) SBA_CUDA(cudaMemcpyAsync(jac, djac.p, nb * 14 * sizeof(double), cudaMemcpyDeviceToHost, st));
  }
  SBA_CUDA(cudaStreamSynchronize(st));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  if (ms_device) *ms_device = ms;
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaStreamDestroy(st);
  return SBA_OK;
}

// =================================== K-CYL-IOU ================================================
namespace sba {
__global__ void __launch_bounds__(256) mask_count_kernel(const uint8_t* __restrict__ masks, size_t HW, int* __restrict__ count) {
  const uint8_t* m = masks + (size_t)blockIdx.x * HW;
  int c = 0;
  for (size_t j = threadIdx.x; j < HW; j += 256) c += m[j] != 0;
  __shared__ int sm[8];
  for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) { int t = 0; for (int i = 0; i < 8; ++i) t += sm[i]; count[blockIdx.x] = t; }
}

struct CylView {
  int H, W;
  const double* q; const double* t; const double* cam;  // per image
  const uint8_t* masks; const int* pos_count;
  const double* cq; const double* ct; const double* cr; const double* ch;  // per cylinder
  const int* b_img; const int* b_cyl;
  double rel;
};

// One CTA per (residual block, evaluation): evaluation 0 is the block itself, 1+2j / 2+2j are
// the +/- perturbations of ambient coordinate j (16 of them).  vals[k*33 + e] = 1 - IoU.
__global__ void __launch_bounds__(256) cylinder_iou_kernel(CylView V, int n_eval, double* __restrict__ vals,
                                                           double* __restrict__ deltas) {
  const int64_t k = blockIdx.x / n_eval;
  const int e = blockIdx.x - (int)(k * n_eval);
  const int im = V.b_img[k], cy = V.b_cyl[k];
  __shared__ QuadRaster R;
  __shared__ int ok;
  __shared__ int red[8][4];
  if (threadIdx.x == 0) {
    double x[16];
    for (int j = 0; j < 4; ++j) { x[j] = V.q[4 * im + j]; x[7 + j] = V.cq[4 * cy + j]; }
    for (int j = 0; j < 3; ++j) { x[4 + j] = V.t[3 * im + j]; x[11 + j] = V.ct[3 * cy + j]; }
    x[14] = V.cr[cy]; x[15] = V.ch[cy];
    if (e > 0) {
      const int j = (e - 1) >> 1;
      const double delta = NumDiffDelta(x[j], V.rel);
      x[j] = ((e - 1) & 1) ? x[j] - delta : x[j] + delta;
      if (!((e - 1) & 1)) deltas[k * 16 + j] = delta;
    }
    double pts[8];
    ok = CylinderQuad(x, V.cam + 3 * im, pts) ? 1 : 0;
    if (ok) BuildRaster(pts, V.H, V.W, &R);
  }
  __syncthreads();
  if (!ok) {
    if (threadIdx.x == 0) vals[k * n_eval + e] = 1.0;
    return;
  }
  const uint8_t* sem = V.masks + (size_t)im * V.H * V.W;
  int inside_pos = 0, tp = 0, fp = 0, snm = 0;
  const int n_px = R.bb.w * R.bb.h;
  for (int idx = threadIdx.x; idx < n_px; idx += 256) {
    const int yy = idx / R.bb.w, xx = idx - yy * R.bb.w;
    const int x = R.bb.x + xx, y = R.bb.y + yy;
    const bool s = sem[(size_t)y * V.W + x] != 0;
    const bool m = RasterInside(R, x, y);
    inside_pos += s; tp += (s && m); fp += (!s && m); snm += (s && !m);
  }
  int v[4] = {inside_pos, tp, fp, snm};
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    for (int o = 16; o > 0; o >>= 1) v[a] += __shfl_down_sync(0xffffffffu, v[a], o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][a] = v[a];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int t[4] = {0, 0, 0, 0};
    for (int w = 0; w < 8; ++w) for (int a = 0; a < 4; ++a) t[a] += red[w][a];
    const int fn = (V.pos_count[im] - t[0]) + t[3];
    const double iou = (double)t[1] / (t[1] * 1. + t[2] + fn);
    vals[k * n_eval + e] = 1.0 - iou;
  }
}

__global__ void cylinder_jac_kernel(int64_t n_blocks, const double* __restrict__ vals, const double* __restrict__ deltas,
                                    double* __restrict__ jac) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_blocks * 16) return;
  const int64_t k = t / 16;
  const int j = (int)(t - k * 16);
  const double fp = vals[k * 33 + 1 + 2 * j], fm = vals[k * 33 + 2 + 2 * j];
  jac[t] = (fp - fm) * ((1.0 / deltas[t]) / 2.0);
}

}  // namespace sba

extern "C" int sba_cylinder_evaluate(const sba_cylinder_desc* d, double* residuals, double* jac, double* ms_device) {
  if (!d || d->num_images <= 0 || d->height <= 0 || d->width <= 0 || d->num_cylinders < 0 || d->num_blocks < 0 ||
      !d->image_qvec || !d->image_tvec || !d->camera_params || !d->masks ||
      (d->num_blocks > 0 && (!d->block_image || !d->block_cylinder || !d->cyl_qvec || !d->cyl_tvec || !d->cyl_radius || !d->cyl_height))) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "bad cylinder description");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  const int64_t nb = d->num_blocks;
  for (int64_t k = 0; k < nb; ++k)
    if (d->block_image[k] < 0 || d->block_image[k] >= d->num_images || d->block_cylinder[k] < 0 ||
        d->block_cylinder[k] >= d->num_cylinders) {
      SetError(SBA_ERR_INVALID_ARGUMENT, "cylinder block index out of range");
      return SBA_ERR_INVALID_ARGUMENT;
    }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    return SBA_ERR_NO_DEVICE;
  }
  cudaStream_t st;
  SBA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  const size_t HW = (size_t)d->height * d->width, n_img = d->num_images, n_cyl = std::max(1, d->num_cylinders);
  DevBuf<double> dq, dt, dc, dcq, dct, dcr, dch, dvals, ddel, djac;
  DevBuf<uint8_t> dmask;
  DevBuf<int> dpos, dbi, dbc;
  auto up = [&](auto& buf, const auto* src, size_t n) -> cudaError_t {
    cudaError_t e = buf.Alloc(n);
    if (e != cudaSuccess || !src || n == 0) return e;
    return cudaMemcpyAsync(buf.p, src, n * sizeof(*src), cudaMemcpyHostToDevice, st);
  };
  SBA_CUDA(up(dq, d->image_qvec, n_img * 4)); SBA_CUDA(up(dt, d->image_tvec, n_img * 3));
  SBA_CUDA(up(dc, d->camera_params, n_img * 3)); SBA_CUDA(up(dmask, d->masks, n_img * HW));
  SBA_CUDA(up(dcq, d->cyl_qvec, (size_t)d->num_cylinders * 4)); SBA_CUDA(up(dct, d->cyl_tvec, (size_t)d->num_cylinders * 3));
  SBA_CUDA(up(dcr, d->cyl_radius, (size_t)d->num_cylinders)); SBA_CUDA(up(dch, d->cyl_height, (size_t)d->num_cylinders));
  (void)n_cyl;
  SBA_CUDA(up(dbi, d->block_image, nb)); SBA_CUDA(up(dbc, d->block_cylinder, nb));
  SBA_CUDA(dpos.Alloc(n_img));
  const int n_eval = jac ? 33 : 1;
  SBA_CUDA(dvals.Alloc(std::max<int64_t>(1, nb * n_eval))); SBA_CUDA(ddel.Alloc(std::max<int64_t>(1, nb * 16)));
  if (jac) SBA_CUDA(djac.Alloc(std::max<int64_t>(1, nb * 16)));
  CylView V;
  V.H = d->height; V.W = d->width; V.q = dq.p; V.t = dt.p; V.cam = dc.p; V.masks = dmask.p; V.pos_count = dpos.p;
  V.cq = dcq.p; V.ct = dct.p; V.cr = dcr.p; V.ch = dch.p; V.b_img = dbi.p; V.b_cyl = dbc.p; V.rel = d->relative_step_size;
  cudaEvent_t e0, e1;
  SBA_CUDA(cudaEventCreate(&e0)); SBA_CUDA(cudaEventCreate(&e1));
  SBA_CUDA(cudaEventRecord(e0, st));
  mask_count_kernel<<<(int)n_img, 256, 0, st>>>(dmask.p, HW, dpos.p);
  if (nb > 0) {
    cylinder_iou_kernel<<<(unsigned)(nb * n_eval), 256, 0, st>>>(V, n_eval, dvals.p, ddel.p);
    if (jac) cylinder_jac_kernel<<<DivUp(nb * 16, 256), 256, 0, st>>>(nb, dvals.p, ddel.p, djac.p);
  }
  SBA_CUDA(cudaEventRecord(e1, st));
  SBA_CUDA(cudaGetLastError());
  if (nb > 0) {
    if (residuals) SBA_CUDA(cudaMemcpy2DAsync(residuals, sizeof(double), dvals.p, n_eval * sizeof(double), sizeof(double), nb,
                                              cudaMemcpyDeviceToHost, st));
    if (jac) SBA_CUDA(cudaMemcpyAsync(jac, djac.p, nb * 16 * sizeof(double), cudaMemcpyDeviceToHost, st));
  }
  SBA_CUDA(cudaStreamSynchronize(st));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  if (ms_device) *ms_device = ms;
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaStreamDestroy(st);
  return SBA_OK;
}

// ---- parity surface of the two halves of the cylinder residual ---------------------------------------------------
namespace sba {
__global__ void cyl_quad_kernel(int64_t n_blocks, CylView V, double* __restrict__ pts, int* __restrict__ ok) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_blocks) return;
  const int im = V.b_img[k], cy = V.b_cyl[k];
  double x[16];
  for (int j = 0; j < 4; ++j) { x[j] = V.q[4 * im + j]; x[7 + j] = V.cq[4 * cy + j]; }
  for (int j = 0; j < 3; ++j) { x[4 + j] = V.t[3 * im + j]; x[11 + j] = V.ct[3 * cy + j]; }
  x[14] = V.cr[cy]; x[15] = V.ch[cy];
  double p8[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  ok[k] = CylinderQuad(x, V.cam + 3 * im, p8) ? 1 : 0;
  for (int j = 0; j < 8; ++j) pts[k * 8 + j] = p8[j];
}
__global__ void __launch_bounds__(256) quad_draw_kernel(const double* __restrict__ pts8, int H, int W, uint8_t* __restrict__ mask) {
  __shared__ QuadRaster R;
  if (threadIdx.x == 0) BuildRaster(pts8, H, W, &R);
  __syncthreads();
  for (int idx = blockIdx.x * 256 + threadIdx.x; idx < H * W; idx += gridDim.x * 256) {
    const int y = idx / W, x = idx - y * W;
    mask[idx] = (InBox(R.bb, x, y) && RasterInside(R, x, y)) ? 1 : 0;
  }
}
}  // namespace sba

extern "C" int sba_cylinder_quadrilateral(const sba_cylinder_desc* d, double* pts8, int32_t* ok) {
  if (!d || !pts8 || !ok || d->num_blocks < 0) { SetError(SBA_ERR_INVALID_ARGUMENT, "bad arguments"); return SBA_ERR_INVALID_ARGUMENT; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback"); return SBA_ERR_NO_DEVICE; }
  const int64_t nb = d->num_blocks;
  if (nb == 0) return SBA_OK;
  cudaStream_t st = nullptr;
  DevBuf<double> dq, dt, dc, dcq, dct, dcr, dch, dpts;
  DevBuf<int> dbi, dbc, dok;
  auto up = [&](auto& buf, const auto* src, size_t n) -> cudaError_t {
    cudaError_t e = buf.Alloc(n);
    if (e != cudaSuccess || !src || n == 0) return e;
    return cudaMemcpy(buf.p, src, n * sizeof(*src), cudaMemcpyHostToDevice);
  };
  const size_t n_img = d->num_images, n_cyl = d->num_cylinders;
  SBA_CUDA(up(dq, d->image_qvec, n_img * 4)); SBA_CUDA(up(dt, d->image_tvec, n_img * 3)); SBA_CUDA(up(dc, d->camera_params, n_img * 3));
  SBA_CUDA(up(dcq, d->cyl_qvec, n_cyl * 4)); SBA_CUDA(up(dct, d->cyl_tvec, n_cyl * 3));
  SBA_CUDA(up(dcr, d->cyl_radius, n_cyl)); SBA_CUDA(up(dch, d->cyl_height, n_cyl));
  SBA_CUDA(up(dbi, d->block_image, nb)); SBA_CUDA(up(dbc, d->block_cylinder, nb));
  SBA_CUDA(dpts.Alloc(nb * 8)); SBA_CUDA(dok.Alloc(nb));
  CylView V;
  V.H = d->height; V.W = d->width; V.q = dq.p; V.t = dt.p; V.cam = dc.p; V.masks = nullptr; V.pos_count = nullptr;
  V.cq = dcq.p; V.ct = dct.p; V.cr = dcr.p; V.ch = dch.p; V.b_img = dbi.p; V.b_cyl = dbc.p; V.rel = d->relative_step_size;
  cyl_quad_kernel<<<DivUp(nb, 128), 128, 0, st>>>(nb, V, dpts.p, dok.p);
  SBA_CUDA(cudaGetLastError());
  SBA_CUDA(cudaMemcpy(pts8, dpts.p, nb * 8 * sizeof(double), cudaMemcpyDeviceToHost));
  SBA_CUDA(cudaMemcpy(ok, dok.p, nb * sizeof(int), cudaMemcpyDeviceToHost));
  return SBA_OK;
}

extern "C" int sba_draw_quadrilateral(const double* pts8, int32_t height, int32_t width, uint8_t* mask) {
  if (!pts8 || !mask || height <= 0 || width <= 0) { SetError(SBA_ERR_INVALID_ARGUMENT, "bad arguments"); return SBA_ERR_INVALID_ARGUMENT; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback"); return SBA_ERR_NO_DEVICE; }
  DevBuf<double> dp;
  DevBuf<uint8_t> dm;
  SBA_CUDA(dp.Alloc(8)); SBA_CUDA(dm.Alloc((size_t)height * width));
  SBA_CUDA(cudaMemcpy(dp.p, pts8, 8 * sizeof(double), cudaMemcpyHostToDevice));
  quad_draw_kernel<<<64, 256>>>(dp.p, height, width, dm.p);
  SBA_CUDA(cudaGetLastError());
  SBA_CUDA(cudaMemcpy(mask, dm.p, (size_t)height * width, cudaMemcpyDeviceToHost));
  return SBA_OK;
}
