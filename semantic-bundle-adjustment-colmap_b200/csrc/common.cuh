// Shared declarations of the B200-native bundle-adjustment hot path.
//
// Reference boundary this library sits behind:
//   BundleAdjuster::Solve   /root/reference/src/optim/bundle_adjustment.cc:258-320
//   cost functors           /root/reference/src/base/cost_functions.h:44-152
// Everything the reference hands to ceres::Solve (bundle_adjustment.cc:306) is
// evaluated and solved on the device by the kernels declared here.
//
// Device data layout (all fp64):
//   images  : pose record  double[8]  = q(w,x,y,z) t(x,y,z) pad     (64 B aligned gather unit)
//   cameras : params       double[12]                               (96 B)
//   points  : X            double[3]  packed
//   observations are kept in POINT-SORTED order (stable by reference order);
//   ref2pt / pt2ref permutations recover the reference residual order.
//   per observation k the linearisation is two records
//     ER = double[8]  : E (2x3 d r/d X, row-major) + r(2)
//     F  = double2[DC]: slot j -> (d r0/d s_j, d r1/d s_j); slots 0..2 quaternion tangent, 3..5 tvec,
//                       6.. free intrinsics
//   stored TILE-BLOCKED in 32-byte units (D4), units PAIRED into 64-byte pieces: observations are grouped in tiles
//   of 32; inside a tile, the pair (2m, 2m+1) of units of all 32 observations is contiguous (2 KB, lane-major), and an
//   odd last unit is a 1 KB row of its own (PairedUnit below):
//     F  unit jj (slots 2jj, 2jj+1) of observation k : F [PairedUnit(k, jj, DC/2)]
//     ER unit u  (0: E0..E3, 1: E4 E5 r0 r1)         : ER[PairedUnit(k, u, 2)]
//   A warp that walks consecutive observations (the point-sorted passes) moves the two halves of a 2 KB piece with two
//   256-bit accesses per lane, every sector fully used.  A pass that gathers observations in image order (the Schur
//   build, the F^T reductions) touches whole 64 B pieces: HBM delivers 64 B bursts, and with single 32 B units per
//   row (round 1) such a gather fetched twice what it used (ncu: 1.99 GB of DRAM reads for 0.96 GB in the C4 build).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <functional>
#include <string>
#include <vector>

#include "../../include/sba_b200.h"

namespace sba {

constexpr int kMaxDC = 18;  // 6 pose slots + 12 intrinsics slots
constexpr int kTile = 32;   // observations per warp tile

// 32-byte unit of the tile-blocked Jacobian storage; moved with one 256-bit LDG / STG (sm_100)
struct __align__(32) D4 { double x, y, z, w; };
#ifdef __CUDACC__
__device__ __forceinline__ D4 Ld256(const D4* p) {
  D4 v;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void St256(D4* p, const D4& v) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
}
#endif
__host__ __device__ inline size_t PairedUnit(int k, int u, int n) {
  const size_t base = (size_t)(k >> 5) * n * 32;
  const int lane = k & 31;
#ifdef SBA_UNPAIRED
  return base + (size_t)u * 32 + lane;
#endif
  return u < (n & ~1) ? base + (size_t)(u >> 1) * 64 + lane * 2 + (u & 1) : base + (size_t)(n - 1) * 32 + lane;
}
__host__ __device__ inline size_t FUnit(int k, int jj, int dch) { return PairedUnit(k, jj, dch); }
__host__ __device__ inline size_t EUnit(int k, int u) { return PairedUnit(k, u, 2); }
inline size_t TilePad(size_t n_obs) { return (n_obs + 31) / 32 * 32; }
// ---- third residual row (extension X1: the bilinear semantic row) ----
// A problem created with semantic maps carries THREE residual rows per observation (nr = 3).  The two reprojection
// rows keep the layout above; the third row lives in two streams of its own, so that a problem without maps
// (nr = 2) moves not one byte more than before:
//   ES[k]           = {d r_s/d X (3), r_s}                       one 32 B unit per observation, linear
//   FS unit q of k  = d r_s / d s_j for slots j = 4q .. 4q+3     tile-blocked like F, FSUnits(DC) units per observation
__host__ __device__ constexpr int FSUnits(int dc) { return (dc + 3) / 4; }
__host__ __device__ inline size_t FSUnit(int k, int q, int nq) { return PairedUnit(k, q, nq); }
#ifdef __CUDACC__
// (d r0 / d s, d r1 / d s) of one slot of one observation: half a unit
__device__ __forceinline__ double2 LoadSlot(const D4* __restrict__ F, int k, int slot, int dch) {
  return reinterpret_cast<const double2*>(F + FUnit(k, slot >> 1, dch))[slot & 1];
}
// d r_s / d s of one slot of one observation (third-row stream)
__device__ __forceinline__ double LoadSlotS(const D4* __restrict__ FS, int k, int slot, int nq) {
  return reinterpret_cast<const double*>(FS + FSUnit(k, slot >> 2, nq))[slot & 3];
}
#endif

struct Err {
  int code;
  std::string msg;
};
void SetError(int code, const std::string& msg);
#define SBA_CUDA(expr)                                                                     \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      ::sba::SetError(SBA_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));   \
      return SBA_ERR_CUDA;                                                                 \
    }                                                                                      \
  } while (0)

// Views passed by value to kernels --------------------------------------------------
struct ParamsView {
  const double* pose;  // [n_img*8]
  const double* cam;   // [n_cam*12]
  const double* X;     // [n_pt*3]
};

struct StructView {
  int n_img, n_cam, n_pt, n_obs;
  int dc;                     // padded slots per observation (6, 8, 10, 14, 18)
  int kmax;                   // max free intrinsics per camera
  int n_cdim;                 // compact camera-side dimension
  int n_pvar;                 // variable points
  int nr;                     // residual rows per observation: 2, or 3 with the semantic row
  const int* img_cam;         // [n_img]
  const int* model;           // [n_cam]
  const int* col_img;         // [n_img*6]  compact column of pose slot or -1
  const int* col_cam;         // [n_cam*12] compact column of intrinsics slot j (j<k_dim) or -1
  const int* k_cols;          // [n_cam*12] ambient param index of intrinsics slot j
  const int* k_slot;          // [n_cam*12] intrinsics slot of ambient param a, or -1 (constant)
  const int* k_dim;           // [n_cam]
  const int* p_off;           // [n_pt] variable-point index or -1
  const int* obs_img;         // [n_obs] point-sorted
  const int* obs_pt;          // [n_obs]
  const double* obs_xy;       // [n_obs*2]
  const uint8_t* obs_active;  // [n_obs] residual block kept in the reduced program
  const uint8_t* pose_const;  // [n_img]
};

// the third-row streams (null when the problem has two rows), the per-image maps they are sampled from
struct Row3View {
  D4* ES;             // [n_obs]
  D4* FS;             // [TilePad(n_obs) * FSUnits(dc)]
  const float* maps;  // [n_img * MapTiledSize(H, W)] top-down, in 8 x 4-texel tiles of 128 B (MapTexel below)
  int H, W;
  double weight;      // r_s = weight * bilinear(map_img, x, y)   (Bitmap::InterpolateBilinear convention, bitmap.cc:238-269)
};

// Label-map layout in HBM.  A bilinear sample reads a 2 x 2 texel footprint; in a row-major map its two rows are W * 4
// bytes apart, i.e. two DRAM bursts per observation.  Tiles of 8 (x) by 4 (rows) texels = 128 B keep the footprint in
// one burst for most samples (rows {0,1} and {2,3} of a tile share a 64 B half).
constexpr int kMapTileW = 8, kMapTileH = 4;
__host__ __device__ __forceinline__ size_t MapTiledSize(int H, int W) {
  return (size_t)((H + kMapTileH - 1) / kMapTileH) * ((W + kMapTileW - 1) / kMapTileW) * (kMapTileW * kMapTileH);
}
__host__ __device__ __forceinline__ size_t MapTexel(int row, int x, int tiles_per_row) {
  return ((size_t)(row >> 2) * tiles_per_row + (x >> 3)) * 32 + ((row & 3) << 3) + (x & 7);
}

struct LossView {
  int type;
  double scale;
  int apply;
};

// warp tiles over the point-sorted observation array: whole points packed into
// <=32 observations; a point longer than 32 gets a tile of its own.
struct TileView {
  int n_tiles;
  const int* tile_start;  // [n_tiles+1]
};

// camera-side chunks: runs of <=256 slots of the image-sorted index list that
// never straddle an image.
struct ChunkView {
  int n_chunks;
  const int* chunk_start;   // [n_chunks+1] into cam_order
  const int* chunk_img;     // [n_chunks]
  const int* cam_order;     // [n_obs] observation ids (point-sorted numbering), image-major
  const int* cam_po;        // [n_obs] variable-point offset of the same observations (-1: constant point)
  const int* img_chunk_ptr; // [n_img+1] chunk range of each image
  const int* cam_img_ptr;   // [n_cam+1] CSR: images of each camera
  const int* cam_img_list;  // [n_img]
};

// Host worker threads for the integer set-up work (sorting, pair lists, staging copies).
int HostThreads();
void ParallelFor(int n_tasks, const std::function<void(int)>& fn);  // fn(task), tasks spread over HostThreads()

// host phase timer, printed when SBA_TIMING is set (diagnostics only)
struct HostTimer {
  const char* what; double t0; bool on;
  static double Now();
  explicit HostTimer(const char* w);
  void Lap(const char* label);
};

inline int DivUp(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

}  // namespace sba
