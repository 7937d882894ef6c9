// Host-side flattening of the problem the reference builds in
//   BundleAdjuster::SetUp / AddImageToProblem / AddPointToProblem /
//   ParameterizeCameras / ParameterizePoints
//   (/root/reference/src/optim/bundle_adjustment.cc:326-530)
// into the device layout of common.cuh: reduced-program structure (which blocks
// are variable, tangent sizes), point-sorted observation order, warp tiles,
// image-major chunk lists, and the per-rank shard (points + their observations).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <atomic>
#include <mutex>
#include <numeric>
#include <thread>

#include "device_math.cuh"
#include "solver.h"

namespace sba {

static thread_local Err g_err = {0, ""};
void SetError(int code, const std::string& msg) { g_err.code = code; g_err.msg = msg; }
const char* LastErrorMessage() { return g_err.msg.c_str(); }

double HostTimer::Now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
HostTimer::HostTimer(const char* w) : what(w), t0(Now()) { static const bool e = getenv("SBA_TIMING") != nullptr; on = e; }
void HostTimer::Lap(const char* label) {
  if (!on) return;
  const double t = Now();
  fprintf(stderr, "[sba timing] %s / %s: %.3f ms\n", what, label, (t - t0) * 1e3);
  t0 = t;
}

int HostThreads() {
  static const int n = [] {
    const char* e = getenv("SBA_HOST_THREADS");
    int v = e ? atoi(e) : (int)std::thread::hardware_concurrency();
    return std::max(1, std::min(v, 16));
  }();
  return n;
}
void ParallelFor(int n_tasks, const std::function<void(int)>& fn) {
  const int nt = std::min(HostThreads(), n_tasks);
  if (nt <= 1) { for (int t = 0; t < n_tasks; ++t) fn(t); return; }
  std::atomic<int> next(0);
  auto work = [&]() { for (int t; (t = next.fetch_add(1)) < n_tasks;) fn(t); };
  std::vector<std::thread> th;
  for (int i = 1; i < nt; ++i) th.emplace_back(work);
  work();
  for (auto& t : th) t.join();
}

cudaStream_t AllocStream() {
  static cudaStream_t streams[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (!streams[dev]) {
    cudaStreamCreateWithFlags(&streams[dev], cudaStreamNonBlocking);
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
      uint64_t keep = ~0ull;  // never hand freed blocks back to the OS: the next handle reuses them
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
  }
  return streams[dev];
}

// A handle's stream and its pinned scalar block are recycled across handles of the same device:
// cudaMallocHost alone costs about a millisecond, which is a tenth of a whole C2-sized solve.
namespace {
struct Ctx { int device; cudaStream_t stream; DeviceScalars* pinned; };
std::mutex g_ctx_mu;
std::vector<Ctx> g_ctx_free;
}  // namespace
static cudaError_t AcquireCtx(int device, cudaStream_t* stream, DeviceScalars** pinned) {
  {
    std::lock_guard<std::mutex> lk(g_ctx_mu);
    for (size_t i = 0; i < g_ctx_free.size(); ++i)
      if (g_ctx_free[i].device == device) {
        *stream = g_ctx_free[i].stream; *pinned = g_ctx_free[i].pinned;
        g_ctx_free.erase(g_ctx_free.begin() + i);
        return cudaSuccess;
      }
  }
  cudaError_t e = cudaStreamCreateWithFlags(stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) return e;
  return cudaMallocHost((void**)pinned, sizeof(DeviceScalars));
}

// Large host -> device copies from the caller's PAGEABLE arrays (label maps: 61 MB at C2, 6 GB at C5).
// A plain cudaMemcpyAsync stages them through the driver on one thread (~8 GB/s measured); here four host
// threads copy 2 MB chunks into process-lifetime pinned buffers (two per thread) and queue the DMA of each
// chunk themselves, so the staging memcpy of one chunk overlaps the DMA of the others.
namespace {
constexpr size_t kStageChunk = 2u << 20;
constexpr int kStageThreads = 4, kStageDepth = 2;
struct StageBufs {
  char* buf[kStageThreads][kStageDepth] = {};
  cudaEvent_t ev[kStageThreads][kStageDepth] = {};
  bool ok = false;
};
std::mutex g_stage_mu;  // one large upload at a time per process
StageBufs g_stage;
}  // namespace
cudaError_t UploadLarge(void* dst, const void* src, size_t bytes, cudaStream_t st) {
  if (bytes < 4 * kStageChunk) return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st);
  std::lock_guard<std::mutex> lk(g_stage_mu);
  if (!g_stage.ok) {
    for (int t = 0; t < kStageThreads; ++t)
      for (int d = 0; d < kStageDepth; ++d) {
        cudaError_t e = cudaMallocHost((void**)&g_stage.buf[t][d], kStageChunk);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g_stage.ev[t][d], cudaEventDisableTiming);
        if (e != cudaSuccess) return e;
      }
    g_stage.ok = true;
  }
  const size_t n_chunks = (bytes + kStageChunk - 1) / kStageChunk;
  std::atomic<size_t> next(0);
  std::atomic<int> err((int)cudaSuccess);
  int dev = 0;
  cudaGetDevice(&dev);
  auto work = [&](int t) {
    cudaSetDevice(dev);
    int d = 0;
    for (size_t c; (c = next.fetch_add(1)) < n_chunks; d ^= 1) {
      const size_t off = c * kStageChunk, len = std::min(kStageChunk, bytes - off);
      cudaEventSynchronize(g_stage.ev[t][d]);  // the DMA that last read this buffer has finished
      std::memcpy(g_stage.buf[t][d], (const char*)src + off, len);
      cudaError_t e = cudaMemcpyAsync((char*)dst + off, g_stage.buf[t][d], len, cudaMemcpyHostToDevice, st);
      if (e == cudaSuccess) e = cudaEventRecord(g_stage.ev[t][d], st);
      if (e != cudaSuccess) err.store((int)e);
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < kStageThreads; ++t) th.emplace_back(work, t);
  work(0);
  for (auto& t : th) t.join();
  return (cudaError_t)err.load();
}

Problem::~Problem() {
  if (stream) cudaStreamSynchronize(stream);  // buffers go back to the pool only once nothing uses them
  CommDestroy(this);
  if (stream && h_scalars) {
    std::lock_guard<std::mutex> lk(g_ctx_mu);
    g_ctx_free.push_back(Ctx{device, stream, h_scalars});
  } else {
    if (h_scalars) cudaFreeHost(h_scalars);
    if (stream) cudaStreamDestroy(stream);
  }
}

ParamsView Problem::Params(bool candidate) const {
  ParamsView v;
  v.pose = candidate ? d_pose2.p : d_pose.p;
  v.cam = candidate ? d_cam2.p : d_cam.p;
  v.X = candidate ? d_X2.p : d_X.p;
  return v;
}
StructView Problem::Struct() const {
  StructView s;
  s.n_img = n_img; s.n_cam = n_cam; s.n_pt = n_pt; s.n_obs = n_obs;
  s.dc = dc; s.kmax = kmax; s.n_cdim = n_cdim; s.n_pvar = n_pvar;
  s.img_cam = d_img_cam.p; s.model = d_model.p; s.col_img = d_col_img.p; s.col_cam = d_col_cam.p;
  s.k_cols = d_k_cols.p; s.k_slot = d_k_slot.p; s.k_dim = d_k_dim.p; s.p_off = d_p_off.p;
  s.obs_img = d_obs_img.p; s.obs_pt = d_obs_pt.p; s.obs_xy = d_obs_xy.p;
  s.obs_active = d_obs_active.p; s.pose_const = d_pose_const.p;
  return s;
}
TileView Problem::Tiles() const { return TileView{n_tiles, d_tile_start.p}; }
ChunkView Problem::Chunks() const {
  ChunkView c;
  c.n_chunks = n_chunks; c.chunk_start = d_chunk_start.p; c.chunk_img = d_chunk_img.p;
  c.cam_order = d_cam_order.p; c.img_chunk_ptr = d_img_chunk_ptr.p;
  c.cam_img_ptr = d_cam_img_ptr.p; c.cam_img_list = d_cam_img_list.p;
  return c;
}

static int Validate(const sba_problem_desc* d) {
  if (!d) { SetError(SBA_ERR_INVALID_ARGUMENT, "null problem description"); return SBA_ERR_INVALID_ARGUMENT; }
  if (d->num_images < 0 || d->num_cameras < 0 || d->num_points < 0 || d->num_observations < 0 ||
      d->num_observations > 2000000000LL) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "negative or oversized counts"); return SBA_ERR_INVALID_ARGUMENT;
  }
  if (d->num_observations > 0 &&
      (!d->image_qvec || !d->image_tvec || !d->image_camera || !d->camera_model || !d->camera_params ||
       !d->points || !d->obs_image || !d->obs_point || !d->obs_xy)) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "null array in problem description"); return SBA_ERR_INVALID_ARGUMENT;
  }
  for (int c = 0; c < d->num_cameras; ++c)
    if (ModelNumParams(d->camera_model[c]) < 0) {
      SetError(SBA_ERR_UNKNOWN_CAMERA_MODEL, "camera model does not exist"); return SBA_ERR_UNKNOWN_CAMERA_MODEL;
    }
  for (int i = 0; i < d->num_images; ++i)
    if (d->image_camera[i] < 0 || d->image_camera[i] >= d->num_cameras) {
      SetError(SBA_ERR_INVALID_ARGUMENT, "image_camera out of range"); return SBA_ERR_INVALID_ARGUMENT;
    }
  for (int64_t k = 0; k < d->num_observations; ++k)
    if (d->obs_image[k] < 0 || d->obs_image[k] >= d->num_images || d->obs_point[k] < 0 ||
        d->obs_point[k] >= d->num_points) {
      SetError(SBA_ERR_INVALID_ARGUMENT, "observation index out of range"); return SBA_ERR_INVALID_ARGUMENT;
    }
  if (d->shard_count < 0 || (d->shard_count > 0 && (d->shard_rank < 0 || d->shard_rank >= d->shard_count))) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "bad shard_rank / shard_count"); return SBA_ERR_INVALID_ARGUMENT;
  }
  return SBA_OK;
}

// Global (un-sharded) reduced-program structure. Mirrors what Ceres derives
// from the parameter blocks / manifolds the reference sets up
// (bundle_adjustment.cc:417-425, 480-530); counts follow Solver::Summary.
struct GlobalStructure {
  std::vector<int> col_img, col_cam, k_cols, k_dim, t_dim, p_off_global, blk_off, blk_dim;
  std::vector<uint8_t> obs_active;
  int n_cdim = 0, kmax = 0, n_pvar_global = 0, uniform_model = -1;
  bool shared_intrinsics = false;
  sba_summary counts;
};

static void Analyse(const sba_problem_desc& d, GlobalStructure* g) {
  const int n_img = d.num_images, n_cam = d.num_cameras, n_pt = d.num_points;
  const int64_t n_obs = d.num_observations;
  std::vector<uint8_t> img_used(n_img, 0), cam_used(n_cam, 0), pt_used(n_pt, 0);
  std::vector<int> cam_img_count(n_cam, 0);
  for (int64_t k = 0; k < n_obs; ++k) {
    img_used[d.obs_image[k]] = 1; pt_used[d.obs_point[k]] = 1;
  }
  for (int i = 0; i < n_img; ++i) if (img_used[i]) { cam_used[d.image_camera[i]] = 1; cam_img_count[d.image_camera[i]]++; }
  std::memset(&g->counts, 0, sizeof(g->counts));
  sba_summary& s = g->counts;
  g->col_img.assign((size_t)n_img * 6, -1); g->col_cam.assign((size_t)n_cam * 12, -1);
  g->k_cols.assign((size_t)n_cam * 12, 0); g->k_dim.assign(n_cam, 0); g->t_dim.assign(n_img, 0);
  int off = 0;
  for (int i = 0; i < n_img; ++i) {
    const bool pc = d.image_pose_constant && d.image_pose_constant[i];
    if (!img_used[i] || pc) continue;
    for (int j = 0; j < 3; ++j) g->col_img[6 * (size_t)i + j] = off + j;
    g->blk_off.push_back(off); g->blk_dim.push_back(3); off += 3;
    s.num_parameter_blocks_reduced++; s.num_parameters_reduced += 4; s.num_effective_parameters_reduced += 3;
    const int mask = d.image_tvec_constant_mask ? d.image_tvec_constant_mask[i] : 0;
    int td = 0;
    for (int a = 0; a < 3; ++a) if (!((mask >> a) & 1)) g->col_img[6 * (size_t)i + 3 + a] = off + td++;
    g->t_dim[i] = td;
    if (td > 0) {
      g->blk_off.push_back(off); g->blk_dim.push_back(td); off += td;
      s.num_parameter_blocks_reduced++; s.num_parameters_reduced += 3; s.num_effective_parameters_reduced += td;
    }
  }
  int first_model = -2;
  for (int c = 0; c < n_cam; ++c) {
    if (!cam_used[c]) continue;
    if (first_model == -2) first_model = d.camera_model[c];
    else if (first_model != d.camera_model[c]) first_model = -1;
    const int P = ModelNumParams(d.camera_model[c]);
    const int mask = d.camera_param_constant_mask ? d.camera_param_constant_mask[c] : 0;
    int kd = 0;
    for (int j = 0; j < P; ++j) if (!((mask >> j) & 1)) { g->k_cols[12 * (size_t)c + kd] = j; g->col_cam[12 * (size_t)c + kd] = off + kd; kd++; }
    g->k_dim[c] = kd;
    if (kd > 0) {
      g->blk_off.push_back(off); g->blk_dim.push_back(kd); off += kd;
      s.num_parameter_blocks_reduced++; s.num_parameters_reduced += P; s.num_effective_parameters_reduced += kd;
      g->kmax = std::max(g->kmax, kd);
      if (cam_img_count[c] > 1) g->shared_intrinsics = true;
    }
  }
  g->uniform_model = first_model < 0 ? -1 : first_model;
  g->n_cdim = off;
  g->p_off_global.assign(n_pt, -1);
  for (int p = 0; p < n_pt; ++p) {
    if (!pt_used[p] || (d.point_constant && d.point_constant[p])) continue;
    g->p_off_global[p] = g->n_pvar_global++;
    s.num_parameter_blocks_reduced++; s.num_parameters_reduced += 3; s.num_effective_parameters_reduced += 3;
  }
  g->obs_active.assign(n_obs, 0);
  s.num_residual_blocks = (int)n_obs; s.num_residuals = (int)(2 * n_obs);
  for (int64_t k = 0; k < n_obs; ++k) {
    const int i = d.obs_image[k], c = d.image_camera[i], p = d.obs_point[k];
    const bool act = g->col_img[6 * (size_t)i] >= 0 || g->k_dim[c] > 0 || g->p_off_global[p] >= 0;
    g->obs_active[k] = act;
    if (act) s.num_residuals_reduced += 2;
  }
}

int ProblemCounts(const sba_problem_desc* d, sba_summary* s) {
  int rc = Validate(d);
  if (rc) return rc;
  GlobalStructure g;
  Analyse(*d, &g);
  *s = g.counts;
  return SBA_OK;
}

static int DcTier(int kmax) {
  if (kmax <= 0) return 6;
  if (kmax <= 2) return 8;
  if (kmax <= 4) return 10;
  if (kmax <= 8) return 14;
  return 18;
}

// contiguous point ranges balanced by observation count
void ShardPointRange(const std::vector<int64_t>& pt_obs_prefix, int nranks, int rank, int* p0, int* p1) {
  const int n_pt = (int)pt_obs_prefix.size() - 1;
  const int64_t total = pt_obs_prefix[n_pt];
  auto cut = [&](int r) -> int {
    if (r <= 0) return 0;
    if (r >= nranks) return n_pt;
    const int64_t target = total * r / nranks;
    return (int)(std::lower_bound(pt_obs_prefix.begin(), pt_obs_prefix.end(), target) - pt_obs_prefix.begin());
  };
  *p0 = std::min(cut(rank), n_pt);
  *p1 = std::min(std::max(cut(rank + 1), *p0), n_pt);
}

int BuildProblem(const sba_problem_desc* dp, Problem** out, bool host_only) {
  HostTimer tm("create");
  int rc = Validate(dp);
  if (rc) return rc;
  tm.Lap("validate");
  const sba_problem_desc& d = *dp;
  GlobalStructure g;
  Analyse(d, &g);
  tm.Lap("analyse");
  Problem* p = new Problem();
  p->counts = g.counts;
  p->n_img = d.num_images; p->n_cam = d.num_cameras;
  p->n_pt_global = d.num_points; p->n_obs_global = d.num_observations;
  p->kmax = g.kmax; p->dc = DcTier(g.kmax); p->n_cdim = g.n_cdim;
  p->uniform_model = g.uniform_model; p->shared_intrinsics = g.shared_intrinsics;
  p->shard_count = d.shard_count > 0 ? d.shard_count : 1;
  p->shard_rank = d.shard_count > 0 ? d.shard_rank : 0;
  p->h_img_cam.assign(d.image_camera, d.image_camera + p->n_img);
  p->h_model.assign(d.camera_model, d.camera_model + p->n_cam);
  p->h_col_img = g.col_img; p->h_col_cam = g.col_cam; p->h_k_cols = g.k_cols; p->h_k_dim = g.k_dim;
  p->h_k_slot.assign((size_t)p->n_cam * 12, -1);
  for (int c = 0; c < p->n_cam; ++c)
    for (int j = 0; j < g.k_dim[c]; ++j) p->h_k_slot[12 * (size_t)c + g.k_cols[12 * (size_t)c + j]] = j;
  p->h_blk_off = g.blk_off; p->h_blk_dim = g.blk_dim; p->h_t_dim = g.t_dim;
  p->h_pose_const.assign(p->n_img, 0);
  for (int i = 0; i < p->n_img; ++i) p->h_pose_const[i] = g.col_img[6 * (size_t)i] < 0 ? 1 : 0;

  // ---- shard: contiguous global point range balanced by observation count ----
  const int n_pt_g = d.num_points;
  const int64_t n_obs_g = d.num_observations;
  std::vector<int64_t> prefix(n_pt_g + 1, 0);
  for (int64_t k = 0; k < n_obs_g; ++k) prefix[d.obs_point[k] + 1]++;
  for (int q = 0; q < n_pt_g; ++q) prefix[q + 1] += prefix[q];
  int p0 = 0, p1 = n_pt_g;
  ShardPointRange(prefix, p->shard_count, p->shard_rank, &p0, &p1);
  p->n_pt = p1 - p0;
  p->n_obs = (int)(prefix[p1] - prefix[p0]);
  p->h_pt_local2global.resize(p->n_pt);
  std::iota(p->h_pt_local2global.begin(), p->h_pt_local2global.end(), p0);

  // ---- point-sorted local observations (stable in reference order) ----
  const int n_obs = p->n_obs;
  p->h_local2ref.resize(n_obs);
  {
    std::vector<int64_t> fill(prefix.begin() + p0, prefix.begin() + p1);
    const int64_t base = prefix[p0];
    for (int64_t k = 0; k < n_obs_g; ++k) {
      const int q = d.obs_point[k];
      if (q < p0 || q >= p1) continue;
      p->h_local2ref[fill[q - p0]++ - base] = k;
    }
  }
  p->h_obs_img.resize(n_obs); p->h_obs_pt.resize(n_obs);
  std::vector<double> h_xy((size_t)n_obs * 2);
  std::vector<uint8_t> h_active(n_obs);
  for (int k = 0; k < n_obs; ++k) {
    const int64_t r = p->h_local2ref[k];
    p->h_obs_img[k] = d.obs_image[r];
    p->h_obs_pt[k] = d.obs_point[r] - p0;
    h_xy[2 * (size_t)k] = d.obs_xy[2 * r]; h_xy[2 * (size_t)k + 1] = d.obs_xy[2 * r + 1];
    h_active[k] = g.obs_active[r];
  }
  p->h_p_off.assign(p->n_pt, -1);
  p->n_pvar = 0;
  for (int q = 0; q < p->n_pt; ++q) if (g.p_off_global[p0 + q] >= 0) p->h_p_off[q] = p->n_pvar++;
  std::vector<int>& h_pt_ptr = p->h_pt_ptr;
  h_pt_ptr.resize(p->n_pt + 1);
  for (int q = 0; q <= p->n_pt; ++q) h_pt_ptr[q] = (int)(prefix[p0 + q] - prefix[p0]);

  tm.Lap("shard + point sort");
  // ---- warp tiles: whole points packed into <= kTile observations ----
  std::vector<int> tile_start;
  {
    int cur = 0, fill = 0;
    tile_start.push_back(0);
    for (int q = 0; q < p->n_pt; ++q) {
      const int len = h_pt_ptr[q + 1] - h_pt_ptr[q];
      if (len == 0) continue;
      if (fill > 0 && fill + len > kTile) { tile_start.push_back(cur); fill = 0; }
      cur += len; fill += len;
      if (fill >= kTile) { tile_start.push_back(cur); fill = 0; }
    }
    if (tile_start.back() != cur) tile_start.push_back(cur);
  }
  p->n_tiles = (int)tile_start.size() - 1;

  // ---- image-major index list + chunks of <= 256 ----
  std::vector<int>& img_ptr = p->h_img_ptr; std::vector<int>& cam_order = p->h_cam_order;
  img_ptr.assign(p->n_img + 1, 0); cam_order.resize(n_obs);
  for (int k = 0; k < n_obs; ++k) img_ptr[p->h_obs_img[k] + 1]++;
  for (int i = 0; i < p->n_img; ++i) img_ptr[i + 1] += img_ptr[i];
  {
    std::vector<int> fill(img_ptr.begin(), img_ptr.end() - 1);
    for (int k = 0; k < n_obs; ++k) cam_order[fill[p->h_obs_img[k]]++] = k;
  }
  std::vector<int> chunk_start, chunk_img, img_chunk_ptr(p->n_img + 1, 0);
  for (int i = 0; i < p->n_img; ++i) {
    img_chunk_ptr[i] = (int)chunk_img.size();
    for (int s = img_ptr[i]; s < img_ptr[i + 1]; s += 256) { chunk_start.push_back(s); chunk_img.push_back(i); }
  }
  img_chunk_ptr[p->n_img] = (int)chunk_img.size();
  chunk_start.push_back(n_obs);
  p->n_chunks = (int)chunk_img.size();
  std::vector<int> cam_img_ptr(p->n_cam + 1, 0), cam_img_list(p->n_img);
  for (int i = 0; i < p->n_img; ++i) cam_img_ptr[p->h_img_cam[i] + 1]++;
  for (int c = 0; c < p->n_cam; ++c) cam_img_ptr[c + 1] += cam_img_ptr[c];
  {
    std::vector<int> fill(cam_img_ptr.begin(), cam_img_ptr.end() - 1);
    for (int i = 0; i < p->n_img; ++i) cam_img_list[fill[p->h_img_cam[i]]++] = i;
  }
  p->n_blk = (int)p->h_blk_off.size();
  std::vector<int> col_blk(std::max(1, p->n_cdim), 0);
  for (int b = 0; b < p->n_blk; ++b)
    for (int j = 0; j < p->h_blk_dim[b]; ++j) col_blk[p->h_blk_off[b] + j] = b;

  tm.Lap("tiles + chunks");
  *out = p;
  if (host_only) return SBA_OK;

  // ---- device ----
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    delete p; *out = nullptr;
    SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    return SBA_ERR_NO_DEVICE;
  }
  cudaGetDevice(&p->device);
  if (cudaDeviceGetAttribute(&p->n_sm, cudaDevAttrMultiProcessorCount, p->device) != cudaSuccess || p->n_sm <= 0) p->n_sm = 148;
#define UP(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { SetError(SBA_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); delete p; *out = nullptr; return SBA_ERR_CUDA; } } while (0)
  UP(AcquireCtx(p->device, &p->stream, &p->h_scalars));
  std::memset(p->h_scalars, 0, sizeof(DeviceScalars));
  cudaStream_t st = p->stream;
  std::vector<double> pose((size_t)p->n_img * 8, 0.0), cam((size_t)p->n_cam * 12, 0.0), X((size_t)p->n_pt * 3);
  for (int i = 0; i < p->n_img; ++i) {
    for (int j = 0; j < 4; ++j) pose[8 * (size_t)i + j] = d.image_qvec[4 * (size_t)i + j];
    for (int j = 0; j < 3; ++j) pose[8 * (size_t)i + 4 + j] = d.image_tvec[3 * (size_t)i + j];
  }
  for (size_t j = 0; j < cam.size(); ++j) cam[j] = d.camera_params[j];
  for (int q = 0; q < p->n_pt; ++q)
    for (int a = 0; a < 3; ++a) X[3 * (size_t)q + a] = d.points[3 * (size_t)(p0 + q) + a];
  tm.Lap("stream + pinned + pack");
  UP(p->d_pose.Upload(pose, st)); UP(p->d_cam.Upload(cam, st)); UP(p->d_X.Upload(X, st));
  UP(p->d_pose2.Upload(pose, st)); UP(p->d_cam2.Upload(cam, st)); UP(p->d_X2.Upload(X, st));
  UP(p->d_pose0.Upload(pose, st)); UP(p->d_cam0.Upload(cam, st)); UP(p->d_X0.Upload(X, st));
  UP(p->d_img_cam.Upload(p->h_img_cam, st)); UP(p->d_model.Upload(p->h_model, st));
  UP(p->d_col_img.Upload(p->h_col_img, st)); UP(p->d_col_cam.Upload(p->h_col_cam, st));
  UP(p->d_k_cols.Upload(p->h_k_cols, st)); UP(p->d_k_slot.Upload(p->h_k_slot, st)); UP(p->d_k_dim.Upload(p->h_k_dim, st));
  UP(p->d_p_off.Upload(p->h_p_off, st));
  UP(p->d_obs_img.Upload(p->h_obs_img, st)); UP(p->d_obs_pt.Upload(p->h_obs_pt, st));
  UP(p->d_obs_xy.Upload(h_xy, st)); UP(p->d_obs_active.Upload(h_active, st));
  UP(p->d_pose_const.Upload(p->h_pose_const, st));
  UP(p->d_tile_start.Upload(tile_start, st));
  UP(p->d_chunk_start.Upload(chunk_start, st)); UP(p->d_chunk_img.Upload(chunk_img, st));
  UP(p->d_cam_order.Upload(cam_order, st)); UP(p->d_img_chunk_ptr.Upload(img_chunk_ptr, st));
  UP(p->d_cam_img_ptr.Upload(cam_img_ptr, st)); UP(p->d_cam_img_list.Upload(cam_img_list, st));
  UP(p->d_pt_ptr.Upload(h_pt_ptr, st));
  UP(p->d_blk_off.Upload(p->h_blk_off, st)); UP(p->d_blk_dim.Upload(p->h_blk_dim, st));
  UP(p->d_col_blk.Upload(col_blk, st));
  tm.Lap("uploads");
  const size_t nc = std::max(1, p->n_cdim), np3 = std::max(1, 3 * p->n_pvar);
  UP(p->d_ER.Alloc(TilePad(std::max(1, n_obs)) * 2)); UP(p->d_F.Alloc(TilePad(std::max(1, n_obs)) * (p->dc / 2)));
  p->n_partial = std::max(std::max(DivUp(std::max(1, n_obs), 128), DivUp(std::max(1, p->n_tiles), 4)),
                          DivUp(p->n_img + p->n_cam + p->n_pt, 128)) + 8;
  UP(p->d_cost_partial.Alloc(4 * (size_t)p->n_partial));
  UP(p->d_xi.Alloc((size_t)std::max(1, p->n_img) * 6)); UP(p->d_xc.Alloc((size_t)std::max(1, p->n_cam) * 12));
  UP(p->d_scale_i.Alloc((size_t)std::max(1, p->n_img) * 6)); UP(p->d_scale_k.Alloc((size_t)std::max(1, p->n_cam) * 12));
  {  // 1 / 0 masks of the free slots (the scale of the unscaled evaluation) and the packed slot -> parameter map
    std::vector<double> mi((size_t)std::max(1, p->n_img) * 6, 0.0), mk((size_t)std::max(1, p->n_cam) * 12, 0.0);
    std::vector<unsigned long long> pack(std::max(1, p->n_cam), 0ull);
    for (size_t j = 0; j < (size_t)p->n_img * 6; ++j) mi[j] = p->h_col_img[j] >= 0 ? 1.0 : 0.0;
    for (int c = 0; c < p->n_cam; ++c)
      for (int j = 0; j < p->h_k_dim[c]; ++j) {
        mk[12 * (size_t)c + j] = p->h_col_cam[12 * (size_t)c + j] >= 0 ? 1.0 : 0.0;
        pack[c] |= (unsigned long long)(p->h_k_cols[12 * (size_t)c + j] & 15) << (4 * j);
      }
    UP(p->d_mask_i.Upload(mi, st)); UP(p->d_mask_k.Upload(mk, st)); UP(p->d_k_pack.Upload(pack, st));
    UP(cudaStreamSynchronize(st));  // the vectors above go out of scope
  }
  UP(p->d_red.Alloc(2048));
  UP(p->d_scale_c.Alloc(nc)); UP(p->d_scale_p.Alloc(np3));
  UP(p->d_sq_c.Alloc(nc)); UP(p->d_sq_p.Alloc(np3));
  UP(p->d_diag_c.Alloc(nc)); UP(p->d_diag_p.Alloc(np3)); UP(p->d_Dc.Alloc(nc)); UP(p->d_Dp.Alloc(np3));
  UP(p->d_Cinv.Alloc((size_t)std::max(1, p->n_pvar) * 6)); UP(p->d_gp.Alloc(np3));
  UP(p->d_chunk_partial.Alloc((size_t)std::max(1, p->n_chunks) * kMaxDC * 2));
  UP(p->d_g2.Alloc((size_t)std::max(1, n_obs) * 2));
  UP(p->d_rhs.Alloc(nc)); UP(p->d_yc.Alloc(nc)); UP(p->d_yp.Alloc(np3)); UP(p->d_step_c.Alloc(nc));
  UP(p->d_grad_c.Alloc(nc)); UP(p->d_grad_p.Alloc(np3)); UP(p->d_tmp_c.Alloc(nc));
  UP(p->d_cg_r.Alloc(nc)); UP(p->d_cg_z.Alloc(nc)); UP(p->d_cg_p.Alloc(nc)); UP(p->d_cg_q.Alloc(nc)); UP(p->d_cg_x.Alloc(nc));
  UP(p->d_scalars.Alloc(1)); UP(p->d_scalars.Zero(st));
  UP(p->d_tickets.Alloc(8)); UP(p->d_tickets.Zero(st));
  tm.Lap("allocs");
  if (d.sem_maps && d.sem_height > 0 && d.sem_width > 0) {
    const size_t n_map = (size_t)p->n_img * d.sem_height * d.sem_width;
    UP(p->d_sem_maps.Alloc(n_map));
    UP(UploadLarge(p->d_sem_maps.p, d.sem_maps, n_map * sizeof(float), st));
    p->sem_h = d.sem_height; p->sem_w = d.sem_width; p->sem_weight = d.sem_weight;
  }
  UP(cudaStreamSynchronize(st));
  tm.Lap("maps + sync");
#undef UP
  return SBA_OK;
}

}  // namespace sba
