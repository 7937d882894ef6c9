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
#include <condition_variable>
#include <map>
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
// Persistent worker pool: the set-up passes below are a few hundred microseconds each, so spawning threads per
// pass (tens of microseconds per thread) would eat the gain.  Workers live for the life of the process.
namespace {
class Pool {
 public:
  explicit Pool(int n_workers) {
    for (int i = 0; i < n_workers; ++i) std::thread([this] { Loop(); }).detach();
  }
  void Run(int n_tasks, const std::function<void(int)>& fn) {
    {
      std::lock_guard<std::mutex> lk(mu_);
      fn_ = &fn; n_tasks_ = n_tasks; next_.store(0); active_ = 0; ++gen_;
    }
    cv_.notify_all();
    for (int t; (t = next_.fetch_add(1)) < n_tasks;) fn(t);
    std::unique_lock<std::mutex> lk(mu_);
    fn_ = nullptr;                                   // late wakers find no job
    done_.wait(lk, [this] { return active_ == 0; });
  }
 private:
  void Loop() {
    uint64_t seen = 0;
    for (;;) {
      const std::function<void(int)>* fn;
      int n;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return gen_ != seen; });
        seen = gen_;
        if (!fn_) continue;
        fn = fn_; n = n_tasks_; ++active_;
      }
      for (int t; (t = next_.fetch_add(1)) < n;) (*fn)(t);
      {
        std::lock_guard<std::mutex> lk(mu_);
        --active_;
      }
      done_.notify_all();
    }
  }
  std::mutex mu_;
  std::condition_variable cv_, done_;
  const std::function<void(int)>* fn_ = nullptr;
  int n_tasks_ = 0, active_ = 0;
  uint64_t gen_ = 0;
  std::atomic<int> next_{0};
};
std::mutex g_pool_mu;  // one parallel pass at a time per process; a second caller runs its pass inline
}  // namespace
void ParallelFor(int n_tasks, const std::function<void(int)>& fn) {
  const int nt = std::min(HostThreads(), n_tasks);
  std::unique_lock<std::mutex> lk(g_pool_mu, std::try_to_lock);
  if (nt <= 1 || !lk.owns_lock()) { for (int t = 0; t < n_tasks; ++t) fn(t); return; }
  static Pool* pool = new Pool(HostThreads() - 1);  // never destroyed: workers may outlive static destruction
  pool->Run(n_tasks, fn);
}
// [begin, end) of slice `t` of `n` items cut into `parts` contiguous slices
static inline void Slice(int64_t n, int parts, int t, int64_t* b, int64_t* e) { *b = n * t / parts; *e = n * (t + 1) / parts; }

// Side stream for the large map upload, so that its DMA overlaps the host set-up and the small uploads.
static cudaStream_t CopyStream() {
  static cudaStream_t streams[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (!streams[dev]) cudaStreamCreateWithFlags(&streams[dev], cudaStreamNonBlocking);
  return streams[dev];
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

namespace {
struct BlockCache {
  std::mutex mu;
  std::multimap<size_t, void*> free_blocks[64];
  size_t cached_bytes[64] = {};
};
BlockCache g_blocks;
}  // namespace

// SBA_POISON=1 (diagnostics): every block handed out is filled with 0xFF bytes (a NaN in every double, -1 in every
// int) first, so that a kernel reading memory nobody wrote fails the parity tests loudly instead of depending on
// what the previous owner left there.
// SBA_POISON_FROM / SBA_POISON_TO restrict it to the allocations with ordinal in [from, to) (bisecting for the buffer a
// result depends on); SBA_POISON_LOG=1 prints ordinal and size of every allocation.
static cudaError_t Poison(void* ptr, size_t bytes) {
  static const bool on = getenv("SBA_POISON") != nullptr;
  if (!on) return cudaSuccess;
  static const long from = getenv("SBA_POISON_FROM") ? atol(getenv("SBA_POISON_FROM")) : 0;
  static const long to = getenv("SBA_POISON_TO") ? atol(getenv("SBA_POISON_TO")) : (1l << 60);
  static const bool log = getenv("SBA_POISON_LOG") != nullptr;
  static std::atomic<long> ordinal{0};
  const long k = ordinal.fetch_add(1);
  if (log) fprintf(stderr, "[sba poison] alloc %ld: %zu bytes%s\n", k, bytes, (k >= from && k < to) ? " (poisoned)" : "");
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaMemset(ptr, (k >= from && k < to) ? 0xFF : 0x00, bytes);
  return e == cudaSuccess ? cudaDeviceSynchronize() : e;
}

cudaError_t CachedAlloc(void** ptr, size_t bytes, size_t* cap) {
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  bytes = (bytes + 511) / 512 * 512;
  {
    std::lock_guard<std::mutex> lk(g_blocks.mu);
    auto& fb = g_blocks.free_blocks[dev];
    auto it = fb.lower_bound(bytes);
    if (it != fb.end() && it->first <= bytes + bytes / 8 + 65536) {  // a block at most ~12 % larger
      *ptr = it->second; *cap = it->first;
      g_blocks.cached_bytes[dev] -= it->first;
      fb.erase(it);
      return Poison(*ptr, *cap);
    }
  }
  *cap = bytes;
  cudaError_t e = cudaMallocAsync(ptr, bytes, AllocStream());
  if (e != cudaSuccess) {  // out of memory with blocks parked in the cache: give them back and retry once
    cudaGetLastError();
    ReleaseCachedBlocks();
    e = cudaMallocAsync(ptr, bytes, AllocStream());
  }
  // the allocation stream carries no kernels, so this returns at once; it makes the block usable from any stream
  if (e == cudaSuccess) e = cudaStreamSynchronize(AllocStream());
  return e == cudaSuccess ? Poison(*ptr, bytes) : e;
}

// Contract (DevBuf): the stream that used the block has been synchronised, so any stream may reuse it at once.
void CachedFree(void* ptr, size_t cap) {
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  size_t free_b = 0, total_b = 0;
  static size_t limit[64] = {};
  if (!limit[dev] && cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) limit[dev] = total_b / 3;
  std::lock_guard<std::mutex> lk(g_blocks.mu);
  if (cap == 0 || g_blocks.cached_bytes[dev] + cap > limit[dev]) { cudaFreeAsync(ptr, AllocStream()); return; }
  g_blocks.free_blocks[dev].emplace(cap, ptr);
  g_blocks.cached_bytes[dev] += cap;
}

void ReleaseCachedBlocks() {
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  std::lock_guard<std::mutex> lk(g_blocks.mu);
  for (auto& kv : g_blocks.free_blocks[dev]) cudaFreeAsync(kv.second, AllocStream());
  g_blocks.free_blocks[dev].clear();
  g_blocks.cached_bytes[dev] = 0;
  cudaStreamSynchronize(AllocStream());
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
  {  // a caller that keeps its arrays in page-locked memory (sba_host_alloc, cudaHostRegister) needs no staging
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeHost)
      return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st);
    cudaGetLastError();
  }
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

// The way back: device -> the caller's PAGEABLE array (sba_get_parameters: 24 MB of points at C4, 48 MB at C5).  The
// DMA of each 2 MB chunk lands in a pinned buffer and a host thread copies it out while the next chunks are in flight;
// a plain cudaMemcpy into pageable memory measured 14 ms for the 24 MB.  Synchronous: returns when dst is complete.
cudaError_t DownloadLarge(void* dst, const void* src, size_t bytes, cudaStream_t st) {
  if (bytes == 0) return cudaSuccess;
  {
    cudaPointerAttributes at;
    if (bytes < 4 * kStageChunk || (cudaPointerGetAttributes(&at, dst) == cudaSuccess && at.type == cudaMemoryTypeHost)) {
      cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st);
      return e == cudaSuccess ? cudaStreamSynchronize(st) : e;
    }
    cudaGetLastError();
  }
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
  cudaError_t e0 = cudaStreamSynchronize(st);  // src is complete; the chunk copies below use per-thread side streams
  if (e0 != cudaSuccess) return e0;
  const size_t n_chunks = (bytes + kStageChunk - 1) / kStageChunk;
  std::atomic<size_t> next(0);
  std::atomic<int> err((int)cudaSuccess);
  int dev = 0;
  cudaGetDevice(&dev);
  auto work = [&](int t) {
    cudaSetDevice(dev);
    cudaStream_t cs;
    if (cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking) != cudaSuccess) { err.store((int)cudaErrorUnknown); return; }
    size_t pending[kStageDepth] = {}; bool has[kStageDepth] = {};
    auto drain = [&](int d) {
      if (!has[d]) return;
      cudaError_t e = cudaEventSynchronize(g_stage.ev[t][d]);
      if (e != cudaSuccess) err.store((int)e);
      const size_t off = pending[d] * kStageChunk, len = std::min(kStageChunk, bytes - off);
      std::memcpy((char*)dst + off, g_stage.buf[t][d], len);
      has[d] = false;
    };
    int d = 0;
    for (size_t c; (c = next.fetch_add(1)) < n_chunks; d = (d + 1) % kStageDepth) {
      drain(d);  // the buffer's previous chunk has been copied out
      const size_t off = c * kStageChunk, len = std::min(kStageChunk, bytes - off);
      cudaError_t e = cudaMemcpyAsync(g_stage.buf[t][d], (const char*)src + off, len, cudaMemcpyDeviceToHost, cs);
      if (e == cudaSuccess) e = cudaEventRecord(g_stage.ev[t][d], cs);
      if (e != cudaSuccess) err.store((int)e);
      pending[d] = c; has[d] = true;
    }
    for (int k = 0; k < kStageDepth; ++k) drain((d + k) % kStageDepth);
    cudaStreamDestroy(cs);
  };
  std::vector<std::thread> th;
  for (int t = 1; t < kStageThreads; ++t) th.emplace_back(work, t);
  work(0);
  for (auto& t : th) t.join();
  return (cudaError_t)err.load();
}

void Problem::DropGraphs() {
  for (auto& g : graph_exec) { if (g) cudaGraphExecDestroy(g); g = nullptr; }
  graph_opts_valid = false;
}

Problem::~Problem() {
  if (stream) cudaStreamSynchronize(stream);  // buffers go back to the pool only once nothing uses them
  DropGraphs();
  CommDestroy(this);
  if (shared_maps) { MapSetRelease(shared_maps); shared_maps = nullptr; }
  if (stream2) { cudaStreamSynchronize(stream2); cudaStreamDestroy(stream2); }
  if (stream3) { cudaStreamSynchronize(stream3); cudaStreamDestroy(stream3); }
  for (auto& e : ev_chol) if (e) cudaEventDestroy(e);
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
  s.dc = dc; s.kmax = kmax; s.n_cdim = n_cdim; s.n_pvar = n_pvar; s.nr = nr;
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
  c.cam_order = d_cam_order.p; c.cam_po = d_cam_po.p; c.img_chunk_ptr = d_img_chunk_ptr.p;
  c.cam_img_ptr = d_cam_img_ptr.p; c.cam_img_list = d_cam_img_list.p;
  return c;
}

// check_observations = false: the caller checks the index ranges of the observation list on the device (sba_create: a
// multi-threaded host pass over obs_image / obs_point leaves their cache lines spread over the cores' private caches,
// and the DMA that follows then reads them at 10-15 GB/s instead of 55 -- 3.4 ms of a 12.9 ms create at C4)
static int Validate(const sba_problem_desc* d, bool check_observations = true) {
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
  if (check_observations) {
    const int64_t n = d->num_observations;
    const int NT = std::max(1, std::min(HostThreads(), (int)(n / 16384) + 1));
    std::atomic<int> bad(0);
    ParallelFor(NT, [&](int t) {
      int64_t kb, ke; Slice(n, NT, t, &kb, &ke);
      const unsigned ni = (unsigned)d->num_images, np = (unsigned)d->num_points;
      int b = 0;
      for (int64_t k = kb; k < ke; ++k) b |= ((unsigned)d->obs_image[k] >= ni) | ((unsigned)d->obs_point[k] >= np);
      if (b) bad.store(1);
    });
    if (bad.load()) { SetError(SBA_ERR_INVALID_ARGUMENT, "observation index out of range"); return SBA_ERR_INVALID_ARGUMENT; }
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
  std::vector<int> col_img, col_cam, k_cols, k_dim, t_dim, blk_off, blk_dim;
  // per point of the whole problem: offset among the variable points (or -1) and the same as a byte mask; in the
  // caller's buffers (the page-locked scratch of sba_create) or in the two vectors
  int* p_off = nullptr;
  uint8_t* pt_live = nullptr;
  std::vector<int> p_off_store;
  std::vector<uint8_t> pt_live_store;
  int n_cdim = 0, kmax = 0, n_pvar_global = 0, uniform_model = -1;
  bool shared_intrinsics = false;
  sba_summary counts;
};

// ---- set-up passes over the observation list, made on the device from the raw copy of the caller's arrays ----
// (at C4 the three host passes they replace -- usage marks, counts per point, counts per image -- were 28 of the 38 ms
//  of sba_create)
// (both histogram kernels also validate the index ranges: *bad becomes 1, nothing is counted out of bounds)
__global__ void __launch_bounds__(256) hist_point_kernel(int64_t n, const int* __restrict__ pt, int n_pt, int* __restrict__ cnt,
                                                         int* __restrict__ bad) {
  const int64_t k = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (k >= n) return;
  const int q = pt[k];
  if ((unsigned)q < (unsigned)n_pt) atomicAdd(cnt + q, 1);
  else *bad = 1;
}
// per-image counts of the observations whose point lies in [p0, p1); a CTA counts into shared memory first when the
// image table fits (n_img <= 8192), so the global atomics are one per image per CTA
__global__ void __launch_bounds__(256) hist_image_kernel(int64_t n, const int* __restrict__ img, const int* __restrict__ pt,
                                                         int p0, int p1, int n_img, int* __restrict__ cnt, int* __restrict__ bad) {
  extern __shared__ int sh[];
  const bool use_sh = n_img <= 8192;
  if (use_sh) { for (int i = threadIdx.x; i < n_img; i += 256) sh[i] = 0; __syncthreads(); }
  for (int64_t k = (int64_t)blockIdx.x * 256 + threadIdx.x; k < n; k += (int64_t)gridDim.x * 256) {
    const int q = pt[k];
    const int i = img[k];
    if ((unsigned)i >= (unsigned)n_img) { if (bad) *bad = 1; continue; }
    if (q < p0 || q >= p1) continue;
    if (use_sh) atomicAdd(sh + i, 1); else atomicAdd(cnt + i, 1);
  }
  if (use_sh) {
    __syncthreads();
    for (int i = threadIdx.x; i < n_img; i += 256) if (sh[i]) atomicAdd(cnt + i, sh[i]);
  }
}
// residual blocks kept in the reduced program: some parameter block of the observation is variable
__global__ void __launch_bounds__(256) count_active_kernel(int64_t n, const int* __restrict__ img, const int* __restrict__ pt,
                                                           const uint8_t* __restrict__ img_live, const uint8_t* __restrict__ pt_live,
                                                           unsigned long long* __restrict__ out) {
  unsigned c = 0;
  for (int64_t k = (int64_t)blockIdx.x * 256 + threadIdx.x; k < n; k += (int64_t)gridDim.x * 256)
    c += (img_live[img[k]] | pt_live[pt[k]]) ? 1u : 0u;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, (unsigned long long)c);
}

static cudaError_t DeviceHistogramImages(cudaStream_t st, int64_t n, const int* d_img, const int* d_pt, int p0, int p1,
                                         int n_img, int* d_cnt, int n_sm, int* d_bad = nullptr) {
  cudaError_t e = cudaMemsetAsync(d_cnt, 0, (size_t)std::max(1, n_img) * sizeof(int), st);
  if (e != cudaSuccess || n == 0) return e;
  const int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)n_sm * 8);
  hist_image_kernel<<<blocks, 256, n_img <= 8192 ? (size_t)n_img * sizeof(int) : 0, st>>>(n, d_img, d_pt, p0, p1, n_img, d_cnt, d_bad);
  return cudaGetLastError();
}

// residual rows per block: the two reprojection rows, plus the semantic row when the problem carries maps (X1)
static int RowsOf(const sba_problem_desc& d, const MapSet* shared = nullptr) {
  if (shared) return 3;
  return (d.sem_maps && d.sem_height > 0 && d.sem_width > 0 && d.num_images > 0) ? 3 : 2;
}

// pt_count / img_count: observations per point / per image of the WHOLE problem.  With a device they come from two
// histogram kernels over the caller's arrays, which are on the device anyway (DeviceHistograms); the host-only entry
// point (sba_problem_counts) counts here.  count_active: also count the residual blocks of the reduced program on the
// host (a pass over the observation list); with a device that count is made there too (CountActiveOnDevice).
static void Analyse(const sba_problem_desc& d, GlobalStructure* g, const int* pt_count, const int* img_count,
                    bool count_active, const MapSet* shared = nullptr, int* p_off_out = nullptr, uint8_t* pt_live_out = nullptr) {
  const int n_img = d.num_images, n_cam = d.num_cameras, n_pt = d.num_points;
  const int64_t n_obs = d.num_observations;
  std::vector<uint8_t> img_used(n_img, 0), cam_used(n_cam, 0), pt_used;
  std::vector<int> cam_img_count(n_cam, 0);
  const int NT = std::max(1, std::min(HostThreads(), (int)(n_obs / 16384) + 1));
  if (pt_count && img_count) {
    for (int i = 0; i < n_img; ++i) img_used[i] = img_count[i] > 0;  // (the points: one pass below, straight from pt_count)
  } else {
    pt_used.assign(n_pt, 0);
    // every writer stores the same value 1 into a byte: the passes can share the mark arrays
    ParallelFor(NT, [&](int t) {
      int64_t kb, ke; Slice(n_obs, NT, t, &kb, &ke);
      for (int64_t k = kb; k < ke; ++k) {
        uint8_t& iu = img_used[d.obs_image[k]]; if (!iu) iu = 1;
        uint8_t& pu = pt_used[d.obs_point[k]]; if (!pu) pu = 1;
      }
    });
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
  // one pass over the points (2 M at C5: every extra host pass is ~2 ms of sba_create): offsets of the variable ones and the
  // byte mask the device's residual count reads
  if (p_off_out && pt_live_out) { g->p_off = p_off_out; g->pt_live = pt_live_out; }
  else {
    g->p_off_store.resize(std::max(1, n_pt)); g->pt_live_store.resize(std::max(1, n_pt));
    g->p_off = g->p_off_store.data(); g->pt_live = g->pt_live_store.data();
  }
  {
    const uint8_t* pc = d.point_constant;
    const uint8_t* used = pt_used.empty() ? nullptr : pt_used.data();
    int nv = 0;
    for (int p = 0; p < n_pt; ++p) {
      const bool var = (used ? used[p] != 0 : pt_count[p] > 0) && !(pc && pc[p]);
      g->p_off[p] = var ? nv++ : -1;
      g->pt_live[p] = var ? 1 : 0;
    }
    g->n_pvar_global = nv;
    s.num_parameter_blocks_reduced += nv; s.num_parameters_reduced += 3 * nv; s.num_effective_parameters_reduced += 3 * nv;
  }
  const int nr = RowsOf(d, shared);
  s.num_residual_blocks = (int)n_obs; s.num_residuals = (int)(nr * n_obs);
  if (!count_active) return;
  std::vector<int64_t> n_act(NT, 0);
  ParallelFor(NT, [&](int t) {
    int64_t kb, ke, na = 0; Slice(n_obs, NT, t, &kb, &ke);
    for (int64_t k = kb; k < ke; ++k) {
      const int i = d.obs_image[k], c = d.image_camera[i], p = d.obs_point[k];
      na += g->col_img[6 * (size_t)i] >= 0 || g->k_dim[c] > 0 || g->p_off[p] >= 0;
    }
    n_act[t] = na;
  });
  for (int t = 0; t < NT; ++t) s.num_residuals_reduced += (int)(nr * n_act[t]);
}

int ProblemCounts(const sba_problem_desc* d, sba_summary* s) {
  int rc = Validate(d);
  if (rc) return rc;
  GlobalStructure g;
  Analyse(*d, &g, nullptr, nullptr, true);
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

// row-major maps -> the tiled layout of common.cuh (MapTexel); padding texels are never sampled and stay 0
__global__ void retile_maps_kernel(const float* __restrict__ src, float* __restrict__ dst, int n_img, int H, int W) {
  const int tpr = (W + kMapTileW - 1) / kMapTileW;
  const size_t per = MapTiledSize(H, W), total = per * n_img;
  for (size_t o = blockIdx.x * (size_t)blockDim.x + threadIdx.x; o < total; o += (size_t)gridDim.x * blockDim.x) {
    const size_t img = o / per, w = o - img * per;
    const size_t tile = w >> 5;
    const int in = (int)(w & 31);
    const int row = (int)(tile / tpr) * kMapTileH + (in >> 3), x = (int)(tile % tpr) * kMapTileW + (in & 7);
    dst[o] = (row < H && x < W) ? __ldg(src + (img * H + row) * (size_t)W + x) : 0.0f;
  }
}

// ---- label maps kept on the device across problems (sba_maps_create) ----
// The reference's semantic adjusters read every image's map once and keep it for all their cost functors
// (SemanticBundleAdjuster::ReadDepthAndSemanticMaps, /root/reference/src/optim/semantic_bundle_adjustment.cc:1020-1066;
// semantic_maps_, semantic_bundle_adjustment.h:260); the maps do not change between the mapper's adjustments.  At C5
// they are 6.5 GB, and moving them host -> device is ~110 of the 220 ms of a create + solve + read-back.
void MapSetRelease(MapSet* m) {
  if (m && m->refs.fetch_sub(1) == 1) {
    int prev = -1;
    const int dev = m->device;
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    cudaDeviceSynchronize();  // no kernel of a destroyed handle may still read the block when it returns to the pool
    delete m;
    if (prev >= 0 && prev != dev) cudaSetDevice(prev);
  }
}

int BuildMapSet(int n_img, int h, int w, const float* maps, MapSet** out) {
  *out = nullptr;
  if (n_img <= 0 || h <= 0 || w <= 0 || !maps) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "sba_maps_create: num_images, height, width must be positive and maps non-null");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    return SBA_ERR_NO_DEVICE;
  }
  MapSet* m = new MapSet();
  m->n_img = n_img; m->h = h; m->w = w;
  cudaGetDevice(&m->device);
  const size_t n_map = (size_t)n_img * h * w, n_out = (size_t)n_img * MapTiledSize(h, w);
  DevBuf<float> stage;
  cudaStream_t cs = CopyStream();
  cudaError_t e = stage.Alloc(n_map);
  if (e == cudaSuccess) e = m->tiled.Alloc(n_out);
  if (e == cudaSuccess) e = UploadLarge(stage.p, maps, n_map * sizeof(float), cs);
  if (e == cudaSuccess) {
    retile_maps_kernel<<<(unsigned)std::min<size_t>(DivUp(n_out, 256), 148 * 64), 256, 0, cs>>>(stage.p, m->tiled.p, n_img, h, w);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(cs);
  else cudaStreamSynchronize(cs);
  if (e != cudaSuccess) {
    SetError(SBA_ERR_CUDA, std::string("sba_maps_create: ") + cudaGetErrorString(e));
    delete m;
    return SBA_ERR_CUDA;
  }
  *out = m;
  return SBA_OK;
}

int BuildProblem(const sba_problem_desc* dp, Problem** out, bool host_only, MapSet* shared) {
  HostTimer tm("create");
  int rc = Validate(dp, host_only);  // with a device the observation indices are checked there (histogram kernels)
  if (rc) return rc;
  if (shared) {  // label maps already on the device (sba_maps_create): the descriptor must not carry a second set
    int dev = -1;
    cudaGetDevice(&dev);
    if (host_only || dp->sem_maps || shared->n_img != dp->num_images || shared->device != dev) {
      SetError(SBA_ERR_INVALID_ARGUMENT, dp->sem_maps ? "sba_create_with_maps: desc->sem_maps must be null when a resident map set is given"
               : shared->n_img != dp->num_images ? "sba_create_with_maps: the map set holds a different number of images than the problem"
                                                 : "sba_create_with_maps: the map set lives on another device than the current one");
      return SBA_ERR_INVALID_ARGUMENT;
    }
  }
  tm.Lap("validate");
  const sba_problem_desc& d = *dp;
  Problem* p = new Problem();
  p->nr = RowsOf(d, shared);  // 3: the semantic row is part of every residual block (evaluation AND solve)
  // ---- device first: the label maps (61 MB at C2, 6 GB at C5) start moving on a side stream right away, so their
  //      DMA runs under the host set-up below instead of after it ----
  DevBuf<int> raw_img, raw_pt;   // the caller's observation arrays on the device (released after the final sync)
  DevBuf<double> raw_xy;
  DevBuf<float> map_stage;       // the caller's row-major label maps; re-tiled into d_sem_maps after the upload
#define UP(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { SetError(SBA_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); maps.Join(); xy_up.Join(); cudaStreamSynchronize(CopyStream()); if (p->stream) cudaStreamSynchronize(p->stream); delete p; *out = nullptr; return SBA_ERR_CUDA; } } while (0)
  struct MapUpload {
    std::thread th;
    cudaError_t err = cudaSuccess;
    void Join() { if (th.joinable()) th.join(); }
    ~MapUpload() { Join(); }
  } maps, xy_up;  // xy_up: obs_xy out of PAGEABLE memory (staged by host threads) runs under the host's point passes
  if (!host_only) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
      delete p;
      SetError(SBA_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
      return SBA_ERR_NO_DEVICE;
    }
    cudaGetDevice(&p->device);
    {
      cudaError_t e = AcquireCtx(p->device, &p->stream, &p->h_scalars);
      const size_t no = (size_t)std::max<int64_t>(1, d.num_observations);
      if (e == cudaSuccess) e = raw_img.Alloc(no);
      if (e == cudaSuccess) e = raw_pt.Alloc(no);
      if (e == cudaSuccess) e = raw_xy.Alloc(2 * no);
      if (e == cudaSuccess && d.num_observations > 0) {
        // the caller's observation arrays go up as they are; the device sorts them (order.cu)
        e = UploadLarge(raw_pt.p, d.obs_point, no * sizeof(int), p->stream);
        if (e == cudaSuccess) e = UploadLarge(raw_img.p, d.obs_image, no * sizeof(int), p->stream);
        // (obs_xy -- two thirds of the bytes -- follows the histogram kernels below: its DMA runs under the host's point passes)
      }
      if (e != cudaSuccess) { SetError(SBA_ERR_CUDA, std::string("observation upload: ") + cudaGetErrorString(e)); delete p; return SBA_ERR_CUDA; }
      std::memset(p->h_scalars, 0, sizeof(DeviceScalars));
    }
    if (shared) {  // a view: the set is reference-counted and outlives the handle
      shared->refs.fetch_add(1);
      p->shared_maps = shared;
      p->d_sem_maps.View(shared->tiled.p, shared->tiled.n);
      p->sem_h = shared->h; p->sem_w = shared->w; p->sem_weight = d.sem_weight;
    } else if (d.sem_maps && d.sem_height > 0 && d.sem_width > 0 && d.num_images > 0) {
      const size_t n_map = (size_t)d.num_images * d.sem_height * d.sem_width;
      cudaError_t e = map_stage.Alloc(n_map);
      if (e == cudaSuccess) e = p->d_sem_maps.Alloc((size_t)d.num_images * MapTiledSize(d.sem_height, d.sem_width));
      if (e != cudaSuccess) { SetError(SBA_ERR_CUDA, std::string("label maps: ") + cudaGetErrorString(e)); delete p; return SBA_ERR_CUDA; }
      p->sem_h = d.sem_height; p->sem_w = d.sem_width; p->sem_weight = d.sem_weight;
      const int dev = p->device;
      float* dst = map_stage.p;
      const float* src = d.sem_maps;
      cudaStream_t cs = CopyStream();
      maps.th = std::thread([&maps, dev, dst, src, n_map, cs] {
        cudaSetDevice(dev);
        maps.err = UploadLarge(dst, src, n_map * sizeof(float), cs);
      });
    }
  }
  tm.Lap("start map upload");
  // counts per point and per image of the whole problem: device histograms over the raw copy (host loops without a device)
  const int n_pt_g = d.num_points;
  const int64_t n_obs_g = d.num_observations;
  // Per-point host arrays of the whole problem (counts, offsets, live mask) live in a page-locked scratch block kept per
  // process; the CSR over points, the offsets and the tiles are written straight into the arena's staging block.  The
  // stage mutex is held from here until the arena upload has left the block (one create at a time per process through
  // this stretch).
  PinnedStage& stg = PinnedStage::Get();
  std::unique_lock<std::mutex> stage_lock(stg.mu, std::defer_lock);
  const size_t npg = (size_t)std::max(1, n_pt_g), npg4 = (npg * sizeof(int) + 255) / 256 * 256;
  std::vector<int> v_cnt, img_count((size_t)std::max(1, d.num_images), 0);
  int* pt_count = nullptr;
  int* p_off_scratch = nullptr;
  uint8_t* live_scratch = nullptr;
  DevBuf<int> d_pt_count, d_img_count;
  if (!host_only) {
    stage_lock.lock();
    if (!stg.ReserveScratch(2 * npg4 + npg + 1024)) { SetError(SBA_ERR_CUDA, "page-locked scratch for sba_create: out of memory"); maps.Join(); cudaStreamSynchronize(CopyStream()); cudaStreamSynchronize(p->stream); delete p; return SBA_ERR_CUDA; }
    pt_count = reinterpret_cast<int*>(stg.q);
    p_off_scratch = reinterpret_cast<int*>(stg.q + npg4);
    live_scratch = reinterpret_cast<uint8_t*>(stg.q + 2 * npg4);
    if (cudaDeviceGetAttribute(&p->n_sm, cudaDevAttrMultiProcessorCount, p->device) != cudaSuccess || p->n_sm <= 0) p->n_sm = 148;
    if (tm.on) { UP(cudaStreamSynchronize(p->stream)); tm.Lap("(timing only) observation upload done"); }
    DevBuf<int> d_bad;
    UP(d_pt_count.Alloc(npg)); UP(d_img_count.Alloc(img_count.size())); UP(d_bad.Alloc(1));
    UP(d_pt_count.Zero(p->stream)); UP(d_bad.Zero(p->stream));
    if (n_obs_g > 0) hist_point_kernel<<<DivUp(n_obs_g, 256), 256, 0, p->stream>>>(n_obs_g, raw_pt.p, n_pt_g, d_pt_count.p, d_bad.p);
    UP(DeviceHistogramImages(p->stream, n_obs_g, raw_img.p, raw_pt.p, 0, n_pt_g, d.num_images, d_img_count.p, p->n_sm, d_bad.p));
    int* h_bad = reinterpret_cast<int*>(stg.q + 2 * npg4 + (npg + 255) / 256 * 256);
    *h_bad = 0;
    UP(cudaMemcpyAsync(h_bad, d_bad.p, sizeof(int), cudaMemcpyDeviceToHost, p->stream));
    if (tm.on) { UP(cudaStreamSynchronize(p->stream)); tm.Lap("(timing only) histogram kernels done"); }
    UP(cudaMemcpyAsync(pt_count, d_pt_count.p, npg * sizeof(int), cudaMemcpyDeviceToHost, p->stream));
    UP(cudaMemcpyAsync(img_count.data(), d_img_count.p, img_count.size() * sizeof(int), cudaMemcpyDeviceToHost, p->stream));
    {
      cudaEvent_t counted = nullptr;
      UP(cudaEventCreateWithFlags(&counted, cudaEventDisableTiming));
      cudaError_t e = cudaEventRecord(counted, p->stream);
      if (e == cudaSuccess && n_obs_g > 0) {
        const size_t xy_bytes = 2 * (size_t)n_obs_g * sizeof(double);
        cudaPointerAttributes at;
        const bool dma = xy_bytes < 4 * kStageChunk ||
                         (cudaPointerGetAttributes(&at, d.obs_xy) == cudaSuccess && at.type == cudaMemoryTypeHost);
        cudaGetLastError();
        if (dma) {  // page-locked source (or small): one asynchronous copy
          e = UploadLarge(raw_xy.p, d.obs_xy, xy_bytes, p->stream);
        } else {    // pageable: the staging copies keep host threads busy for ~2.5 ms at C4 -- not this one
          const int dev = p->device;
          double* dst = raw_xy.p;
          const double* src = d.obs_xy;
          cudaStream_t s = p->stream;
          MapUpload* up = &xy_up;
          xy_up.th = std::thread([up, dev, dst, src, xy_bytes, s] {
            cudaSetDevice(dev);
            up->err = UploadLarge(dst, src, xy_bytes, s);
          });
        }
      }
      if (e == cudaSuccess) e = cudaEventSynchronize(counted);  // the counts are on the host; obs_xy is still on its way
      cudaEventDestroy(counted);
      UP(e);
    }
    p->launches += 2;
    if (*h_bad) {
      xy_up.Join();
      cudaStreamSynchronize(p->stream);
      SetError(SBA_ERR_INVALID_ARGUMENT, "observation index out of range");
      maps.Join(); cudaStreamSynchronize(CopyStream());
      delete p;
      return SBA_ERR_INVALID_ARGUMENT;
    }
  } else {
    v_cnt.assign(npg, 0);
    pt_count = v_cnt.data();
    for (int64_t k = 0; k < n_obs_g; ++k) { pt_count[d.obs_point[k]]++; img_count[d.obs_image[k]]++; }
  }
  tm.Lap("histograms");
  p->shard_count = d.shard_count > 0 ? d.shard_count : 1;
  p->shard_rank = d.shard_count > 0 ? d.shard_rank : 0;
  // ---- shard: contiguous global point range balanced by observation count (from the counts alone) ----
  // The sort of the observations -- and every per-observation array -- is made on the DEVICE from the caller's arrays
  // (cub stable radix sort, order.cu); the host only walks the points.
  int p0 = 0, p1 = n_pt_g;
  int64_t obs_base = 0;
  if (p->shard_count > 1) {
    std::vector<int64_t> prefix(n_pt_g + 1, 0);
    for (int q = 0; q < n_pt_g; ++q) prefix[q + 1] = prefix[q] + pt_count[q];
    ShardPointRange(prefix, p->shard_count, p->shard_rank, &p0, &p1);
    obs_base = prefix[p0];
  }
  p->n_pt = p1 - p0;
  p->pt_base = p0;
  DeviceArena& A = p->arena;
  DeviceArena::Slot s_pt_ptr{}, s_p_off{}, s_tile{};
  std::vector<int> v_pt_ptr, v_p_off, v_tile;  // host-only: no staging block
  const size_t tile_bound = (size_t)(n_obs_g / kTile) * 2 + 16;  // two consecutive tiles hold >= kTile observations
  int *h_pt_ptr, *h_p_off, *tile_start;
  if (!host_only) {
    if (!stg.Reserve(((size_t)p->n_pt * 2 + 1 + tile_bound) * sizeof(int) + 4 * 256 + (1u << 20))) {
      SetError(SBA_ERR_CUDA, "page-locked staging block for sba_create: out of memory");
      maps.Join(); xy_up.Join(); cudaStreamSynchronize(CopyStream()); cudaStreamSynchronize(p->stream); delete p; return SBA_ERR_CUDA;
    }
    h_pt_ptr = A.Claim<int>((size_t)p->n_pt + 1, &s_pt_ptr);
    h_p_off = A.Claim<int>((size_t)p->n_pt, &s_p_off);
    tile_start = A.Claim<int>(tile_bound, &s_tile);
  } else {
    v_pt_ptr.resize((size_t)p->n_pt + 1); v_p_off.resize(std::max(1, p->n_pt)); v_tile.resize(tile_bound);
    h_pt_ptr = v_pt_ptr.data(); h_p_off = v_p_off.data(); tile_start = v_tile.data();
  }
  // CSR over the shard's points and the warp tiles (whole points packed into <= kTile observations): a sequential scan
  // that needs the counts only -- on a thread of its own, under the analysis of the parameter blocks below
  int n_tile_starts = 0;
  std::thread tiler([&, p0]() {
    int acc = 0, fill = 0, nt = 0;
    double pairs = 0.0;
    const int* cnt = pt_count + p0;
    const int n_pt = p->n_pt;
    h_pt_ptr[0] = 0;
    tile_start[nt++] = 0;
    // (branch-free: the boundary is stored every time and kept only when a tile closes -- the data-dependent branches
    //  of the plain form mispredict on every other point)
    for (int q = 0; q < n_pt; ++q) {
      const int len = cnt[q];
      const int close_before = (fill > 0) & (fill + len > kTile);  // (len == 0 never closes: fill < kTile here)
      tile_start[nt] = acc;
      nt += close_before;
      fill = close_before ? 0 : fill;
      acc += len; fill += len;
      const int close_after = fill >= kTile;
      tile_start[nt] = acc;
      nt += close_after;
      fill = close_after ? 0 : fill;
      pairs += 0.5 * (double)len * (double)(len - 1);
      h_pt_ptr[q + 1] = acc;
    }
    if (tile_start[nt - 1] != acc) tile_start[nt++] = acc;
    p->n_obs = acc;
    p->n_tiles = nt - 1;
    p->pair_estimate = pairs;
    n_tile_starts = nt;
  });
  GlobalStructure g;
  Analyse(d, &g, pt_count, img_count.data(), host_only, shared, p_off_scratch, live_scratch);
  {  // the shard's offsets among its own variable points
    int nv = 0;
    const int* go = g.p_off + p0;
    for (int q = 0; q < p->n_pt; ++q) { const int var = go[q] >= 0; h_p_off[q] = var ? nv : -1; nv += var; }
    p->n_pvar = nv;
  }
  tm.Lap("analyse + offsets");
  p->counts = g.counts;
  p->n_img = d.num_images; p->n_cam = d.num_cameras;
  p->n_pt_global = d.num_points; p->n_obs_global = d.num_observations;
  p->kmax = g.kmax; p->dc = DcTier(g.kmax); p->n_cdim = g.n_cdim;
  p->uniform_model = g.uniform_model; p->shared_intrinsics = g.shared_intrinsics;
  p->h_img_cam.assign(d.image_camera, d.image_camera + p->n_img);
  p->h_model.assign(d.camera_model, d.camera_model + p->n_cam);
  p->h_col_img = g.col_img; p->h_col_cam = g.col_cam; p->h_k_cols = g.k_cols; p->h_k_dim = g.k_dim;
  p->h_k_slot.assign((size_t)p->n_cam * 12, -1);
  for (int c = 0; c < p->n_cam; ++c)
    for (int j = 0; j < g.k_dim[c]; ++j) p->h_k_slot[12 * (size_t)c + g.k_cols[12 * (size_t)c + j]] = j;
  p->h_blk_off = g.blk_off; p->h_blk_dim = g.blk_dim; p->h_t_dim = g.t_dim;
  p->h_pose_const.assign(p->n_img, 0);
  for (int i = 0; i < p->n_img; ++i) p->h_pose_const[i] = g.col_img[6 * (size_t)i] < 0 ? 1 : 0;
  tiler.join();
  if (!host_only) A.ShrinkLast<int>(&s_tile, (size_t)n_tile_starts);
  const int n_obs = p->n_obs;
  tm.Lap("blocks + tile scan joined");

  // ---- image-major chunks of <= 256: sizes from a counting pass by image over the shard's observations ----
  std::vector<int>& img_ptr = p->h_img_ptr;
  img_ptr.assign(p->n_img + 1, 0);
  if (p->shard_count > 1) {  // the shard's own per-image counts: observations whose point lies in [p0, p1)
    if (!host_only) {
      UP(DeviceHistogramImages(p->stream, n_obs_g, raw_img.p, raw_pt.p, p0, p1, p->n_img, d_img_count.p, p->n_sm));
      UP(cudaMemcpyAsync(img_count.data(), d_img_count.p, img_count.size() * sizeof(int), cudaMemcpyDeviceToHost, p->stream));
      UP(cudaStreamSynchronize(p->stream));
      p->launches++;
    } else {
      std::fill(img_count.begin(), img_count.end(), 0);
      for (int64_t k = 0; k < n_obs_g; ++k) { const int q = d.obs_point[k]; if (q >= p0 && q < p1) img_count[d.obs_image[k]]++; }
    }
  }
  for (int i = 0; i < p->n_img; ++i) img_ptr[i + 1] = img_ptr[i] + img_count[i];
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
  tm.Lap("chunks");
  *out = p;
  if (host_only) return SBA_OK;

  // ---- device ----
  cudaStream_t st = p->stream;
  {  // residual blocks of the reduced program (Summary::num_residuals_reduced): counted on the device
    std::vector<uint8_t> img_live((size_t)std::max(1, p->n_img), 0);
    for (int i = 0; i < p->n_img; ++i) img_live[i] = (g.col_img[6 * (size_t)i] >= 0 || g.k_dim[d.image_camera[i]] > 0) ? 1 : 0;
    DevBuf<uint8_t> d_il, d_pl;
    DevBuf<unsigned long long> d_n;
    UP(d_il.Upload(img_live, st)); UP(d_pl.Alloc(npg)); UP(d_n.Alloc(1)); UP(d_n.Zero(st));
    UP(cudaMemcpyAsync(d_pl.p, g.pt_live, npg, cudaMemcpyHostToDevice, st));  // (from the page-locked scratch)
    if (n_obs_g > 0) {
      const int blocks = (int)std::min<int64_t>((n_obs_g + 255) / 256, (int64_t)p->n_sm * 8);
      count_active_kernel<<<blocks, 256, 0, st>>>(n_obs_g, raw_img.p, raw_pt.p, d_il.p, d_pl.p, d_n.p);
      p->launches++;
    }
    unsigned long long n_act = 0;
    UP(cudaMemcpyAsync(&n_act, d_n.p, sizeof(n_act), cudaMemcpyDeviceToHost, st));
    UP(cudaStreamSynchronize(st));  // (also: the temporaries above are released on scope exit)
    p->counts.num_residuals_reduced = (int)(p->nr * (int64_t)n_act);
  }
  tm.Lap("active count");
  // parameters: one upload into the creation copies, device-to-device into the current / candidate buffers
  // (these three triples swap pointers during the solve, so they own their memory)
  std::vector<double> pose((size_t)p->n_img * 8, 0.0), cam((size_t)p->n_cam * 12, 0.0);
  for (int i = 0; i < p->n_img; ++i) {
    for (int j = 0; j < 4; ++j) pose[8 * (size_t)i + j] = d.image_qvec[4 * (size_t)i + j];
    for (int j = 0; j < 3; ++j) pose[8 * (size_t)i + 4 + j] = d.image_tvec[3 * (size_t)i + j];
  }
  for (size_t j = 0; j < cam.size(); ++j) cam[j] = d.camera_params[j];
  const size_t nX = (size_t)std::max(1, p->n_pt) * 3;
  UP(p->d_pose0.Upload(pose, st)); UP(p->d_cam0.Upload(cam, st));
  UP(p->d_X0.Alloc(nX));
  if (p->n_pt > 0) UP(UploadLarge(p->d_X0.p, d.points + 3 * (size_t)p0, (size_t)p->n_pt * 3 * sizeof(double), st));
  UP(p->d_pose.Alloc(pose.size())); UP(p->d_pose2.Alloc(pose.size()));
  UP(p->d_cam.Alloc(cam.size())); UP(p->d_cam2.Alloc(cam.size()));
  UP(p->d_X.Alloc(nX)); UP(p->d_X2.Alloc(nX));
  for (double* dst : {p->d_pose.p, p->d_pose2.p})
    if (!pose.empty()) UP(cudaMemcpyAsync(dst, p->d_pose0.p, pose.size() * sizeof(double), cudaMemcpyDeviceToDevice, st));
  for (double* dst : {p->d_cam.p, p->d_cam2.p})
    if (!cam.empty()) UP(cudaMemcpyAsync(dst, p->d_cam0.p, cam.size() * sizeof(double), cudaMemcpyDeviceToDevice, st));
  for (double* dst : {p->d_X.p, p->d_X2.p})
    if (p->n_pt > 0) UP(cudaMemcpyAsync(dst, p->d_X0.p, (size_t)p->n_pt * 3 * sizeof(double), cudaMemcpyDeviceToDevice, st));
  tm.Lap("parameters");

  // structure arrays: one packed block, one copy
  std::vector<double> mi((size_t)std::max(1, p->n_img) * 6, 0.0), mk((size_t)std::max(1, p->n_cam) * 12, 0.0);
  std::vector<unsigned long long> pack(std::max(1, p->n_cam), 0ull);
  for (size_t j = 0; j < (size_t)p->n_img * 6; ++j) mi[j] = p->h_col_img[j] >= 0 ? 1.0 : 0.0;
  for (int c = 0; c < p->n_cam; ++c)
    for (int j = 0; j < p->h_k_dim[c]; ++j) {
      mk[12 * (size_t)c + j] = p->h_col_cam[12 * (size_t)c + j] >= 0 ? 1.0 : 0.0;
      pack[c] |= (unsigned long long)(p->h_k_cols[12 * (size_t)c + j] & 15) << (4 * j);
    }
  const auto s_img_cam = A.Push(p->h_img_cam), s_model = A.Push(p->h_model), s_col_img = A.Push(p->h_col_img),
             s_col_cam = A.Push(p->h_col_cam), s_k_cols = A.Push(p->h_k_cols), s_k_slot = A.Push(p->h_k_slot),
             s_k_dim = A.Push(p->h_k_dim), s_pose_const = A.Push(p->h_pose_const),
             s_chunk_start = A.Push(chunk_start), s_chunk_img = A.Push(chunk_img),
             s_img_chunk = A.Push(img_chunk_ptr), s_cam_img_ptr = A.Push(cam_img_ptr), s_cam_img_list = A.Push(cam_img_list),
             s_blk_off = A.Push(p->h_blk_off), s_blk_dim = A.Push(p->h_blk_dim),
             s_col_blk = A.Push(col_blk), s_mi = A.Push(mi), s_mk = A.Push(mk), s_pack = A.Push(pack);
  UP(A.Upload(st, true));  // (the per-point arrays were written in place; the stage mutex is ours since the histograms)
  stage_lock.unlock();
  A.Bind(&p->d_img_cam, s_img_cam); A.Bind(&p->d_model, s_model); A.Bind(&p->d_col_img, s_col_img);
  A.Bind(&p->d_col_cam, s_col_cam); A.Bind(&p->d_k_cols, s_k_cols); A.Bind(&p->d_k_slot, s_k_slot);
  A.Bind(&p->d_k_dim, s_k_dim); A.Bind(&p->d_p_off, s_p_off); A.Bind(&p->d_pose_const, s_pose_const);
  A.Bind(&p->d_tile_start, s_tile); A.Bind(&p->d_chunk_start, s_chunk_start); A.Bind(&p->d_chunk_img, s_chunk_img);
  A.Bind(&p->d_img_chunk_ptr, s_img_chunk); A.Bind(&p->d_cam_img_ptr, s_cam_img_ptr);
  A.Bind(&p->d_cam_img_list, s_cam_img_list); A.Bind(&p->d_pt_ptr, s_pt_ptr); A.Bind(&p->d_blk_off, s_blk_off);
  A.Bind(&p->d_blk_dim, s_blk_dim); A.Bind(&p->d_col_blk, s_col_blk); A.Bind(&p->d_mask_i, s_mi);
  A.Bind(&p->d_mask_k, s_mk); A.Bind(&p->d_k_pack, s_pack);
  tm.Lap("structure block");

  // point-sorted observations and the image-major list, sorted on the device from the raw copies
  {
    xy_up.Join();  // (pageable obs_xy: every chunk has been queued on the stream)
    UP(xy_up.err);
    const int orc = BuildDeviceOrder(p, p0, obs_base, n_obs_g, n_pt_g, raw_img.p, raw_pt.p, raw_xy.p);
    if (orc) { maps.Join(); cudaStreamSynchronize(CopyStream()); delete p; *out = nullptr; return orc; }
  }
  tm.Lap("device order");
  const size_t nc = std::max(1, p->n_cdim), np3 = std::max(1, 3 * p->n_pvar);
  UP(p->d_ER.Alloc(TilePad(std::max(1, n_obs)) * 2)); UP(p->d_F.Alloc(TilePad(std::max(1, n_obs)) * (p->dc / 2)));
  if (p->nr == 3) {
    UP(p->d_ES.Alloc(TilePad(std::max(1, n_obs)))); UP(p->d_FS.Alloc(TilePad(std::max(1, n_obs)) * FSUnits(p->dc)));
    UP(p->d_g1.Alloc((size_t)std::max(1, n_obs)));
  }
  p->n_partial = std::max(std::max(DivUp(std::max(1, n_obs), 128), DivUp(std::max(1, p->n_tiles), 4)),
                          DivUp(p->n_img + p->n_cam + p->n_pt, 128)) + 8;
  UP(p->d_cost_partial.Alloc(4 * (size_t)p->n_partial));
  UP(p->d_xi.Alloc((size_t)std::max(1, p->n_img) * 6)); UP(p->d_xc.Alloc((size_t)std::max(1, p->n_cam) * 12));
  UP(p->d_scale_i.Alloc((size_t)std::max(1, p->n_img) * 6)); UP(p->d_scale_k.Alloc((size_t)std::max(1, p->n_cam) * 12));
  UP(p->d_red.Alloc(2048));
  UP(p->d_scale_c.Alloc(nc)); UP(p->d_scale_p.Alloc(np3));
  UP(p->d_sq_p.Alloc(np3));
  UP(p->d_diag_c.Alloc(nc)); UP(p->d_diag_p.Alloc(np3)); UP(p->d_Dc.Alloc(nc)); UP(p->d_Dp.Alloc(np3));
  UP(p->d_Cinv.Alloc((size_t)std::max(1, p->n_pvar) * 6)); UP(p->d_gp.Alloc(np3));
  UP(p->d_Cfac.Alloc((size_t)std::max(1, p->n_pvar) * 6)); UP(p->d_EtE.Alloc((size_t)std::max(1, p->n_pvar) * 6));
  UP(p->d_chunk_partial.Alloc((size_t)std::max(1, p->n_chunks) * kMaxDC * 2));
  UP(p->d_g2.Alloc((size_t)std::max(1, n_obs) * 2));
  UP(p->d_rhs.Alloc(nc)); UP(p->d_yc.Alloc(nc)); UP(p->d_yp.Alloc(np3)); UP(p->d_step_c.Alloc(nc));
  UP(p->d_grad_c.Alloc(2 * nc + 2)); p->d_sq_c.View(p->d_grad_c.p + p->n_cdim, nc);  // adjacent: Gradient(with_norms) all-reduces both at once
  UP(p->d_grad_p.Alloc(np3)); UP(p->d_tmp_c.Alloc(nc));
  UP(p->d_cg_r.Alloc(nc)); UP(p->d_cg_z.Alloc(nc)); UP(p->d_cg_p.Alloc(nc)); UP(p->d_cg_q.Alloc(nc)); UP(p->d_cg_x.Alloc(nc));
  UP(p->d_scalars.Alloc(1)); UP(p->d_scalars.Zero(st));
  UP(p->d_tickets.Alloc(8)); UP(p->d_tickets.Zero(st));
  tm.Lap("allocs");
  maps.Join();  // every chunk of the map upload has been queued on the side stream
  UP(maps.err);
  if (map_stage.p) {
    const size_t n_out = (size_t)p->n_img * MapTiledSize(p->sem_h, p->sem_w);
    retile_maps_kernel<<<(unsigned)std::min<size_t>(DivUp(n_out, 256), 148 * 64), 256, 0, CopyStream()>>>(
        map_stage.p, p->d_sem_maps.p, p->n_img, p->sem_h, p->sem_w);
    UP(cudaGetLastError());
    p->launches += 1;
  }
  UP(cudaStreamSynchronize(CopyStream()));
  UP(cudaStreamSynchronize(st));
  tm.Lap("maps + sync");
#undef UP
  return SBA_OK;
}

}  // namespace sba
