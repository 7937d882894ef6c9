// Host-side state of one bundle-adjustment problem resident on one B200.
// The drop-in boundary is include/sba_b200.h; this struct is what sba_handle points at.
// Reference path: BundleAdjuster::SetUp + ceres::Solve,
//   /root/reference/src/optim/bundle_adjustment.cc:326-530, :306.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <mutex>
#include <atomic>
#include <vector>

#include "common.cuh"

namespace sba {

// Device memory comes from CUDA's stream-ordered pool (cudaMallocAsync) on one long-lived allocation stream
// per device, with the pool's release threshold lifted: the ~70 buffers of a problem cost microseconds to
// create and destroy after the first handle, instead of a cudaMalloc / synchronising cudaFree each.
// Contract: a buffer is freed only after the stream that used it has been synchronised (sba_destroy does).
cudaStream_t AllocStream();  // problem.cu
// Blocks released by a handle are kept in a per-device size-keyed cache and handed to the next request of (nearly)
// the same size: repeated problems of one shape then make no driver allocation call at all -- the pool's own reuse
// costs tens of milliseconds now and then once blocks reach hundreds of MB.  `*cap` receives the true block size.
cudaError_t CachedAlloc(void** ptr, size_t bytes, size_t* cap);
void CachedFree(void* ptr, size_t cap);
void ReleaseCachedBlocks();  // current device: hand every cached block back to the pool

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  DevBuf() {}
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { Free(); }
  bool owned = true;  // false: a view into a DeviceArena block
  size_t cap_bytes = 0;  // true size of the owned block (CachedAlloc may hand out a slightly larger one)
  void Free() {
    if (p && owned) CachedFree(p, cap_bytes);
    p = nullptr; n = 0; owned = true; cap_bytes = 0;
  }
  void View(T* ptr, size_t count) { Free(); p = ptr; n = count; owned = false; }
  cudaError_t Alloc(size_t count) {
    if (count <= n && p) return cudaSuccess;
    if (p) cudaDeviceSynchronize();  // growing a live buffer: nothing in flight may still read the old one
    Free();
    n = count;
    return CachedAlloc((void**)&p, (count ? count : 1) * sizeof(T), &cap_bytes);
  }
  cudaError_t Upload(const std::vector<T>& h, cudaStream_t s) {
    cudaError_t e = Alloc(h.size());
    if (e != cudaSuccess) return e;
    if (h.empty()) return cudaSuccess;
    return cudaMemcpyAsync(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, s);
  }
  cudaError_t Zero(cudaStream_t s) { return cudaMemsetAsync(p, 0, (n ? n : 1) * sizeof(T), s); }
};

// The ~30 small structure arrays of a problem travel in ONE host -> device copy: they are packed (256 B aligned)
// into a host block, copied once, and the DevBufs become views into the device block.
// The ~22 small structure arrays of a problem travel as ONE block.  The host side of the block is a page-locked staging
// buffer kept per process (grown on demand, guarded by a mutex for the duration of one create): arrays are copied into it
// once and the upload is one DMA -- a growing std::vector re-allocated, zero-filled and page-faulted 12 MB per create at C4.
struct PinnedStage {
  char* p = nullptr; size_t cap = 0;   // staging block of the arena upload
  char* q = nullptr; size_t qcap = 0;  // scratch of one create: per-point counts, offsets and live mask of the whole problem
  std::mutex mu;
  static PinnedStage& Get() { static PinnedStage s; return s; }
  static bool Grow(char** blk, size_t* have, size_t bytes, size_t keep) {
    if (bytes <= *have) return true;
    const size_t want = std::max<size_t>(bytes + bytes / 2, 1u << 20);
    char* nb = nullptr;
    if (cudaMallocHost((void**)&nb, want) != cudaSuccess) { cudaGetLastError(); return false; }
    if (*blk && keep) std::memcpy(nb, *blk, std::min(keep, *have));
    if (*blk) cudaFreeHost(*blk);
    *blk = nb; *have = want;
    return true;
  }
  bool Reserve(size_t bytes, size_t keep = 0) { return Grow(&p, &cap, bytes, keep); }
  bool ReserveScratch(size_t bytes) { return Grow(&q, &qcap, bytes, 0); }
};
struct DeviceArena {
  DevBuf<char> dev;
  struct Slot { size_t off, bytes; const void* src; };  // src == nullptr: written in place (Claim)
  std::vector<Slot> slots;
  size_t size = 0;
  template <typename T>
  Slot Push(const std::vector<T>& v) {  // (v must stay alive until Upload)
    const size_t off = (size + 255) / 256 * 256, bytes = v.size() * sizeof(T);
    size = off + std::max<size_t>(bytes, 1);
    slots.push_back(Slot{off, bytes, v.data()});
    return slots.back();
  }
  // An array written straight into the staging block: no host vector (12 MB of them were allocated, zero-filled and
  // page-faulted per create at C4) and no copy at Upload.  The caller holds PinnedStage::mu from the first Claim until
  // Upload(s, true) returns and has Reserved the staging block for every Claim it makes.
  template <typename T>
  T* Claim(size_t count, Slot* sl) {
    const size_t off = (size + 255) / 256 * 256, bytes = count * sizeof(T);
    size = off + std::max<size_t>(bytes, 1);
    slots.push_back(Slot{off, bytes, nullptr});
    *sl = slots.back();
    return reinterpret_cast<T*>(PinnedStage::Get().p + off);
  }
  template <typename T>
  void ShrinkLast(Slot* sl, size_t count) {  // the most recent Claim turned out smaller than its bound
    Slot& b = slots.back();
    b.bytes = count * sizeof(T);
    size = b.off + std::max<size_t>(b.bytes, 1);
    *sl = b;
  }
  cudaError_t Upload(cudaStream_t s, bool locked = false) {
    cudaError_t e = dev.Alloc(std::max<size_t>(size, 1));
    if (e != cudaSuccess) return e;
    PinnedStage& st = PinnedStage::Get();
    std::unique_lock<std::mutex> lk(st.mu, std::defer_lock);
    if (!locked) lk.lock();
    if (!st.Reserve(size, locked ? size : 0)) return cudaErrorMemoryAllocation;  // (growing keeps what was written in place)
    for (const Slot& sl : slots)
      if (sl.bytes && sl.src) std::memcpy(st.p + sl.off, sl.src, sl.bytes);
    e = cudaMemcpyAsync(dev.p, st.p, size, cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) return e;
    e = cudaStreamSynchronize(s);  // the staging buffer is shared: the copy has left it before the lock is released
    slots.clear();
    return e;
  }
  template <typename T>
  void Bind(DevBuf<T>* b, const Slot& s) { b->View(reinterpret_cast<T*>(dev.p + s.off), s.bytes / sizeof(T)); }
};

// scalar block shared between kernels and polled by the host once per LM
// iteration (trust-region bookkeeping stays on the device).
struct DeviceScalars {
  double cost;              // 0.5 sum rho (active blocks), current linearisation point
  double fixed_cost;        // blocks with only constant parameters
  double cand_cost;         // candidate cost
  double model_cost_change;
  double step2_p, x2_p;     // point-side step / parameter square norms (sharded); with the two above: ONE all-reduce of 4
  double radius;
  double decrease_factor;
  double relative_decrease;
  double step_norm;
  double x_norm;
  double gradient_max_norm;
  double cost_change;
  int step_valid;
  int step_accepted;
  int reuse_diagonal;
  int solver_failed;
  // PCG scalars
  double cg_rho, cg_last_rho, cg_alpha, cg_beta, cg_pq, cg_Q0, cg_Q1, cg_zeta, cg_norm_b;
  int cg_iter;
  int cg_state;  // 0 running, 1 converged, 2 no-convergence stop, 3 failure
  // Cholesky
  int chol_fail;
  int converged;  // 1 parameter tolerance, 2 function tolerance reached in this iteration
  // squared norms for step_norm / x_norm: camera-side (replicated) and point-side (sharded)
  double step2_c, x2_c;
};

class Comm;  // NCCL wrapper (comm.cu)

// pairs (a,b) of observations of one point, sorted by (image(a), image(b)) key, in segments
struct PairList {
  DevBuf<int> seg_ptr, key_seg_ptr, key_gi, key_gj;
  DevBuf<unsigned long long> ab;  // packed pair: a in the high word, b in the low word (pairs.cu)
  DevBuf<double> partial;  // [n_seg*dc*dc]
  int n_seg = 0, n_keys = 0, seg_pairs = 32;
  int64_t n_pairs = 0;
  bool built = false;
};

// Explicit block-sparse reduced camera matrix for the iterative solver (bsr.cu): blocks live in SLOT space (dc x dc
// per camera pair), block i < n_img = diagonal block of image i, block n_img + k = global camera-pair key k (gi > gj).
struct BlockSchur {
  bool decided = false, enabled = false;
  int n_keys = 0;                    // global camera-pair keys (same on every rank)
  int64_t n_blocks = 0, n_entries = 0;
  DevBuf<double> blocks;             // [n_blocks * dc * dc]
  DevBuf<int> row_ptr;               // [n_img + 1] into entries
  DevBuf<int2> entries;              // {block, other image | transposed << 31}: diagonal, lower, then upper neighbours
  DevBuf<int> slot_col;              // [n_img * dc] compact column of a slot or -1
  DevBuf<int> key_block;             // multi-GPU: local key -> block index (single GPU: n_img + key)
  int pcg_grid = 0;                  // CTAs of the persistent PCG kernel (pcg_blocks.cu), co-resident by construction
  int pcg_rows = 0;                  // block rows per CTA
  DevBuf<double> pcg_partials;       // [3 * pcg_grid]
  DevBuf<double> minv_s;             // [n_img * dc * dc] SCHUR_JACOBI inverse in slot space (zero outside a parameter block)
  DevBuf<double> p_s;                // slot-space CG vectors other CTAs read: p [2][n_img*dc] | z | x (one GPU; N > 1: the peer-mapped area)
  bool pcg_dist = false;             // N > 1: block rows split across the ranks (pcg_blocks.cu DIST)
};

struct Problem {
  // ---- sizes ----
  int n_img = 0, n_cam = 0, n_pt = 0;  // global images / cameras, LOCAL points
  int n_obs = 0;                       // LOCAL observations
  int64_t n_obs_global = 0;
  int n_pt_global = 0;
  int dc = 6, kmax = 0, n_cdim = 0, n_pvar = 0;
  int uniform_model = -1;  // camera model shared by all used cameras, else -1
  int shard_rank = 0, shard_count = 1;
  bool shared_intrinsics = false;  // some camera with free intrinsics is used by >1 image

  // ---- host copies of structure (small) ----
  std::vector<int> h_img_cam, h_model, h_col_img, h_col_cam, h_k_cols, h_k_slot, h_k_dim;
  std::vector<int> h_blk_off, h_blk_dim;       // SCHUR_JACOBI parameter blocks (compact cols)
  std::vector<int> h_obs_img, h_obs_pt;        // local, point-sorted
  std::vector<int> h_img_ptr, h_cam_order;  // image-major list of local observation ids
  double pair_estimate = 0.0;               // sum over points of k (k - 1) / 2: the size of the pair list (bsr.cu)
  std::vector<int64_t> h_local2ref;            // local point-sorted position -> reference residual index
  int pt_base = 0;                             // local point q is global point pt_base + q (shards are contiguous ranges)
  std::vector<uint8_t> h_pose_const;
  std::vector<int> h_t_dim;
  sba_summary counts;                          // structural counts (global problem)
  std::vector<sba_iteration_summary> trace;

  DeviceArena arena;                           // backing store of the small structure arrays below
  bool host_order_ready = false;               // h_local2ref / h_obs_img fetched from the device (parity surface only)

  // ---- device: structure ----
  DevBuf<unsigned> d_local2ref;                // [n_obs] local point-sorted position -> reference residual index
  DevBuf<int> d_img_cam, d_model, d_col_img, d_col_cam, d_k_cols, d_k_slot, d_k_dim, d_p_off;
  DevBuf<int> d_obs_img, d_obs_pt;
  DevBuf<double> d_obs_xy;
  DevBuf<uint8_t> d_obs_active, d_pose_const;
  DevBuf<int> d_tile_start;  int n_tiles = 0;
  DevBuf<int> d_chunk_start, d_chunk_img, d_cam_order, d_img_chunk_ptr, d_cam_img_ptr, d_cam_img_list;
  DevBuf<int> d_cam_po;      // [n_obs] p_off of the point of the k-th image-major observation
  DevBuf<int> d_cam_pos;     // [n_obs] inverse of d_cam_order: image-major position of an observation (rows of d_W / d_H)
  int n_chunks = 0;
  DevBuf<int> d_pt_ptr;      // [n_pt+1] CSR over point-sorted observations
  DevBuf<int> d_blk_off, d_blk_dim, d_col_blk;  int n_blk = 0;

  // observation-pair lists for the off-diagonal Schur contributions (schur.cu)
  PairList pairs_dense, pairs_diag;
  BlockSchur bsr;
  DevBuf<int> d_blk_moff;  int m_size = 0;      // offsets of the SCHUR_JACOBI blocks inside d_M / d_Minv
  DevBuf<double> d_M;                           // block diagonal of S before inversion
  DevBuf<double> d_chunk_block;                 // [n_chunks*dc*dc] per-chunk diagonal block partials
  DevBuf<double> d_W, d_H;                      // [n_obs*dc*3] F^T E and F^T E C^{-1}, rows in IMAGE-MAJOR order (d_cam_pos)

  // ---- device: parameters (current / candidate) ----
  DevBuf<double> d_pose, d_cam, d_X, d_pose2, d_cam2, d_X2;
  DevBuf<double> d_pose0, d_cam0, d_X0;  // parameters at creation (sba_reset)

  // ---- device: linearisation ----
  DevBuf<D4> d_ER;       // [TilePad(n_obs)*2]       tile-blocked (common.cuh)
  DevBuf<D4> d_F;        // [TilePad(n_obs)*dc/2]    tile-blocked
  int nr = 2;            // residual rows per observation: 3 when the problem carries the semantic row (extension X1)
  DevBuf<D4> d_ES;       // [n_obs]                          third row: {E_s (3), r_s}
  DevBuf<D4> d_FS;       // [TilePad(n_obs)*FSUnits(dc)]     third row of F, tile-blocked
  DevBuf<double> d_g1;   // [n_obs] third component of the per-observation scratch vector (d_g2 holds the first two)
  DevBuf<double> d_cost_partial;  // [blocks*2]
  DevBuf<double> d_scale_c, d_scale_p;      // Jacobi scaling
  DevBuf<double> d_diag_c, d_diag_p;        // clamped column square norms
  DevBuf<double> d_Dc, d_Dp;                // LM diagonal
  DevBuf<double> d_Cinv, d_gp;              // [n_pvar*6] sym inverse, [n_pvar*3]
  DevBuf<double> d_EtE;                     // [n_pvar*6] E_p^T E_p of the current linearisation (without D_p^2): the LM diagonal
                                            // changes every iteration, E only with an accepted step (PointBlocks, linalg.cu)
  bool ete_valid = false;                   // d_EtE and d_gp belong to the current d_ER
  DevBuf<double> d_Cfac;                    // [n_pvar*6] M = L^{-1} (lower) with C_p^{-1} = M^T M: U_a = (F_a^T E_a) M^T (schur.cu)
  bool rhs_from_schur = false;              // the explicit Schur build of this iteration also leaves the reduced rhs in d_rhs
  DevBuf<double> d_chunk_partial;           // [n_chunks*dc] (or dc*dc for block sums)
  DevBuf<double> d_g2;                      // [n_obs*2] per-observation 2-vector scratch
  DevBuf<double> d_rhs, d_yc, d_yp, d_step_c, d_grad_c, d_grad_p, d_tmp_c;
  DevBuf<double> d_S;                       // dense reduced camera matrix [n_cdim^2]
  DevBuf<double> d_Linv;                    // inverted diagonal blocks of the Cholesky factor
  DevBuf<int> chol_flags;  int chol_epoch = 0;  bool chol_add_diag = false;  // chol_small.cu tile epochs
  DevBuf<double> d_Minv;                    // SCHUR_JACOBI inverse blocks
  DevBuf<double> d_cg_r, d_cg_z, d_cg_p, d_cg_q, d_cg_x;
  DevBuf<double> d_xi, d_xc;                // padded camera-space vector (n_img*6 | n_cam*12)
  DevBuf<double> d_scale_i, d_scale_k;      // padded Jacobi scale (0 at constant slots)
  DevBuf<double> d_mask_i, d_mask_k;        // padded 1 / 0 masks: the unscaled evaluation uses these as its scale
  DevBuf<unsigned long long> d_k_pack;      // [n_cam] nibble s = ambient parameter index of intrinsics slot s
  DevBuf<double> d_sq_c, d_sq_p;            // raw column square norms
  DevBuf<double> d_red;                     // small reduction scratch (>= 1024 doubles)
  int n_partial = 0;                        // capacity (pairs) of d_cost_partial
  DevBuf<unsigned> d_tickets;               // GridSum tickets (reduce.cuh), zero between launches
  DevBuf<DeviceScalars> d_scalars;
  DeviceScalars* h_scalars = nullptr;       // pinned

  // CUDA graphs of the LM iteration's launch sequence (lm.cu): [0..3] step segment by (parity, refresh), [4..5]
  // re-linearisation segment by parity; parity = which of the two parameter buffers is current
  cudaGraphExec_t graph_exec[6] = {};
  int64_t graph_launches[6] = {};
  sba_solver_options graph_opts;             // the options the graphs were captured with
  bool graph_opts_valid = false;
  int parity = 0;
  int solve_count = 0;                       // sba_solve calls on this handle
  bool graphs_unusable = false;              // a capture failed once (e.g. a launch kind the driver cannot capture)
  void DropGraphs();

  cudaStream_t stream = nullptr;
  cudaStream_t stream2 = nullptr, stream3 = nullptr;  // side streams of the blocked Cholesky (cholesky.cu), forked from `stream`
  cudaEvent_t ev_chol[8] = {};
  int device = 0;
  int n_sm = 148;                           // multiprocessors of `device`
  int64_t launches = 0;
  int64_t collectives = 0;
  // extension X1: per-image semantic maps (row-major float), sampled bilinearly at the projection
  DevBuf<float> d_sem_maps;  int sem_h = 0, sem_w = 0;  double sem_weight = 0.0;
  struct MapSet* shared_maps = nullptr;     // d_sem_maps is a view into this reference-counted set (sba_create_with_maps)
  bool solved = false;
  bool linearized = false;
  bool cost_unreduced = false;              // multi-GPU: scalars->cost / fixed_cost are partial sums; Gradient(with_norms) carries them
  bool rhs_unreduced = false;               // multi-GPU: d_rhs holds this rank's partial sum; the Schur build's all-reduce carries it
  bool colnorm_current = false;             // d_sq_c holds the camera column norms of the CURRENT Jacobian (Gradient(with_norms))
  int cg_poll_every = 4;                    // host polls the CG state every this many iterations
  double* jq_out = nullptr;                 // parity surface: raw 2x4 quaternion block per observation
  DevBuf<char> d_flush;                     // L2 flush scratch for sba_bench_stage
  Comm* comm = nullptr;

  // stage timers
  double t_eval = 0, t_elim = 0, t_solve = 0, t_update = 0;

  ~Problem();
  ParamsView Params(bool candidate = false) const;
  Row3View Row3() const { return Row3View{d_ES.p, d_FS.p, d_sem_maps.p, sem_h, sem_w, sem_weight}; }
  StructView Struct() const;
  TileView Tiles() const;
  ChunkView Chunks() const;
};

// problem.cu
// label maps resident on one device across problems (sba_maps_create): tiled like Problem::d_sem_maps
struct MapSet {
  DevBuf<float> tiled;
  int device = 0, n_img = 0, h = 0, w = 0;
  std::atomic<int> refs{1};  // the caller's reference + one per live handle created over it
};
int BuildMapSet(int n_img, int h, int w, const float* maps, MapSet** out);
void MapSetRelease(MapSet* m);
int BuildProblem(const sba_problem_desc* d, Problem** out, bool host_only, MapSet* shared = nullptr);
int FetchHostOrder(Problem* p);  // order.cu
int BuildDeviceOrder(Problem* p, int p0, int64_t obs_base, int64_t n_obs_g, int n_pt_g, const int* d_raw_img,
                     const int* d_raw_pt, const double* d_raw_xy);
cudaError_t UploadLarge(void* dst, const void* src, size_t bytes, cudaStream_t st);
cudaError_t DownloadLarge(void* dst, const void* src, size_t bytes, cudaStream_t st);  // device -> (pageable) host, synchronous  // fills h_local2ref / h_obs_img / h_obs_pt on first use (sba_evaluate)
int ProblemCounts(const sba_problem_desc* d, sba_summary* s);
// eval.cu
int LaunchEvaluate(Problem* p, const LossView& loss, bool use_scale, bool candidate, bool defer_allreduce = false);
int LaunchCostOnly(Problem* p, const LossView& loss, bool candidate, double* d_out_cost /* &scalars->cand_cost */,
                   bool with_step_scalars = false);
// linalg.cu (column norms, point blocks, rhs, matvec, back-substitution, update)
int ColumnNorms(Problem* p, double* d_nc, double* d_np, bool camera_part = true);
int Gradient(Problem* p, bool with_norms = false);
int FillOnes(Problem* p);
int AllReduceMax(Problem* p, double* d_buf, size_t n);
int InitJacobiScale(Problem* p);
int PrepareLMDiagonal(Problem* p, const sba_solver_options& o);
int PointBlocks(Problem* p, bool grad_only);
int FinalizeCameraVector(Problem* p, const double* d_partial, double* d_out);
int ReducedRhs(Problem* p, bool defer_allreduce = false);  // deferred: see Problem::rhs_unreduced, AllReduceSumWithRhs
int SchurMatVec(Problem* p, const double* d_x, double* d_y, const int* skip = nullptr);  // *skip != 0: the big kernels return at once
int BackSubstitute(Problem* p, bool defer_allreduce = false);
int ModelCostChange(Problem* p);
int ApplyStep(Problem* p, bool defer_allreduce = false);
int TrustRegionUpdate(Problem* p, const sba_solver_options& o, int phase, bool sync);
int AllReduceSum(Problem* p, double* d_buf, size_t n);
// sums d_buf[0..n) and, when rhs_unreduced, d_rhs in the same collective (d_buf needs n + n_cdim doubles of room)
int AllReduceSumWithRhs(Problem* p, double* d_buf, size_t n);
int AllReduceMaxInt(Problem* p, int* d_buf, size_t n);
// pairs.cu
int BuildPairLists(Problem* p, bool dense);
// pairs per segment: the lane-per-pair kernels (dc <= 8) amortise one reduction over 4 pairs per lane
inline int PairSegmentSize(int dc) { return dc <= 8 ? 128 : 32; }
// schur.cu
// with_rhs: the build's pass over the observations also forms the reduced right-hand side F^T (r - E C^{-1} g) into d_rhs
// (dc <= 8: the lane kernels read everything it needs anyway); SchurBuildsRhs tells whether it will.  Otherwise the
// caller runs ReducedRhs first.
bool SchurBuildsRhs(const Problem* p);
int BuildDenseSchur(Problem* p, bool with_rhs = false);
int BuildSchurJacobi(Problem* p);
int BuildExplicitSchur(Problem* p, bool with_rhs = false);   // iterative solver: block-sparse S + its SCHUR_JACOBI blocks (when bsr.enabled)
// bsr.cu
int DecideBlockSchur(Problem* p);     // builds the pair lists and the block structure; sets bsr.enabled
int BlockSchurMatVec(Problem* p, const double* d_x, double* d_y, const int* skip);
// cholesky.cu
int CholeskySolve(Problem* p);  // d_S, d_rhs -> d_yc ; sets scalars.chol_fail
// chol_small.cu
constexpr int kCholSmallMax = 1024;
int CholeskySolveSmall(Problem* p);  // n_cdim <= kCholSmallMax: one dataflow kernel
// leading dimension of the dense reduced camera matrix d_S ((n + 1) rows: the last one is the right-hand side): n, rounded
// up to even for the blocked factorisation so that every row starts on a 16-byte boundary (cp.async / double2 moves)
inline int DenseLd(int n) { return n > kCholSmallMax ? (n + 1) / 2 * 2 : n; }
// pcg.cu
int SolvePCG(Problem* p, const sba_solver_options& o, int* iters);
int SolvePCGBlocks(Problem* p, const sba_solver_options& o);  // pcg_blocks.cu: one persistent kernel on the block-sparse S
// lm.cu
int SolveLM(Problem* p, const sba_solver_options* o, sba_iteration_callback cb, void* user, sba_summary* s);
// comm.cu
int CommGetUniqueId(void* id128);
int CostNoAlloc(Problem* p, const LossView& loss, double* cost_host, double* cost_device);  // eval.cu: static-scratch cost (sba_cost_no_alloc)
int CommInit(Problem* p, int nranks, int rank, const void* id128);
int CommPcgArea(Problem* p, size_t bytes, void** area, int* nranks, int* rank);  // peer-mapped vectors of the row-sharded PCG (collective)
bool CommHasPeerMemory(const Problem* p);
void CommDestroy(Problem* p);
bool CommAcquire(Problem* p);   // one collective-issuing API call at a time per communicator
void CommRelease(Problem* p);
struct CommScope {              // RAII: ok == false -> the caller returns SBA_ERR_COMM
  Problem* p; bool ok;
  explicit CommScope(Problem* q) : p(q), ok(CommAcquire(q)) {
    if (!ok) SetError(SBA_ERR_COMM, "the communicator is in use by another handle of the same group (one solve at a time per group)");
  }
  ~CommScope() { if (ok) CommRelease(p); }
  CommScope(const CommScope&) = delete;
  CommScope& operator=(const CommScope&) = delete;
};

}  // namespace sba
