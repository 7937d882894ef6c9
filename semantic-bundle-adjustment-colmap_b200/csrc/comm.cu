// Multi-GPU plumbing: one process per GPU, NCCL over NVLink 5 / NVSwitch.
//
// The reference has no distributed path at all; its only
// parallelism is Ceres' num_threads (/root/reference/src/optim/bundle_adjustment.cc:288-301).
// Here observations are sharded by point ownership (problem.cu) and camera-space
// quantities -- cost scalars, gradient / reduced right-hand side, the reduced camera
// matrix or its SCHUR_JACOBI block diagonal, and one camera-space vector per PCG
// product -- are summed across ranks: operands of up to 1 MB by the peer-memory all-reduce kernel
// below, larger ones by fp64 ncclAllReduce.  lm.cu lets partial sums that meet at the same point share one collective.
//
// NCCL is resolved at run time with dlopen("libnccl.so.2"): inside a PyTorch process
// this binds to the copy torch already loaded (one NCCL per process); in a plain C++
// host it binds to the system library.  Nothing here is needed for single-GPU use.
#include <dlfcn.h>
#include <nccl.h>

#include <atomic>
#include <mutex>

#include "solver.h"

namespace sba {

namespace {
struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool Load() {
    if (lib) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (lib) break;
    }
    if (!lib) return false;
    GetUniqueId = (decltype(GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    CommInitRank = (decltype(CommInitRank))dlsym(lib, "ncclCommInitRank");
    CommDestroy = (decltype(CommDestroy))dlsym(lib, "ncclCommDestroy");
    AllReduce = (decltype(AllReduce))dlsym(lib, "ncclAllReduce");
    GetErrorString = (decltype(GetErrorString))dlsym(lib, "ncclGetErrorString");
    return GetUniqueId && CommInitRank && CommDestroy && AllReduce && GetErrorString;
  }
};
NcclApi g_nccl;
}  // namespace

// ---- small all-reduces over NVLink peer memory --------------------------------------------------------------------
// The LM iteration's collectives are short (a few scalars, a camera-space vector, the reduced camera matrix of a
// small problem) and latency-bound: at 8 GPUs ncclAllReduce measured ~17 us more per call than this kernel (C2 weak, 30
// collectives per solve: 3.47 vs 2.95 ms, profiles/r01_bench_c2_8gpu_weak*.json).  For up to kP2PCapacity
// doubles every rank instead owns one cudaMalloc'ed exchange area that all peers map through CUDA IPC.  One launch:
// CTA b copies its slice of the operand into its slot of the local area, raises flag (rank, b) on every peer with a
// system-scope release store, waits for the same flag from every peer, then reads slot b of every rank over NVLink and
// combines them IN RANK ORDER -- so every rank computes bit-identical results, which the replicated camera-side state
// relies on.  Slots are double-buffered on the parity of a per-CTA epoch kept in device memory (a peer can run at most
// one collective ahead: it cannot finish collective e+1 without this rank's flag e+1), and nothing in the protocol is a
// kernel argument that changes between launches, so the launch replays inside a CUDA graph.
constexpr int kP2PCtas = 128, kP2PThreads = 512, kP2PSlot = 1024, kP2PMaxRanks = 8;
constexpr size_t kP2PCapacity = (size_t)kP2PCtas * kP2PSlot;                         // doubles per collective
constexpr size_t kP2PDataBytes = 2 * kP2PCapacity * sizeof(double);                  // two parities
constexpr size_t kP2PAreaBytes = kP2PDataBytes + (size_t)(kP2PMaxRanks + 1) * kP2PCtas * sizeof(unsigned);

struct P2PView {
  double* data[kP2PMaxRanks];     // every rank's exchange area as mapped here ([rank] is the local one)
  unsigned* flags[kP2PMaxRanks];  // every rank's flag array [kP2PMaxRanks][kP2PCtas]
  unsigned* epoch;                // local: collectives completed, per CTA
  int nranks, rank;
};

__device__ __forceinline__ void StoreReleaseSys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned LoadAcquireSys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double LoadRelaxedSys(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ double2 LoadRelaxedSys2(const double* p) {  // p is 16-byte aligned
  double2 v;
  asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
  return v;
}

template <bool MAX>
__global__ void __launch_bounds__(kP2PThreads) p2p_allreduce_kernel(P2PView V, double* __restrict__ buf, int n) {
  const int b = blockIdx.x, t = threadIdx.x;
  __shared__ unsigned s_epoch;
  if (t == 0) s_epoch = V.epoch[b] + 1;
  __syncthreads();
  const unsigned e = s_epoch;
  const int per = (n + gridDim.x - 1) / gridDim.x, lo = b * per, cnt = max(0, min(per, n - lo));
  const size_t slot = ((size_t)(e & 1) * kP2PCtas + b) * kP2PSlot;
  double* mine = V.data[V.rank] + slot;
  for (int i = t; i < cnt; i += kP2PThreads) mine[i] = buf[lo + i];
  __syncthreads();
  if (t < V.nranks && t != V.rank) {
    // (the release is cumulative over the barrier above: the whole CTA's slot writes are visible before the flag)
    StoreReleaseSys(V.flags[t] + (size_t)V.rank * kP2PCtas + b, e);
    const unsigned* f = V.flags[V.rank] + (size_t)t * kP2PCtas + b;
    // a peer that died or left the sequence must not leave this GPU spinning for ever: after 30 s the kernel traps and
    // every later CUDA call of this process reports the failure
    unsigned long long t0 = 0;
    for (unsigned spins = 0; (int)(LoadAcquireSys(f) - e) < 0; ++spins) {
      if ((spins & 0xfff) != 0xfff) continue;
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 30000000000ull) __trap();
    }
  }
  __syncthreads();
  // all ranks' loads are issued before the first use (an NVLink round trip is ~2 us: a load-add chain would pay it per rank)
  for (int i = 2 * t; i < cnt; i += 2 * kP2PThreads) {
    double2 v[kP2PMaxRanks];
#pragma unroll
    for (int r = 0; r < kP2PMaxRanks; ++r)
      if (r < V.nranks) v[r] = LoadRelaxedSys2(V.data[r] + slot + i);  // (i + 1 may lie past cnt, never past the slot)
    double2 a = v[0];
#pragma unroll
    for (int r = 1; r < kP2PMaxRanks; ++r)
      if (r < V.nranks) {
        a.x = MAX ? fmax(a.x, v[r].x) : a.x + v[r].x;
        a.y = MAX ? fmax(a.y, v[r].y) : a.y + v[r].y;
      }
    buf[lo + i] = a.x;
    if (i + 1 < cnt) buf[lo + i + 1] = a.y;
  }
  if (t == 0) V.epoch[b] = e;
}

class Comm {
 public:
  ncclComm_t comm = nullptr;
  int nranks = 1, rank = 0;
  ncclUniqueId id{};               // the group this communicator was created for (cache key, with the shape)
  std::atomic_flag in_use = ATOMIC_FLAG_INIT;  // held by the one API call that is issuing collectives (CommAcquire)
  bool p2p = false;  // the exchange areas are mapped on every rank
  P2PView view{};
  // second peer-mapped area, sized by the first problem that needs it: the vectors of the row-sharded PCG (pcg_blocks.cu)
  void* pcg_area[kP2PMaxRanks] = {nullptr};
  size_t pcg_bytes = 0;
};

// cudaMalloc's `bytes` (zeroed) on this rank and maps every rank's block here through CUDA IPC.  Collective over the
// communicator (NCCL on the default stream, with a consensus after each step): true on every rank or false on every rank.
static bool MapPeerArea(Comm* c, size_t bytes, void** peer) {
  const int nr = c->nranks, hw = (int)(sizeof(cudaIpcMemHandle_t) / sizeof(int)), words = nr * (hw + 1);
  int ok = 1;
  void* area = nullptr;
  if (cudaMalloc(&area, bytes) != cudaSuccess) { ok = 0; area = nullptr; }
  if (ok && cudaMemset(area, 0, bytes) != cudaSuccess) ok = 0;
  cudaIpcMemHandle_t mine;
  std::memset(&mine, 0, sizeof(mine));
  if (ok && cudaIpcGetMemHandle(&mine, area) != cudaSuccess) ok = 0;
  cudaGetLastError();
  std::vector<int> h((size_t)words, 0);
  std::memcpy(h.data() + (size_t)c->rank * hw, &mine, sizeof(mine));
  h[(size_t)nr * hw + c->rank] = ok;
  int* d = nullptr;
  auto gather = [&]() -> bool {  // a sum over slots only their owner fills = an all-gather
    if (cudaMemcpy(d, h.data(), words * sizeof(int), cudaMemcpyHostToDevice) != cudaSuccess) return false;
    if (g_nccl.AllReduce(d, d, words, ncclInt32, ncclSum, c->comm, nullptr) != ncclSuccess) return false;
    if (cudaStreamSynchronize(nullptr) != cudaSuccess) return false;
    return cudaMemcpy(h.data(), d, words * sizeof(int), cudaMemcpyDeviceToHost) == cudaSuccess;
  };
  auto all_ok = [&]() { for (int r = 0; r < nr; ++r) if (h[(size_t)nr * hw + r] != 1) return false; return true; };
  if (cudaMalloc(&d, words * sizeof(int)) != cudaSuccess) { cudaGetLastError(); if (area) cudaFree(area); return false; }  // (cannot even vote: the peers' gather fails loudly)
  bool good = gather() && all_ok();
  for (int r = 0; r < nr; ++r) peer[r] = nullptr;
  if (good) {
    for (int r = 0; r < nr; ++r) {
      if (r == c->rank) { peer[r] = area; continue; }
      cudaIpcMemHandle_t hr;
      std::memcpy(&hr, h.data() + (size_t)r * hw, sizeof(hr));
      if (cudaIpcOpenMemHandle(&peer[r], hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; peer[r] = nullptr; }
    }
    cudaGetLastError();
    std::fill(h.begin(), h.end(), 0);
    h[(size_t)nr * hw + c->rank] = ok;
    good = gather() && all_ok();  // (also a barrier: every area is zeroed and mapped before the first flag is raised)
  }
  cudaFree(d);
  if (!good) {
    for (int r = 0; r < nr; ++r) if (r != c->rank && peer[r]) cudaIpcCloseMemHandle(peer[r]);
    if (area) cudaFree(area);
    for (int r = 0; r < nr; ++r) peer[r] = nullptr;
    cudaGetLastError();
    return false;
  }
  return true;
}

static int NcclFail(const char* what, ncclResult_t r) {
  SetError(SBA_ERR_COMM, std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "nccl error"));
  return SBA_ERR_COMM;
}

// Maps every rank's exchange area into this process.  Every step is followed by a consensus over NCCL, so either all
// ranks switch to the peer-memory path or none does (SBA_NO_P2P_ALLREDUCE=1 keeps every collective on NCCL).
static void SetupP2P(Comm* c) {
  if (getenv("SBA_NO_P2P_ALLREDUCE") || c->nranks > kP2PMaxRanks) return;
  void* peer[kP2PMaxRanks] = {nullptr};
  if (!MapPeerArea(c, kP2PAreaBytes, peer)) return;
  const int nr = c->nranks;
  P2PView& V = c->view;
  V.nranks = nr; V.rank = c->rank;
  for (int r = 0; r < nr; ++r) {
    V.data[r] = static_cast<double*>(peer[r]);
    V.flags[r] = reinterpret_cast<unsigned*>(static_cast<char*>(peer[r]) + kP2PDataBytes);
  }
  V.epoch = V.flags[c->rank] + (size_t)kP2PMaxRanks * kP2PCtas;
  c->p2p = true;
}

// The peer-mapped area of the row-sharded PCG: every rank's block of >= `bytes`, as mapped here (area[rank] is the local
// one).  Collective: every rank of the group calls it at the same point of its first solve.  SBA_ERR_COMM when the
// group has no peer mapping (the caller then keeps the replicated solve).
int CommPcgArea(Problem* p, size_t bytes, void** area, int* nranks, int* rank) {
  Comm* c = p->comm;
  if (!c || !c->p2p) return SBA_ERR_COMM;
  if (c->pcg_bytes < bytes) {
    SBA_CUDA(cudaDeviceSynchronize());
    if (c->pcg_bytes) {
      for (int r = 0; r < c->nranks; ++r) {
        if (r != c->rank && c->pcg_area[r]) cudaIpcCloseMemHandle(c->pcg_area[r]);
        else if (c->pcg_area[r]) cudaFree(c->pcg_area[r]);
        c->pcg_area[r] = nullptr;
      }
      c->pcg_bytes = 0;
    }
    const size_t want = (bytes + (1u << 20) - 1) & ~((size_t)(1u << 20) - 1);
    if (!MapPeerArea(c, want, c->pcg_area)) { SetError(SBA_ERR_COMM, "peer mapping of the PCG exchange area failed"); return SBA_ERR_COMM; }
    c->pcg_bytes = want;
  }
  for (int r = 0; r < c->nranks; ++r) area[r] = c->pcg_area[r];
  *nranks = c->nranks; *rank = c->rank;
  return SBA_OK;
}

int CommGetUniqueId(void* id128) {
  static_assert(sizeof(ncclUniqueId) == SBA_COMM_ID_BYTES, "ncclUniqueId size");
  if (!g_nccl.Load()) { SetError(SBA_ERR_COMM, "libnccl.so.2 could not be loaded"); return SBA_ERR_COMM; }
  ncclUniqueId id;
  const ncclResult_t r = g_nccl.GetUniqueId(&id);
  if (r != ncclSuccess) return NcclFail("ncclGetUniqueId", r);
  std::memcpy(id128, &id, sizeof(id));
  return SBA_OK;
}

int CommInit(Problem* p, int nranks, int rank, const void* id128) {
  if (nranks <= 1) return SBA_OK;
  if (nranks != p->shard_count || rank != p->shard_rank) {
    SetError(SBA_ERR_INVALID_ARGUMENT, "communicator shape differs from the problem's shard_count / shard_rank");
    return SBA_ERR_INVALID_ARGUMENT;
  }
  if (!g_nccl.Load()) { SetError(SBA_ERR_COMM, "libnccl.so.2 could not be loaded"); return SBA_ERR_COMM; }
  // One communicator per process and GROUP (unique id + shape), shared by every handle that names that group (creating a
  // communicator costs far more than a solve); it lives until the process exits.  A new id is a new group: it gets a
  // communicator and exchange areas of its own, never the mappings of an older group.
  // All handles of a group share one exchange area and one epoch sequence, so their collectives must not interleave:
  // the API calls that issue collectives take the communicator for their duration (CommAcquire) and a second handle
  // trying to solve at the same time is refused with SBA_ERR_COMM instead of corrupting the sums.
  static std::vector<Comm*> cache;
  static std::mutex cache_mu;
  std::lock_guard<std::mutex> lk(cache_mu);
  ncclUniqueId id;
  std::memcpy(&id, id128, sizeof(id));
  for (Comm* c : cache)
    if (c->nranks == nranks && c->rank == rank && std::memcmp(&c->id, &id, sizeof(id)) == 0) { p->comm = c; return SBA_OK; }
  Comm* c = new Comm();
  c->nranks = nranks; c->rank = rank; c->id = id;
  const ncclResult_t r = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
  if (r != ncclSuccess) { delete c; return NcclFail("ncclCommInitRank", r); }
  SetupP2P(c);
  cache.push_back(c);
  p->comm = c;
  return SBA_OK;
}

bool CommHasPeerMemory(const Problem* p) { return p->comm && p->comm->p2p; }
void CommDestroy(Problem* p) { p->comm = nullptr; }  // communicators are process-wide (see CommInit)

// Exclusive use of the handle's communicator for one API call (no-op on a single GPU).  false: another handle of the
// same group is inside a call that issues collectives.
bool CommAcquire(Problem* p) { return !p->comm || !p->comm->in_use.test_and_set(std::memory_order_acquire); }
void CommRelease(Problem* p) { if (p->comm) p->comm->in_use.clear(std::memory_order_release); }

static int AllReduce(Problem* p, double* d_buf, size_t n, ncclRedOp_t op) {
  if (!p->comm || n == 0) return SBA_OK;
  if (p->comm->p2p && n <= kP2PCapacity && (op == ncclSum || op == ncclMax)) {  // latency-bound sizes: peer memory, one launch
    // (a CTA's epoch advances only in the collectives it takes part in; the same CTAs take part on every rank)
    const int grid = (int)std::min<size_t>(kP2PCtas, (n + 1023) / 1024);
    if (op == ncclSum) p2p_allreduce_kernel<false><<<grid, kP2PThreads, 0, p->stream>>>(p->comm->view, d_buf, (int)n);
    else p2p_allreduce_kernel<true><<<grid, kP2PThreads, 0, p->stream>>>(p->comm->view, d_buf, (int)n);
    SBA_CUDA(cudaGetLastError());
    p->collectives++; p->launches++;
    return SBA_OK;
  }
  const ncclResult_t r = g_nccl.AllReduce(d_buf, d_buf, n, ncclDouble, op, p->comm->comm, p->stream);
  if (r != ncclSuccess) return NcclFail("ncclAllReduce", r);
  p->collectives++;
  return SBA_OK;
}
int AllReduceSum(Problem* p, double* d_buf, size_t n) { return AllReduce(p, d_buf, n, ncclSum); }
int AllReduceSumWithRhs(Problem* p, double* d_buf, size_t n) {
  if (!p->comm || !p->rhs_unreduced || p->n_cdim == 0) { p->rhs_unreduced = false; return AllReduceSum(p, d_buf, n); }
  const size_t nb = (size_t)p->n_cdim * sizeof(double);
  SBA_CUDA(cudaMemcpyAsync(d_buf + n, p->d_rhs.p, nb, cudaMemcpyDeviceToDevice, p->stream));
  const int rc = AllReduceSum(p, d_buf, n + p->n_cdim);
  if (rc) return rc;
  SBA_CUDA(cudaMemcpyAsync(p->d_rhs.p, d_buf + n, nb, cudaMemcpyDeviceToDevice, p->stream));
  p->rhs_unreduced = false;
  return SBA_OK;
}
int AllReduceMax(Problem* p, double* d_buf, size_t n) { return AllReduce(p, d_buf, n, ncclMax); }
int AllReduceMaxInt(Problem* p, int* d_buf, size_t n) {
  if (!p->comm || n == 0) return SBA_OK;
  const ncclResult_t r = g_nccl.AllReduce(d_buf, d_buf, n, ncclInt32, ncclMax, p->comm->comm, p->stream);
  if (r != ncclSuccess) return NcclFail("ncclAllReduce", r);
  p->collectives++;
  return SBA_OK;
}

}  // namespace sba
