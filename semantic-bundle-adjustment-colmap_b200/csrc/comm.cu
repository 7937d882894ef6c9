// Multi-GPU plumbing: one process per GPU, NCCL over NVLink 5 / NVSwitch.
//
// The reference has no distributed path at all; its only
// parallelism is Ceres' num_threads (/root/reference/src/optim/bundle_adjustment.cc:288-301).
// Here observations are sharded by point ownership (problem.cu) and camera-space
// quantities -- cost scalars, gradient / reduced right-hand side, the reduced camera
// matrix or its SCHUR_JACOBI block diagonal, and one camera-space vector per PCG
// product -- are summed across ranks with fp64 ncclAllReduce.
//
// NCCL is resolved at run time with dlopen("libnccl.so.2"): inside a PyTorch process
// this binds to the copy torch already loaded (one NCCL per process); in a plain C++
// host it binds to the system library.  Nothing here is needed for single-GPU use.
#include <dlfcn.h>
#include <nccl.h>

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

class Comm {
 public:
  ncclComm_t comm = nullptr;
  int nranks = 1, rank = 0;
};

static int NcclFail(const char* what, ncclResult_t r) {
  SetError(SBA_ERR_COMM, std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "nccl error"));
  return SBA_ERR_COMM;
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
  // One communicator per process and shape, shared by every handle (creating a communicator costs
  // far more than a solve); it lives until the process exits.
  static std::vector<Comm*> cache;
  for (Comm* c : cache)
    if (c->nranks == nranks && c->rank == rank) { p->comm = c; return SBA_OK; }
  ncclUniqueId id;
  std::memcpy(&id, id128, sizeof(id));
  Comm* c = new Comm();
  c->nranks = nranks; c->rank = rank;
  const ncclResult_t r = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
  if (r != ncclSuccess) { delete c; return NcclFail("ncclCommInitRank", r); }
  cache.push_back(c);
  p->comm = c;
  return SBA_OK;
}

void CommDestroy(Problem* p) { p->comm = nullptr; }  // communicators are process-wide (see CommInit)

static int AllReduce(Problem* p, double* d_buf, size_t n, ncclRedOp_t op) {
  if (!p->comm || n == 0) return SBA_OK;
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
