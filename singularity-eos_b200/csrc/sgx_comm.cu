// The one exchange step of the path: the optional all-reduce of the convergence-statistics
// vector (uint64[SGX_PTE_NSTATS]) over the ranks that own the shards.  NCCL is
// bound at run time (dlopen "libnccl.so.2": the copy the host application / PyTorch already
// loaded, else the system one), so the library carries no link-time NCCL dependency and a
// single-GPU user never touches it.
#include <dlfcn.h>

#include <cstring>

#include "sgx_internal.hpp"

namespace {

struct NcclUniqueId {
  char internal[128];
};
using ncclComm_t = void *;
enum { kNcclSuccess = 0, kNcclUint64 = 5, kNcclSum = 0 };

struct NcclApi {
  int (*GetUniqueId)(NcclUniqueId *) = nullptr;
  int (*CommInitRank)(ncclComm_t *, int, NcclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
  bool ok = false;
};

const NcclApi &nccl() {
  static const NcclApi api = [] {
    NcclApi a;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return a;
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(h, "ncclAllReduce"));
    a.GetErrorString =
        reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
    a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllReduce;
    return a;
  }();
  return api;
}

int nccl_fail(const char *what, int rc) {
  const NcclApi &a = nccl();
  return sgx::fail(SGX_ERR_CUDA, "%s failed: %s", what,
                   a.GetErrorString ? a.GetErrorString(rc) : "NCCL error");
}

} // namespace

struct sgx_comm_s {
  ncclComm_t comm = nullptr;
  bool owned = false;
};

using namespace sgx;

extern "C" {

int sgx_comm_unique_id(char id[SGX_COMM_ID_BYTES]) {
  if (!id) return fail(SGX_ERR_INVALID, "null id buffer");
  const NcclApi &a = nccl();
  if (!a.ok) return fail(SGX_ERR_UNSUPPORTED, "libnccl.so.2 could not be loaded");
  NcclUniqueId u;
  const int rc = a.GetUniqueId(&u);
  if (rc != kNcclSuccess) return nccl_fail("ncclGetUniqueId", rc);
  static_assert(SGX_COMM_ID_BYTES == sizeof(NcclUniqueId), "ncclUniqueId is 128 bytes");
  memcpy(id, u.internal, sizeof(u.internal));
  return SGX_OK;
}

int sgx_comm_create(const char id[SGX_COMM_ID_BYTES], int nranks, int rank, sgx_comm_t *out) {
  if (!id || !out) return fail(SGX_ERR_INVALID, "null argument");
  if (nranks < 1 || rank < 0 || rank >= nranks)
    return fail(SGX_ERR_INVALID, "bad rank %d of %d", rank, nranks);
  const NcclApi &a = nccl();
  if (!a.ok) return fail(SGX_ERR_UNSUPPORTED, "libnccl.so.2 could not be loaded");
  NcclUniqueId u;
  memcpy(u.internal, id, sizeof(u.internal));
  ncclComm_t c = nullptr;
  const int rc = a.CommInitRank(&c, nranks, u, rank); // on the calling thread's current device
  if (rc != kNcclSuccess) return nccl_fail("ncclCommInitRank", rc);
  sgx_comm_s *h = new sgx_comm_s;
  h->comm = c;
  h->owned = true;
  *out = h;
  return SGX_OK;
}

int sgx_comm_from_nccl(void *nccl_comm, sgx_comm_t *out) {
  if (!nccl_comm || !out) return fail(SGX_ERR_INVALID, "null argument");
  if (!nccl().ok) return fail(SGX_ERR_UNSUPPORTED, "libnccl.so.2 could not be loaded");
  sgx_comm_s *h = new sgx_comm_s;
  h->comm = nccl_comm;
  h->owned = false;
  *out = h;
  return SGX_OK;
}

int sgx_comm_destroy(sgx_comm_t c) {
  if (!c) return SGX_OK;
  if (c->owned && c->comm && nccl().ok) nccl().CommDestroy(c->comm);
  delete c;
  return SGX_OK;
}

int sgx_stats_allreduce(sgx_comm_t c, uint64_t *stats, int n, sgx_stream_t stream) {
  if (!c || !c->comm) return fail(SGX_ERR_INVALID, "null communicator");
  if (!stats || n < 0) return fail(SGX_ERR_INVALID, "bad statistics vector");
  if (n == 0) return SGX_OK;
  const int rc = nccl().AllReduce(stats, stats, (size_t)n, kNcclUint64, kNcclSum, c->comm,
                                  (cudaStream_t)stream);
  if (rc != kNcclSuccess) return nccl_fail("ncclAllReduce", rc);
  return SGX_OK;
}

} // extern "C"
