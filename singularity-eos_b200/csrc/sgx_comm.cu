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

// ---- FP64 pipe peak: the denominator for the FP64-bound kernels ---------
// 8 independent accumulators per thread; fused = DFMA chain, else DMUL + DADD pairs (the
// library is compiled with -fmad=false, so its kernels run on the non-fused rate for a*b+c).
template <int FUSED>
__global__ void __launch_bounds__(256) fp64_peak_kernel(double *out, const int iters,
                                                       const double a, const double b) {
  double x[8];
#pragma unroll
  for (int k = 0; k < 8; ++k)
    x[k] = threadIdx.x * 1e-3 + k;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (FUSED) {
        x[k] = __fma_rn(x[k], a, b);
      } else {
        x[k] = __dadd_rn(__dmul_rn(x[k], a), b);
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k)
    s += x[k];
  if (s == 12345.678) out[0] = s; // keep the chain alive
}

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

int sgx_microbench_fp64(int fused, double *tflops) {
  if (!tflops) return fail(SGX_ERR_INVALID, "null result");
  double *d = nullptr;
  SGX_CUDA_OK(cudaMalloc(&d, sizeof(double)));
  const int iters = 4096, threads = 256;
  const int blocks = sm_count() * 8;
  cudaEvent_t e0, e1;
  SGX_CUDA_OK(cudaEventCreate(&e0));
  SGX_CUDA_OK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    SGX_CUDA_OK(cudaEventRecord(e0));
    if (fused) {
      fp64_peak_kernel<1><<<blocks, threads>>>(d, iters, 1.0000001, 1e-9);
    } else {
      fp64_peak_kernel<0><<<blocks, threads>>>(d, iters, 1.0000001, 1e-9);
    }
    SGX_CUDA_OK(cudaEventRecord(e1));
    SGX_CUDA_OK(cudaEventSynchronize(e1));
    float ms = 0;
    SGX_CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && ms < best) best = ms;
  }
  count_launch(4);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  // 2 flops per element-iteration either way (one multiply and one add)
  *tflops = 2.0 * 8 * iters * (double)threads * blocks / (best * 1e-3) / 1e12;
  return SGX_OK;
}

} // extern "C"
