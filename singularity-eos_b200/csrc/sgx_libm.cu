// Test hook for sgx_libm.cuh: evaluates the device exp / log / pow / log10 / cbrt over an
// array so tests/test_libm_parity.py can compare them bit for bit with the host libm of the
// GPU box (the libm the reference links).  Declared in include/sgx_b200.h.
#include "sgx_internal.hpp"

namespace sgx {
namespace {
__global__ void libm_kernel(const int which, const double *__restrict__ x,
                            const double *__restrict__ y, double *__restrict__ out,
                            const int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double a = x[i];
    double r;
    switch (which) {
    case SGX_LIBM_EXP: r = libm::exp(a); break;
    case SGX_LIBM_LOG: r = libm::log(a); break;
    case SGX_LIBM_LOG10: r = libm::log10(a); break;
    case SGX_LIBM_CBRT: r = libm::cbrt(a); break;
    default: r = libm::pow(a, y[i]); break;
    }
    out[i] = r;
  }
}
} // namespace
} // namespace sgx

extern "C" int sgx_libm_eval(int which, const double *x, const double *y, double *out,
                             int64_t n, sgx_stream_t stream) {
  using namespace sgx;
  if (n <= 0) return n == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative n");
  if (which < SGX_LIBM_EXP || which > SGX_LIBM_POW) return fail(SGX_ERR_INVALID, "bad function id");
  if (!x || !out || (which == SGX_LIBM_POW && !y)) return fail(SGX_ERR_INVALID, "null array");
  const LaunchCfg c = stream_cfg(n, 256, 8);
  libm_kernel<<<c.blocks, c.threads, 0, (cudaStream_t)stream>>>(which, x, y, out, n);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}
