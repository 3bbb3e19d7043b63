// Helmholtz EOS: table loader + FillEos kernels (eos/eos_helmholtz.hpp).
#include "sgx_internal.hpp"
namespace sgx {
int build_helmholtz(const char *, const sgx_helm_options *, sgx_eos_s *) {
  return fail(SGX_ERR_UNSUPPORTED, "Helmholtz not built yet");
}
int launch_helm_fill(sgx_eos_s *, double *, double *, double *, double *, double *, double *,
                     int64_t, unsigned long, double *, int, cudaStream_t) {
  return fail(SGX_ERR_UNSUPPORTED, "Helmholtz not built yet");
}
int launch_helm_eval(sgx_eos_s *, int, const double *, const double *, double *, int64_t,
                     double *, int, cudaStream_t) {
  return fail(SGX_ERR_UNSUPPORTED, "Helmholtz not built yet");
}
} // namespace sgx
