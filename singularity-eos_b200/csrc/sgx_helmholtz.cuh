// Helmholtz EOS on device (eos/eos_helmholtz.hpp).  Filled in by sgx_helm.cu.
#ifndef SGX_HELMHOLTZ_CUH_
#define SGX_HELMHOLTZ_CUH_
#include "sgx_math.cuh"
namespace sgx {
struct HelmDev {
  const double *tables = nullptr;
  int options = 0;
};
} // namespace sgx
#endif
