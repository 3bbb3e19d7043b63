// PTESolverFixedT over mixtures with tabulated materials (see sgx_pte_tab.cu).
#include "sgx_internal.hpp"
#include "sgx_pte_kernel.cuh"
#include "sgx_pte_tab.cuh"

namespace sgx {

#define SGX_TAB_PARAMS                                                                     \
  const EosSet &set, const MixParamsDev &prm, double *rho, double *vfrac, double *sie,     \
      double *temp, double *press, const double *sie_tot, const double *vfrac_tot,         \
      const double *Tguess, uint64_t *status, double *residual, uint64_t *stats,           \
      int64_t ncell, cudaStream_t s, double *lambda
#define SGX_TAB_ARGS                                                                       \
  set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, status, residual,    \
      stats, ncell, s, lambda

int launch_pte_tab_fixedt(SGX_TAB_PARAMS) {
  switch (set.nmat) {
  case 1:
    return launch_pte_percell<1, TabEos, SGX_PTE_FIXED_T>(SGX_TAB_ARGS);
  case 2:
    return launch_pte_percell<2, TabEos, SGX_PTE_FIXED_T>(SGX_TAB_ARGS);
  case 3:
    return launch_pte_percell<3, TabEos, SGX_PTE_FIXED_T>(SGX_TAB_ARGS);
  case 4:
    return launch_pte_percell<4, TabEos, SGX_PTE_FIXED_T>(SGX_TAB_ARGS);
  default:
    return fail(SGX_ERR_UNSUPPORTED, "PTESolverFixedT with tabulated materials: nmat = %d "
                "(1..4 instantiated)", set.nmat);
  }
}

} // namespace sgx
