// PTESolverRhoT over mixtures with tabulated materials and NO modifier layer: the per-cell kernel
// on TabEosT<false> (sgx_pte_tab.cuh), 1..4 materials.  A translation unit of its own so that it
// compiles next to the modifier-carrying instantiations of sgx_pte_tab.cu, not after them.
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

int launch_pte_tab_rhot_plain(SGX_TAB_PARAMS) {
  switch (set.nmat) {
  case 1:
    return launch_pte_percell<1, TabEosPlain, SGX_PTE_RHOT>(SGX_TAB_ARGS);
  case 2:
    return launch_pte_percell<2, TabEosPlain, SGX_PTE_RHOT>(SGX_TAB_ARGS);
  case 3:
    return launch_pte_percell<3, TabEosPlain, SGX_PTE_RHOT>(SGX_TAB_ARGS);
  default:
    return launch_pte_percell<4, TabEosPlain, SGX_PTE_RHOT>(SGX_TAB_ARGS);
  }
}

} // namespace sgx
