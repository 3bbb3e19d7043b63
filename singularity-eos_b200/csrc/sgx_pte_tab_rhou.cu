// PTESolverRhoU (closure/mixed_cell_models.hpp:1934-2183) over mixtures with tabulated or modified materials (see sgx_pte_tab.cu): the
// per-cell kernel on the TabEos accessor, lambda caches in the CacheAccessor layout.
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

int launch_pte_tab_rhou(SGX_TAB_PARAMS) {
  switch (set.nmat) {
  case 2:
    return launch_pte_percell<2, TabEos, SGX_PTE_RHOU>(SGX_TAB_ARGS);
  case 3:
    return launch_pte_percell<3, TabEos, SGX_PTE_RHOU>(SGX_TAB_ARGS);
  default:
    return fail(SGX_ERR_UNSUPPORTED, "PTESolverRhoU with tabulated materials: nmat = %d "
                "(2..3 instantiated)", set.nmat);
  }
}

} // namespace sgx
