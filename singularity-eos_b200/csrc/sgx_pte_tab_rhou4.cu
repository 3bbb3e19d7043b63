// PTESolverRhoU (closure/mixed_cell_models.hpp:1934-2183), the 8 x 8 system, over FOUR-material mixtures with tabulated or
// modified materials (see sgx_pte_tab.cu): the per-cell kernel on the TabEos accessor.  A unit of
// its own so that the 2..3-material units keep their build time.
#include "sgx_internal.hpp"
#include "sgx_pte_kernel.cuh"
#include "sgx_pte_tab.cuh"

namespace sgx {

int launch_pte_tab_rhou4(const EosSet &set, const MixParamsDev &prm, double *rho, double *vfrac,
                         double *sie, double *temp, double *press, const double *sie_tot,
                         const double *vfrac_tot, const double *Tguess, uint64_t *status,
                         double *residual, uint64_t *stats, int64_t ncell, cudaStream_t s,
                         double *lambda) {
  return launch_pte_percell<4, TabEos, SGX_PTE_RHOU>(set, prm, rho, vfrac, sie, temp, press,
                                                    sie_tot, vfrac_tot, Tguess, status, residual,
                                                    stats, ncell, s, lambda);
}

} // namespace sgx
