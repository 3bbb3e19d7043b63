// PTESolverRhoT over mixtures with tabulated materials (modifier layers) through the lane-refill
// state machine (sgx_pte_sm.cuh) on the TabEos accessor (sgx_pte_tab.cuh), 1..4 materials, with
// the per-material lambda caches held per lane.  Same arithmetic per cell as the per-cell kernels
// (sgx_pte_tab*.cu), which stay reachable with SGX_PTE_KERNEL=percell.
#include "sgx_pte_rhot_launch.cuh"
#include "sgx_pte_tab.cuh"

namespace sgx {

int launch_pte_tab_rhot_sm_mod(const EosSet &set, const MixParamsDev &prm, double *rho, double *vfrac, double *sie,
         double *temp, double *press, const double *sie_tot, const double *vfrac_tot,
         const double *Tguess, uint64_t *status, double *residual, uint64_t *stats,
         int64_t ncell, cudaStream_t s, double *lambda, const unsigned char *skip) {
#define SGX_ARGS                                                                           \
  set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, status, residual,    \
      stats, ncell, s, lambda, skip
  raise_stack_limit_once();
  switch (set.nmat) {
  case 1:
    return launch_sm<1, TabEos>(SGX_ARGS);
  case 2:
    return launch_sm<2, TabEos>(SGX_ARGS);
  case 3:
    return launch_sm<3, TabEos>(SGX_ARGS);
  default:
    return launch_sm<4, TabEos>(SGX_ARGS);
  }
#undef SGX_ARGS
}

} // namespace sgx
