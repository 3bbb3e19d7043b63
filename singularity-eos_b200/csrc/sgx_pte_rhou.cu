// PTESolverRhoU launchers (closure/mixed_cell_models.hpp:1934-2183), per-cell kernel; a translation
// unit of its own so the solver families compile in parallel.
#include "sgx_internal.hpp"
#include "sgx_pte_kernel.cuh"

namespace sgx {

int launch_pte_rhou(const EosSet &set, bool bench_mix, const MixParamsDev &prm, double *rho,
         double *vfrac, double *sie, double *temp, double *press, const double *sie_tot,
         const double *vfrac_tot, const double *Tguess, uint64_t *status, double *residual,
         uint64_t *stats, int64_t ncell, cudaStream_t s) {
#define SGX_PTE_ARGS                                                                     \
  set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, status, residual,   \
      stats, ncell, s
  if (bench_mix) return launch_pte_percell<3, Mix3GrDrDp, SGX_PTE_RHOU>(SGX_PTE_ARGS);
  switch (set.nmat) {
  case 2:
    return launch_pte_percell<2, DynEos, SGX_PTE_RHOU>(SGX_PTE_ARGS);
  case 3:
    return launch_pte_percell<3, DynEos, SGX_PTE_RHOU>(SGX_PTE_ARGS);
  default: // 4 materials: the 8 x 8 system through the same static-index pivoted Householder
    return launch_pte_percell<4, DynEos, SGX_PTE_RHOU>(SGX_PTE_ARGS);
  }
#undef SGX_PTE_ARGS
}

} // namespace sgx
