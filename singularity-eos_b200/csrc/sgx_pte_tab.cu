// PTE solves for mixtures that hold tabulated materials (SpinerEOSDependsRhoT / RhoSie next to
// the analytic models): the per-cell kernels instantiated on the TabEos accessor
// (sgx_pte_tab.cuh), with the per-material lambda caches of the reference's CacheAccessor.
// PTESolverRhoT, PTESolverFixedT, PTESolverFixedP (1..4 materials), PTESolverPT and
// PTESolverRhoU (2..4); every solver family compiles in a translation unit of its own, the
// four-material PT / RhoU / FixedP instantiations in a further one each (this one: the
// dispatcher and RhoT).
#include <cstdlib>
#include <cstring>

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

int launch_pte_tab_fixedt(SGX_TAB_PARAMS);
int launch_pte_tab_pt(SGX_TAB_PARAMS);
int launch_pte_tab_rhou(SGX_TAB_PARAMS);
int launch_pte_tab_fixedp(SGX_TAB_PARAMS);
int launch_pte_tab_pt4(SGX_TAB_PARAMS);
int launch_pte_tab_rhou4(SGX_TAB_PARAMS);
int launch_pte_tab_fixedp4(SGX_TAB_PARAMS);
int launch_pte_tab_rhot_plain(SGX_TAB_PARAMS);
int launch_pte_tab_rhot_sm_plain(SGX_TAB_PARAMS, const unsigned char *skip);
int launch_pte_tab_rhot_sm_mod(SGX_TAB_PARAMS, const unsigned char *skip);

int launch_pte_tab(int solver, const EosSet &set, const MixParamsDev &prm, double *rho,
                   double *vfrac, double *sie, double *temp, double *press,
                   const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                   uint64_t *status, double *residual, uint64_t *stats, int64_t ncell,
                   cudaStream_t s, double *lambda) {
  if (solver == SGX_PTE_FIXED_T) return launch_pte_tab_fixedt(SGX_TAB_ARGS);
  if (set.nmat == 4) { // sgx_pte_tab_{pt,rhou,fixedp}4.cu
    if (solver == SGX_PTE_PT) return launch_pte_tab_pt4(SGX_TAB_ARGS);
    if (solver == SGX_PTE_RHOU) return launch_pte_tab_rhou4(SGX_TAB_ARGS);
    if (solver == SGX_PTE_FIXED_P) return launch_pte_tab_fixedp4(SGX_TAB_ARGS);
  }
  if (solver == SGX_PTE_PT) return launch_pte_tab_pt(SGX_TAB_ARGS);
  if (solver == SGX_PTE_RHOU) return launch_pte_tab_rhou(SGX_TAB_ARGS);
  if (solver == SGX_PTE_FIXED_P) return launch_pte_tab_fixedp(SGX_TAB_ARGS);
  bool modified = false;
  for (int m = 0; m < set.nmat; ++m)
    modified = modified || set.mod[m].flags != 0;
  // RhoT: the lane-refill state machine (sgx_pte_tab_sm*.cu); SGX_PTE_KERNEL=percell keeps the
  // thread-per-cell kernels (sgx_pte_tab_plain.cu and below), bit-identical results
  const char *variant = getenv("SGX_PTE_KERNEL"); // read per call: the tests switch it
  const bool percell = variant && strcmp(variant, "percell") == 0;
  if (solver == SGX_PTE_RHOT && !percell && set.nmat >= 1 && set.nmat <= 4)
    return modified ? launch_pte_tab_rhot_sm_mod(SGX_TAB_ARGS, nullptr)
                    : launch_pte_tab_rhot_sm_plain(SGX_TAB_ARGS, nullptr);
  if (solver == SGX_PTE_RHOT && !modified && set.nmat >= 1 && set.nmat <= 4)
    return launch_pte_tab_rhot_plain(SGX_TAB_ARGS);
  if (solver == SGX_PTE_RHOT) {
    switch (set.nmat) {
    case 1:
      return launch_pte_percell<1, TabEos, SGX_PTE_RHOT>(SGX_TAB_ARGS);
    case 2:
      return launch_pte_percell<2, TabEos, SGX_PTE_RHOT>(SGX_TAB_ARGS);
    case 3:
      return launch_pte_percell<3, TabEos, SGX_PTE_RHOT>(SGX_TAB_ARGS);
    case 4:
      return launch_pte_percell<4, TabEos, SGX_PTE_RHOT>(SGX_TAB_ARGS);
    default:
      break;
    }
  }
  return fail(SGX_ERR_UNSUPPORTED,
              "PTE with tabulated materials: solver %d with %d materials is not instantiated "
              "(RhoT / FixedT / FixedP: 1..4, PT / RhoU: 2..4)", solver, set.nmat);
}

} // namespace sgx
