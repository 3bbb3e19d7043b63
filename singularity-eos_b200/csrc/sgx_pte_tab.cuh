// Mixed cells that hold tabulated materials: the runtime-dispatched EOS
// accessor of the per-cell PTE kernels, extended to SpinerEOSDependsRhoT / RhoSie, with each
// material's lambda cache slice (closure/mixed_cell_models.hpp:307-324 CacheAccessor:
// cache[m] = scratch + m * MAX_NUM_LAMBDAS, MAX_NUM_LAMBDAS = 3, base/constants.hpp:37).
//
// The tabulated EOS structs are too big to travel by value in EosSet (kMaxMat x ~700 bytes): the
// handle keeps a device copy and AnyEos carries its address.  Tabulated models prefer (rho, T)
// (eos_spiner_rho_temp.hpp:269-272), so PTESolverBase takes the P(rho, T) branch of
// GetPressureFromPreferred (:583-597) and the e / cv (rho, T) branch of GetSieCvFromTAndPreferred
// (:601-620) for them, and their RhoPmin(T) is the table's rho_at_pmin curve: the non-trivial
// branch of SetVfracFromT (:423-498) and the density guard of ScaleDx (:1124-1141).
#ifndef SGX_PTE_TAB_CUH_
#define SGX_PTE_TAB_CUH_

#include "sgx_pte.cuh"
#include "sgx_spiner.cuh"

namespace sgx {

constexpr int kMaxNumLambdas = 3; // MAX_NUM_LAMBDAS

#ifdef __CUDACC__
struct TabEos {
  const EosSet &set;
  double *lam = nullptr; // this cell's cache: [nmat][kMaxNumLambdas], or null (NullIndexer)
  // participating-material compaction (get_sg_eos: EOSAccessor_, eos/get_sg_eos.hpp:38-47):
  // solver material m is set.e[idx[m]]; null = identity
  const int *idx = nullptr;
  __device__ TabEos(const EosSet &s) : set(s) {}
  SGX_D const AnyEos &mat(const int m) const { return set.e[idx ? idx[m] : m]; }
  SGX_D double *cache(const int m) const { return lam ? lam + kMaxNumLambdas * m : nullptr; }
  SGX_D bool tab(const int m) const { return mat(m).kind >= SGX_EOS_SPINER_RHOT; }
  SGX_D unsigned long preferred(int m) const {
    if (tab(m)) return TQ_DENSITY | TQ_TEMPERATURE;
    return mat(m).kind == 0 ? IdealGas::preferred_input : (TQ_DENSITY | TQ_SIE);
  }
#define SGX_TAB_DISPATCH(m, CALL, TABCALL)                                                 \
  switch (mat(m).kind) {                                                                 \
  case 0:                                                                                  \
    return mat(m).ideal.CALL;                                                            \
  case 1:                                                                                  \
    return mat(m).gr.CALL;                                                               \
  case 2:                                                                                  \
    return mat(m).dr.CALL;                                                               \
  case 3:                                                                                  \
    return mat(m).dp.CALL;                                                               \
  case 4:                                                                                  \
    return mat(m).rhot->TABCALL;                                                         \
  default:                                                                                 \
    return mat(m).rhosie->TABCALL;                                                       \
  }
  SGX_D double e_from_rho_T(int m, double rho, double T) const {
    SGX_TAB_DISPATCH(m, e_from_rho_T(rho, T), e_from_rho_T(rho, T, cache(m)))
  }
  SGX_D double P_from_rho_e(int m, double rho, double e) const {
    SGX_TAB_DISPATCH(m, P_from_rho_e(rho, e), P_from_rho_e(rho, e, cache(m)))
  }
  SGX_D double T_from_rho_e(int m, double rho, double e) const {
    SGX_TAB_DISPATCH(m, T_from_rho_e(rho, e), T_from_rho_e(rho, e, cache(m)))
  }
  SGX_D double P_from_rho_T(int m, double rho, double T) const {
    SGX_TAB_DISPATCH(m, P_from_rho_T(rho, T), P_from_rho_T(rho, T, cache(m)))
  }
  SGX_D double cv_from_rho_T(int m, double rho, double T) const {
    SGX_TAB_DISPATCH(m, cv_from_rho_T(rho, T), cv_from_rho_T(rho, T, cache(m)))
  }
  SGX_D double cv_from_rho_e(int m, double rho, double e) const {
    SGX_TAB_DISPATCH(m, cv_from_rho_e(rho, e), cv_from_rho_e(rho, e, cache(m)))
  }
  SGX_D double bmod_from_rho_T(int m, double rho, double T) const {
    SGX_TAB_DISPATCH(m, bmod_from_rho_T(rho, T), bmod_from_rho_T(rho, T, cache(m)))
  }
  SGX_D double gamma_from_rho_T(int m, double rho, double T) const {
    SGX_TAB_DISPATCH(m, gamma_from_rho_T(rho, T), gamma_from_rho_T(rho, T, cache(m)))
  }
  SGX_D double RhoPmin(int m, double T) const { SGX_TAB_DISPATCH(m, RhoPmin(T), RhoPmin(T)) }
  SGX_D double MaximumDensity(int m) const {
    SGX_TAB_DISPATCH(m, MaximumDensity(), MaximumDensity())
  }
  SGX_D double MinimumTemperature(int m) const {
    SGX_TAB_DISPATCH(m, MinimumTemperature(), MinimumTemperature())
  }
  SGX_D double MinimumPressure(int m) const {
    SGX_TAB_DISPATCH(m, MinimumPressure(), MinimumPressure())
  }
  SGX_D double MaximumPressureAtTemperature(int m, double T) const {
    SGX_TAB_DISPATCH(m, MaximumPressureAtTemperature(T), MaximumPressureAtTemperature(T))
  }
  __device__ inline void rho_e_from_P_T(int m, double P, double T, double &rho,
                                        double &sie) const {
    switch (mat(m).kind) {
    case 0:
      mat(m).ideal.rho_e_from_P_T(P, T, nullptr, rho, sie);
      return;
    case 1:
      mat(m).gr.rho_e_from_P_T(P, T, nullptr, rho, sie);
      return;
    case 2:
      mat(m).dr.rho_e_from_P_T(P, T, nullptr, rho, sie);
      return;
    case 3:
      mat(m).dp.rho_e_from_P_T(P, T, nullptr, rho, sie);
      return;
    case 4:
      mat(m).rhot->rho_e_from_P_T(P, T, cache(m), rho, sie);
      return;
    default:
      mat(m).rhosie->rho_e_from_P_T(P, T, cache(m), rho, sie);
      return;
    }
  }
  __device__ inline void rho_e_from_P_T_pressure_preferred(int m, double P, double T,
                                                           double &rho, double &sie) const {
    if (mat(m).kind == 0) mat(m).ideal.rho_e_from_P_T(P, T, nullptr, rho, sie);
  }
#undef SGX_TAB_DISPATCH
};
SGX_D void set_cell_cache(TabEos &eos, double *p) { eos.lam = p; }
#endif // __CUDACC__

} // namespace sgx
#endif // SGX_PTE_TAB_CUH_
