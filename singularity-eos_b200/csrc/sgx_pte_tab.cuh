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
// MODS = false: an accessor for mixtures without modifier layers (every `flags &` test below
// folds away; measured: the unmodified tabulated mixture lost 17 % when it carried them).
template <bool MODS>
struct TabEosT {
  const EosSet &set;
  double *lam = nullptr; // this cell's cache: [nmat][kMaxNumLambdas], or null (NullIndexer)
  // participating-material compaction (get_sg_eos: EOSAccessor_, eos/get_sg_eos.hpp:38-47):
  // solver material m is set.e[idx[m]]; null = identity
  const int *idx = nullptr;
  __device__ TabEosT(const EosSet &s) : set(s) {}
  SGX_D int flags_of(const ModParams &q) const { return MODS ? q.flags : 0; }
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
  // ---- the staged interface of the lane-refill state machine (sgx_pte_sm.cuh) ----
  // Tabulated models have no density-only part worth caching (two log transforms); the
  // adapters go through the same layered methods as the per-cell kernels, so both paths do
  // the same arithmetic per cell.
  static constexpr int kCacheSlots = kMaxNumLambdas;
  SGX_D void rho_part(int, double, RhoPart &p) const { p = RhoPart{1.0, 0.0, 0.0, 0.0}; }
  SGX_D double e_from_part(int m, const RhoPart &, double rho, double T) const {
    return e_from_rho_T(m, rho, T);
  }
  // GetPressureFromPreferred(..., do_e_lookup = false) (:583-597)
  SGX_D double P_pref_from_part(int m, const RhoPart &, double rho, double sie,
                                double T) const {
    if (preferred(m) & TQ_SIE) return P_from_rho_e(m, rho, sie);
    return P_from_rho_T(m, rho, T);
  }
  // ---- the unmodified model (the innermost T of BilinearRamp(Scaled(Shifted(T)))) ----
  SGX_D double b_e_from_rho_T(int m, double rho, double T) const {
    SGX_TAB_DISPATCH(m, e_from_rho_T(rho, T), e_from_rho_T(rho, T, cache(m)))
  }
  SGX_D double b_P_from_rho_e(int m, double rho, double e) const {
    SGX_TAB_DISPATCH(m, P_from_rho_e(rho, e), P_from_rho_e(rho, e, cache(m)))
  }
  SGX_D double b_T_from_rho_e(int m, double rho, double e) const {
    SGX_TAB_DISPATCH(m, T_from_rho_e(rho, e), T_from_rho_e(rho, e, cache(m)))
  }
  SGX_D double b_P_from_rho_T(int m, double rho, double T) const {
    SGX_TAB_DISPATCH(m, P_from_rho_T(rho, T), P_from_rho_T(rho, T, cache(m)))
  }
  SGX_D double b_cv_from_rho_T(int m, double rho, double T) const {
    SGX_TAB_DISPATCH(m, cv_from_rho_T(rho, T), cv_from_rho_T(rho, T, cache(m)))
  }
  SGX_D double b_cv_from_rho_e(int m, double rho, double e) const {
    SGX_TAB_DISPATCH(m, cv_from_rho_e(rho, e), cv_from_rho_e(rho, e, cache(m)))
  }
  SGX_D double b_bmod_from_rho_T(int m, double rho, double T) const {
    SGX_TAB_DISPATCH(m, bmod_from_rho_T(rho, T), bmod_from_rho_T(rho, T, cache(m)))
  }
  SGX_D double b_bmod_from_rho_e(int m, double rho, double e) const {
    SGX_TAB_DISPATCH(m, bmod_from_rho_e(rho, e), bmod_from_rho_e(rho, e, cache(m)))
  }
  SGX_D double b_gamma_from_rho_T(int m, double rho, double T) const {
    SGX_TAB_DISPATCH(m, gamma_from_rho_T(rho, T), gamma_from_rho_T(rho, T, cache(m)))
  }
  SGX_D double b_gamma_from_rho_e(int m, double rho, double e) const {
    SGX_TAB_DISPATCH(m, gamma_from_rho_e(rho, e), gamma_from_rho_e(rho, e, cache(m)))
  }
  // ---- modifier layers (sgx_pte.cuh ModParams) ----
  SGX_D const ModParams &mod(const int m) const { return set.mod[idx ? idx[m] : m]; }
  // density and energy as the unmodified model sees them: ScaledEOS hands scale * rho and
  // inv_scale * sie down (scaled_eos.hpp:71-75), ShiftedEOS sie - shift (shifted_eos.hpp:69)
  SGX_D double in_rho(const ModParams &q, const double rho) const {
    return (flags_of(q) & kModScale) ? q.scale * rho : rho;
  }
  SGX_D double in_sie(const ModParams &q, double sie) const {
    if (flags_of(q) & kModScale) sie = q.inv_scale * sie;
    if (flags_of(q) & kModShift) sie = sie - q.shift;
    return sie;
  }
  // BilinearRampEOS::get_ramp_pressure / get_ramp_dpdrho (ramps_eos.hpp:104-124)
  SGX_D double ramp_pressure(const ModParams &q, const double rho) const {
    const double r_r0 = ratio(rho, q.r0);
    return rho < q.r0 ? 0.0 : (rho < q.rmid ? q.a * (r_r0 - 1.0) : q.b * (r_r0 - q.c));
  }
  SGX_D double ramp_dpdrho(const ModParams &q, const double rho) const {
    return rho < q.rmid ? ratio(q.a, q.r0) : ratio(q.b, q.r0);
  }
  SGX_D double e_from_rho_T(int m, double rho, double T) const {
    const ModParams &q = mod(m);
    // sic: ScaledEOS::InternalEnergyFromDensityTemperature passes rho on UNSCALED (:76-80)
    double e = b_e_from_rho_T(m, rho, T);
    if (flags_of(q) & kModShift) e = e + q.shift;
    if (flags_of(q) & kModScale) e = q.scale * e;
    return e;
  }
  SGX_D double P_from_rho_e(int m, double rho, double e) const {
    const ModParams &q = mod(m);
    const double p_eos = b_P_from_rho_e(m, in_rho(q, rho), in_sie(q, e));
    if (!(flags_of(q) & kModRamp)) return p_eos;
    const double p_ramp = ramp_pressure(q, rho);
    return p_eos < p_ramp ? p_ramp : p_eos;
  }
  SGX_D double T_from_rho_e(int m, double rho, double e) const {
    const ModParams &q = mod(m);
    return b_T_from_rho_e(m, in_rho(q, rho), in_sie(q, e));
  }
  SGX_D double P_from_rho_T(int m, double rho, double T) const {
    const ModParams &q = mod(m);
    const double p_eos = b_P_from_rho_T(m, in_rho(q, rho), T);
    if (!(flags_of(q) & kModRamp)) return p_eos;
    const double p_ramp = ramp_pressure(q, rho);
    return p_eos < p_ramp ? p_ramp : p_eos;
  }
  SGX_D double cv_from_rho_T(int m, double rho, double T) const {
    return b_cv_from_rho_T(m, in_rho(mod(m), rho), T);
  }
  SGX_D double cv_from_rho_e(int m, double rho, double e) const {
    const ModParams &q = mod(m);
    return b_cv_from_rho_e(m, in_rho(q, rho), in_sie(q, e));
  }
  SGX_D double bmod_from_rho_T(int m, double rho, double T) const {
    const ModParams &q = mod(m);
    if (flags_of(q) & kModRamp) { // ramps_eos.hpp:196-204
      const double p_ramp = ramp_pressure(q, rho);
      const double p_eos = b_P_from_rho_T(m, in_rho(q, rho), T);
      if (p_eos < p_ramp) return rho * ramp_dpdrho(q, rho);
    }
    return b_bmod_from_rho_T(m, in_rho(q, rho), T);
  }
  SGX_D double bmod_from_rho_e(int m, double rho, double e) const {
    const ModParams &q = mod(m);
    if (flags_of(q) & kModRamp) { // ramps_eos.hpp:160-168
      const double p_ramp = ramp_pressure(q, rho);
      const double p_eos = b_P_from_rho_e(m, in_rho(q, rho), in_sie(q, e));
      if (p_eos < p_ramp) return rho * ramp_dpdrho(q, rho);
    }
    return b_bmod_from_rho_e(m, in_rho(q, rho), in_sie(q, e));
  }
  SGX_D double gamma_from_rho_T(int m, double rho, double T) const {
    return b_gamma_from_rho_T(m, in_rho(mod(m), rho), T);
  }
  SGX_D double gamma_from_rho_e(int m, double rho, double e) const {
    const ModParams &q = mod(m);
    return b_gamma_from_rho_e(m, in_rho(q, rho), in_sie(q, e));
  }
  // bounds: ShiftedEOS and BilinearRampEOS forward, ScaledEOS scales the densities
  // (scaled_eos.hpp:253-269)
  SGX_D double b_RhoPmin(int m, double T) const { SGX_TAB_DISPATCH(m, RhoPmin(T), RhoPmin(T)) }
  SGX_D double b_MaximumDensity(int m) const {
    SGX_TAB_DISPATCH(m, MaximumDensity(), MaximumDensity())
  }
  SGX_D double RhoPmin(int m, double T) const {
    const ModParams &q = mod(m);
    return (flags_of(q) & kModScale) ? q.inv_scale * b_RhoPmin(m, T) : b_RhoPmin(m, T);
  }
  SGX_D double MaximumDensity(int m) const {
    const ModParams &q = mod(m);
    return (flags_of(q) & kModScale) ? q.inv_scale * b_MaximumDensity(m) : b_MaximumDensity(m);
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
  // DensityEnergyFromPressureTemperature: shifted_eos.hpp:398-403, scaled_eos.hpp:245-251,
  // ramps_eos.hpp:522-525 (forwards)
  __device__ inline void rho_e_from_P_T(int m, double P, double T, double &rho,
                                        double &sie) const {
    const ModParams &q = mod(m);
    b_rho_e_from_P_T(m, P, T, rho, sie);
    if (flags_of(q) & kModShift) sie = sie + q.shift;
    if (flags_of(q) & kModScale) {
      rho = rho * q.inv_scale;
      sie = sie * q.scale;
    }
  }
  __device__ inline void b_rho_e_from_P_T(int m, double P, double T, double &rho,
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
    // only IdealGas prefers pressure; through the modifier layers like every other call
    if (mat(m).kind == 0) {
      const ModParams &q = mod(m);
      mat(m).ideal.rho_e_from_P_T(P, T, nullptr, rho, sie);
      if (flags_of(q) & kModShift) sie = sie + q.shift;
      if (flags_of(q) & kModScale) {
        rho = rho * q.inv_scale;
        sie = sie * q.scale;
      }
    }
  }
#undef SGX_TAB_DISPATCH
};
using TabEos = TabEosT<true>;       // any mixture
using TabEosPlain = TabEosT<false>; // no material carries a modifier
template <bool MODS>
SGX_D void set_cell_cache(TabEosT<MODS> &eos, double *p) {
  eos.lam = p;
}
#endif // __CUDACC__

} // namespace sgx
#endif // SGX_PTE_TAB_CUH_
