// PTE mixed-cell closure on device: PTESolverRhoT, PTESolverPT, the PTESolver Newton /
// line-search driver, ApproxTemperatureFromRhoMatU and the small dense solve.  One thread
// owns one cell; all solver scratch (the reference's 39 / 18 doubles per cell) lives in
// registers / local memory, never in HBM.
//
// Reference: singularity-eos/closure/mixed_cell_models.hpp
//   :43-78    MixParams                   :80-87     SolverStatus
//   :161-278  check_nans, solve_Ax_b_wscr :287-305   sum_neumaier
//   :329-786  PTESolverBase               :790-847   ApproxTemperatureFromRhoMatU
//   :883-1193 PTESolverRhoT               :1207-1446 PTESolverPT
//   :2188-2291 PTESolver
// Sequential order of every floating-point reduction is preserved (Neumaier sums, Jacobian
// assembly, line-search acceptance) so converged states and flags match the reference.
#ifndef SGX_PTE_CUH_
#define SGX_PTE_CUH_

#include "sgx_analytic.cuh"
#include "sgx_linsolve.cuh"
#include "sgx_math.cuh"

namespace sgx {

constexpr int kMaxMat = 8;

struct MixParamsDev {
  int iterate_t_guess;
  double derivative_eps;
  double pte_rel_tolerance_p, pte_rel_tolerance_t, pte_rel_tolerance_e, pte_rel_tolerance_v;
  double pte_abs_tolerance_p, pte_abs_tolerance_t;
  double pte_slow_convergence_thresh;
  double pte_rel_tolerance_p_sufficient, pte_rel_tolerance_t_sufficient,
      pte_rel_tolerance_e_sufficient, pte_rel_tolerance_v_sufficient;
  double pte_abs_tolerance_p_sufficient, pte_abs_tolerance_t_sufficient;
  unsigned long long pte_max_iter_per_mat;
  double line_search_alpha;
  unsigned long long line_search_max_iter;
  double line_search_fac;
  double vfrac_safety_fac;
  double temperature_limit;
  double default_tguess;
  double min_dtde;
  unsigned long long pte_small_step_tries;
  double pte_small_step_thresh;
  double pte_max_dpdv;
  long long pte_reduced_system_exclude_idx;
};

struct SolverStatusDev {
  bool converged = false;
  bool slow_convergence_detected = false;
  unsigned long long max_niter = 0;
  unsigned long long max_line_niter = 0;
  unsigned long long small_step_iters = 0;
  double residual;
};

struct SpinerRhoT;
struct SpinerRhoSie;
// One material = tagged union of the analytic models (the Variant of the reference); the
// tabulated models travel as the address of the handle's device copy (sgx_pte_tab.cuh).
struct AnyEos {
  int kind; // enum sgx_eos_kind
  union {
    IdealGas ideal;
    Gruneisen gr;
    DavisReactants dr;
    DavisProducts dp;
    const SpinerRhoT *rhot;
    const SpinerRhoSie *rhosie;
  };
};
// The modifiers the reference's default variant wraps around every model
// (eos/eos_type_lists.hpp:69-94), always nested BilinearRamp(Scaled(Shifted(T))); each layer is
// optional.  eos/modifiers/shifted_eos.hpp, scaled_eos.hpp, ramps_eos.hpp.  Applied by the
// run-time accessor TabEos (sgx_pte_tab.cuh).
enum { kModShift = 1, kModScale = 2, kModRamp = 4 };
struct ModParams {
  int flags = 0;
  double shift = 0.0;                    // ShiftedEOS::shift_
  double scale = 1.0, inv_scale = 1.0;   // ScaledEOS::scale_, inv_scale_ = 1 / scale_
  double r0 = 0.0, a = 0.0, b = 0.0, c = 0.0, rmid = 0.0, Pmid = 0.0; // BilinearRampEOS
};
struct EosSet {
  int nmat;
  AnyEos e[kMaxMat];
  ModParams mod[kMaxMat];
};

#ifdef __CUDACC__

#define SGX_EOS_DISPATCH(m, CALL)                                                        \
  switch (set.e[m].kind) {                                                               \
  case 0:                                                                                \
    return set.e[m].ideal.CALL;                                                          \
  case 1:                                                                                \
    return set.e[m].gr.CALL;                                                             \
  case 2:                                                                                \
    return set.e[m].dr.CALL;                                                             \
  default:                                                                               \
    return set.e[m].dp.CALL;                                                             \
  }

// Runtime-dispatched accessor (eos[m].X(...) through the Variant's visit,
// eos/eos_variant.hpp:236-267).
struct DynEos {
  const EosSet &set;
  __device__ DynEos(const EosSet &s) : set(s) {}
  SGX_D unsigned long preferred(int m) const {
    switch (set.e[m].kind) {
    case 0:
      return IdealGas::preferred_input;
    default:
      return TQ_DENSITY | TQ_SIE;
    }
  }
  SGX_D double e_from_rho_T(int m, double rho, double T) const {
    SGX_EOS_DISPATCH(m, e_from_rho_T(rho, T))
  }
  SGX_D double P_from_rho_e(int m, double rho, double e) const {
    SGX_EOS_DISPATCH(m, P_from_rho_e(rho, e))
  }
  SGX_D double T_from_rho_e(int m, double rho, double e) const {
    SGX_EOS_DISPATCH(m, T_from_rho_e(rho, e))
  }
  SGX_D double P_from_rho_T(int m, double rho, double T) const {
    SGX_EOS_DISPATCH(m, P_from_rho_T(rho, T))
  }
  SGX_D double cv_from_rho_T(int m, double rho, double T) const {
    SGX_EOS_DISPATCH(m, cv_from_rho_T(rho, T))
  }
  SGX_D double cv_from_rho_e(int m, double rho, double e) const {
    SGX_EOS_DISPATCH(m, cv_from_rho_e(rho, e))
  }
  SGX_D double RhoPmin(int m, double T) const { SGX_EOS_DISPATCH(m, RhoPmin(T)) }
  SGX_D double MaximumDensity(int m) const { SGX_EOS_DISPATCH(m, MaximumDensity()) }
  SGX_D double MinimumTemperature(int m) const { SGX_EOS_DISPATCH(m, MinimumTemperature()) }
  SGX_D double MinimumPressure(int m) const { SGX_EOS_DISPATCH(m, MinimumPressure()) }
  SGX_D double MaximumPressureAtTemperature(int m, double T) const {
    SGX_EOS_DISPATCH(m, MaximumPressureAtTemperature(T))
  }
  __device__ inline void rho_e_from_P_T(int m, double P, double T, double &rho,
                                        double &sie) const {
    switch (set.e[m].kind) {
    case 0:
      set.e[m].ideal.rho_e_from_P_T(P, T, nullptr, rho, sie);
      return;
    case 1:
      set.e[m].gr.rho_e_from_P_T(P, T, nullptr, rho, sie);
      return;
    case 2:
      set.e[m].dr.rho_e_from_P_T(P, T, nullptr, rho, sie);
      return;
    default:
      set.e[m].dp.rho_e_from_P_T(P, T, nullptr, rho, sie);
      return;
    }
  }
  // staged evaluation (sgx_analytic.cuh RhoPart)
  __device__ inline void rho_part(int m, double rho, RhoPart &p) const {
    switch (set.e[m].kind) {
    case 0:
      set.e[m].ideal.rho_part(rho, p);
      return;
    case 1:
      set.e[m].gr.rho_part(rho, p);
      return;
    case 2:
      set.e[m].dr.rho_part(rho, p);
      return;
    default:
      set.e[m].dp.rho_part(rho, p);
      return;
    }
  }
  SGX_D double e_from_part(int m, const RhoPart &p, double rho, double T) const {
    SGX_EOS_DISPATCH(m, e_from_part(p, rho, T))
  }
  SGX_D double P_from_part(int m, const RhoPart &p, double rho, double sie) const {
    SGX_EOS_DISPATCH(m, P_from_part(p, rho, sie))
  }
  // GetPressureFromPreferred(..., do_e_lookup = false) (:583-597) for the state machine: every
  // model this accessor holds has sie among its preferred inputs, so T is not used
  SGX_D double P_pref_from_part(int m, const RhoPart &p, double rho, double sie, double) const {
    return P_from_part(m, p, rho, sie);
  }
  static constexpr int kCacheSlots = 0; // no lambda cache
  SGX_D double cv_from_part(int m, const RhoPart &p, double rho, double sie) const {
    SGX_EOS_DISPATCH(m, cv_from_part(p, rho, sie))
  }
  // same, restricted to the models whose PreferredInput() includes pressure (only those
  // reach it through GetSieCvFromTAndPreferred, :603-612): keeps the solids' nested root
  // finders out of the RhoT kernels
  __device__ inline void rho_e_from_P_T_pressure_preferred(int m, double P, double T,
                                                           double &rho, double &sie) const {
    if (set.e[m].kind == 0) set.e[m].ideal.rho_e_from_P_T(P, T, nullptr, rho, sie);
  }
};

// The benchmark mixture of test/pte_test_3mat_analytic.hpp:38-50 with the material types
// resolved at compile time: with `m` a literal after unrolling, each call inlines the one
// model it needs and the Variant switch disappears.
struct Mix3GrDrDp {
  const EosSet &set;
  __device__ Mix3GrDrDp(const EosSet &s) : set(s) {}
  SGX_D unsigned long preferred(int) const { return TQ_DENSITY | TQ_SIE; }
#define SGX_MIX3(CALL)                                                                   \
  if (m == 0) return set.e[0].gr.CALL;                                                   \
  if (m == 1) return set.e[1].dr.CALL;                                                   \
  return set.e[2].dp.CALL;
  SGX_D double e_from_rho_T(int m, double rho, double T) const {
    SGX_MIX3(e_from_rho_T(rho, T))
  }
  SGX_D double P_from_rho_e(int m, double rho, double e) const {
    SGX_MIX3(P_from_rho_e(rho, e))
  }
  SGX_D double T_from_rho_e(int m, double rho, double e) const {
    SGX_MIX3(T_from_rho_e(rho, e))
  }
  SGX_D double P_from_rho_T(int m, double rho, double T) const {
    SGX_MIX3(P_from_rho_T(rho, T))
  }
  SGX_D double cv_from_rho_T(int m, double rho, double T) const {
    SGX_MIX3(cv_from_rho_T(rho, T))
  }
  SGX_D double cv_from_rho_e(int m, double rho, double e) const {
    SGX_MIX3(cv_from_rho_e(rho, e))
  }
  SGX_D double RhoPmin(int, double) const { return 0.0; }
  SGX_D double MaximumDensity(int m) const { return m == 0 ? set.e[0].gr.rho_max : 1e100; }
  SGX_D double MinimumTemperature(int) const { return 0; }
  SGX_D double MinimumPressure(int m) const {
    return m == 0 ? set.e[0].gr.MinimumPressure() : 0.0;
  }
  SGX_D double MaximumPressureAtTemperature(int m, double T) const {
    return m == 0 ? set.e[0].gr.MaximumPressureAtTemperature(T) : 1e100;
  }
  __device__ inline void rho_e_from_P_T(int m, double P, double T, double &rho,
                                        double &sie) const {
    if (m == 0) {
      set.e[0].gr.rho_e_from_P_T(P, T, nullptr, rho, sie);
    } else if (m == 1) {
      set.e[1].dr.rho_e_from_P_T(P, T, nullptr, rho, sie);
    } else {
      set.e[2].dp.rho_e_from_P_T(P, T, nullptr, rho, sie);
    }
  }
  SGX_D void rho_e_from_P_T_pressure_preferred(int, double, double, double &, double &) const {}
  __device__ inline void rho_part(int m, double rho, RhoPart &p) const {
    if (m == 0) {
      set.e[0].gr.rho_part(rho, p);
    } else if (m == 1) {
      set.e[1].dr.rho_part(rho, p);
    } else {
      set.e[2].dp.rho_part(rho, p);
    }
  }
  SGX_D double e_from_part(int m, const RhoPart &p, double rho, double T) const {
    SGX_MIX3(e_from_part(p, rho, T))
  }
  SGX_D double P_from_part(int m, const RhoPart &p, double rho, double sie) const {
    SGX_MIX3(P_from_part(p, rho, sie))
  }
  SGX_D double P_pref_from_part(int m, const RhoPart &p, double rho, double sie, double) const {
    return P_from_part(m, p, rho, sie);
  }
  static constexpr int kCacheSlots = 0;
  SGX_D double cv_from_part(int m, const RhoPart &p, double rho, double sie) const {
    SGX_MIX3(cv_from_part(p, rho, sie))
  }
#undef SGX_MIX3
};

namespace mix {

// ApproxTemperatureFromRhoMatU (:790-847); lambdas are NullIndexer here.
template <int NM, class Eos>
__device__ inline double ApproxTemperatureFromRhoMatU(const Eos &eos, const double u_tot,
                                                      const double *rho, const double *vfrac,
                                                      const double Tguess = 0.0) {
  constexpr double minimum_temperature = 1.e-9;
  constexpr double maximum_temperature = 1.e9;
  auto ufunc = [&](const double T) {
    double usum = 0.0;
#pragma unroll
    for (int m = 0; m < NM; m++) {
      usum += rho[m] * vfrac[m] * eos.e_from_rho_T(m, rho[m], T);
    }
    return usum;
  };
  double ulo = ufunc(minimum_temperature);
  if (u_tot < ulo) return minimum_temperature;
  double uhi = ufunc(maximum_temperature);
  if (u_tot > uhi) return maximum_temperature;
  double lTlo = fastmath::lg(minimum_temperature);
  double lThi = fastmath::lg(maximum_temperature);
  if (Tguess > minimum_temperature && Tguess < maximum_temperature) {
    const double ug = ufunc(Tguess);
    if (ug < u_tot) {
      lTlo = fastmath::lg(Tguess);
      ulo = ug;
    } else {
      lThi = fastmath::lg(Tguess);
      uhi = ug;
    }
  }
  int iter = 0;
  constexpr int max_iter = 10;
  while (lThi - lTlo > 0.01 && iter < max_iter) {
    const double lT = 0.5 * (lTlo + lThi);
    const double uT = ufunc(fastmath::pow2(lT));
    if (uT < u_tot) {
      lTlo = lT;
      ulo = uT;
    } else {
      lThi = lT;
      uhi = uT;
    }
    iter++;
  }
  const double alpha = ratio((u_tot - ulo), (uhi - ulo));
  return fastmath::pow2((1.0 - alpha) * lTlo + alpha * lThi);
}

// ---------------------------------------------------------------------------------------
// PTESolverBase (:329-786) -- state shared by the RhoT and PT systems
// ---------------------------------------------------------------------------------------
template <int NM, int NEQ, class Eos>
struct Base {
  static constexpr int nmat = NM;
  static constexpr int neq = NEQ;
  const Eos &eos;
  const MixParamsDev &params_;
  int niter;
  double vfrac_total, sie_total;
  double rho[NM], vfrac[NM], sie[NM], temp[NM], press[NM];
  double jacobian[NEQ * NEQ], dx[NEQ], residual[NEQ], u[NM], rhobar[NM];
  double rho_total, uscale, utotal_scale, Tnorm;

  __device__ Base(const Eos &e, const MixParamsDev &p) : eos(e), params_(p), niter(0) {}

  SGX_D void InitRhoBarandRho() {
#pragma unroll
    for (int m = 0; m < NM; ++m)
      rhobar[m] = rho[m] * vfrac[m];
    rho_total = sum_neumaier(rhobar, NM);
  }
  SGX_D void NormalizeVfrac() {
    const double vfrac_sum = sum_neumaier(vfrac, NM);
#pragma unroll
    for (int m = 0; m < NM; ++m)
      vfrac[m] *= ratio(vfrac_total, vfrac_sum);
  }
  SGX_D double get_vfrac_max(const double T, const int m) const {
    constexpr double safety = 0.9;
    const double rho_min = eos.RhoPmin(m, T);
    return stdmin(safety * ratio(rhobar[m], rho_min), 1.0);
  }
  // SetVfracFromT (:423-498); returns true on FAIL like the reference
  __device__ inline bool SetVfracFromT(const double T) {
    constexpr double min_dv = 1e-12;
    double vfrac_max_sum = 0;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      const double vfrac_max = get_vfrac_max(T, m);
      if (vfrac[m] <= 0) vfrac[m] = vfrac_max;
      vfrac_max_sum += vfrac_max;
    }
    if (vfrac_max_sum < vfrac_total) return true;
    NormalizeVfrac();
    for (int it = 0; it < NM; ++it) {
      double remaining_vfrac = 0;
      int n_uncapped_materials = NM;
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        const double vfrac_max = get_vfrac_max(T, m);
        const double vnew = stdmin(vfrac_max, vfrac[m]);
        const double dv = vfrac[m] - vnew;
        vfrac[m] -= dv;
        remaining_vfrac += dv;
        n_uncapped_materials -= (dv > min_dv);
      }
      if (fabs(remaining_vfrac) < min_dv) break;
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        const bool uncapped = ((get_vfrac_max(T, m) - vfrac[m]) > min_dv);
        if (uncapped) {
          // robust::ratio(Real, std::size_t): sgn(unsigned) == 1 and SMALL<size_t>() == 0,
          // so this is a plain division (robust_utils.hpp:27-62)
          vfrac[m] += remaining_vfrac / (double)(unsigned long long)n_uncapped_materials;
        }
      }
    }
    NormalizeVfrac();
    return false;
  }
  // GetPressureFromPreferred (:583-597)
  SGX_D double GetPressureFromPreferred(const int m, const double rho_, const double T,
                                        double sie_, const bool do_e_lookup) const {
    if (eos.preferred(m) & TQ_SIE) {
      if (do_e_lookup) sie_ = eos.e_from_rho_T(m, rho_, T);
      return eos.P_from_rho_e(m, rho_, sie_);
    }
    return eos.P_from_rho_T(m, rho_, T);
  }
  // GetSieCvFromTAndPreferred (:601-620)
  SGX_D void GetSieCvFromTAndPreferred(const int m, double rho_, const double P,
                                       const double T, double &sie_, double &cv) const {
    if (eos.preferred(m) & TQ_PRESSURE) {
      const double Pmin = eos.MinimumPressure(m);
      const double Pmax = eos.MaximumPressureAtTemperature(m, T);
      double Pguess;
      if (bad_value(P)) {
        Pguess = 0.5 * (Pmin + Pmax);
      } else {
        Pguess = stdmax(Pmin, stdmin(Pmax, P));
      }
      // only models whose PreferredInput() includes pressure get here
      eos.rho_e_from_P_T_pressure_preferred(m, Pguess, T, rho_, sie_);
      cv = eos.cv_from_rho_e(m, rho_, sie_);
    } else {
      sie_ = eos.e_from_rho_T(m, rho_, T);
      cv = eos.cv_from_rho_T(m, rho_, T);
    }
  }
  // GetTguess (:500-581)
  __device__ inline double GetTguess() {
    double Tguess;
    if (Tnorm > 0.0) {
      Tguess = Tnorm;
    } else {
      Tguess = params_.default_tguess;
#pragma unroll
      for (int m = 0; m < NM; ++m)
        Tguess = stdmax(Tguess, temp[m]);
    }
    auto newton_step = [this](const double T) {
      double dudt = 0;
      double usum = 0;
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        const double rho_max = eos.MaximumDensity(m);
        double sie_, cv;
        GetSieCvFromTAndPreferred(m, stdmin(rho[m], rho_max), press[m], T, sie_, cv);
        usum += rhobar[m] * sie_;
        dudt += rhobar[m] * cv;
      }
      return T - ratio((usum - utotal_scale * uscale), dudt);
    };
    const double Tfactor = 10.0;
    bool bad_vfrac_guess = false;
    for (int i = 0; i < 3; i++) {
      if (params_.iterate_t_guess) {
        Tguess = stdmin(params_.temperature_limit, stdmax(Tguess, newton_step(Tguess)));
      }
#pragma unroll
      for (int m = 0; m < NM; ++m)
        Tguess = stdmax(eos.MinimumTemperature(m), Tguess);
      bad_vfrac_guess = SetVfracFromT(Tguess);
      if (bad_vfrac_guess) {
        Tguess *= Tfactor;
        continue;
      }
#pragma unroll
      for (int m = 0; m < NM; ++m)
        rho[m] = ratio(rhobar[m], vfrac[m]);
      bad_vfrac_guess = false;
      for (int m = 0; m < NM; ++m) {
        const double rho_min = eos.RhoPmin(m, Tguess);
        if (rho[m] < rho_min) {
          bad_vfrac_guess = true;
          Tguess *= Tfactor;
          break;
        }
      }
      if (!bad_vfrac_guess) break;
    }
    return Tguess;
  }
  // InitBase (:627-662)
  __device__ inline void InitBase() {
    InitRhoBarandRho();
    const double utotal = rho_total * sie_total;
    uscale = stdmax(fabs(utotal), 1.0e-14);
    utotal_scale = ratio(utotal, uscale);
    const double Tguess = GetTguess();
    Tnorm = Tguess;
#pragma unroll
    for (int m = 0; m < NM; m++) {
      temp[m] = 1.0;
      sie[m] = eos.e_from_rho_T(m, rho[m], Tguess);
      const unsigned long prefinput = eos.preferred(m);
      if ((prefinput & TQ_PRESSURE) && !(bad_value(press[m]))) {
        press[m] = ratio(press[m], uscale);
      } else {
        press[m] = ratio(GetPressureFromPreferred(m, rho[m], Tguess, sie[m], false), uscale);
      }
    }
#pragma unroll
    for (int m = 0; m < NM; ++m)
      u[m] = sie[m] * ratio(rhobar[m], uscale);
  }
  SGX_D double ResidualNorm() const {
    return 0.5 * sum_neumaier(residual, NEQ, 0, -1, [](const double x) { return x * x; });
  }
  SGX_D bool Solve() {
#pragma unroll
    for (int m = 0; m < NEQ; m++)
      dx[m] = residual[m];
    return solve_Ax_b_static<NEQ>(jacobian, dx);
  }
};

// ---------------------------------------------------------------------------------------
// PTESolverRhoT (:883-1193): unknowns T and the NM-1 non-excluded volume fractions
// ---------------------------------------------------------------------------------------
template <int NM, class Eos>
struct RhoT : public Base<NM, NM, Eos> {
  using B = Base<NM, NM, Eos>;
  using B::dx;
  using B::eos;
  using B::jacobian;
  using B::params_;
  using B::press;
  using B::residual;
  using B::rho;
  using B::rhobar;
  using B::sie;
  using B::temp;
  using B::Tnorm;
  using B::u;
  using B::uscale;
  using B::utotal_scale;
  using B::vfrac;
  double dpdv[NM], dedv[NM], dpdT[NM], vtemp[NM];
  double Tequil, Ttemp;
  int ms;

  __device__ RhoT(const Eos &e, const MixParamsDev &p) : B(e, p) {}

  __device__ inline double Init() {
    B::InitBase();
    if (params_.pte_reduced_system_exclude_idx >= 0) {
      ms = (int)params_.pte_reduced_system_exclude_idx;
    } else {
      ms = 0;
      double xmax = 0;
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        const double x = ratio(rhobar[m], B::rho_total);
        if (x > xmax) {
          xmax = x;
          ms = m;
        }
      }
    }
    Residual();
    Tequil = temp[0];
    return B::ResidualNorm();
  }
  SGX_D void Fixup() {
    if (params_.pte_reduced_system_exclude_idx < 0) {
      double alphamax = 0;
      ms = 0;
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        if (vfrac[m] > alphamax) {
          alphamax = vfrac[m];
          ms = m;
        }
      }
      Residual();
    }
  }
  SGX_D void Finalize() {
#pragma unroll
    for (int m = 0; m < NM; m++) {
      temp[m] *= Tnorm;
      u[m] *= uscale;
      press[m] *= uscale;
    }
  }
  SGX_D void Residual() {
    const double esum = sum_neumaier(u, NM);
    residual[0] = utotal_scale - esum;
    // residual[k], k = 1..NM-1, belongs to the k-th material other than ms (static indices:
    // a runtime ires++ would force the solver object into local memory)
#pragma unroll
    for (int k = 1; k < NM; ++k)
      residual[k] = press_at(ms) - press_at(other(k));
  }
  // the material of unknown / equation k >= 1 of the reduced system
  SGX_D int other(const int k) const { return (k - 1 < ms) ? k - 1 : k; }
  SGX_D double press_at(const int i) const {
    double v = press[0];
#pragma unroll
    for (int m = 1; m < NM; ++m)
      if (m == i) v = press[m];
    return v;
  }
  SGX_D void CheckPTE(bool &converged, bool &acceptable) const {
    double mean_p = 0;
#pragma unroll
    for (int m = 0; m < NM; ++m)
      mean_p += vfrac[m] * press[m];
    const double error_p =
        sqrt(sum_neumaier(residual, NM - 1, 1, -1, [](const double x) { return x * x; }));
    const double error_u = fabs(residual[0]);
    const bool converged_p = (error_p < params_.pte_rel_tolerance_p * fabs(mean_p) ||
                              error_p < params_.pte_abs_tolerance_p);
    const bool converged_u = (error_u < params_.pte_rel_tolerance_e);
    const bool acceptable_p =
        (error_p < params_.pte_rel_tolerance_p_sufficient * fabs(mean_p) ||
         error_p < params_.pte_abs_tolerance_p_sufficient);
    const bool acceptable_u = (error_u < params_.pte_rel_tolerance_e_sufficient);
    converged = converged_p && converged_u;
    acceptable = acceptable_p && acceptable_u;
  }
  __device__ inline void Jacobian() {
    double dedT_sum = 0.0;
#pragma unroll
    for (int m = 0; m < NM; m++) {
      const double dv = (vfrac[m] < 0.5 ? 1.0 : -1.0) * vfrac[m] * params_.derivative_eps;
      const double vf_pert = vfrac[m] + dv;
      const double rho_pert = ratio(rhobar[m], vf_pert);
      double e_pert = eos.e_from_rho_T(m, rho_pert, Tnorm * Tequil);
      double p_pert = ratio(
          B::GetPressureFromPreferred(m, rho_pert, Tnorm * Tequil, e_pert, false), uscale);
      dpdv[m] = stdmin(ratio((p_pert - press[m]), dv), -fabs(params_.pte_max_dpdv));
      dedv[m] = ratio(rhobar[m] * ratio(e_pert, uscale) - u[m], dv);
      const double dT = Tequil * params_.derivative_eps;
      e_pert = eos.e_from_rho_T(m, rho[m], Tnorm * (Tequil + dT));
      p_pert = ratio(
          B::GetPressureFromPreferred(m, rho[m], Tnorm * (Tequil + dT), e_pert, false),
          uscale);
      dpdT[m] = ratio((p_pert - press[m]), dT);
      dedT_sum += ratio(rhobar[m] * ratio(e_pert, uscale) - u[m], dT);
    }
    AssembleJacobian(dedT_sum);
  }
  // Fill the reduced (T, alpha_{m != ms}) system from dpdv/dedv/dpdT (:1083-1104)
  SGX_D void AssembleJacobian(const double dedT_sum) {
#pragma unroll
    for (int i = 0; i < NM * NM; ++i)
      jacobian[i] = 0.0;
    const double dedv_s = pick(dedv, ms), dpdT_s = pick(dpdT, ms), dpdv_s = pick(dpdv, ms);
    jacobian[0] = dedT_sum;
#pragma unroll
    for (int k = 1; k < NM; ++k)
      jacobian[k] = pick(dedv, other(k)) - dedv_s;
#pragma unroll
    for (int row = 1; row < NM; ++row) {
      const int m = other(row);
      jacobian[row * NM + 0] = pick(dpdT, m) - dpdT_s;
#pragma unroll
      for (int c = 1; c < NM; ++c)
        jacobian[row * NM + c] = (c == row) ? pick(dpdv, m) + dpdv_s : dpdv_s;
    }
  }
  SGX_D static double pick(const double *a, const int i) {
    double v = a[0];
#pragma unroll
    for (int m = 1; m < NM; ++m)
      if (m == i) v = a[m];
    return v;
  }
  __device__ inline double ScaleDx() {
    double scale = 1.0;
    const double dalphaskip = -sum_neumaier(dx, NM - 1, 1);
    int idx = 1;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      const double mydx = (m == ms) ? dalphaskip : pick(dx, idx++);
      if (scale * mydx < -params_.vfrac_safety_fac * vfrac[m]) {
        scale = stdmin(scale, fabs(params_.vfrac_safety_fac * ratio(vfrac[m], mydx)));
      }
    }
    const double Tnew = Tequil + scale * dx[0];
    idx = 1;
#pragma unroll
    for (int m = 0; m < NM; m++) {
      const double rho_min =
          stdmax(eos.RhoPmin(m, Tnorm * Tequil), eos.RhoPmin(m, Tnorm * Tnew));
      const double alpha_max = stdmin(1.0, ratio(rhobar[m], rho_min));
      if (alpha_max < vfrac[m]) continue; // NB: skips idx++ exactly like the reference
      const double mydx = (m == ms) ? dalphaskip : pick(dx, idx++);
      if (scale * mydx > 0.5 * (alpha_max - vfrac[m])) {
        scale = stdmin(ratio(0.5 * (alpha_max - vfrac[m]), mydx), scale);
      }
    }
    if (scale * dx[0] < -0.95 * Tequil) {
      scale = stdmin(scale, ratio(-0.95 * Tequil, dx[0]));
    }
#pragma unroll
    for (int i = 0; i < NM; ++i)
      dx[i] *= scale;
    return scale;
  }
  __device__ inline double TestUpdate(const double scale, const bool cache_state = false) {
    if (cache_state) {
      Ttemp = Tequil;
#pragma unroll
      for (int m = 0; m < NM; ++m)
        vtemp[m] = vfrac[m];
    }
    Tequil = Ttemp + scale * dx[0];
    int idx = 1;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      if (ms == m) continue;
      vfrac[m] = vtemp[m] + scale * pick(dx, idx++);
    }
    const double vsum_others = sum_neumaier(vfrac, NM, 0, ms);
#pragma unroll
    for (int m = 0; m < NM; ++m)
      if (m == ms) vfrac[m] = 1 - vsum_others;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      rho[m] = ratio(rhobar[m], vfrac[m]);
      u[m] = rhobar[m] * eos.e_from_rho_T(m, rho[m], Tnorm * Tequil);
      sie[m] = ratio(u[m], rhobar[m]);
      u[m] = ratio(u[m], uscale);
      temp[m] = Tequil;
      press[m] =
          ratio(B::GetPressureFromPreferred(m, rho[m], Tnorm * Tequil, sie[m], false), uscale);
    }
    Residual();
    return B::ResidualNorm();
  }
};

// ---------------------------------------------------------------------------------------
// PTESolverPT (:1207-1446): unknowns (T, P); every state update inverts each material's
// P(rho,T) with the EOS's own DensityEnergyFromPressureTemperature.
// ---------------------------------------------------------------------------------------
template <int NM, class Eos>
struct PT : public Base<NM, 2, Eos> {
  using B = Base<NM, 2, Eos>;
  using B::dx;
  using B::eos;
  using B::jacobian;
  using B::params_;
  using B::press;
  using B::residual;
  using B::rho;
  using B::rhobar;
  using B::sie;
  using B::temp;
  using B::Tnorm;
  using B::u;
  using B::uscale;
  using B::utotal_scale;
  using B::vfrac;
  double Pequil, Tequil, Ptemp, Ttemp;

  __device__ PT(const Eos &e, const MixParamsDev &p) : B(e, p) {}

  __device__ inline double Init() {
    B::InitBase();
    Residual();
    Pequil = 0;
#pragma unroll
    for (int m = 0; m < NM; ++m)
      Pequil += fabs(press[m]) * vfrac[m];
    const double vsum = sum_neumaier(vfrac, NM);
    Pequil /= vsum;
    Tequil = 1;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      eos.rho_e_from_P_T(m, Pequil * uscale, Tequil * Tnorm, rho[m], sie[m]);
      vfrac[m] = ratio(rhobar[m], rho[m]);
      u[m] = ratio(sie[m] * rhobar[m], uscale);
      temp[m] = Tequil;
    }
    Residual();
    return B::ResidualNorm();
  }
  SGX_D void Fixup() {}
  SGX_D void Finalize() {
#pragma unroll
    for (int m = 0; m < NM; m++) {
      press[m] = Pequil * uscale;
      temp[m] = Tequil * Tnorm;
      u[m] *= uscale;
    }
  }
  SGX_D void Residual() {
    const double vsum = sum_neumaier(vfrac, NM);
    const double esum = sum_neumaier(u, NM);
    residual[0] = B::vfrac_total - vsum;
    residual[1] = utotal_scale - esum;
  }
  SGX_D void CheckPTE(bool &converged, bool &acceptable) const {
    const double error_v = fabs(residual[0]);
    const double error_u = fabs(residual[1]);
    const bool converged_u = error_u < params_.pte_rel_tolerance_e;
    const bool converged_v = error_v < params_.pte_rel_tolerance_v;
    const bool acceptable_u = error_u < params_.pte_rel_tolerance_e_sufficient;
    const bool acceptable_v = error_v < params_.pte_rel_tolerance_v_sufficient;
    converged = converged_v && converged_u;
    acceptable = acceptable_u && acceptable_v;
  }
  __device__ inline void Jacobian() {
    double dudT_P_sum = 0.0, dudP_T_sum = 0.0, rbor2_dr_dT_P_sum = 0.0,
           rbor2_dr_dP_T_sum = 0.0;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      double r_pert = rho[m];
      double e_pert = sie[m];
      const double dp = -Pequil * params_.derivative_eps;
      eos.rho_e_from_P_T(m, uscale * (Pequil + dp), Tnorm * Tequil, r_pert, e_pert);
      const double drdp = ratio(r_pert - rho[m], dp);
      const double dudp = ratio(ratio(rhobar[m] * e_pert, uscale) - u[m], dp);
      rbor2_dr_dP_T_sum += ratio(rhobar[m], rho[m] * rho[m]) * drdp;
      dudP_T_sum += dudp;
      const double dT = Tequil * params_.derivative_eps;
      eos.rho_e_from_P_T(m, uscale * Pequil, Tnorm * (Tequil + dT), r_pert, e_pert);
      const double drdT = ratio(r_pert - rho[m], dT);
      const double dudT = ratio(ratio(rhobar[m] * e_pert, uscale) - u[m], dT);
      rbor2_dr_dT_P_sum += ratio(rhobar[m], rho[m] * rho[m]) * drdT;
      dudT_P_sum += dudT;
    }
    jacobian[0] = -rbor2_dr_dT_P_sum;
    jacobian[1] = -rbor2_dr_dP_T_sum;
    jacobian[2] = dudT_P_sum;
    jacobian[3] = dudP_T_sum;
  }
  __device__ inline double ScaleDx() {
    double scale = 1.0;
    if (scale * dx[0] < -0.95 * Tequil) {
      scale = ratio(-0.95 * Tequil, dx[0]);
    }
    auto bounded = [this](const double Pbound, const double delta) {
      return ratio(ratio(Pbound, uscale) - Pequil, delta);
    };
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      const double Pmin = eos.MinimumPressure(m);
      scale = stdmin(fabs(scale), fabs(0.95 * bounded(Pmin, dx[1])));
    }
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      const double Ttest = (Tequil + scale * dx[0]) * Tnorm;
      const double Pmax = eos.MaximumPressureAtTemperature(m, Ttest);
      scale = stdmin(fabs(scale), fabs(0.95 * bounded(Pmax, dx[0])));
    }
    dx[0] *= scale;
    dx[1] *= scale;
    return scale;
  }
  __device__ inline double TestUpdate(const double scale, const bool cache_state = false) {
    if (cache_state) {
      Ttemp = Tequil;
      Ptemp = Pequil;
    }
    Tequil = Ttemp + scale * dx[0];
    Pequil = Ptemp + scale * dx[1];
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      eos.rho_e_from_P_T(m, Pequil * uscale, Tequil * Tnorm, rho[m], sie[m]);
      vfrac[m] = ratio(rhobar[m], rho[m]);
      u[m] = ratio(sie[m] * rhobar[m], uscale);
      temp[m] = Tequil;
      press[m] = Pequil;
    }
    Residual();
    return B::ResidualNorm();
  }
};

// ---------------------------------------------------------------------------------------
// PTESolverFixedT (:1462-1672): temperature given, unknowns are the NM volume fractions;
// equations: volume sum, and pressure equality of consecutive materials.
// ---------------------------------------------------------------------------------------
template <int NM, class Eos>
struct FixedT : public Base<NM, NM, Eos> {
  using B = Base<NM, NM, Eos>;
  using B::dx;
  using B::eos;
  using B::jacobian;
  using B::params_;
  using B::press;
  using B::residual;
  using B::rho;
  using B::rhobar;
  using B::sie;
  using B::temp;
  using B::Tnorm;
  using B::u;
  using B::uscale;
  using B::vfrac;
  double dpdv[NM], vtemp[NM];
  double Tequil;

  __device__ FixedT(const Eos &e, const MixParamsDev &p) : B(e, p) {}

  // Init (:1523-1551).  The kernel puts T_true into Tnorm; the constructor (:1498-1510) keeps
  // it as Tequil and resets Tnorm to 1.
  __device__ inline double Init() {
    Tequil = Tnorm;
    Tnorm = 1.0;
    B::InitRhoBarandRho();
    B::SetVfracFromT(Tequil);
    uscale = 0.0;
#pragma unroll
    for (int m = 0; m < NM; m++) {
      rho[m] = ratio(rhobar[m], vfrac[m]);
      sie[m] = eos.e_from_rho_T(m, rho[m], Tequil);
      uscale += fabs(sie[m]) * rho[m];
      press[m] = eos.P_from_rho_T(m, rho[m], Tequil);
    }
    uscale = stdmax(uscale, 1.0e-14);
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      press[m] = ratio(press[m], uscale);
      u[m] = sie[m] * ratio(rhobar[m], uscale);
    }
    Residual();
    return B::ResidualNorm();
  }
  SGX_D void Fixup() {}
  SGX_D void Finalize() { // base Finalize (:351-357)
#pragma unroll
    for (int m = 0; m < NM; m++) {
      temp[m] *= Tnorm;
      u[m] *= uscale;
      press[m] *= uscale;
    }
  }
  SGX_D void Residual() {
    const double vsum = sum_neumaier(vfrac, NM);
    residual[0] = B::vfrac_total - vsum;
#pragma unroll
    for (int m = 0; m < NM - 1; ++m)
      residual[1 + m] = press[m] - press[m + 1];
  }
  SGX_D void CheckPTE(bool &converged, bool &acceptable) const {
    double mean_p = vfrac[0] * press[0];
    double error_p = 0;
#pragma unroll
    for (int m = 1; m < NM; ++m) {
      mean_p += vfrac[m] * press[m];
      error_p += residual[m] * residual[m];
    }
    error_p = sqrt(error_p);
    const double error_v = fabs(residual[0]);
    const bool converged_p = (error_p < params_.pte_rel_tolerance_p * fabs(mean_p) ||
                              error_p < params_.pte_abs_tolerance_p);
    const bool converged_v = error_v < params_.pte_rel_tolerance_v;
    const bool acceptable_p =
        (error_p < params_.pte_rel_tolerance_p_sufficient * fabs(mean_p) ||
         error_p < params_.pte_abs_tolerance_p_sufficient);
    const bool acceptable_v = error_v < params_.pte_rel_tolerance_v_sufficient;
    converged = converged_p && converged_v;
    acceptable = acceptable_p && acceptable_v;
  }
  __device__ inline void Jacobian() {
#pragma unroll
    for (int m = 0; m < NM; m++) {
      const double dv = (vfrac[m] < 0.5 ? 1.0 : -1.0) * vfrac[m] * params_.derivative_eps;
      const double vf_pert = vfrac[m] + dv;
      const double rho_pert = ratio(rhobar[m], vf_pert);
      const double p_pert = ratio(eos.P_from_rho_T(m, rho_pert, Tequil), uscale);
      dpdv[m] = stdmin(ratio((p_pert - press[m]), dv), -fabs(params_.pte_max_dpdv));
    }
#pragma unroll
    for (int i = 0; i < NM * NM; ++i)
      jacobian[i] = 0.0;
#pragma unroll
    for (int m = 0; m < NM; ++m)
      jacobian[m] = 1.0;
#pragma unroll
    for (int m = 0; m < NM - 1; m++) {
      jacobian[(m + 1) * NM + m] = -dpdv[m];
      jacobian[(m + 1) * NM + m + 1] = dpdv[m + 1];
    }
  }
  __device__ inline double ScaleDx() {
    double scale = 1.0;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      if (scale * dx[m] < -params_.vfrac_safety_fac * vfrac[m]) {
        scale = ratio(-params_.vfrac_safety_fac * vfrac[m], dx[m]);
      }
    }
#pragma unroll
    for (int m = 0; m < NM; m++) {
      const double rho_min = eos.RhoPmin(m, Tequil);
      const double alpha_max = ratio(rhobar[m], rho_min);
      if (alpha_max < vfrac[m]) continue;
      if (scale * dx[m] > 0.5 * (alpha_max - vfrac[m])) {
        scale = ratio(0.5 * (alpha_max - vfrac[m]), dx[m]);
      }
    }
#pragma unroll
    for (int i = 0; i < NM; ++i)
      dx[i] *= scale;
    return scale;
  }
  __device__ inline double TestUpdate(const double scale, const bool cache_state = false) {
    if (cache_state) {
#pragma unroll
      for (int m = 0; m < NM; ++m)
        vtemp[m] = vfrac[m];
    }
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      vfrac[m] = vtemp[m] + scale * dx[m];
      rho[m] = ratio(rhobar[m], vfrac[m]);
      u[m] = rhobar[m] * eos.e_from_rho_T(m, rho[m], Tequil);
      sie[m] = ratio(u[m], rhobar[m]);
      u[m] = ratio(u[m], uscale);
      press[m] = ratio(eos.P_from_rho_T(m, rho[m], Tequil), uscale);
    }
    Residual();
    return B::ResidualNorm();
  }
};

// ---------------------------------------------------------------------------------------
// PTESolverFixedP (:1674-1932): pressure given; unknowns are the NM volume fractions and the
// temperature; equations: each material's pressure equals P, and the volume sum.
// Init calls GetTguess, whose optional Newton step (MixParams::iterate_t_guess) reads uscale and
// utotal_scale BEFORE this solver's Init sets them (:1747-1755 vs :536-538; the base constructor
// :371-387 leaves both uninitialised), so the reference's answer is only defined with
// iterate_t_guess == false: the C ABI rejects the other setting.
// ---------------------------------------------------------------------------------------
template <int NM, class Eos>
struct FixedP : public Base<NM, NM + 1, Eos> {
  static constexpr int NEQ = NM + 1;
  using B = Base<NM, NM + 1, Eos>;
  using B::dx;
  using B::eos;
  using B::jacobian;
  using B::params_;
  using B::press;
  using B::residual;
  using B::rho;
  using B::rhobar;
  using B::sie;
  using B::temp;
  using B::Tnorm;
  using B::u;
  using B::uscale;
  using B::vfrac;
  double dpdv[NM], dpdT[NM], vtemp[NM];
  double Tequil, Ttemp, Pequil;

  __device__ FixedP(const Eos &e, const MixParamsDev &p) : B(e, p) {}

  // Init (:1741-1774).  The kernel puts P into Pequil and 0 into Tnorm (ctor :1731-1739).
  __device__ inline double Init() {
    B::InitRhoBarandRho();
    const double Tguess = B::GetTguess();
    Tnorm = Tguess;
    uscale = stdmax(fabs(Pequil), 1.0e-14);
#pragma unroll
    for (int m = 0; m < NM; m++) {
      temp[m] = 1.0;
      sie[m] = eos.e_from_rho_T(m, rho[m], Tguess);
    }
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      u[m] = ratio(sie[m] * rhobar[m], uscale);
      press[m] = ratio(eos.P_from_rho_T(m, rho[m], Tguess), uscale);
    }
    Residual();
    Tequil = temp[0];
    Ttemp = Tequil;
    return B::ResidualNorm();
  }
  SGX_D void Fixup() {}
  SGX_D void Finalize() { // base Finalize (:351-357)
#pragma unroll
    for (int m = 0; m < NM; m++) {
      temp[m] *= Tnorm;
      u[m] *= uscale;
      press[m] *= uscale;
    }
  }
  SGX_D void Residual() {
    const double vsum = sum_neumaier(vfrac, NM);
#pragma unroll
    for (int m = 0; m < NM; ++m)
      residual[m] = ratio(Pequil, uscale) - press[m];
    residual[NM] = B::vfrac_total - vsum;
  }
  SGX_D void CheckPTE(bool &converged, bool &acceptable) const {
    double error_p = 0;
#pragma unroll
    for (int m = 0; m < NM; ++m)
      error_p += residual[m] * residual[m];
    error_p = sqrt(error_p);
    error_p *= uscale;
    const double error_v = fabs(residual[NEQ - 1]);
    const bool converged_p =
        (error_p < params_.pte_rel_tolerance_p || error_p < params_.pte_abs_tolerance_p);
    const bool converged_v = error_v < params_.pte_rel_tolerance_v;
    const bool acceptable_p = (error_p < params_.pte_rel_tolerance_p_sufficient ||
                               error_p < params_.pte_abs_tolerance_p_sufficient);
    const bool acceptable_v = error_v < params_.pte_rel_tolerance_v_sufficient;
    converged = converged_p && converged_v;
    acceptable = acceptable_p && acceptable_v;
  }
  __device__ inline void Jacobian() {
#pragma unroll
    for (int m = 0; m < NM; m++) {
      const double dv = (vfrac[m] < 0.5 ? 1.0 : -1.0) * vfrac[m] * params_.derivative_eps;
      const double vf_pert = vfrac[m] + dv;
      const double rho_pert = ratio(rhobar[m], vf_pert);
      double e_pert = 0.0;
      double p_pert = ratio(
          B::GetPressureFromPreferred(m, rho_pert, Tnorm * Tequil, e_pert, true), uscale);
      dpdv[m] = stdmin(ratio((p_pert - press[m]), dv), -fabs(params_.pte_max_dpdv));
      const double dT = Tequil * params_.derivative_eps;
      p_pert = ratio(
          B::GetPressureFromPreferred(m, rho[m], Tnorm * (Tequil + dT), e_pert, true), uscale);
      dpdT[m] = ratio((p_pert - press[m]), dT);
    }
#pragma unroll
    for (int i = 0; i < NEQ * NEQ; ++i)
      jacobian[i] = 0.0;
#pragma unroll
    for (int m = 0; m < NM; m++) {
      jacobian[m * NEQ + m] = dpdv[m];
      jacobian[m * NEQ + NM] = dpdT[m];
    }
#pragma unroll
    for (int m = 0; m < NEQ - 1; ++m)
      jacobian[(NEQ - 1) * NEQ + m] = 1.0;
  }
  __device__ inline double ScaleDx() {
    double scale = 1.0;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      if (scale * dx[m] < -params_.vfrac_safety_fac * vfrac[m]) {
        scale = -params_.vfrac_safety_fac * ratio(vfrac[m], dx[m]);
      }
    }
    const double Tnew = Tequil + scale * dx[NM];
#pragma unroll
    for (int m = 0; m < NM; m++) {
      const double rho_min =
          stdmax(eos.RhoPmin(m, Tnorm * Tequil), eos.RhoPmin(m, Tnorm * Tnew));
      const double alpha_max = ratio(rhobar[m], rho_min);
      if (alpha_max < vfrac[m]) continue;
      if (scale * dx[m] > 0.5 * (alpha_max - vfrac[m])) {
        scale = 0.5 * ratio(alpha_max - vfrac[m], dx[m]);
      }
    }
    if (scale * dx[NM] < -0.95 * Tequil) {
      scale = -0.95 * ratio(Tequil, dx[NM]);
    }
#pragma unroll
    for (int i = 0; i < NEQ; ++i)
      dx[i] *= scale;
    return scale;
  }
  __device__ inline double TestUpdate(const double scale, const bool cache_state = false) {
    if (cache_state) {
      Ttemp = Tequil;
#pragma unroll
      for (int m = 0; m < NM; ++m)
        vtemp[m] = vfrac[m];
    }
    Tequil = Ttemp + scale * dx[NM];
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      vfrac[m] = vtemp[m] + scale * dx[m];
      rho[m] = ratio(rhobar[m], vfrac[m]);
      u[m] = rhobar[m] * eos.e_from_rho_T(m, rho[m], Tnorm * Tequil);
      press[m] = ratio(eos.P_from_rho_T(m, rho[m], Tnorm * Tequil), uscale);
      sie[m] = ratio(u[m], rhobar[m]);
      u[m] = ratio(u[m], uscale);
      temp[m] = Tequil;
    }
    Residual();
    return B::ResidualNorm();
  }
};

// ---------------------------------------------------------------------------------------
// PTESolverRhoU (:1934-2183): unknowns are the NM volume fractions and the NM energy
// densities; equations: volume sum, energy sum, pressure and temperature equality of
// consecutive materials.  Init tries the ideal-gas PTE guess (TryIdealPTE, :692-763).
// ---------------------------------------------------------------------------------------
template <int NM, class Eos>
struct RhoU : public Base<NM, 2 * NM, Eos> {
  static constexpr int NEQ = 2 * NM;
  using B = Base<NM, 2 * NM, Eos>;
  using B::dx;
  using B::eos;
  using B::jacobian;
  using B::params_;
  using B::press;
  using B::residual;
  using B::rho;
  using B::rhobar;
  using B::sie;
  using B::temp;
  using B::Tnorm;
  using B::u;
  using B::uscale;
  using B::utotal_scale;
  using B::vfrac;
  double dpdv[NM], dtdv[NM], dpde[NM], dtde[NM], vtemp[NM], utemp[NM];

  __device__ RhoU(const Eos &e, const MixParamsDev &p) : B(e, p) {}

  __device__ inline double Init() {
    B::InitBase();
    Residual();
    TryIdealPTE();
    return B::ResidualNorm();
  }
  SGX_D void Fixup() {}
  SGX_D void Finalize() { // base Finalize (:351-357)
#pragma unroll
    for (int m = 0; m < NM; m++) {
      temp[m] *= Tnorm;
      u[m] *= uscale;
      press[m] *= uscale;
    }
  }
  SGX_D void Residual() {
    const double vsum = sum_neumaier(vfrac, NM);
    const double esum = sum_neumaier(u, NM);
    residual[0] = B::vfrac_total - vsum;
    residual[1] = utotal_scale - esum;
#pragma unroll
    for (int m = 0; m < NM - 1; ++m)
      residual[2 + m] = press[m + 1] - press[m];
#pragma unroll
    for (int m = NM + 1; m < NEQ; m++)
      residual[m] = temp[m - NM] - temp[m - NM - 1];
  }
  // GetIdealPTE (:676-688) + TryIdealPTE (:692-763)
  __device__ inline void TryIdealPTE() {
    double rhoBsum = 0.0;
    double Asum = 0.0;
#pragma unroll
    for (int m = 0; m < NM; m++) {
      Asum += vfrac[m] * ratio(press[m], temp[m]);
      rhoBsum += rho[m] * vfrac[m] * ratio(sie[m], temp[m]);
    }
    Asum *= uscale / Tnorm;
    rhoBsum /= Tnorm;
    const double Tideal = uscale / rhoBsum / Tnorm;
    const double Pideal = ratio(Tnorm * Tideal * Asum, uscale * B::vfrac_total);
    double etemp[NM], ptemp[NM], vsave[NM], rtemp[NM], res[NEQ];
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      etemp[m] = u[m];
      ptemp[m] = press[m];
      vsave[m] = vfrac[m];
      rtemp[m] = rho[m];
    }
    double res_norm_old = 0.0;
#pragma unroll
    for (int m = 0; m < NEQ; ++m) {
      res[m] = residual[m];
      res_norm_old += res[m] * res[m];
    }
    const double alpha = ratio(Pideal, Tideal);
    bool aborted = false;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      if (!aborted) {
        vfrac[m] *= ratio(press[m], (temp[m] * alpha));
        if (Tnorm * Tideal < 0 ||
            ratio(rhobar[m], vfrac[m]) < eos.RhoPmin(m, Tnorm * Tideal)) {
#pragma unroll
          for (int n = 0; n < NM; n++)
            if (n <= m) vfrac[n] = vsave[n];
          aborted = true;
        }
      }
    }
    if (aborted) return;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      rho[m] = ratio(rhobar[m], vfrac[m]);
      const double sie_m = eos.e_from_rho_T(m, rho[m], Tnorm * Tideal);
      u[m] = rhobar[m] * ratio(sie_m, uscale);
      press[m] =
          ratio(B::GetPressureFromPreferred(m, rho[m], Tnorm * Tideal, sie_m, false), uscale);
    }
    Residual();
    double res_norm_new = 0.0;
#pragma unroll
    for (int m = 0; m < NEQ; m++)
      res_norm_new += residual[m] * residual[m];
    if (res_norm_new > res_norm_old) {
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        vfrac[m] = vsave[m];
        rho[m] = rtemp[m];
        u[m] = etemp[m];
        press[m] = ptemp[m];
      }
#pragma unroll
      for (int m = 0; m < NEQ; ++m)
        residual[m] = res[m];
    } else {
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        temp[m] = Tideal;
        sie[m] = uscale * ratio(u[m], rhobar[m]);
      }
    }
  }
  SGX_D void CheckPTE(bool &converged, bool &acceptable) const {
    double mean_p = vfrac[0] * press[0];
    double mean_t = rhobar[0] * temp[0];
    double error_p = 0.0;
    double error_t = 0.0;
#pragma unroll
    for (int m = 1; m < NM; ++m) {
      mean_p += vfrac[m] * press[m];
      mean_t += rhobar[m] * temp[m];
      error_p += residual[m + 1] * residual[m + 1];
      error_t += residual[m + NM] * residual[m + NM];
    }
    mean_t = ratio(mean_t, B::rho_total);
    error_p = sqrt(error_p);
    error_t = sqrt(error_t);
    const bool converged_p = (error_p < params_.pte_rel_tolerance_p * fabs(mean_p) ||
                              error_p < params_.pte_abs_tolerance_p);
    const bool converged_t = (error_t < params_.pte_rel_tolerance_t * mean_t ||
                              error_t < params_.pte_abs_tolerance_t);
    const bool acceptable_p =
        (error_p < params_.pte_rel_tolerance_p_sufficient * fabs(mean_p) ||
         error_p < params_.pte_abs_tolerance_p_sufficient);
    const bool acceptable_t = (error_t < params_.pte_rel_tolerance_t_sufficient * mean_t ||
                               error_t < params_.pte_abs_tolerance_t_sufficient);
    converged = converged_p && converged_t;
    acceptable = acceptable_p && acceptable_t;
  }
  __device__ inline void Jacobian() {
#pragma unroll
    for (int m = 0; m < NM; m++) {
      const double dv = (vfrac[m] < 0.5 ? 1.0 : -1.0) * vfrac[m] * params_.derivative_eps;
      const double vf_pert = vfrac[m] + dv;
      const double rho_pert = ratio(rhobar[m], vf_pert);
      double t_pert = ratio(eos.T_from_rho_e(m, rho_pert, sie[m]), Tnorm);
      double p_pert = ratio(
          B::GetPressureFromPreferred(m, rho_pert, Tnorm * t_pert, sie[m], false), uscale);
      dpdv[m] = stdmin(ratio(p_pert - press[m], dv), -fabs(params_.pte_max_dpdv));
      dtdv[m] = ratio(t_pert - temp[m], dv);
      const double de = fabs(u[m]) * params_.derivative_eps;
      const double e_pert = ratio(u[m] + de, rhobar[m]);
      t_pert = ratio(eos.T_from_rho_e(m, rho[m], uscale * e_pert), Tnorm);
      p_pert = ratio(
          B::GetPressureFromPreferred(m, rho[m], Tnorm * t_pert, uscale * e_pert, false),
          uscale);
      dpde[m] = ratio(p_pert - press[m], de);
      dtde[m] = ratio(t_pert - temp[m], de);
      if (fabs(dtde[m]) < params_.min_dtde) dtde[m] = params_.derivative_eps;
    }
#pragma unroll
    for (int i = 0; i < NEQ * NEQ; ++i)
      jacobian[i] = 0.0;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      jacobian[m] = 1.0;
      jacobian[2 * NM + NM + m] = 1.0;
    }
    // the reference writes these eight entries unconditionally (:2076-2083); for NM == 1 they
    // index rows 2 and 3 of a 2 x 2 matrix, i.e. past the Jacobian, so they are guarded here
    if constexpr (NM > 1) {
      jacobian[2 * 2 * NM] = dpdv[0];
      jacobian[NM * 2 * NM + NM - 1] = -dpdv[NM - 1];
      jacobian[(NM + 1) * 2 * NM] = dtdv[0];
      jacobian[(2 * NM - 1) * 2 * NM + NM - 1] = -dtdv[NM - 1];
      jacobian[2 * 2 * NM + NM] = dpde[0];
      jacobian[NM * 2 * NM + 2 * NM - 1] = -dpde[NM - 1];
      jacobian[(NM + 1) * 2 * NM + NM] = dtde[0];
      jacobian[(2 * NM - 1) * 2 * NM + 2 * NM - 1] = -dtde[NM - 1];
    }
#pragma unroll
    for (int m = 1; m < NM - 1; ++m) {
      jacobian[(1 + m) * 2 * NM + m] = -dpdv[m];
      jacobian[(2 + m) * 2 * NM + m] = dpdv[m];
      jacobian[(NM + m) * 2 * NM + m] = -dtdv[m];
      jacobian[(NM + m + 1) * 2 * NM + m] = dtdv[m];
      jacobian[(1 + m) * 2 * NM + NM + m] = -dpde[m];
      jacobian[(2 + m) * 2 * NM + NM + m] = dpde[m];
      jacobian[(NM + m) * 2 * NM + NM + m] = -dtde[m];
      jacobian[(NM + m + 1) * 2 * NM + NM + m] = dtde[m];
    }
  }
  __device__ inline double ScaleDx() {
    double scale = 1.0;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      if (scale * dx[m] < -0.1 * vfrac[m]) {
        scale = -0.1 * ratio(vfrac[m], dx[m]);
      }
      const double dt = (dtdv[m] * dx[m] + dtde[m] * dx[m + NM]);
      if (scale * dt < -0.1 * temp[m]) {
        scale = -0.1 * ratio(temp[m], dt);
      }
      const double tt = temp[m] + scale * dt;
      const double rho_min = stdmax(eos.RhoPmin(m, Tnorm * temp[m]), eos.RhoPmin(m, Tnorm * tt));
      const double alpha_max = ratio(rhobar[m], rho_min);
      if (alpha_max < vfrac[m]) continue;
      if (scale * dx[m] > 0.5 * (alpha_max - vfrac[m])) {
        scale = 0.5 * ratio(alpha_max - vfrac[m], dx[m]);
      }
    }
#pragma unroll
    for (int i = 0; i < NEQ; ++i)
      dx[i] *= scale;
    return scale;
  }
  __device__ inline double TestUpdate(const double scale, const bool cache_state = false) {
    if (cache_state) {
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        vtemp[m] = vfrac[m];
        utemp[m] = u[m];
      }
    }
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      vfrac[m] = vtemp[m] + scale * dx[m];
      rho[m] = ratio(rhobar[m], vfrac[m]);
      u[m] = utemp[m] + scale * dx[NM + m];
      sie[m] = uscale * ratio(u[m], rhobar[m]);
      temp[m] = ratio(eos.T_from_rho_e(m, rho[m], sie[m]), Tnorm);
      press[m] = ratio(
          B::GetPressureFromPreferred(m, rho[m], Tnorm * temp[m], sie[m], false), uscale);
    }
    Residual();
    return B::ResidualNorm();
  }
};

// ---------------------------------------------------------------------------------------
// PTESolver (:2188-2291)
// ---------------------------------------------------------------------------------------
template <class System>
__device__ inline SolverStatusDev PTESolver(System &s) {
  SolverStatusDev status;
  double &err = status.residual;
  bool &converged = status.converged;
  bool close_enough = false;

  err = s.Init();
  const MixParamsDev &params = s.params_;
  converged = false;
  const unsigned long long pte_max_iter =
      (unsigned long long)System::nmat * params.pte_max_iter_per_mat;
  int &niter = s.niter;
  unsigned long long &small_step_iters = status.small_step_iters;
  for (niter = 0; (unsigned long long)niter < pte_max_iter; ++niter) {
    status.max_niter =
        status.max_niter > (unsigned long long)niter ? status.max_niter : niter;
    s.CheckPTE(converged, close_enough);
    if (converged) break;
    s.Jacobian();
    const bool success = s.Solve();
    if (!success) {
      converged = false;
      break;
    }
    double scale = stdmin(1.0, s.ScaleDx());
    if (fabs(scale) < params.pte_small_step_thresh) small_step_iters++;
    if (small_step_iters >= params.pte_small_step_tries) {
      converged = false;
      break;
    }
    const double gradfdx = -2.0 * scale * err;
    scale = 1.0;
    const double err_old = err;
    err = s.TestUpdate(scale, true);
    if (err > err_old + params.line_search_alpha * gradfdx) {
      scale = 0.5;
      const double err_mid = s.TestUpdate(scale);
      if (err_mid < err && err_mid < err_old) {
        scale = 0.75 + 0.5 * ratio(err_mid - err, err - 2.0 * err_mid + err_old);
      }
      for (unsigned long long line_iter = 0; line_iter < params.line_search_max_iter;
           line_iter++) {
        err = s.TestUpdate(scale);
        if (err < err_old + params.line_search_alpha * scale * gradfdx) break;
        scale *= params.line_search_fac;
        status.max_line_niter =
            line_iter > status.max_line_niter ? line_iter : status.max_line_niter;
      }
    }
    s.Fixup();
    if (((err > params.pte_slow_convergence_thresh * err_old) ||
         (unsigned long long)niter == pte_max_iter - 1) &&
        close_enough) {
      converged = true;
      status.slow_convergence_detected = true;
      break;
    }
  }
  s.Fixup();
  s.Finalize();
  return status;
}

SGX_D unsigned long long pack_status(const SolverStatusDev &s, const int niter) {
  auto clamp = [](unsigned long long v, unsigned long long mx) { return v > mx ? mx : v; };
  unsigned long long w = 0;
  w |= s.converged ? 1ull : 0ull;
  w |= s.slow_convergence_detected ? 2ull : 0ull;
  w |= clamp((unsigned long long)niter, 0xffff) << 8;
  w |= clamp(s.max_line_niter, 0xff) << 24;
  w |= clamp(s.small_step_iters, 0xffff) << 32;
  w |= clamp(s.max_niter, 0xffff) << 48;
  return w;
}

} // namespace mix
#endif // __CUDACC__
} // namespace sgx
#endif // SGX_PTE_CUH_
