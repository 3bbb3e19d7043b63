// get_sg_eos, the reference's batch mixed-cell driver (eos/get_sg_eos.cpp:36-280), for the
// (rho, e) input conditions (input_int 0 and 1: get_sg_eos_rho_e.cpp:22-90), the (rho, T)
// condition (input_int -3: get_sg_eos_rho_t.cpp:22-101, PTESolverFixedT) and the (P, T)
// condition (input_int -1: get_sg_eos_p_t.cpp:22-96, one DensityEnergyFromPressureTemperature
// per material, no solver), with init_functor / final_functor of
// get_sg_eos_functors.hpp:27-394.  The (rho, P) condition (-2) is refused: the reference runs
// PTESolverFixedP with iterate_t_guess = true there, whose result is undefined.
// For (rho, e), per cell
//   * mass fractions normalised, materials below `mass_frac_cutoff` dropped, the participating
//     ones compacted (pte_mats / pte_idxs) and given consistent volume-fraction / density /
//     energy guesses;
//   * more than one participant: PTESolverRhoT over the compacted set with the library-default
//     MixParams and per-material lambda caches (CacheAccessor, zero-initialised like the
//     reference's solver scratch); a single participant: the pure-cell fast path
//     T(rho, sie) then P(rho, T);
//   * cell pressure / temperature as volume-fraction-weighted sums, per-material energies and
//     volumes, bulk modulus, specific heat (per eV) and dP/de by the reference's mixing rules,
//     pmax.
// One thread owns one cell; the participant count is a per-cell run-time value, so the solver is
// entered through a switch over its compile-time sizes (2..4 participants).
// Array layout is the reference's: per-cell arrays [cell_dim], per-material arrays
// LayoutLeft (cell_dim, nmat): frac(i, m) = p[i + cell_dim * m]; offsets are 1-based (Fortran).
// This header holds the kernel template; one translation unit per input mode instantiates it
// (sgx_get_sg_eos_rho_e.cu, _rho_t.cu, _p_t.cu) so that the three compile in parallel -- as one
// unit they were the longest compile of the library (9.5 min).
#ifndef SGX_GET_SG_EOS_CUH_
#define SGX_GET_SG_EOS_CUH_
#include "sgx_internal.hpp"
#include "sgx_pte_tab.cuh"


namespace sgx {
namespace sg {

constexpr double kEv2K = 1.160451930280894026e4; // eos/get_sg_eos.hpp:68

struct SgArgs {
  int nmat, ncell, cell_dim;
  const int *offsets;     // [ncell], 1-based cell index
  int eos_idx[kMaxMat];   // eos_offsets[m] - 1
  double *press, *pmax, *vol, *spvol, *sie, *temp, *bmod, *dpde, *cv;
  double *frac_mass, *frac_vol, *frac_ie, *frac_bmod, *frac_dpde, *frac_cv;
  double mass_frac_cutoff;
  int *nfail; // cells whose PTE solve did not converge (or whose participant count is not built)
  // (rho, e) mode, staged form: cells in which EVERY material participates are solved by the
  // lane-refill state machine (sgx_pte_tab_sm_mod.cu) between a prep and this kernel; pre_skip
  // [iloop] != 0 marks the cells that were not (solved here as before).  Null: solve all here.
  const unsigned char *pre_skip = nullptr;
  const double *pre_rho = nullptr, *pre_vfrac = nullptr, *pre_sie = nullptr, *pre_temp = nullptr,
               *pre_press = nullptr; // [ncell][nmat] by loop index, as the solver left them
  const unsigned long long *pre_status = nullptr;
};

// Stages the full cells of the (rho, e) mode for the state machine: init_functor's arithmetic
// (get_sg_eos_functors.hpp:66-121 with t_mult = 0, s_mult = 1, p_mult = 0) on LOCAL copies --
// the caller's arrays are written by the finishing kernel only, exactly once.
static __global__ void __launch_bounds__(256)
    get_sg_eos_prep_kernel(const __grid_constant__ SgArgs A, unsigned char *__restrict__ skip,
                           double *__restrict__ s_rho, double *__restrict__ s_vfrac,
                           double *__restrict__ s_sie, double *__restrict__ s_temp,
                           double *__restrict__ s_press, double *__restrict__ sie_tot,
                           double *__restrict__ vfrac_tot) {
  const int nmat = A.nmat;
  const int64_t cd = A.cell_dim;
  for (int iloop = blockIdx.x * blockDim.x + threadIdx.x; iloop < A.ncell;
       iloop += gridDim.x * blockDim.x) {
    const int64_t i = A.offsets[iloop] - 1;
    double fm[kMaxMat];
    double mass_sum = 0.0;
    for (int m = 0; m < nmat; ++m)
      mass_sum += A.frac_mass[i + cd * m];
    int npte = 0;
    for (int m = 0; m < nmat; ++m) {
      fm[m] = A.frac_mass[i + cd * m] / mass_sum;
      if (fm[m] > A.mass_frac_cutoff) npte += 1;
    }
    if (npte != nmat) {
      skip[iloop] = 1;
      continue;
    }
    skip[iloop] = 0;
    double vfrac_sum = 0.;
    const int64_t base = (int64_t)iloop * nmat;
    for (int m = 0; m < nmat; ++m) {
      const double vf = fm[m];
      vfrac_sum += vf;
      s_vfrac[base + m] = vf;
      s_rho[base + m] = fm[m] / A.spvol[i] / vf;
      s_temp[base + m] = A.temp[i] * kEv2K * 0.0;
      s_press[base + m] = A.press[i] * 0.0;
      s_sie[base + m] = A.sie[i] * fm[m] * 1.0;
    }
    sie_tot[iloop] = A.sie[i];
    vfrac_tot[iloop] = vfrac_sum;
  }
}

template <int NM>
__device__ bool solve_rhot(TabEos &eos, const MixParamsDev &prm, const double vfrac_sum,
                           const double sie_cell, double *rho, double *vfrac, double *sie,
                           double *temp, double *press) {
  mix::RhoT<NM, TabEos> s(eos, prm);
#pragma unroll
  for (int m = 0; m < NM; ++m) {
    s.rho[m] = rho[m];
    s.vfrac[m] = vfrac[m];
    s.sie[m] = sie[m];
    s.temp[m] = temp[m];
    s.press[m] = press[m];
  }
  s.sie_total = sie_cell;
  s.vfrac_total = vfrac_sum;
  s.Tnorm = 0.0; // no guess: GetTguess decides (PTESolverRhoT default argument, :927)
  const SolverStatusDev st = mix::PTESolver(s);
#pragma unroll
  for (int m = 0; m < NM; ++m) {
    rho[m] = s.rho[m];
    vfrac[m] = s.vfrac[m];
    sie[m] = s.sie[m];
    temp[m] = s.temp[m];
    press[m] = s.press[m];
  }
  return st.converged;
}

// final_functor (get_sg_eos_functors.hpp:207-283): cell averages and per-material outputs
__device__ inline void finalise(const SgArgs &A, TabEos &eos, const int64_t i, const int64_t cd,
                                const int nmat, const int npte, const int *mats,
                                const double mass_sum, const double t_mult, const double s_mult,
                                const double p_mult, const double *rho_pte,
                                const double *vfrac_pte, const double *sie_pte,
                                const double *temp_pte, const double *press_pte) {
  auto F = [&](double *p, const int m) -> double & { return p[i + cd * m]; };
  const bool do_t = t_mult == 1.0, do_s = s_mult == 1.0, do_p = p_mult == 1.0;
  if (do_t) A.temp[i] = 0.0;
  if (do_p) A.press[i] = 0.0;
  if (do_s) A.sie[i] = 0.0;
  A.bmod[i] = 0.0;
  A.cv[i] = 0.0;
  A.dpde[i] = 0.0;
  for (int mp = 0; mp < npte; ++mp) {
    const int m = mats[mp];
    A.press[i] += press_pte[mp] * vfrac_pte[mp] * p_mult;
    A.temp[i] += temp_pte[mp] * vfrac_pte[mp] * t_mult;
    const double ie_m = sie_pte[mp] * F(A.frac_mass, m) * mass_sum;
    A.sie[i] += ie_m * s_mult;
    F(A.frac_ie, m) = ie_m;
    F(A.frac_vol, m) = vfrac_pte[mp] * A.vol[i];
    const double bmod_m = eos.bmod_from_rho_T(mp, rho_pte[mp], temp_pte[mp]);
    A.bmod[i] += bmod_m * vfrac_pte[mp];
    const double cv_m = kEv2K * eos.cv_from_rho_T(mp, rho_pte[mp], temp_pte[mp]);
    A.cv[i] += cv_m * F(A.frac_mass, m);
    const double dpde_m = eos.gamma_from_rho_T(mp, rho_pte[mp], temp_pte[mp]);
    A.dpde[i] += dpde_m * vfrac_pte[mp];
    if (A.frac_bmod) F(A.frac_bmod, m) = bmod_m;
    if (A.frac_cv) F(A.frac_cv, m) = cv_m;
    if (A.frac_dpde) F(A.frac_dpde, m) = dpde_m;
  }
  if (do_t) A.temp[i] /= kEv2K;
  if (do_s) A.sie[i] /= mass_sum;
  for (int m = 0; m < nmat; ++m)
    F(A.frac_mass, m) *= mass_sum;
}

// PTESolverFixedT over the compacted set (get_sg_eos_rho_t.cpp:66-70): T_true = temp_pte[0]
template <int NM>
__device__ bool solve_fixed_t(TabEos &eos, const MixParamsDev &prm, const double vfrac_sum,
                              const double T_true, double *rho, double *vfrac, double *sie,
                              double *temp, double *press) {
  mix::FixedT<NM, TabEos> s(eos, prm);
#pragma unroll
  for (int m = 0; m < NM; ++m) {
    s.rho[m] = rho[m];
    s.vfrac[m] = vfrac[m];
    s.sie[m] = sie[m];
    s.temp[m] = temp[m];
    s.press[m] = press[m];
  }
  s.sie_total = 1.0; // the FixedT constructor hands the base sie_tot = 1 (:1504-1506)
  s.vfrac_total = vfrac_sum;
  s.Tnorm = T_true;
  const SolverStatusDev st = mix::PTESolver(s);
#pragma unroll
  for (int m = 0; m < NM; ++m) {
    rho[m] = s.rho[m];
    vfrac[m] = s.vfrac[m];
    sie[m] = s.sie[m];
    temp[m] = s.temp[m];
    press[m] = s.press[m];
  }
  return st.converged;
}

constexpr int kRhoE = 1, kRhoT = -3, kPT = -1; // input_condition, eos/get_sg_eos.hpp:50-56

template <int MODE>
__global__ void __launch_bounds__(128)
    get_sg_eos_kernel(const __grid_constant__ EosSet set,
                            const __grid_constant__ MixParamsDev prm,
                            const __grid_constant__ SgArgs A) {
  TabEos eos(set);
  const int nmat = A.nmat;
  const int64_t cd = A.cell_dim;
  for (int iloop = blockIdx.x * blockDim.x + threadIdx.x; iloop < A.ncell;
       iloop += gridDim.x * blockDim.x) {
    const int64_t i = A.offsets[iloop] - 1;
    auto F = [&](double *p, const int m) -> double & { return p[i + cd * m]; };
    int idxs[kMaxMat], mats[kMaxMat];
    double vfrac_pte[kMaxMat], sie_pte[kMaxMat], temp_pte[kMaxMat], press_pte[kMaxMat],
        rho_pte[kMaxMat];
    double cache[kMaxMat * kMaxNumLambdas];
    // the reference zeroes the solver scratch that the cache accessor lives in
    for (int k = 0; k < kMaxMat * kMaxNumLambdas; ++k)
      cache[k] = 0.0;
    eos.idx = idxs;
    eos.lam = cache;
    if constexpr (MODE == kPT) { // get_sg_eos_p_t.cpp:40-90: every material takes part
      double mass_sum = 0.0;
      for (int m = 0; m < nmat; ++m) {
        mass_sum += F(A.frac_mass, m);
        idxs[m] = A.eos_idx[m];
        mats[m] = m;
        temp_pte[m] = A.temp[i] * kEv2K;
        press_pte[m] = A.press[i];
      }
      for (int m = 0; m < nmat; ++m)
        F(A.frac_mass, m) /= mass_sum;
      double vfrac_tot = 0.0, sie_tot = 0.0;
      for (int m = 0; m < nmat; ++m) {
        eos.rho_e_from_P_T(m, press_pte[m], temp_pte[m], rho_pte[m], sie_pte[m]);
        vfrac_pte[m] = F(A.frac_mass, m) / rho_pte[m] * mass_sum; // a physical volume
        vfrac_tot += vfrac_pte[m];
        sie_tot += sie_pte[m] * F(A.frac_mass, m);
      }
      A.sie[i] = sie_tot;
      A.vol[i] = vfrac_tot;
      A.spvol[i] = A.vol[i] / mass_sum;
      for (int m = 0; m < nmat; ++m)
        vfrac_pte[m] /= vfrac_tot;
      finalise(A, eos, i, cd, nmat, nmat, mats, mass_sum, 0.0, 0.0, 0.0, rho_pte, vfrac_pte,
               sie_pte, temp_pte, press_pte);
      A.pmax[i] = A.press[i] > A.pmax[i] ? A.press[i] : A.pmax[i];
      continue;
    }
    // ---- init_functor(i, tid, mass_sum, npte, vfrac_sum, t_mult, s_mult, p_mult):
    //      (0, 1, 0) for (rho, e) (get_sg_eos_rho_e.cpp:45), (1, 0, 0) for (rho, T) (_rho_t:43)
    const double t_mult = MODE == kRhoT ? 1.0 : 0.0, s_mult = MODE == kRhoT ? 0.0 : 1.0,
                 p_mult = 0.0;
    double mass_sum = 0.0;
    for (int m = 0; m < nmat; ++m) {
      mass_sum += F(A.frac_mass, m);
      idxs[m] = A.eos_idx[m];
      F(A.frac_vol, m) = F(A.frac_mass, m);
      F(A.frac_ie, m) = A.sie[i] * F(A.frac_mass, m) * s_mult;
    }
    for (int m = 0; m < nmat; ++m)
      F(A.frac_mass, m) /= mass_sum;
    int npte = 0;
    for (int m = 0; m < nmat; ++m) {
      if (F(A.frac_mass, m) > A.mass_frac_cutoff) {
        idxs[npte] = A.eos_idx[m];
        mats[npte] = m;
        npte += 1;
      }
      vfrac_pte[m] = 0.0;
      sie_pte[m] = 0.0;
      temp_pte[m] = 0.0;
      press_pte[m] = 0.0;
    }
    double vfrac_sum = 0.;
    for (int mp = 0; mp < npte; ++mp) {
      const int m = mats[mp];
      vfrac_pte[mp] = F(A.frac_mass, m);
      vfrac_sum += vfrac_pte[mp];
      rho_pte[mp] = F(A.frac_mass, m) / A.spvol[i] / vfrac_pte[mp];
      temp_pte[mp] = A.temp[i] * kEv2K * t_mult;
      press_pte[mp] = A.press[i] * p_mult;
      sie_pte[mp] = A.sie[i] * F(A.frac_mass, m) * s_mult;
    }
    bool pte_converged = true;
    if constexpr (MODE == kRhoT) { // get_sg_eos_rho_t.cpp:44-93
      double sie_tot_true = 0.0;
      if (npte > 1) {
        const double T_true = temp_pte[0];
        switch (npte) {
        case 2:
          pte_converged = solve_fixed_t<2>(eos, prm, vfrac_sum, T_true, rho_pte, vfrac_pte,
                                           sie_pte, temp_pte, press_pte);
          break;
        case 3:
          pte_converged = solve_fixed_t<3>(eos, prm, vfrac_sum, T_true, rho_pte, vfrac_pte,
                                           sie_pte, temp_pte, press_pte);
          break;
        case 4:
          pte_converged = solve_fixed_t<4>(eos, prm, vfrac_sum, T_true, rho_pte, vfrac_pte,
                                           sie_pte, temp_pte, press_pte);
          break;
        default:
          pte_converged = false;
          break;
        }
        if (!pte_converged && A.nfail) atomicAdd(A.nfail, 1);
        for (int mp = 0; mp < npte; ++mp)
          sie_tot_true += sie_pte[mp] * F(A.frac_mass, mats[mp]);
      } else if (npte == 1) { // pure cell
        sie_pte[0] = eos.e_from_rho_T(0, rho_pte[0], temp_pte[0]);
        sie_tot_true = sie_pte[0];
        press_pte[0] = eos.P_from_rho_T(0, rho_pte[0], temp_pte[0]);
      }
      A.sie[i] = sie_tot_true;
      finalise(A, eos, i, cd, nmat, npte, mats, mass_sum, 0.0, 0.0, 1.0, rho_pte, vfrac_pte,
               sie_pte, temp_pte, press_pte);
      A.pmax[i] = A.press[i] > A.pmax[i] ? A.press[i] : A.pmax[i];
      continue;
    }
    if (npte > 1 && A.pre_skip && !A.pre_skip[iloop]) { // solved by the state machine
      const int64_t pb = (int64_t)iloop * nmat;
      for (int mp = 0; mp < npte; ++mp) {
        rho_pte[mp] = A.pre_rho[pb + mp];
        vfrac_pte[mp] = A.pre_vfrac[pb + mp];
        sie_pte[mp] = A.pre_sie[pb + mp];
        temp_pte[mp] = A.pre_temp[pb + mp];
        press_pte[mp] = A.pre_press[pb + mp];
      }
      pte_converged = (A.pre_status[iloop] & 1ull) != 0;
      if (!pte_converged && A.nfail) atomicAdd(A.nfail, 1);
    } else if (npte > 1) {
      switch (npte) {
      case 2:
        pte_converged = solve_rhot<2>(eos, prm, vfrac_sum, A.sie[i], rho_pte, vfrac_pte, sie_pte,
                                      temp_pte, press_pte);
        break;
      case 3:
        pte_converged = solve_rhot<3>(eos, prm, vfrac_sum, A.sie[i], rho_pte, vfrac_pte, sie_pte,
                                      temp_pte, press_pte);
        break;
      case 4:
        pte_converged = solve_rhot<4>(eos, prm, vfrac_sum, A.sie[i], rho_pte, vfrac_pte, sie_pte,
                                      temp_pte, press_pte);
        break;
      default:
        pte_converged = false; // more participants than this build instantiates
        break;
      }
      if (!pte_converged && A.nfail) atomicAdd(A.nfail, 1);
    } else if (npte == 1) { // pure cell
      temp_pte[0] = eos.T_from_rho_e(0, rho_pte[0], sie_pte[0]);
      press_pte[0] = eos.P_from_rho_T(0, rho_pte[0], temp_pte[0]);
    }
    finalise(A, eos, i, cd, nmat, npte, mats, mass_sum, 1.0, 0.0, 1.0, rho_pte, vfrac_pte,
             sie_pte, temp_pte, press_pte);
    A.pmax[i] = A.press[i] > A.pmax[i] ? A.press[i] : A.pmax[i];
  }
}


template <int MODE>
inline void launch_get_sg_eos_mode(const EosSet &set, const MixParamsDev &prm, const SgArgs &A,
                                   const LaunchCfg &c, cudaStream_t st) {
  get_sg_eos_kernel<MODE><<<c.blocks, c.threads, 0, st>>>(set, prm, A);
}

} // namespace sg
// defined in sgx_get_sg_eos_rho_e.cu / _rho_t.cu / _p_t.cu
void launch_get_sg_eos_rho_e(const EosSet &set, const MixParamsDev &prm, const sg::SgArgs &A,
                             const LaunchCfg &c, cudaStream_t st);
void launch_get_sg_eos_rho_t(const EosSet &set, const MixParamsDev &prm, const sg::SgArgs &A,
                             const LaunchCfg &c, cudaStream_t st);
void launch_get_sg_eos_p_t(const EosSet &set, const MixParamsDev &prm, const sg::SgArgs &A,
                           const LaunchCfg &c, cudaStream_t st);
} // namespace sgx
#endif // SGX_GET_SG_EOS_CUH_
