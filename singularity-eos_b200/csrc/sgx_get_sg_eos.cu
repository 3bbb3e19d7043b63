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
#include "sgx_internal.hpp"
#include "sgx_pte_tab.cuh"

namespace sgx {
namespace {

constexpr double kEv2K = 1.160451930280894026e4; // eos/get_sg_eos.hpp:68

struct SgArgs {
  int nmat, ncell, cell_dim;
  const int *offsets;     // [ncell], 1-based cell index
  int eos_idx[kMaxMat];   // eos_offsets[m] - 1
  double *press, *pmax, *vol, *spvol, *sie, *temp, *bmod, *dpde, *cv;
  double *frac_mass, *frac_vol, *frac_ie, *frac_bmod, *frac_dpde, *frac_cv;
  double mass_frac_cutoff;
  int *nfail; // cells whose PTE solve did not converge (or whose participant count is not built)
};

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
    if (npte > 1) {
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

} // namespace
} // namespace sgx

using namespace sgx;

extern "C" {

int sgx_get_sg_eos_device(int nmat, int ncell, int cell_dim, int input_int,
                          const int *eos_offsets, const sgx_eos_t *eos, const int *offsets,
                          double *press, double *pmax, double *vol, double *spvol, double *sie,
                          double *temp, double *bmod, double *dpde, double *cv,
                          double *frac_mass, double *frac_vol, double *frac_ie,
                          double *frac_bmod, double *frac_dpde, double *frac_cv,
                          double mass_frac_cutoff, int *nfail, sgx_stream_t stream) {
  DeviceGuard guard;
  if (input_int == -2)
    return fail(SGX_ERR_UNSUPPORTED, "get_sg_eos: the (rho, P) input condition runs "
                "PTESolverFixedP with iterate_t_guess = true in the reference, whose result is "
                "undefined (uninitialised uscale); not built");
  if (input_int != 0 && input_int != 1 && input_int != -3 && input_int != -1)
    return fail(SGX_ERR_INVALID, "get_sg_eos: unknown input_int %d", input_int);
  if (nmat < 1 || nmat > kMaxMat)
    return fail(SGX_ERR_UNSUPPORTED, "get_sg_eos: nmat = %d, this build holds 1..%d", nmat, kMaxMat);
  if (ncell < 0 || cell_dim < ncell) return fail(SGX_ERR_INVALID, "get_sg_eos: bad sizes");
  if (!eos_offsets || !eos || !offsets || !press || !pmax || !vol || !spvol || !sie || !temp ||
      !bmod || !dpde || !cv || !frac_mass || !frac_vol || !frac_ie)
    return fail(SGX_ERR_INVALID, "get_sg_eos: null array");
  if (ncell == 0) return SGX_OK;
  int max_off = 0;
  for (int m = 0; m < nmat; ++m)
    max_off = eos_offsets[m] > max_off ? eos_offsets[m] : max_off;
  if (max_off > kMaxMat)
    return fail(SGX_ERR_UNSUPPORTED, "get_sg_eos: %d EOS objects, this build holds %d", max_off,
                kMaxMat);
  EosSet set;
  set.nmat = max_off;
  SgArgs A;
  for (int m = 0; m < nmat; ++m) {
    const int k = eos_offsets[m] - 1;
    if (k < 0 || !eos[k]) return fail(SGX_ERR_INVALID, "get_sg_eos: bad eos_offsets[%d]", m);
    A.eos_idx[m] = k;
  }
  int device = -1;
  for (int k = 0; k < max_off; ++k) {
    if (!eos[k]) { // an index no material refers to
      set.e[k].kind = SGX_EOS_IDEAL;
      set.e[k].ideal = IdealGas{1.0, 1.0};
      set.mod[k] = ModParams();
      continue;
    }
    if (device < 0) device = eos[k]->device;
    if (eos[k]->device != device) return fail(SGX_ERR_INVALID, "EOS %d lives on another device", k);
    if (!set_material(eos[k], set, k))
      return fail(SGX_ERR_UNSUPPORTED, "get_sg_eos: EOS %d is kind %d", k, eos[k]->kind);
  }
  SGX_CUDA_OK(cudaSetDevice(device));
  sgx_mix_params dflt;
  sgx_mix_params_default(&dflt);
  const MixParamsDev prm = mix_params_to_dev(&dflt);
  A.nmat = nmat;
  A.ncell = ncell;
  A.cell_dim = cell_dim;
  A.offsets = offsets;
  A.press = press, A.pmax = pmax, A.vol = vol, A.spvol = spvol, A.sie = sie, A.temp = temp;
  A.bmod = bmod, A.dpde = dpde, A.cv = cv;
  A.frac_mass = frac_mass, A.frac_vol = frac_vol, A.frac_ie = frac_ie;
  A.frac_bmod = frac_bmod, A.frac_dpde = frac_dpde, A.frac_cv = frac_cv;
  A.mass_frac_cutoff = mass_frac_cutoff;
  A.nfail = nfail;
  const LaunchCfg c = stream_cfg(ncell, 128, 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (input_int == kRhoT)
    get_sg_eos_kernel<kRhoT><<<c.blocks, c.threads, 0, st>>>(set, prm, A);
  else if (input_int == kPT)
    get_sg_eos_kernel<kPT><<<c.blocks, c.threads, 0, st>>>(set, prm, A);
  else
    get_sg_eos_kernel<kRhoE><<<c.blocks, c.threads, 0, st>>>(set, prm, A);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

// The reference's signature and memory contract: every array is a HOST array; inputs are
// mirrored to the device, results copied back exactly as get_sg_eos.cpp:236-278 does.
int sgx_get_sg_eos(int nmat, int ncell, int cell_dim, int input_int, int *eos_offsets,
                   const sgx_eos_t *eos, int *offsets, double *press, double *pmax, double *vol,
                   double *spvol, double *sie, double *temp, double *bmod, double *dpde,
                   double *cv, double *frac_mass, double *frac_vol, double *frac_ie,
                   double *frac_bmod, double *frac_dpde, double *frac_cv,
                   double mass_frac_cutoff) {
  DeviceGuard guard;
  if (input_int == -2)
    return fail(SGX_ERR_UNSUPPORTED, "get_sg_eos: the (rho, P) input condition runs "
                "PTESolverFixedP with iterate_t_guess = true in the reference, whose result is "
                "undefined (uninitialised uscale); not built");
  if (input_int != 0 && input_int != 1 && input_int != -3 && input_int != -1)
    return fail(SGX_ERR_INVALID, "get_sg_eos: unknown input_int %d", input_int);
  if (!eos || !eos_offsets || nmat < 1 || !eos[eos_offsets[0] - 1])
    return fail(SGX_ERR_INVALID, "get_sg_eos: null EOS");
  if (ncell <= 0) return ncell == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative ncell");
  SGX_CUDA_OK(cudaSetDevice(eos[eos_offsets[0] - 1]->device));
  const bool do_bmod = frac_bmod != nullptr, do_dpde = frac_dpde != nullptr;
  const bool do_cv = frac_dpde != nullptr; // sic: get_sg_eos.cpp:66 tests frac_dpde for frac_cv
  const size_t nc = (size_t)cell_dim, nf = (size_t)cell_dim * nmat;
  // one device allocation: 9 per-cell + 6 per-material arrays, offsets, failure counter
  const size_t doubles = 9 * nc + 6 * nf;
  char *d = nullptr;
  const size_t bytes = doubles * sizeof(double) + (size_t)ncell * sizeof(int) + 256;
  SGX_CUDA_OK(cudaMalloc(&d, bytes));
  double *p = reinterpret_cast<double *>(d);
  double *dv[9], *df[6];
  double *hv[9] = {press, pmax, vol, spvol, sie, temp, bmod, dpde, cv};
  double *hf[6] = {frac_mass, frac_vol, frac_ie, frac_bmod, frac_dpde, frac_cv};
  for (int k = 0; k < 9; ++k, p += nc)
    dv[k] = p;
  for (int k = 0; k < 6; ++k, p += nf)
    df[k] = p;
  int *d_offsets = reinterpret_cast<int *>(p);
  int *d_nfail = d_offsets + ncell + ((ncell & 1) ? 1 : 0);
  cudaError_t err = cudaSuccess;
  auto up = [&](void *dst, const void *src, size_t n) {
    if (err == cudaSuccess && src) err = cudaMemcpy(dst, src, n, cudaMemcpyHostToDevice);
  };
  for (int k = 0; k < 9; ++k)
    up(dv[k], hv[k], nc * sizeof(double));
  for (int k = 0; k < 6; ++k)
    up(df[k], hf[k], nf * sizeof(double));
  up(d_offsets, offsets, (size_t)ncell * sizeof(int));
  if (err == cudaSuccess) err = cudaMemset(d_nfail, 0, sizeof(int));
  int rc = SGX_OK;
  if (err == cudaSuccess) {
    rc = sgx_get_sg_eos_device(nmat, ncell, cell_dim, input_int, eos_offsets, eos, d_offsets,
                               dv[0], dv[1], dv[2], dv[3], dv[4], dv[5], dv[6], dv[7], dv[8],
                               df[0], df[1], df[2], do_bmod ? df[3] : nullptr,
                               do_dpde ? df[4] : nullptr, do_cv ? df[5] : nullptr,
                               mass_frac_cutoff, d_nfail, nullptr);
    if (rc == SGX_OK) err = cudaDeviceSynchronize();
  }
  auto down = [&](void *dst, const void *src, size_t n) {
    if (err == cudaSuccess && rc == SGX_OK && dst)
      err = cudaMemcpy(dst, src, n, cudaMemcpyDeviceToHost);
  };
  // get_sg_eos.cpp:236-262: a cell quantity is copied back unless it is an input of the mode
  const bool r_inp = input_int != kPT, t_inp = input_int == kRhoT || input_int == kPT,
             p_inp = input_int == kPT, s_inp = input_int == 0 || input_int == 1;
  if (!p_inp) down(press, dv[0], nc * sizeof(double));
  down(pmax, dv[1], nc * sizeof(double));
  down(vol, dv[2], nc * sizeof(double));
  if (!r_inp) down(spvol, dv[3], nc * sizeof(double));
  if (!s_inp) down(sie, dv[4], nc * sizeof(double));
  if (!t_inp) down(temp, dv[5], nc * sizeof(double));
  down(bmod, dv[6], nc * sizeof(double));
  down(dpde, dv[7], nc * sizeof(double));
  down(cv, dv[8], nc * sizeof(double));
  down(frac_vol, df[1], nf * sizeof(double));
  down(frac_ie, df[2], nf * sizeof(double));
  if (do_bmod) down(frac_bmod, df[3], nf * sizeof(double));
  if (do_dpde) down(frac_dpde, df[4], nf * sizeof(double));
  if (do_cv) down(frac_cv, df[5], nf * sizeof(double));
  cudaFree(d);
  if (rc != SGX_OK) return rc;
  if (err != cudaSuccess) return fail(SGX_ERR_CUDA, "get_sg_eos: %s", cudaGetErrorString(err));
  return SGX_OK; // like the reference, whose failure count is never copied back (:237-240)
}

} // extern "C"
