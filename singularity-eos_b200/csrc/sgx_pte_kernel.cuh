// The straight transcription kernel: one thread walks one cell through PTESolver().
// Used for PTESolverPT (nested root finds defeat lane-level regrouping) and, for the
// benchmark mixture, as the comparison variant of the RhoT state-machine kernel.
#ifndef SGX_PTE_KERNEL_CUH_
#define SGX_PTE_KERNEL_CUH_

#include "sgx_internal.hpp"

namespace sgx {

// per-cell lambda cache (CacheAccessor, closure/mixed_cell_models.hpp:307-324): only the
// accessor of sgx_pte_tab.cuh has one
template <class Acc>
SGX_D void set_cell_cache(Acc &, double *) {}

template <int NM, class Acc, int SOLVER>
__global__ void __launch_bounds__(128)
    pte_kernel(const __grid_constant__ EosSet set, const __grid_constant__ MixParamsDev prm,
               double *__restrict__ rho, double *__restrict__ vfrac, double *__restrict__ sie,
               double *__restrict__ temp, double *__restrict__ press,
               const double *__restrict__ sie_tot_in, const double *__restrict__ vfrac_tot_in,
               const double *__restrict__ Tguess_in, unsigned long long *__restrict__ status,
               double *__restrict__ residual, unsigned long long *__restrict__ stats,
               const int64_t ncell, double *__restrict__ lambda = nullptr) {
  __shared__ unsigned int s_hist[SGX_PTE_NSTATS];
  if (stats) {
    for (int k = threadIdx.x; k < SGX_PTE_NSTATS; k += blockDim.x)
      s_hist[k] = 0;
    __syncthreads();
  }
  Acc eos(set);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = tid; t < ncell; t += nthr) {
    set_cell_cache(eos, lambda ? lambda + t * (NM * 3) : nullptr);
    using System = typename std::conditional<
        SOLVER == SGX_PTE_RHOT, mix::RhoT<NM, Acc>,
        typename std::conditional<
            SOLVER == SGX_PTE_PT, mix::PT<NM, Acc>,
            typename std::conditional<
                SOLVER == SGX_PTE_FIXED_T, mix::FixedT<NM, Acc>,
                typename std::conditional<SOLVER == SGX_PTE_FIXED_P, mix::FixedP<NM, Acc>,
                                          mix::RhoU<NM, Acc>>::type>::type>::type>::type;
    System s(eos, prm);
    const int64_t base = t * NM;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      s.rho[m] = rho[base + m];
      s.vfrac[m] = vfrac[base + m];
      s.sie[m] = sie[base + m];
      s.temp[m] = temp[base + m];
      s.press[m] = press[base + m];
    }
    // test/test_pte.cpp:188-195
    double sie_tot = 0.0, rho_tot = 0.0;
#pragma unroll
    for (int m = 0; m < NM; m++) {
      const double rhobar = s.rho[m] * s.vfrac[m];
      rho_tot += rhobar;
      sie_tot += rhobar * s.sie[m];
    }
    sie_tot /= rho_tot;
    if (sie_tot_in) sie_tot = sie_tot_in[t];
    s.sie_total = sie_tot;
    s.vfrac_total = vfrac_tot_in ? vfrac_tot_in[t] : 1.0;
    if constexpr (SOLVER == SGX_PTE_FIXED_T) {
      s.Tnorm = Tguess_in[t]; // T_true (PTESolverFixedT ctor, :1498-1510)
    } else if constexpr (SOLVER == SGX_PTE_FIXED_P) {
      s.Pequil = Tguess_in[t]; // P; the ctor hands the base sie_tot = 1, Tnorm = 0 (:1731-1739)
      s.sie_total = 1.0;
      s.Tnorm = 0.0;
    } else if (Tguess_in) {
      s.Tnorm = Tguess_in[t];
    } else {
      s.Tnorm = mix::ApproxTemperatureFromRhoMatU<NM>(eos, rho_tot * sie_tot, s.rho, s.vfrac);
    }
    const SolverStatusDev st = mix::PTESolver(s);
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      rho[base + m] = s.rho[m];
      vfrac[base + m] = s.vfrac[m];
      sie[base + m] = s.sie[m];
      temp[base + m] = s.temp[m];
      press[base + m] = s.press[m];
    }
    if (status) status[t] = mix::pack_status(st, s.niter);
    if (residual) residual[t] = st.residual;
    if (stats) {
      atomicAdd(&s_hist[SGX_PTE_STAT_NCELL], 1u);
      if (st.converged) atomicAdd(&s_hist[SGX_PTE_STAT_CONVERGED], 1u);
      if (st.slow_convergence_detected) atomicAdd(&s_hist[SGX_PTE_STAT_SLOW], 1u);
      if (!st.converged) atomicAdd(&s_hist[SGX_PTE_STAT_FAILED], 1u);
      const int bin = s.niter < 31 ? s.niter : 31;
      atomicAdd(&s_hist[SGX_PTE_STAT_HIST0 + bin], 1u);
    }
  }
  if (stats) {
    __syncthreads();
    for (int k = threadIdx.x; k < SGX_PTE_NSTATS; k += blockDim.x)
      if (s_hist[k]) atomicAdd(&stats[k], (unsigned long long)s_hist[k]);
  }
}

template <int NM, class Acc, int SOLVER>
static int launch_pte_percell(const EosSet &set, const MixParamsDev &prm, double *rho,
                              double *vfrac, double *sie, double *temp, double *press,
                              const double *sie_tot, const double *vfrac_tot,
                              const double *Tguess, uint64_t *status, double *residual,
                              uint64_t *stats, int64_t ncell, cudaStream_t s,
                              double *lambda = nullptr) {
  const LaunchCfg c = stream_cfg(ncell, 128, 8);
  pte_kernel<NM, Acc, SOLVER><<<c.blocks, c.threads, 0, s>>>(
      set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess,
      reinterpret_cast<unsigned long long *>(status), residual,
      reinterpret_cast<unsigned long long *>(stats), ncell, lambda);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

} // namespace sgx
#endif // SGX_PTE_KERNEL_CUH_
