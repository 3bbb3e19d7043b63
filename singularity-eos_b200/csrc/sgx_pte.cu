// PTE mixed-cell solve kernels + C ABI (sgx_pte_solve).  Replaces the user-side cell loop of
// test/test_pte.cpp:177-265 around PTESolverRhoT / PTESolverPT + PTESolver
// (closure/mixed_cell_models.hpp:883-1446, 2188-2291).
//
// Mapping: one thread per mixed cell, grid-stride over cells, 128-thread blocks, the whole
// Newton state in registers / local memory (no per-cell scratch in HBM: the reference's
// 39-double scratch would add 312 B/cell of traffic).  All lanes of a warp execute the same
// material's EOS at the same time (the material loop is uniform), so the FP64 pipe sees
// full warps; divergence only comes from per-cell iteration counts.
// Convergence statistics are accumulated in a shared-memory histogram per block and flushed
// with one atomicAdd per bin per block.
#include "sgx_internal.hpp"

namespace sgx {

template <int NM, class Acc, int SOLVER>
__global__ void __launch_bounds__(128)
    pte_kernel(const __grid_constant__ EosSet set, const __grid_constant__ MixParamsDev prm,
               double *__restrict__ rho, double *__restrict__ vfrac, double *__restrict__ sie,
               double *__restrict__ temp, double *__restrict__ press,
               const double *__restrict__ sie_tot_in, const double *__restrict__ vfrac_tot_in,
               const double *__restrict__ Tguess_in, unsigned long long *__restrict__ status,
               double *__restrict__ residual, unsigned long long *__restrict__ stats,
               const int64_t ncell) {
  __shared__ unsigned int s_hist[SGX_PTE_NSTATS];
  if (stats) {
    for (int k = threadIdx.x; k < SGX_PTE_NSTATS; k += blockDim.x)
      s_hist[k] = 0;
    __syncthreads();
  }
  const Acc eos(set);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = tid; t < ncell; t += nthr) {
    using System =
        typename std::conditional<SOLVER == SGX_PTE_RHOT, mix::RhoT<NM, Acc>, mix::PT<NM, Acc>>::type;
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
    if (Tguess_in) {
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

static MixParamsDev to_dev(const sgx_mix_params *p) {
  sgx_mix_params d;
  if (!p) {
    sgx_mix_params_default(&d);
    p = &d;
  }
  MixParamsDev m;
  m.iterate_t_guess = p->iterate_t_guess;
  m.derivative_eps = p->derivative_eps;
  m.pte_rel_tolerance_p = p->pte_rel_tolerance_p;
  m.pte_rel_tolerance_t = p->pte_rel_tolerance_t;
  m.pte_rel_tolerance_e = p->pte_rel_tolerance_e;
  m.pte_rel_tolerance_v = p->pte_rel_tolerance_v;
  m.pte_abs_tolerance_p = p->pte_abs_tolerance_p;
  m.pte_abs_tolerance_t = p->pte_abs_tolerance_t;
  m.pte_slow_convergence_thresh = p->pte_slow_convergence_thresh;
  m.pte_rel_tolerance_p_sufficient = p->pte_rel_tolerance_p_sufficient;
  m.pte_rel_tolerance_t_sufficient = p->pte_rel_tolerance_t_sufficient;
  m.pte_rel_tolerance_e_sufficient = p->pte_rel_tolerance_e_sufficient;
  m.pte_rel_tolerance_v_sufficient = p->pte_rel_tolerance_v_sufficient;
  m.pte_abs_tolerance_p_sufficient = p->pte_abs_tolerance_p_sufficient;
  m.pte_abs_tolerance_t_sufficient = p->pte_abs_tolerance_t_sufficient;
  m.pte_max_iter_per_mat = p->pte_max_iter_per_mat;
  m.line_search_alpha = p->line_search_alpha;
  m.line_search_max_iter = p->line_search_max_iter;
  m.line_search_fac = p->line_search_fac;
  m.vfrac_safety_fac = p->vfrac_safety_fac;
  m.temperature_limit = p->temperature_limit;
  m.default_tguess = p->default_tguess;
  m.min_dtde = p->min_dtde;
  m.pte_small_step_tries = p->pte_small_step_tries;
  m.pte_small_step_thresh = p->pte_small_step_thresh;
  m.pte_max_dpdv = p->pte_max_dpdv;
  m.pte_reduced_system_exclude_idx = p->pte_reduced_system_exclude_idx;
  return m;
}

template <int NM, class Acc>
static int launch_pte(int solver, const EosSet &set, const MixParamsDev &prm, double *rho,
                      double *vfrac, double *sie, double *temp, double *press,
                      const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                      uint64_t *status, double *residual, uint64_t *stats, int64_t ncell,
                      cudaStream_t s) {
  const LaunchCfg c = stream_cfg(ncell, 128, 8);
  auto *st = reinterpret_cast<unsigned long long *>(status);
  auto *sx = reinterpret_cast<unsigned long long *>(stats);
  if (solver == SGX_PTE_RHOT) {
    pte_kernel<NM, Acc, SGX_PTE_RHOT><<<c.blocks, c.threads, 0, s>>>(
        set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, st, residual, sx,
        ncell);
  } else {
    pte_kernel<NM, Acc, SGX_PTE_PT><<<c.blocks, c.threads, 0, s>>>(
        set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, st, residual, sx,
        ncell);
  }
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

} // namespace sgx

using namespace sgx;

extern "C" {

void sgx_mix_params_default(sgx_mix_params *p) {
  if (!p) return;
  // closure/mixed_cell_models.hpp:43-78
  p->verbose = 0;
  p->iterate_t_guess = 1;
  p->derivative_eps = 3.0e-6;
  p->pte_rel_tolerance_p = 1.0e-6;
  p->pte_rel_tolerance_t = 1.0e-4;
  p->pte_rel_tolerance_e = 1.0e-6;
  p->pte_rel_tolerance_v = 1.0e-6;
  p->pte_abs_tolerance_p = 1.0e-6;
  p->pte_abs_tolerance_t = 1.0e-4;
  p->pte_slow_convergence_thresh = 0.99;
  p->pte_rel_tolerance_p_sufficient = 1.e2 * p->pte_rel_tolerance_p;
  p->pte_rel_tolerance_t_sufficient = 10 * p->pte_rel_tolerance_t;
  p->pte_rel_tolerance_e_sufficient = 1.e2 * p->pte_rel_tolerance_e;
  p->pte_rel_tolerance_v_sufficient = 1.e2 * p->pte_rel_tolerance_v;
  p->pte_abs_tolerance_p_sufficient = 1.e2 * p->pte_abs_tolerance_p;
  p->pte_abs_tolerance_t_sufficient = 10 * p->pte_abs_tolerance_t;
  p->pte_max_iter_per_mat = 128;
  p->line_search_alpha = 1.e-2;
  p->line_search_max_iter = 6;
  p->line_search_fac = 0.5;
  p->vfrac_safety_fac = 0.95;
  p->temperature_limit = 1.0e15;
  p->default_tguess = 300.;
  p->min_dtde = 1.0e-16;
  p->pte_small_step_tries = 2;
  p->pte_small_step_thresh = 1e-16;
  p->pte_max_dpdv = -1e-8;
  p->pte_reduced_system_exclude_idx = -1;
}

int sgx_pte_required_scratch(int solver, int nmat) {
  if (solver == SGX_PTE_RHOT) return nmat * nmat + 4 * nmat + 6 * nmat; // :873-878
  return 2 * 2 + 4 * 2 + 2 * nmat;                                       // :1198-1203
}

int sgx_pte_solve(int solver, const sgx_eos_t *eos, int nmat, int64_t ncell, double *rho,
                  double *vfrac, double *sie, double *temp, double *press,
                  const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                  const sgx_mix_params *params, uint64_t *status, double *residual,
                  uint64_t *stats, sgx_stream_t stream) {
  if (!eos) return fail(SGX_ERR_INVALID, "null EOS array");
  if (solver != SGX_PTE_RHOT && solver != SGX_PTE_PT)
    return fail(SGX_ERR_INVALID, "unknown PTE solver %d", solver);
  if (nmat < 1 || nmat > 4)
    return fail(SGX_ERR_UNSUPPORTED, "nmat=%d: this build instantiates 1..4 materials", nmat);
  if (ncell < 0) return fail(SGX_ERR_INVALID, "negative ncell");
  if (!rho || !vfrac || !sie || !temp || !press)
    return fail(SGX_ERR_INVALID, "null state array");
  EosSet set;
  set.nmat = nmat;
  for (int m = 0; m < nmat; ++m) {
    if (!eos[m]) return fail(SGX_ERR_INVALID, "null EOS handle for material %d", m);
    if (eos[m]->kind > SGX_EOS_DAVIS_PRODUCTS)
      return fail(SGX_ERR_UNSUPPORTED,
                  "PTE with tabulated materials is not built yet (material %d kind %d)", m,
                  eos[m]->kind);
    set.e[m] = eos[m]->analytic;
  }
  cudaError_t err = cudaSetDevice(eos[0]->device);
  if (err != cudaSuccess)
    return fail(SGX_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(err));
  if (ncell == 0) return SGX_OK;
  const MixParamsDev prm = to_dev(params);
  cudaStream_t s = (cudaStream_t)stream;
  const bool is_bench_mix = nmat == 3 && set.e[0].kind == SGX_EOS_GRUNEISEN &&
                            set.e[1].kind == SGX_EOS_DAVIS_REACTANTS &&
                            set.e[2].kind == SGX_EOS_DAVIS_PRODUCTS;
#define SGX_PTE_ARGS                                                                     \
  solver, set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, status,     \
      residual, stats, ncell, s
  if (is_bench_mix) return launch_pte<3, Mix3GrDrDp>(SGX_PTE_ARGS);
  switch (nmat) {
  case 1:
    return launch_pte<1, DynEos>(SGX_PTE_ARGS);
  case 2:
    return launch_pte<2, DynEos>(SGX_PTE_ARGS);
  case 3:
    return launch_pte<3, DynEos>(SGX_PTE_ARGS);
  default:
    return launch_pte<4, DynEos>(SGX_PTE_ARGS);
  }
#undef SGX_PTE_ARGS
}

int sgx_pte_solve_host(int solver, const sgx_eos_t *eos, int nmat, int64_t ncell,
                       double *rho, double *vfrac, double *sie, double *temp, double *press,
                       const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                       const sgx_mix_params *params, uint64_t *status, double *residual) {
  if (!eos || !eos[0]) return fail(SGX_ERR_INVALID, "null EOS array");
  if (ncell <= 0) return ncell == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative ncell");
  sgx_eos_s *e0 = eos[0];
  cudaError_t err = cudaSetDevice(e0->device);
  if (err != cudaSuccess)
    return fail(SGX_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(err));
  const size_t nb = (size_t)ncell * nmat * sizeof(double);
  const size_t cb = (size_t)ncell * sizeof(double);
  int rc = ensure_staging(e0, 0, 5 * nb + 5 * cb);
  if (rc) return rc;
  cudaStream_t s = e0->own_stream;
  char *base = (char *)e0->dscratch;
  double *d[5];
  double *h[5] = {rho, vfrac, sie, temp, press};
  for (int k = 0; k < 5; ++k) {
    d[k] = (double *)(base + k * nb);
    SGX_CUDA_OK(cudaMemcpyAsync(d[k], h[k], nb, cudaMemcpyHostToDevice, s));
  }
  double *d_sie_tot = (double *)(base + 5 * nb);
  double *d_vf_tot = d_sie_tot + ncell, *d_Tg = d_vf_tot + ncell, *d_res = d_Tg + ncell;
  uint64_t *d_status = (uint64_t *)(d_res + ncell);
  if (sie_tot) SGX_CUDA_OK(cudaMemcpyAsync(d_sie_tot, sie_tot, cb, cudaMemcpyHostToDevice, s));
  if (vfrac_tot) SGX_CUDA_OK(cudaMemcpyAsync(d_vf_tot, vfrac_tot, cb, cudaMemcpyHostToDevice, s));
  if (Tguess) SGX_CUDA_OK(cudaMemcpyAsync(d_Tg, Tguess, cb, cudaMemcpyHostToDevice, s));
  rc = sgx_pte_solve(solver, eos, nmat, ncell, d[0], d[1], d[2], d[3], d[4],
                     sie_tot ? d_sie_tot : nullptr, vfrac_tot ? d_vf_tot : nullptr,
                     Tguess ? d_Tg : nullptr, params, status ? d_status : nullptr,
                     residual ? d_res : nullptr, nullptr, s);
  if (rc) return rc;
  for (int k = 0; k < 5; ++k)
    SGX_CUDA_OK(cudaMemcpyAsync(h[k], d[k], nb, cudaMemcpyDeviceToHost, s));
  if (status) SGX_CUDA_OK(cudaMemcpyAsync(status, d_status, cb, cudaMemcpyDeviceToHost, s));
  if (residual) SGX_CUDA_OK(cudaMemcpyAsync(residual, d_res, cb, cudaMemcpyDeviceToHost, s));
  SGX_CUDA_OK(cudaStreamSynchronize(s));
  return SGX_OK;
}

} // extern "C"
