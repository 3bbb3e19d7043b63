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
#include "sgx_pte_kernel.cuh"

namespace sgx {

MixParamsDev mix_params_to_dev(const sgx_mix_params *p) {
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
  if (solver == SGX_PTE_FIXED_T) return nmat * nmat + 4 * nmat + 4 * nmat; // :1450-1456
  if (solver == SGX_PTE_FIXED_P)                                          // :1674-1680
    return (nmat + 1) * (nmat + 1) + 4 * (nmat + 1) + 2 * nmat + 3 * nmat;
  if (solver == SGX_PTE_RHOU) return 4 * nmat * nmat + 8 * nmat + 8 * nmat; // :1924-1929
  return 2 * 2 + 4 * 2 + 2 * nmat;                                       // :1198-1203
}

int sgx_pte_solve(int solver, const sgx_eos_t *eos, int nmat, int64_t ncell, double *rho,
                  double *vfrac, double *sie, double *temp, double *press,
                  const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                  const sgx_mix_params *params, uint64_t *status, double *residual,
                  uint64_t *stats, sgx_stream_t stream) {
  return sgx_pte_solve_cached(solver, eos, nmat, ncell, rho, vfrac, sie, temp, press, sie_tot,
                              vfrac_tot, Tguess, params, status, residual, stats, nullptr,
                              stream);
}

int sgx_pte_solve_cached(int solver, const sgx_eos_t *eos, int nmat, int64_t ncell, double *rho,
                         double *vfrac, double *sie, double *temp, double *press,
                         const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                         const sgx_mix_params *params, uint64_t *status, double *residual,
                         uint64_t *stats, double *lambda, sgx_stream_t stream) {
  DeviceGuard guard;
  if (!eos) return fail(SGX_ERR_INVALID, "null EOS array");
  if (solver != SGX_PTE_RHOT && solver != SGX_PTE_PT && solver != SGX_PTE_FIXED_T &&
      solver != SGX_PTE_RHOU && solver != SGX_PTE_FIXED_P)
    return fail(SGX_ERR_INVALID, "unknown PTE solver %d", solver);
  if (solver == SGX_PTE_RHOU && (nmat < 2 || nmat > 4))
    return fail(SGX_ERR_UNSUPPORTED, "PTESolverRhoU: nmat=%d, this build instantiates 2..4",
                nmat);
  if (solver == SGX_PTE_FIXED_T && !Tguess)
    return fail(SGX_ERR_INVALID, "PTESolverFixedT needs the temperature array (Tguess)");
  if (solver == SGX_PTE_FIXED_P && !Tguess)
    return fail(SGX_ERR_INVALID, "PTESolverFixedP needs the pressure array (Tguess)");
  if (solver == SGX_PTE_FIXED_P && (!params || params->iterate_t_guess))
    return fail(SGX_ERR_UNSUPPORTED,
                "PTESolverFixedP with MixParams::iterate_t_guess: the reference's GetTguess reads "
                "uscale / utotal_scale before FixedP::Init sets them "
                "(mixed_cell_models.hpp:536-538 vs 1747-1755), its answer is undefined; pass "
                "iterate_t_guess = 0");
  if (nmat < 1 || nmat > 4)
    return fail(SGX_ERR_UNSUPPORTED, "nmat=%d: this build instantiates 1..4 materials", nmat);
  if (ncell < 0) return fail(SGX_ERR_INVALID, "negative ncell");
  if (!rho || !vfrac || !sie || !temp || !press)
    return fail(SGX_ERR_INVALID, "null state array");
  EosSet set;
  set.nmat = nmat;
  bool tabulated = false;
  for (int m = 0; m < nmat; ++m) {
    if (!eos[m]) return fail(SGX_ERR_INVALID, "null EOS handle for material %d", m);
    if (eos[m]->device != eos[0]->device)
      return fail(SGX_ERR_INVALID, "material %d lives on another device", m);
    if (needs_tab_accessor(eos[m])) tabulated = true;
    if (!set_material(eos[m], set, m))
      return fail(SGX_ERR_UNSUPPORTED, "PTE: material %d is EOS kind %d (Helmholtz mixtures are "
                  "not built)", m, eos[m]->kind);
  }
  if (lambda && !tabulated) lambda = nullptr; // the analytic models take no lambda
  bool any_lambda_user = false;
  for (int m = 0; m < nmat; ++m)
    any_lambda_user = any_lambda_user || set.e[m].kind >= SGX_EOS_SPINER_RHOT;
  if (!any_lambda_user) lambda = nullptr; // modified analytic materials only
  cudaError_t err = cudaSetDevice(eos[0]->device);
  if (err != cudaSuccess)
    return fail(SGX_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(err));
  if (ncell == 0) return SGX_OK;
  const MixParamsDev prm = mix_params_to_dev(params);
  cudaStream_t s = (cudaStream_t)stream;
  const bool is_bench_mix = nmat == 3 && set.e[0].kind == SGX_EOS_GRUNEISEN &&
                            set.e[1].kind == SGX_EOS_DAVIS_REACTANTS &&
                            set.e[2].kind == SGX_EOS_DAVIS_PRODUCTS;
#define SGX_PTE_TAIL                                                                     \
  prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, status, residual, stats,  \
      ncell, s
#define SGX_PTE_ARGS                                                                     \
  set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, status, residual,   \
      stats, ncell, s
  if (tabulated)
    return launch_pte_tab(solver, set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot,
                          Tguess, status, residual, stats, ncell, s, lambda);
  if (solver == SGX_PTE_RHOT) return launch_pte_rhot(set, is_bench_mix, prm, rho, vfrac, sie, temp, press, sie_tot,
                           vfrac_tot, Tguess, status, residual, stats, ncell, s);
  if (solver == SGX_PTE_RHOU) return launch_pte_rhou(set, is_bench_mix, SGX_PTE_TAIL);
  if (solver == SGX_PTE_FIXED_T) return launch_pte_fixedt(set, is_bench_mix, SGX_PTE_TAIL);
  if (solver == SGX_PTE_FIXED_P) return launch_pte_fixedp(set, is_bench_mix, SGX_PTE_TAIL);
  if (is_bench_mix) return launch_pte_percell<3, Mix3GrDrDp, SGX_PTE_PT>(SGX_PTE_ARGS);
  switch (nmat) {
  case 1:
    return launch_pte_percell<1, DynEos, SGX_PTE_PT>(SGX_PTE_ARGS);
  case 2:
    return launch_pte_percell<2, DynEos, SGX_PTE_PT>(SGX_PTE_ARGS);
  case 3:
    return launch_pte_percell<3, DynEos, SGX_PTE_PT>(SGX_PTE_ARGS);
  default:
    return launch_pte_percell<4, DynEos, SGX_PTE_PT>(SGX_PTE_ARGS);
  }
#undef SGX_PTE_ARGS
#undef SGX_PTE_TAIL
}

int sgx_pte_solve_host(int solver, const sgx_eos_t *eos, int nmat, int64_t ncell,
                       double *rho, double *vfrac, double *sie, double *temp, double *press,
                       const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                       const sgx_mix_params *params, uint64_t *status, double *residual) {
  if (!eos || !eos[0]) return fail(SGX_ERR_INVALID, "null EOS array");
  if (ncell <= 0) return ncell == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative ncell");
  if (nmat < 1 || nmat > 4)
    return fail(SGX_ERR_UNSUPPORTED, "nmat=%d: this build instantiates 1..4 materials", nmat);
  sgx_eos_s *e0 = eos[0];
  DeviceGuard guard;
  cudaError_t err = cudaSetDevice(e0->device);
  if (err != cudaSuccess)
    return fail(SGX_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(err));
  // Chunks of cells flow through kSlots device slots, each on its own stream
  // (H2D -> solve -> D2H), so both PCIe directions overlap the solver.
  constexpr int kSlots = 3;
  constexpr int64_t kChunk = int64_t(1) << 20;
  const int64_t chunk = ncell < kChunk ? ncell : kChunk;
  const size_t nb_max = (size_t)chunk * nmat * sizeof(double);
  const size_t cb_max = (size_t)chunk * sizeof(double);
  const size_t slot_bytes = (5 * nb_max + 5 * cb_max + 255) & ~size_t(255);
  int rc = ensure_staging(e0, 0, kSlots * slot_bytes);
  if (rc) return rc;
  rc = ensure_host_streams(e0);
  if (rc) return rc;
  double *h[5] = {rho, vfrac, sie, temp, press};
  // any failure drains the slot streams before returning (no copy into the caller's buffers
  // may still be pending when the call comes back)
#define SGX_PIPE_TRY(expr)                                                                 \
  do {                                                                                     \
    cudaError_t err__ = (expr);                                                            \
    if (err__ != cudaSuccess) {                                                            \
      rc = fail(SGX_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(err__));                 \
      goto drain;                                                                          \
    }                                                                                      \
  } while (0)
  {
    const size_t tot = (size_t)ncell * nmat * sizeof(double), tc = (size_t)ncell * sizeof(double);
    HostPin pins[5] = {HostPin(h[0], tot), HostPin(h[1], tot), HostPin(h[2], tot),
                       HostPin(h[3], tot), HostPin(h[4], tot)};
    HostPin p_st(sie_tot, tc), p_vt(vfrac_tot, tc), p_tg(Tguess, tc), p_status(status, tc),
        p_res(residual, tc);
    int slot = 0;
    for (int64_t c0 = 0; c0 < ncell; c0 += chunk, slot = (slot + 1) % kSlots) {
      const int64_t n = (ncell - c0) < chunk ? (ncell - c0) : chunk;
      const size_t nb = (size_t)n * nmat * sizeof(double);
      const size_t cb = (size_t)n * sizeof(double);
      cudaStream_t s = e0->host_streams[slot];
      char *base = (char *)e0->dscratch + slot * slot_bytes;
      double *d[5];
      for (int k = 0; k < 5; ++k) {
        d[k] = (double *)(base + k * nb_max);
        SGX_PIPE_TRY(cudaMemcpyAsync(d[k], h[k] + c0 * nmat, nb, cudaMemcpyHostToDevice, s));
      }
      double *d_sie_tot = (double *)(base + 5 * nb_max);
      double *d_vf_tot = d_sie_tot + chunk, *d_Tg = d_vf_tot + chunk, *d_res = d_Tg + chunk;
      uint64_t *d_status = (uint64_t *)(d_res + chunk);
      if (sie_tot)
        SGX_PIPE_TRY(cudaMemcpyAsync(d_sie_tot, sie_tot + c0, cb, cudaMemcpyHostToDevice, s));
      if (vfrac_tot)
        SGX_PIPE_TRY(cudaMemcpyAsync(d_vf_tot, vfrac_tot + c0, cb, cudaMemcpyHostToDevice, s));
      if (Tguess) SGX_PIPE_TRY(cudaMemcpyAsync(d_Tg, Tguess + c0, cb, cudaMemcpyHostToDevice, s));
      rc = sgx_pte_solve(solver, eos, nmat, n, d[0], d[1], d[2], d[3], d[4],
                         sie_tot ? d_sie_tot : nullptr, vfrac_tot ? d_vf_tot : nullptr,
                         Tguess ? d_Tg : nullptr, params, status ? d_status : nullptr,
                         residual ? d_res : nullptr, nullptr, s);
      if (rc) goto drain;
      for (int k = 0; k < 5; ++k)
        SGX_PIPE_TRY(cudaMemcpyAsync(h[k] + c0 * nmat, d[k], nb, cudaMemcpyDeviceToHost, s));
      if (status)
        SGX_PIPE_TRY(cudaMemcpyAsync(status + c0, d_status, cb, cudaMemcpyDeviceToHost, s));
      if (residual)
        SGX_PIPE_TRY(cudaMemcpyAsync(residual + c0, d_res, cb, cudaMemcpyDeviceToHost, s));
    }
  drain:
    for (int k = 0; k < kSlots; ++k) {
      cudaError_t e2 = cudaStreamSynchronize(e0->host_streams[k]);
      if (e2 != cudaSuccess && rc == SGX_OK)
        rc = fail(SGX_ERR_CUDA, "cudaStreamSynchronize: %s", cudaGetErrorString(e2));
    }
  }
#undef SGX_PIPE_TRY
  return rc;
}

} // extern "C"
