// sgx/closure.hpp -- header-only C++17 host side of the batch PTE mixed-cell closure.
//
// The reference has no batch PTE entry point: a host code constructs one
//   PTESolverRhoT<EOSIndexer,RealIndexer,LambdaIndexer>(nmat, eos, vfrac_tot, sie_tot, rho,
//       vfrac, sie, temp, press, lambda, scratch, Tnorm, params)
// per cell inside its own parallel loop and calls `PTESolver(system)`
// (closure/mixed_cell_models.hpp:922-935, 1245-1253, 2188-2291; the loop is user code,
// test/test_pte.cpp:177-265).  Here that loop IS the kernel: one call solves all cells.
// Argument names and order follow the per-cell constructor; arrays are [ncell][nmat]
// row-major device pointers, all in/out (doc/sphinx/src/using-closures.rst:1044-1051).
#ifndef SGX_CLOSURE_HPP_
#define SGX_CLOSURE_HPP_

#include <cstdint>
#include <vector>

#include "eos.hpp"

namespace sgx {

// closure/mixed_cell_models.hpp:43-78: same field names, same defaults
struct MixParams : sgx_mix_params {
  MixParams() { sgx_mix_params_default(this); }
};

// closure/mixed_cell_models.hpp:80-87, unpacked from the per-cell status word
struct SolverStatus {
  bool converged = false;
  bool slow_convergence_detected = false;
  std::size_t max_niter = 0;
  std::size_t max_line_niter = 0;
  std::size_t small_step_iters = 0;
  std::size_t niter = 0; // System::Niter() after the solve (test/test_pte.cpp:263)
  static SolverStatus unpack(std::uint64_t w) {
    SolverStatus s;
    s.converged = SGX_PTE_CONVERGED(w) != 0;
    s.slow_convergence_detected = SGX_PTE_SLOW(w) != 0;
    s.niter = SGX_PTE_NITER(w);
    s.max_line_niter = SGX_PTE_MAX_LINE_NITER(w);
    s.small_step_iters = SGX_PTE_SMALL_STEP_ITERS(w);
    s.max_niter = SGX_PTE_MAX_NITER(w);
    return s;
  }
};

// RequiredScratch (:873-878, 1198-1203): kept for source compatibility; the batch solver
// holds its scratch on chip, so callers need not allocate any.
inline int PTESolverRhoTRequiredScratch(int nmat) {
  return sgx_pte_required_scratch(SGX_PTE_RHOT, nmat);
}
inline int PTESolverPTRequiredScratch(int nmat) {
  return sgx_pte_required_scratch(SGX_PTE_PT, nmat);
}

namespace detail {
inline void pte(int solver, const std::vector<EosBase> &eos, std::int64_t ncell, Real *rho,
                Real *vfrac, Real *sie, Real *temp, Real *press, const Real *sie_tot,
                const Real *vfrac_tot, const Real *Tguess, const MixParams &params,
                std::uint64_t *status, Real *residual, std::uint64_t *stats,
                sgx_stream_t stream) {
  std::vector<sgx_eos_t> h;
  h.reserve(eos.size());
  for (const EosBase &e : eos)
    h.push_back(e.handle());
  check(sgx_pte_solve(solver, h.data(), static_cast<int>(h.size()), ncell, rho, vfrac, sie,
                      temp, press, sie_tot, vfrac_tot, Tguess, &params, status, residual, stats,
                      stream));
}
} // namespace detail

// Solve `ncell` mixed cells with PTESolverRhoT + PTESolver.
//   sie_tot   [ncell] or nullptr => sum(rho vfrac sie) / sum(rho vfrac)  (test_pte.cpp:188-195)
//   vfrac_tot [ncell] or nullptr => 1
//   Tguess    [ncell] or nullptr => ApproxTemperatureFromRhoMatU (:790-847)
//   status    [ncell] packed SolverStatus words (SolverStatus::unpack), may be nullptr
//   stats     uint64[SGX_PTE_NSTATS] accumulated convergence statistics, may be nullptr
inline void PTESolverRhoT(const std::vector<EosBase> &eos, std::int64_t ncell, Real *rho,
                          Real *vfrac, Real *sie, Real *temp, Real *press,
                          const MixParams &params = MixParams(),
                          std::uint64_t *status = nullptr, const Real *sie_tot = nullptr,
                          const Real *vfrac_tot = nullptr, const Real *Tguess = nullptr,
                          Real *residual = nullptr, std::uint64_t *stats = nullptr,
                          sgx_stream_t stream = nullptr) {
  detail::pte(SGX_PTE_RHOT, eos, ncell, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot,
              Tguess, params, status, residual, stats, stream);
}
inline void PTESolverPT(const std::vector<EosBase> &eos, std::int64_t ncell, Real *rho,
                        Real *vfrac, Real *sie, Real *temp, Real *press,
                        const MixParams &params = MixParams(), std::uint64_t *status = nullptr,
                        const Real *sie_tot = nullptr, const Real *vfrac_tot = nullptr,
                        const Real *Tguess = nullptr, Real *residual = nullptr,
                        std::uint64_t *stats = nullptr, sgx_stream_t stream = nullptr) {
  detail::pte(SGX_PTE_PT, eos, ncell, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess,
              params, status, residual, stats, stream);
}

// PTESolverRhoU + PTESolver (closure/mixed_cell_models.hpp:1934-2183); same arguments as
// PTESolverRhoT
inline void PTESolverRhoU(const std::vector<EosBase> &eos, std::int64_t ncell, Real *rho,
                          Real *vfrac, Real *sie, Real *temp, Real *press,
                          const MixParams &params = MixParams(),
                          std::uint64_t *status = nullptr, const Real *sie_tot = nullptr,
                          const Real *vfrac_tot = nullptr, const Real *Tguess = nullptr,
                          Real *residual = nullptr, std::uint64_t *stats = nullptr,
                          sgx_stream_t stream = nullptr) {
  detail::pte(SGX_PTE_RHOU, eos, ncell, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot,
              Tguess, params, status, residual, stats, stream);
}

// PTESolverFixedT + PTESolver (closure/mixed_cell_models.hpp:1462-1672): the temperature of
// every cell is given (T_true [ncell]); unknowns are the volume fractions.
inline void PTESolverFixedT(const std::vector<EosBase> &eos, std::int64_t ncell,
                            const Real *T_true, Real *rho, Real *vfrac, Real *sie, Real *temp,
                            Real *press, const MixParams &params = MixParams(),
                            std::uint64_t *status = nullptr, const Real *vfrac_tot = nullptr,
                            Real *residual = nullptr, std::uint64_t *stats = nullptr,
                            sgx_stream_t stream = nullptr) {
  detail::pte(SGX_PTE_FIXED_T, eos, ncell, rho, vfrac, sie, temp, press, nullptr, vfrac_tot,
              T_true, params, status, residual, stats, stream);
}

// PTESolverFixedP + PTESolver (closure/mixed_cell_models.hpp:1674-1932): the pressure of every
// cell is given (P_true [ncell]); unknowns are the volume fractions and the temperature.
// params.iterate_t_guess must be 0 (include/sgx_b200.h, SGX_PTE_FIXED_P).
inline void PTESolverFixedP(const std::vector<EosBase> &eos, std::int64_t ncell,
                            const Real *P_true, Real *rho, Real *vfrac, Real *sie, Real *temp,
                            Real *press, const MixParams &params,
                            std::uint64_t *status = nullptr, const Real *vfrac_tot = nullptr,
                            Real *residual = nullptr, std::uint64_t *stats = nullptr,
                            sgx_stream_t stream = nullptr) {
  detail::pte(SGX_PTE_FIXED_P, eos, ncell, rho, vfrac, sie, temp, press, nullptr, vfrac_tot,
              P_true, params, status, residual, stats, stream);
}

// PTESolverRhoT over a mixture that holds tabulated materials, with the per-material lambda
// caches of the reference's CacheAccessor (:307-324): lambda = [ncell][nmat][3] or nullptr.
inline void PTESolverRhoTCached(const std::vector<EosBase> &eos, std::int64_t ncell, Real *rho,
                                Real *vfrac, Real *sie, Real *temp, Real *press, Real *lambda,
                                const MixParams &params = MixParams(),
                                std::uint64_t *status = nullptr, const Real *sie_tot = nullptr,
                                const Real *vfrac_tot = nullptr, const Real *Tguess = nullptr,
                                Real *residual = nullptr, sgx_stream_t stream = nullptr) {
  std::vector<sgx_eos_t> h;
  for (const EosBase &e : eos)
    h.push_back(e.handle());
  check(sgx_pte_solve_cached(SGX_PTE_RHOT, h.data(), static_cast<int>(h.size()), ncell, rho,
                             vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, &params,
                             status, residual, nullptr, lambda, stream));
}

// The three materials of test/pte_test_3mat_analytic.hpp:38-50
inline std::vector<EosBase> set_eos_3mat_analytic() {
  return {Gruneisen(394000.0, 1.489, 0.0, 0.0, 2.02, 0.47, 8.93, 297.0, 1.0e6, 0.383e7),
          DavisReactants(1.890, 4.115e10, 1.0e6, 297.0, 1.8e5, 4.6, 0.34, 0.56, 0.0, 0.4265,
                         0.001074e10),
          DavisProducts(0.798311, 0.58, 1.35, 2.66182, 0.75419, 3.2e10, 0.001072e10)};
}

} // namespace sgx
#endif // SGX_CLOSURE_HPP_
