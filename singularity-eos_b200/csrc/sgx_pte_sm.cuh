// PTESolverRhoT as a per-lane state machine ("lane refill" kernel).
//
// Why: the straight transcription in sgx_pte.cuh (one thread walks one cell through
// Init / Jacobian / TestUpdate) inlines the materials' EOS code at ~12 call sites -- 484 KB of
// SASS, instruction-fetch bound (profiles/r1a_pte_details.txt: 13.7 `no_instruction` stall
// cycles per issue) -- and a warp waits for its slowest cell (14.8 of 32 lanes active).
//
// Here every lane still owns one cell and executes exactly the reference's sequence of
// floating-point operations for it (closure/mixed_cell_models.hpp:329-1193, 2188-2291), but the
// solver is cut at each point where it needs the materials' EOS:
//
//   * all EOS work of all phases goes through ONE warp-converged call site, eval_round():
//     "for each material m: e = e(rho_q[m], T_q), then P(rho_q[m], e) or cv";
//   * between two rounds a lane runs its own phase handler (cheap scalar algebra: Jacobian
//     assembly, the 3x3 solve, ScaleDx, line-search bookkeeping);
//   * a lane whose cell is finished stores it and takes the next cell of its warp's contiguous
//     cell range at once, so iteration-count differences between cells no longer idle lanes.
//
// The solve is split in two kernels:
//   pte_rhot_init_kernel  ApproxTemperatureFromRhoMatU (:790-847) and GetTguess (:500-581).
//                         Every cell takes the same path here (2 + 10 bisection steps, one
//                         Newton step), so plain thread-per-cell code is already converged;
//                         the 12 energies of the bisection share one density, so the
//                         density-only EOS terms are computed once (RhoPart).  Leaves
//                         Tguess, vfrac[], rho[] in a 2*NM+1 doubles/cell scratch record.
//   pte_rhot_sm_kernel    everything from InitBase's first EOS call on, as the state machine.
// Evaluation points of the state machine (reference line numbers):
//   INIT_EP              InitBase                      :627-662   (e and P)
//   J                    PTESolverRhoT::Jacobian       :1042-1105 (dv- and dT-perturbed e, P)
//   TU                   PTESolverRhoT::TestUpdate     :1155-1187 (e, then P from sie=u/rhobar)
#ifndef SGX_PTE_SM_CUH_
#define SGX_PTE_SM_CUH_

#include "sgx_pte.cuh"

namespace sgx {
namespace mix {

#ifdef __CUDACC__

enum SmPhase : int {
  PH_IDLE = 0,
  // phases entered with eval_round() results ready
  PH_INIT_EP,
  PH_J,
  PH_TU,
  // pure-compute phases (run back to back inside advance())
  PH_NEWTON_TOP,
  PH_AFTER_LS,
  PH_FINISH
};

enum SmMode : int { MODE_EP = 1, MODE_TU = 3, MODE_J = 4 };

// The one place the materials' EOS code is instantiated.  Evaluation is staged (RhoPart,
// sgx_analytic.cuh): density-only curves first, then the temperature / energy terms.
//   MODE_EP  e, then GetPressureFromPreferred(m, rho, T, e, false)          (:583-597);
//            keeps the density-only terms in `cache` (InitBase: rho is the state density)
//   MODE_TU  TestUpdate: sie = ratio(rhobar*e, rhobar), P from that sie     (:1175-1184);
//            keeps the density-only terms in `cache`
//   MODE_J   Jacobian: (e,P) at the dv-perturbed densities rq and T = Tq, and (e2,P2) at the
//            state densities rho_c -- whose density-only terms are in `cache` since the last
//            TestUpdate / InitBase -- and T = Tq2                          (:1042-1075)
// GetPressureFromPreferred is the accessor's P_pref_from_part: P(rho, e) for the analytic
// models (sie among their preferred inputs), P(rho, T) for the tabulated ones (sgx_pte_tab.cuh).
template <int NM, class Eos>
__device__ __forceinline__ void eval_round(const Eos &eos, const int mode, const double Tq,
                                           const double Tq2, const double (&rq)[NM],
                                           const double (&rho_c)[NM], const double (&rb)[NM],
                                           RhoPart (&cache)[NM], double (&eq)[NM],
                                           double (&pq)[NM], double (&eq2)[NM],
                                           double (&pq2)[NM]) {
#pragma unroll
  for (int m = 0; m < NM; ++m) {
    RhoPart part;
    eos.rho_part(m, rq[m], part);
    const double e = eos.e_from_part(m, part, rq[m], Tq);
    eq[m] = e;
    double sie_for_p = e;
    if (mode == MODE_TU) sie_for_p = ratio(rb[m] * e, rb[m]);
    pq[m] = eos.P_pref_from_part(m, part, rq[m], sie_for_p, Tq);
    if (mode == MODE_J) {
      const double e2 = eos.e_from_part(m, cache[m], rho_c[m], Tq2);
      eq2[m] = e2;
      pq2[m] = eos.P_pref_from_part(m, cache[m], rho_c[m], e2, Tq2);
    } else {
      cache[m] = part;
    }
  }
}

// ApproxTemperatureFromRhoMatU (:790-847) with the density-only terms hoisted: the same
// operations as mix::ApproxTemperatureFromRhoMatU (sgx_pte.cuh), in the same order.
template <int NM, class Eos>
__device__ inline double ApproxTemperatureStaged(const Eos &eos, const double u_tot,
                                                 const double *rho, const double *vfrac) {
  constexpr double minimum_temperature = 1.e-9;
  constexpr double maximum_temperature = 1.e9;
  RhoPart part[NM];
#pragma unroll
  for (int m = 0; m < NM; m++)
    eos.rho_part(m, rho[m], part[m]);
  auto ufunc = [&](const double T) {
    double usum = 0.0;
#pragma unroll
    for (int m = 0; m < NM; m++) {
      usum += rho[m] * vfrac[m] * eos.e_from_part(m, part[m], rho[m], T);
    }
    return usum;
  };
  double ulo = ufunc(minimum_temperature);
  if (u_tot < ulo) return minimum_temperature;
  double uhi = ufunc(maximum_temperature);
  if (u_tot > uhi) return maximum_temperature;
  double lTlo = fastmath::lg(minimum_temperature);
  double lThi = fastmath::lg(maximum_temperature);
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

// One lane's cell: the reference solver object plus the program counter and the few locals
// of PTESolver() / GetTguess() / ApproxTemperatureFromRhoMatU() that live across a cut.
template <int NM, class Eos>
struct RhoTLane {
  RhoT<NM, Eos> s;
  SolverStatusDev st;
  int phase = PH_IDLE;
  int mode = MODE_EP;
  double Tq = 0.0, Tq2 = 0.0;
  double rq[NM];
  RhoPart cache[NM]; // density-only EOS terms at the state densities s.rho
  // PTESolver locals
  double err, err_old, gradfdx, scale, err_first;
  bool close_enough;
  int tu_kind; // 0: full step, 1: half step, 2: line-search candidates
  unsigned long long line_iter, pte_max_iter;

  __device__ RhoTLane(const Eos &e, const MixParamsDev &p) : s(e, p) {}

  // ---- requests ----------------------------------------------------------------------
  SGX_D void request(const int mode_, const double T, const int next_phase) {
    mode = mode_;
    Tq = T;
    phase = next_phase;
  }
  SGX_D void request_state_rho(const int mode_, const double T, const int next_phase) {
#pragma unroll
    for (int m = 0; m < NM; ++m)
      rq[m] = s.rho[m];
    request(mode_, T, next_phase);
  }
  // first half of TestUpdate (:1155-1174): new T and volume fractions, densities to evaluate
  SGX_D void prepare_test_update(const double scale_, const bool cache_state) {
    if (cache_state) {
      s.Ttemp = s.Tequil;
#pragma unroll
      for (int m = 0; m < NM; ++m)
        s.vtemp[m] = s.vfrac[m];
    }
    s.Tequil = s.Ttemp + scale_ * s.dx[0];
    int idx = 1;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      if (s.ms == m) continue;
      s.vfrac[m] = s.vtemp[m] + scale_ * s.pick(s.dx, idx++);
    }
    const double vsum_others = sum_neumaier(s.vfrac, NM, 0, s.ms);
#pragma unroll
    for (int m = 0; m < NM; ++m)
      if (m == s.ms) s.vfrac[m] = 1 - vsum_others;
#pragma unroll
    for (int m = 0; m < NM; ++m)
      s.rho[m] = ratio(s.rhobar[m], s.vfrac[m]);
    request_state_rho(MODE_TU, s.Tnorm * s.Tequil, PH_TU);
  }

  // ---- cell entry ------------------------------------------------------------------------
  // The kernel has loaded the caller's state into s.rho .. s.press.  The init kernel
  // (pte_rhot_init_kernel below) already ran ApproxTemperatureFromRhoMatU and GetTguess for
  // this cell and left Tguess and the volume fractions / densities GetTguess settled on;
  // here: the cell totals (test/test_pte.cpp:188-195) and the scalings of InitBase (:627-640).
  SGX_D void begin(const double sie_tot_in, const bool have_sie_tot, const double vfrac_tot,
                   const double Tguess, const double (&vfrac_init)[NM],
                   const double (&rho_init)[NM]) {
    st = SolverStatusDev();
    s.niter = 0;
    double sie_tot = 0.0, rho_tot = 0.0;
#pragma unroll
    for (int m = 0; m < NM; m++) {
      const double rhobar = s.rho[m] * s.vfrac[m];
      rho_tot += rhobar;
      sie_tot += rhobar * s.sie[m];
    }
    sie_tot /= rho_tot;
    if (have_sie_tot) sie_tot = sie_tot_in;
    s.sie_total = sie_tot;
    s.vfrac_total = vfrac_tot;
    s.InitRhoBarandRho();
    const double utotal = s.rho_total * s.sie_total;
    s.uscale = stdmax(fabs(utotal), 1.0e-14);
    s.utotal_scale = ratio(utotal, s.uscale);
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      s.vfrac[m] = vfrac_init[m];
      s.rho[m] = rho_init[m];
    }
    s.Tnorm = Tguess;
    request_state_rho(MODE_EP, Tguess, PH_INIT_EP);
  }
  SGX_D bool wants_eval() const { return phase != PH_IDLE && phase < PH_NEWTON_TOP; }

  // ---- the solver, cut at its evaluation points -----------------------------------------
  // Called after eval_round() (eq/pq hold the results for a lane in a result phase); runs
  // this lane's handlers until the next evaluation request or the end of the cell.
  __device__ __forceinline__ void advance(const double (&eq)[NM], const double (&pq)[NM],
                                          const double (&eq2)[NM],
                                          const double (&pq2)[NM]) {
    const MixParamsDev &params = s.params_;
    while (true) {
      switch (phase) {
      // ===== rest of InitBase + PTESolverRhoT::Init (:947-984) =========================
      case PH_INIT_EP: {
#pragma unroll
        for (int m = 0; m < NM; m++) {
          s.temp[m] = 1.0;
          s.sie[m] = eq[m];
          if ((s.eos.preferred(m) & TQ_PRESSURE) && !bad_value(s.press[m])) {
            s.press[m] = ratio(s.press[m], s.uscale); // input pressure is the guess (:646-649)
          } else {
            s.press[m] = ratio(pq[m], s.uscale);
          }
        }
#pragma unroll
        for (int m = 0; m < NM; ++m)
          s.u[m] = s.sie[m] * ratio(s.rhobar[m], s.uscale);
        if (params.pte_reduced_system_exclude_idx >= 0) {
          s.ms = (int)params.pte_reduced_system_exclude_idx;
        } else {
          s.ms = 0;
          double xmax = 0;
#pragma unroll
          for (int m = 0; m < NM; ++m) {
            const double x = ratio(s.rhobar[m], s.rho_total);
            if (x > xmax) {
              xmax = x;
              s.ms = m;
            }
          }
        }
        s.Residual();
        s.Tequil = s.temp[0];
        err = s.ResidualNorm();
        // PTESolver prologue (:2188-2203)
        st.converged = false;
        close_enough = false;
        pte_max_iter = (unsigned long long)NM * params.pte_max_iter_per_mat;
        s.niter = 0;
        phase = PH_NEWTON_TOP;
      } break;
      // ===== PTESolver loop head (:2204-2211) ==========================================
      case PH_NEWTON_TOP: {
        if ((unsigned long long)s.niter >= pte_max_iter) {
          phase = PH_FINISH;
          break;
        }
        st.max_niter =
            st.max_niter > (unsigned long long)s.niter ? st.max_niter : s.niter;
        s.CheckPTE(st.converged, close_enough);
        if (st.converged) {
          phase = PH_FINISH;
          break;
        }
        // Jacobian, volume-fraction perturbation (:1046-1054)
#pragma unroll
        for (int m = 0; m < NM; m++) {
          const double dv =
              (s.vfrac[m] < 0.5 ? 1.0 : -1.0) * s.vfrac[m] * params.derivative_eps;
          const double vf_pert = s.vfrac[m] + dv;
          rq[m] = ratio(s.rhobar[m], vf_pert);
        }
        Tq2 = s.Tnorm * (s.Tequil + s.Tequil * params.derivative_eps);
        request(MODE_J, s.Tnorm * s.Tequil, PH_J);
      } break;
      case PH_J: { // :1055-1105, then Solve / ScaleDx / first TestUpdate (:2212-2226)
        const double dT = s.Tequil * params.derivative_eps;
        double dedT_sum = 0.0;
#pragma unroll
        for (int m = 0; m < NM; m++) {
          const double dv =
              (s.vfrac[m] < 0.5 ? 1.0 : -1.0) * s.vfrac[m] * params.derivative_eps;
          double p_pert = ratio(pq[m], s.uscale);
          s.dpdv[m] = stdmin(ratio((p_pert - s.press[m]), dv), -fabs(params.pte_max_dpdv));
          s.dedv[m] = ratio(s.rhobar[m] * ratio(eq[m], s.uscale) - s.u[m], dv);
          p_pert = ratio(pq2[m], s.uscale);
          s.dpdT[m] = ratio((p_pert - s.press[m]), dT);
          dedT_sum += ratio(s.rhobar[m] * ratio(eq2[m], s.uscale) - s.u[m], dT);
        }
        s.AssembleJacobian(dedT_sum);
        const bool success = s.Solve();
        if (!success) {
          st.converged = false;
          phase = PH_FINISH;
          break;
        }
        scale = stdmin(1.0, s.ScaleDx());
        if (fabs(scale) < params.pte_small_step_thresh) st.small_step_iters++;
        if (st.small_step_iters >= params.pte_small_step_tries) {
          st.converged = false;
          phase = PH_FINISH;
          break;
        }
        gradfdx = -2.0 * scale * err;
        scale = 1.0;
        err_old = err;
        tu_kind = 0;
        prepare_test_update(scale, true);
      } break;
      // ===== TestUpdate second half (:1175-1187) + line search (:2226-2251) =============
      case PH_TU: {
#pragma unroll
        for (int m = 0; m < NM; ++m) {
          s.u[m] = s.rhobar[m] * eq[m];
          s.sie[m] = ratio(s.u[m], s.rhobar[m]);
          s.u[m] = ratio(s.u[m], s.uscale);
          s.temp[m] = s.Tequil;
          s.press[m] = ratio(pq[m], s.uscale);
        }
        s.Residual();
        const double err_new = s.ResidualNorm();
        if (tu_kind == 0) {
          err = err_new;
          if (err > err_old + params.line_search_alpha * gradfdx) {
            scale = 0.5;
            tu_kind = 1;
            prepare_test_update(scale, false);
          } else {
            phase = PH_AFTER_LS;
          }
        } else if (tu_kind == 1) {
          const double err_mid = err_new;
          if (err_mid < err && err_mid < err_old) {
            scale = 0.75 + 0.5 * ratio(err_mid - err, err - 2.0 * err_mid + err_old);
          }
          line_iter = 0;
          if (line_iter < params.line_search_max_iter) {
            tu_kind = 2;
            prepare_test_update(scale, false);
          } else {
            phase = PH_AFTER_LS;
          }
        } else {
          err = err_new;
          if (err < err_old + params.line_search_alpha * scale * gradfdx) {
            phase = PH_AFTER_LS;
          } else {
            scale *= params.line_search_fac;
            st.max_line_niter = line_iter > st.max_line_niter ? line_iter : st.max_line_niter;
            line_iter++;
            if (line_iter < params.line_search_max_iter) {
              prepare_test_update(scale, false);
            } else {
              phase = PH_AFTER_LS;
            }
          }
        }
      } break;
      // ===== PTESolver loop tail (:2252-2262) ==========================================
      case PH_AFTER_LS: {
        s.Fixup();
        if (((err > params.pte_slow_convergence_thresh * err_old) ||
             (unsigned long long)s.niter == pte_max_iter - 1) &&
            close_enough) {
          st.converged = true;
          st.slow_convergence_detected = true;
          phase = PH_FINISH;
        } else {
          ++s.niter;
          phase = PH_NEWTON_TOP;
        }
      } break;
      case PH_FINISH: { // :2263-2265
        s.Fixup();
        s.Finalize();
        st.residual = err;
        phase = PH_IDLE;
      } break;
      default:
        break;
      }
      if (phase < PH_NEWTON_TOP) break; // IDLE or waiting for an evaluation
    }
  }
};

#endif // __CUDACC__
} // namespace mix

#ifdef __CUDACC__
// Init kernel: one thread per cell, every lane on the same path.  test/test_pte.cpp:188-198
// (cell totals, Tguess = ApproxTemperatureFromRhoMatU unless the caller supplies one), then
// InitBase up to and including GetTguess (:627-640, :500-581).  The caller's arrays are only
// read; the record {Tguess, vfrac[NM], rho[NM]} goes to `init`.
template <int NM, class Acc>
__global__ void __launch_bounds__(256)
    pte_rhot_init_kernel(const __grid_constant__ EosSet set,
                         const __grid_constant__ MixParamsDev prm,
                         const double *__restrict__ rho, const double *__restrict__ vfrac,
                         const double *__restrict__ sie, const double *__restrict__ temp,
                         const double *__restrict__ press,
                         const double *__restrict__ sie_tot_in,
                         const double *__restrict__ vfrac_tot_in,
                         const double *__restrict__ Tguess_in, double *__restrict__ init,
                         const int64_t ncell) {
  const Acc eos(set);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = tid; t < ncell; t += nthr) {
    mix::Base<NM, NM, Acc> s(eos, prm);
    const int64_t base = t * NM;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      s.rho[m] = rho[base + m];
      s.vfrac[m] = vfrac[base + m];
      s.sie[m] = sie[base + m];
      s.temp[m] = temp[base + m];
      s.press[m] = press[base + m];
    }
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
      s.Tnorm = mix::ApproxTemperatureStaged<NM>(eos, rho_tot * sie_tot, s.rho, s.vfrac);
    }
    s.InitRhoBarandRho();
    const double utotal = s.rho_total * s.sie_total;
    s.uscale = stdmax(fabs(utotal), 1.0e-14);
    s.utotal_scale = ratio(utotal, s.uscale);
    const double Tguess = s.GetTguess();
    double *rec = init + t * (2 * NM + 1);
    rec[0] = Tguess;
#pragma unroll
    for (int m = 0; m < NM; ++m) {
      rec[1 + m] = s.vfrac[m];
      rec[1 + NM + m] = s.rho[m];
    }
  }
}

// One lane per cell.  A lane that finishes its cell stores it and pulls the next unsolved cell
// of the WHOLE launch from a global cursor (one warp-aggregated atomicAdd per refill), so no
// warp, block or SM runs dry before the range is exhausted: the only tail is the last cells'
// own iterations.  (Round 1 gave each warp a static contiguous slice; its lanes, then its
// block, then the grid idled while the slowest slice finished.)
//
// Rounds are block-synchronous (one barrier per round) and the Newton cycle is aligned to
// the block's round counter r: the Jacobian evaluations happen when r % 2 == 0, the
// full-step TestUpdate when r % 2 == 1.  A lane entering the Newton loop out of step (fresh
// from Init, or after line-search rounds) skips one round.  So in any round nearly all lanes of ALL warps of the SM run the same
// handler: the handlers execute with full warps, and the 16 warps walk the same stretch of
// SASS at the same time, which is what lets a 160 KB loop body live with the instruction
// cache (profiles/r1e: unaligned, 75 % of stall samples were instruction fetch and the
// handlers ran with 5 of 32 lanes).
constexpr int kSmThreads = 512;
template <int NM, class Acc, int THREADS = kSmThreads, int MINB = 1>
__global__ void __launch_bounds__(THREADS, MINB)
    pte_rhot_sm_kernel(const __grid_constant__ EosSet set,
                       const __grid_constant__ MixParamsDev prm, double *__restrict__ rho,
                       double *__restrict__ vfrac, double *__restrict__ sie,
                       double *__restrict__ temp, double *__restrict__ press,
                       const double *__restrict__ sie_tot_in,
                       const double *__restrict__ vfrac_tot_in,
                       const double *__restrict__ init,
                       unsigned long long *__restrict__ status, double *__restrict__ residual,
                       unsigned long long *__restrict__ stats, const int64_t ncell,
                       unsigned long long *__restrict__ cursor,
                       double *__restrict__ lambda = nullptr,
                       const unsigned char *__restrict__ skip = nullptr) {
  __shared__ unsigned int s_hist[SGX_PTE_NSTATS];
  if (stats) {
    for (int k = threadIdx.x; k < SGX_PTE_NSTATS; k += blockDim.x)
      s_hist[k] = 0;
    __syncthreads();
  }
  Acc eos(set);
  mix::RhoTLane<NM, Acc> L(eos, prm);
  // Accessors with a lambda cache (tabulated materials, CacheAccessor layout [cell][NM][slots]):
  // the lane keeps its cell's cache here and stores it when the cell retires.  In this solver
  // the tabulated models only WRITE their cache (no call reads the cached lT), so the init
  // kernel runs without one.
  constexpr int kLam = Acc::kCacheSlots > 0 ? NM * Acc::kCacheSlots : 1;
  double lamreg[kLam];
  const unsigned lane = threadIdx.x & 31u;
  bool exhausted = false; // warp-uniform: the cursor has passed the end of the range
  int64_t cell = -1;
  double eq[NM], pq[NM], eq2[NM], pq2[NM];
#pragma unroll
  for (int m = 0; m < NM; ++m) {
    eq[m] = pq[m] = eq2[m] = pq2[m] = 0.0;
    L.s.rho[m] = 1.0;
    L.cache[m] = RhoPart{1.0, 0.0, 0.0, 0.0};
    L.rq[m] = 1.0;
    L.s.rhobar[m] = 1.0;
  }
  for (unsigned r = 0;; ++r) {
    const bool idle = (L.phase == mix::PH_IDLE);
    const unsigned idle_mask = __ballot_sync(0xffffffffu, idle);
    if (idle_mask) {
      if (idle) {
        if (cell >= 0) { // retire the finished cell
          const int64_t base = cell * NM;
#pragma unroll
          for (int m = 0; m < NM; ++m) {
            rho[base + m] = L.s.rho[m];
            vfrac[base + m] = L.s.vfrac[m];
            sie[base + m] = L.s.sie[m];
            temp[base + m] = L.s.temp[m];
            press[base + m] = L.s.press[m];
          }
          if (status) status[cell] = mix::pack_status(L.st, L.s.niter);
          if (residual) residual[cell] = L.st.residual;
          if constexpr (Acc::kCacheSlots > 0) {
            if (lambda)
              for (int k = 0; k < kLam; ++k)
                lambda[cell * kLam + k] = lamreg[k];
          }
          if (stats) {
            atomicAdd(&s_hist[SGX_PTE_STAT_NCELL], 1u);
            if (L.st.converged) atomicAdd(&s_hist[SGX_PTE_STAT_CONVERGED], 1u);
            if (L.st.slow_convergence_detected) atomicAdd(&s_hist[SGX_PTE_STAT_SLOW], 1u);
            if (!L.st.converged) atomicAdd(&s_hist[SGX_PTE_STAT_FAILED], 1u);
            const int bin = L.s.niter < 31 ? L.s.niter : 31;
            atomicAdd(&s_hist[SGX_PTE_STAT_HIST0 + bin], 1u);
          }
          cell = -1;
        }
      }
      // one atomic per warp per refill; lanes take consecutive cells (coalesced loads)
      unsigned long long first = (unsigned long long)ncell;
      if (!exhausted) {
        if (lane == 0) first = atomicAdd(cursor, (unsigned long long)__popc(idle_mask));
        first = __shfl_sync(0xffffffffu, first, 0);
        exhausted = first + (unsigned long long)__popc(idle_mask) >= (unsigned long long)ncell;
      }
      if (idle) {
        const int64_t mine = (int64_t)first + __popc(idle_mask & ((1u << lane) - 1u));
        // skip[]: cells the caller solves elsewhere (get_sg_eos: cells whose participating
        // materials are fewer than NM); the lane stays idle and draws again next round
        if (mine < ncell && !(skip && skip[mine])) {
          cell = mine;
          const int64_t base = cell * NM;
#pragma unroll
          for (int m = 0; m < NM; ++m) {
            L.s.rho[m] = rho[base + m];
            L.s.vfrac[m] = vfrac[base + m];
            L.s.sie[m] = sie[base + m];
            L.s.temp[m] = temp[base + m];
            L.s.press[m] = press[base + m];
          }
          if constexpr (Acc::kCacheSlots > 0) {
            if (lambda)
              for (int k = 0; k < kLam; ++k)
                lamreg[k] = lambda[cell * kLam + k];
            set_cell_cache(eos, lambda ? lamreg : nullptr);
          }
          const double *rec = init + cell * (2 * NM + 1);
          double vfrac_init[NM], rho_init[NM];
#pragma unroll
          for (int m = 0; m < NM; ++m) {
            vfrac_init[m] = rec[1 + m];
            rho_init[m] = rec[1 + NM + m];
          }
          L.begin(sie_tot_in ? sie_tot_in[cell] : 0.0, sie_tot_in != nullptr,
                  vfrac_tot_in ? vfrac_tot_in[cell] : 1.0, rec[0], vfrac_init, rho_init);
        }
      }
    }
    // block-wide round barrier; also the termination vote
    if (!__syncthreads_or(L.phase != mix::PH_IDLE || !exhausted)) break;
    const bool hold = (L.phase == mix::PH_J) && (r % 2u != 0u); // wait for the Newton beat
    if (L.wants_eval() && !hold)
      mix::eval_round<NM>(eos, L.mode, L.Tq, L.Tq2, L.rq, L.s.rho, L.s.rhobar, L.cache, eq,
                          pq, eq2, pq2);
    __syncwarp();
    if (L.phase != mix::PH_IDLE && !hold) L.advance(eq, pq, eq2, pq2);
    __syncwarp();
  }
  if (stats) {
    __syncthreads();
    for (int k = threadIdx.x; k < SGX_PTE_NSTATS; k += blockDim.x)
      if (s_hist[k]) atomicAdd(&stats[k], (unsigned long long)s_hist[k]);
  }
}

// ---------------------------------------------------------------------------------------
// The warp-per-cell variant north_star names, kept for the side-by-side
// measurement (SGX_PTE_KERNEL=warp): ONE warp owns one mixed cell.  Lane 0 runs the strictly
// sequential part of the solve -- the same RhoTLane state machine as above: Neumaier sums,
// pivoted Householder solve, ScaleDx, line-search acceptance -- and the EOS evaluations of each
// round are dealt out to lanes, lane 2m + p = material m, evaluation point p (p = 1 is the
// dT-perturbed point of the Jacobian), requests and results travelling by __shfl_sync.  Results
// are bit-identical to the lane-per-cell kernel (same operations per cell).
// Why it loses (profiles/r2_pte_variants.md): a 3-material cell has at most 6 independent EOS
// evaluations per round, and the three materials are three different code paths, which SIMT
// serialises anyway; 26+ lanes idle through every round, and lane 0's scalar handlers -- a third
// of the instructions -- run at 1/32.
// ---------------------------------------------------------------------------------------
template <int NM, class Acc>
__global__ void __launch_bounds__(256, 2)
    pte_rhot_warp_kernel(const __grid_constant__ EosSet set,
                         const __grid_constant__ MixParamsDev prm, double *__restrict__ rho,
                         double *__restrict__ vfrac, double *__restrict__ sie,
                         double *__restrict__ temp, double *__restrict__ press,
                         const double *__restrict__ sie_tot_in,
                         const double *__restrict__ vfrac_tot_in,
                         const double *__restrict__ init,
                         unsigned long long *__restrict__ status, double *__restrict__ residual,
                         const int64_t ncell) {
  static_assert(2 * NM <= 32, "one lane per (material, evaluation point)");
  const Acc eos(set);
  mix::RhoTLane<NM, Acc> L(eos, prm);
  const unsigned lane = threadIdx.x & 31u;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarp = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int mm = (int)(lane >> 1), pp = (int)(lane & 1u); // my material / evaluation point
  RhoPart my_cache{1.0, 0.0, 0.0, 0.0}; // density-only terms of material mm (lane 2 mm)
  double eq[NM], pq[NM], eq2[NM], pq2[NM];
  for (int64_t cell = warp; cell < ncell; cell += nwarp) {
    if (lane == 0) {
      const int64_t base = cell * NM;
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        L.s.rho[m] = rho[base + m];
        L.s.vfrac[m] = vfrac[base + m];
        L.s.sie[m] = sie[base + m];
        L.s.temp[m] = temp[base + m];
        L.s.press[m] = press[base + m];
      }
      const double *rec = init + cell * (2 * NM + 1);
      double vfrac_init[NM], rho_init[NM];
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        vfrac_init[m] = rec[1 + m];
        rho_init[m] = rec[1 + NM + m];
      }
      L.begin(sie_tot_in ? sie_tot_in[cell] : 0.0, sie_tot_in != nullptr,
              vfrac_tot_in ? vfrac_tot_in[cell] : 1.0, rec[0], vfrac_init, rho_init);
    }
    while (true) {
      const int phase = __shfl_sync(0xffffffffu, L.phase, 0);
      if (phase == mix::PH_IDLE) break;
      const int wants = __shfl_sync(0xffffffffu, (int)L.wants_eval(), 0);
      if (wants) {
        const int mode = __shfl_sync(0xffffffffu, L.mode, 0);
        const double Tq = __shfl_sync(0xffffffffu, L.Tq, 0);
        const double Tq2 = __shfl_sync(0xffffffffu, L.Tq2, 0);
        double my_rq = 1.0, my_rho_c = 1.0, my_rb = 1.0;
#pragma unroll
        for (int m = 0; m < NM; ++m) { // requests: lane 0 -> lanes 2m, 2m+1
          const double a = __shfl_sync(0xffffffffu, L.rq[m], 0);
          const double b = __shfl_sync(0xffffffffu, L.s.rho[m], 0);
          const double c = __shfl_sync(0xffffffffu, L.s.rhobar[m], 0);
          if (mm == m) {
            my_rq = a;
            my_rho_c = b;
            my_rb = c;
          }
        }
        // the density-only terms kept since the last TestUpdate / InitBase live in lane 2 mm;
        // the dT-perturbed point of the Jacobian (lane 2 mm + 1) needs them
        RhoPart nb;
        nb.Ts = __shfl_sync(0xffffffffu, my_cache.Ts, lane & ~1u);
        nb.Es = __shfl_sync(0xffffffffu, my_cache.Es, lane & ~1u);
        nb.Ps = __shfl_sync(0xffffffffu, my_cache.Ps, lane & ~1u);
        nb.rG = __shfl_sync(0xffffffffu, my_cache.rG, lane & ~1u);
        double e = 0.0, P = 0.0;
        if (mm < NM) {
          if (pp == 0) {
            RhoPart part;
            eos.rho_part(mm, my_rq, part);
            e = eos.e_from_part(mm, part, my_rq, Tq);
            double sie_for_p = e;
            if (mode == mix::MODE_TU) sie_for_p = ratio(my_rb * e, my_rb);
            P = eos.P_pref_from_part(mm, part, my_rq, sie_for_p, Tq);
            if (mode != mix::MODE_J) my_cache = part;
          } else if (mode == mix::MODE_J) {
            e = eos.e_from_part(mm, nb, my_rho_c, Tq2);
            P = eos.P_pref_from_part(mm, nb, my_rho_c, e, Tq2);
          }
        }
#pragma unroll
        for (int m = 0; m < NM; ++m) { // results: lanes 2m, 2m+1 -> lane 0
          eq[m] = __shfl_sync(0xffffffffu, e, 2 * m);
          pq[m] = __shfl_sync(0xffffffffu, P, 2 * m);
          eq2[m] = __shfl_sync(0xffffffffu, e, 2 * m + 1);
          pq2[m] = __shfl_sync(0xffffffffu, P, 2 * m + 1);
        }
      }
      if (lane == 0) L.advance(eq, pq, eq2, pq2);
      __syncwarp();
    }
    if (lane == 0) {
      const int64_t base = cell * NM;
#pragma unroll
      for (int m = 0; m < NM; ++m) {
        rho[base + m] = L.s.rho[m];
        vfrac[base + m] = L.s.vfrac[m];
        sie[base + m] = L.s.sie[m];
        temp[base + m] = L.s.temp[m];
        press[base + m] = L.s.press[m];
      }
      if (status) status[cell] = mix::pack_status(L.st, L.s.niter);
      if (residual) residual[cell] = L.st.residual;
    }
  }
}
#endif // __CUDACC__

} // namespace sgx
#endif // SGX_PTE_SM_CUH_
