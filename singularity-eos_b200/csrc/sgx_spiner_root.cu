// SpinerEOSDependsRhoT::TemperatureFromDensityInternalEnergy over an array of cells
// (eos/eos_spiner_rho_temp.hpp:777-784 -> lTFromlRhoSie_ :1192-1250 -> regula_falsi,
// base/root-finding-1d/root_finding.hpp:154-240) as a warp-level flow kernel.
//
// Why not one thread per cell (eval_kernel<T_FROM_RHO_E, SpinerRhoT>): the regula-falsi trip
// count is data dependent, so a warp waits for its slowest cell -- profiles/r1h_root_details:
// 13.3 of 32 lanes active, issue slots 70 % busy, FP64 pipe 44 %.  The trip count cannot be
// changed (the iterate is the parity contract), the idling can:
//
//   * each warp owns a contiguous cell range and runs three kinds of steps on it;
//   * SETUP, a full-warp batch of 32 fresh cells: logs, density-axis weights, table location,
//     bracket selection (2-3 function evaluations).  Off-table cells are finished here; the
//     others become 72-byte iteration records in a per-warp shared-memory ready queue;
//   * ITERATE, one regula-falsi step per live lane.  A lane whose root is found appends
//     (cell, lT) to the warp's done queue and takes the next record from the ready queue in
//     the same round, so all lanes keep iterating whatever their trip counts;
//   * RETIRE, a full-warp batch of 32 finished cells: T = 10^lT - offset, stores, lambda.
//
// MEASURED (B200, 2^24 cells, profiles/r1_tuning.md): lanes active 13.3 -> 24.7, but the step
// chain (weights -> 32-byte gather from L2 -> blend -> FP64 divide) is latency-bound, the
// queues take L1 capacity and ~65 bookkeeping instructions per round, and the kernel ends up
// level with the per-thread one on cold starts (2.59 vs 2.49 ms) and behind it on warm starts
// (0.89 vs 0.61 ms).  It is therefore NOT the default: SGX_ROOT_KERNEL=flow selects it.
//
// Every cell executes exactly the floating-point operations of the per-thread path (same
// bracket logic, same secant / Illinois updates, same loop condition), so results are
// bit-identical to it (tests/test_gpu_parity.py::test_root_flow_equals_per_thread_kernel).
#include <cstdlib>
#include <cstring>

#include "sgx_internal.hpp"

namespace sgx {

namespace {

constexpr int kWarpsPerBlock = 8;
constexpr int kBlocksPerSM = 3;
constexpr int kReadyCap = 64; // iteration records per warp (power of two)
constexpr int kDoneCap = 128; // finished cells per warp (power of two)
constexpr int kMaxIter = 1000; // SECANT_NITER_MAX
constexpr int kStepsPerRound = 4; // regula-falsi steps between two rounds of queue bookkeeping

// per-warp queues in shared memory, structure of arrays (bank-conflict free per field)
struct WarpQueues {
  double a[kReadyCap], b[kReadyCap], ya[kReadyCap], yb[kReadyCap];
  double tgt[kReadyCap], w21[kReadyCap], guess[kReadyCap]; // w20 = 1 - w21 (weights())
  int cell[kReadyCap];   // offset of the cell in the warp's range
  int ixsign[kReadyCap]; // density-axis cell index; bit 30 set when sign = -1
  double done_lT[kDoneCap];
  int done_cell[kDoneCap];
};

} // namespace

template <int STEPS, int BPS>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, BPS)
    rhot_T_from_rho_e_flow_kernel(const __grid_constant__ SpinerRhoT eos,
                                  const double *__restrict__ rho_in,
                                  const double *__restrict__ sie_in, double *__restrict__ T_out,
                                  const int64_t n, double *__restrict__ lambda, const int ls,
                                  const int64_t cells_per_warp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpQueues &q = reinterpret_cast<WarpQueues *>(smem_raw)[threadIdx.x >> 5];
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lt_mask = (1u << lane) - 1u;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t begin = warp * cells_per_warp;
  const int ncells = (int)((begin + cells_per_warp < n ? begin + cells_per_warp : n) - begin);
  if (ncells <= 0) return; // whole warp
  const double *__restrict__ tab = eos.sie;
  const int n1 = eos.numT;
  const double xtol = SpinerRhoT::ROOT_THRESH, ytol = SpinerRhoT::ROOT_THRESH;

  // f(x) - ytarget on the density row fixed by (ix2, w20, w21): the reference's four-corner
  // blend (databox.hpp:471-474) with the rho-axis weights hoisted
  auto func = [&](const double x, const int ix2, const double w20, const double w21,
                  const double tgt) {
    int ix1;
    double w10, w11;
    eos.gT.weights(x, ix1, w10, w11);
    return blend2c(tab, n1, ix2, w20, w21, ix1, w10, w11) - tgt;
  };

  int next = 0;                        // next fresh cell of the range (warp-uniform)
  unsigned rhead = 0, rtail = 0;       // ready queue (monotonic, slot = idx & (cap-1))
  unsigned dhead = 0, dtail = 0;       // done queue
  // this lane's live iteration
  bool live = false;
  double a = 0, b = 0, ya = 0, yb = 0, tgt = 0, w20 = 0, w21 = 0, guess = 0, sign = 1;
  int ix2 = 0, cell = 0, b1 = 0, b2 = 0, it = 0;

  while (true) {
    // ---- SETUP: keep at least 32 records ahead of the iterating lanes ---------------------
    while (next < ncells && (rtail - rhead) <= (unsigned)(kReadyCap - 32) &&
           (dtail - dhead) < 32u) {
      const int c = next + (int)lane;
      const bool have = c < ncells;
      bool to_ready = false, to_done = false;
      double s_a = 0, s_b = 0, s_ya = 0, s_yb = 0, s_tgt = 0, s_w21 = 0, s_guess = 0,
             s_lT = 0;
      int s_ixsign = 0;
      if (have) {
        const double rho = ldg(rho_in + begin + c);
        const double sie = ldg(sie_in + begin + c);
        const double lRho = eos.lRho_(rho);
        const AxisPos pr(eos.gRho, lRho);
        const TableStatus w = eos.locRhoSie(pr, sie);
        double *lam = lambda ? lambda + (begin + c) * ls : nullptr;
        if (w == OffBottom) {
          s_lT = eos.lTMin;
          to_done = true;
        } else if (w == OffTop) {
          const double Cv = interp1d(eos.dEdTMax, pr);
          const double e0 = interp1d(eos.sielTMax, pr);
          const double T = eos.TMax + ratio(sie - e0, Cv);
          s_lT = eos.lT_(T);
          to_done = true;
        } else {
          double lTGuess = eos.reproducible ? eos.lTMin : 0.5 * (eos.lTMin + eos.lTMax);
          if (lam) {
            const double lT_cache = lam[1];
            if ((eos.lTMin <= lT_cache) && (lT_cache <= eos.lTMax)) lTGuess = lT_cache;
          }
          // regula_falsi prologue (root_finding.hpp:163-188)
          s_a = eos.lTMin;
          s_b = eos.lTMax;
          s_ya = func(s_a, pr.ix, pr.w0, pr.w1, sie);
          const double yg = func(lTGuess, pr.ix, pr.w0, pr.w1, sie);
          bool bracketed = true;
          if (root::check_bracket(s_ya, yg)) {
            s_b = lTGuess;
            s_yb = yg;
          } else {
            s_yb = func(s_b, pr.ix, pr.w0, pr.w1, sie);
            if (root::check_bracket(yg, s_yb)) {
              s_a = lTGuess;
              s_ya = yg;
            } else {
              auto f1 = [&](const double x) { return func(x, pr.ix, pr.w0, pr.w1, sie); };
              bracketed = root::set_bracket(f1, s_a, lTGuess, s_b, s_ya, yg, s_yb);
            }
          }
          if (!bracketed) { // Status::FAIL -> lT = guess (:1241)
            s_lT = eos.reproducible ? eos.lTMin : lTGuess;
            to_done = true;
          } else {
            const bool neg = !(s_ya < 0); // sign = (ya < 0 ? 1.0 : -1.0)
            const double sg = neg ? -1.0 : 1.0;
            s_ya *= sg;
            s_yb *= sg;
            s_tgt = sie;
            s_w21 = pr.w1;
            s_guess = lTGuess;
            s_ixsign = pr.ix | (neg ? (1 << 30) : 0);
            to_ready = true;
          }
        }
        if (lam) lam[0] = lRho; // lambda[1] = lT follows at retire
      }
      const unsigned rmask = __ballot_sync(0xffffffffu, to_ready);
      const unsigned dmask = __ballot_sync(0xffffffffu, to_done);
      if (to_ready) {
        const unsigned slot = (rtail + __popc(rmask & lt_mask)) & (kReadyCap - 1);
        q.a[slot] = s_a;
        q.b[slot] = s_b;
        q.ya[slot] = s_ya;
        q.yb[slot] = s_yb;
        q.tgt[slot] = s_tgt;
        q.w21[slot] = s_w21;
        q.guess[slot] = s_guess;
        q.cell[slot] = c;
        q.ixsign[slot] = s_ixsign;
      }
      if (to_done) {
        const unsigned slot = (dtail + __popc(dmask & lt_mask)) & (kDoneCap - 1);
        q.done_lT[slot] = s_lT;
        q.done_cell[slot] = c;
      }
      rtail += __popc(rmask);
      dtail += __popc(dmask);
      next += 32;
      __syncwarp();
    }
    // ---- REFILL: idle lanes take the next records of the ready queue ----------------------
    {
      const unsigned idle_mask = __ballot_sync(0xffffffffu, !live);
      const unsigned avail = rtail - rhead;
      const unsigned rank = __popc(idle_mask & lt_mask);
      if (!live && rank < avail) {
        const unsigned slot = (rhead + rank) & (kReadyCap - 1);
        a = q.a[slot];
        b = q.b[slot];
        ya = q.ya[slot];
        yb = q.yb[slot];
        tgt = q.tgt[slot];
        w21 = q.w21[slot];
        w20 = 1. - w21; // as RegularGrid1D::weights computes it (regular_grid_1d.hpp:86)
        guess = q.guess[slot];
        cell = q.cell[slot];
        const int ixs = q.ixsign[slot];
        ix2 = ixs & ((1 << 30) - 1);
        sign = (ixs >> 30) ? -1.0 : 1.0;
        b1 = 0;
        b2 = 0;
        it = 0;
        live = true;
      }
      const unsigned taken = __popc(idle_mask);
      rhead += taken < avail ? taken : avail;
      __syncwarp();
    }
    const bool exhausted =
        (next >= ncells) && (rtail == rhead) && !__any_sync(0xffffffffu, live);

    // ---- ITERATE: one regula-falsi step per live lane (root_finding.hpp:193-222) ----------
    bool finished = false;
    double lT_fin = 0;
#pragma unroll 1
    for (int step = 0; step < STEPS; ++step) {
      if (live) {
        bool go = b - a > 2.0 * xtol && (fabs(ya) > ytol || fabs(yb) > ytol) && it < kMaxIter;
        if (go) {
          const double c = (a * yb - b * ya) / (yb - ya);
          if (c == a) {
            b = a; // `continue`: the loop condition then fails on b - a
          } else if (c == b) {
            a = b;
          } else {
            const double yc = sign * func(c, ix2, w20, w21, tgt);
            if (yc > 0.0) {
              b = c;
              yb = yc;
              b1++;
              ya *= (b1 > 1 ? 0.5 : 1.0);
              b2 = 0;
            } else if (yc < 0.0) {
              a = c;
              ya = yc;
              b2++;
              yb *= (b2 > 1 ? 0.5 : 1.0);
              b1 = 0;
            } else {
              a = c;
              b = c;
            }
            it++;
          }
          go = b - a > 2.0 * xtol && (fabs(ya) > ytol || fabs(yb) > ytol) && it < kMaxIter;
        }
        if (!go) {
          finished = true;
          // xroot = 0.5*(a+b); FAIL only when the iteration cap was hit (:223-230, :1241)
          lT_fin = (it == kMaxIter) ? (eos.reproducible ? eos.lTMin : guess) : 0.5 * (a + b);
          live = false;
        }
      }
    }
    {
      const unsigned fmask = __ballot_sync(0xffffffffu, finished);
      if (finished) {
        const unsigned slot = (dtail + __popc(fmask & lt_mask)) & (kDoneCap - 1);
        q.done_lT[slot] = lT_fin;
        q.done_cell[slot] = cell;
      }
      dtail += __popc(fmask);
      __syncwarp();
    }

    // ---- RETIRE: full-warp batches of finished cells ------------------------------------
    const bool draining =
        (next >= ncells) && (rtail == rhead) && !__any_sync(0xffffffffu, live);
    while ((dtail - dhead) >= 32u || (draining && dtail != dhead)) {
      const unsigned cnt = dtail - dhead;
      if (lane < cnt) {
        const unsigned slot = (dhead + lane) & (kDoneCap - 1);
        const double lT = q.done_lT[slot];
        const int64_t c = begin + q.done_cell[slot];
        T_out[c] = eos.T_(lT);
        if (lambda) lambda[c * ls + 1] = lT;
      }
      dhead += cnt < 32u ? cnt : 32u;
      __syncwarp();
    }
    if (exhausted) break; // nothing was live this round and everything has been retired
  }
}

template <int STEPS, int BPS>
static int launch_flow(const SpinerRhoT &eos, const double *rho, const double *sie, double *out,
                       int64_t n, double *lambda, int ls, cudaStream_t s) {
  const size_t smem = sizeof(WarpQueues) * kWarpsPerBlock;
  SGX_CUDA_OK(cudaFuncSetAttribute(rhot_T_from_rho_e_flow_kernel<STEPS, BPS>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // BPS blocks of kWarpsPerBlock warps per SM; each warp owns a contiguous range (>= 256 cells)
  int64_t warps = (int64_t)sm_count() * BPS * kWarpsPerBlock;
  const int64_t max_warps = (n + 255) / 256;
  if (warps > max_warps) warps = max_warps;
  const int64_t blocks = (warps + kWarpsPerBlock - 1) / kWarpsPerBlock;
  warps = blocks * kWarpsPerBlock;
  const int64_t cpw = (n + warps - 1) / warps;
  rhot_T_from_rho_e_flow_kernel<STEPS, BPS><<<(unsigned)blocks, kWarpsPerBlock * 32, smem, s>>>(
      eos, rho, sie, out, n, lambda, ls, cpw);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

int launch_rhot_T_from_rho_e_flow(const SpinerRhoT &eos, const double *rho, const double *sie,
                                  double *out, int64_t n, double *lambda, int ls,
                                  cudaStream_t s) {
  return launch_flow<kStepsPerRound, kBlocksPerSM>(eos, rho, sie, out, n, lambda, ls, s);
}

} // namespace sgx
