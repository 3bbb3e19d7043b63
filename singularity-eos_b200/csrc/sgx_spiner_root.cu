// SpinerEOSDependsRhoT::TemperatureFromDensityInternalEnergy over an array of cells
// (eos/eos_spiner_rho_temp.hpp:777-784 -> lTFromlRhoSie_ :1192-1250 -> regula_falsi,
// base/root-finding-1d/root_finding.hpp:154-240): direct inversion where it provably agrees with
// the reference iterate, the reference iterate itself everywhere else.
//
// At fixed lRho the inverted function is the reference's four-corner blend with (ix2, w2)
// frozen: piecewise linear in lT.  SURVEY.md 9-K2 allows replacing the root find by a segment
// search + linear solve where the two agree to 1e-12.  Measured against the reference on the
// config-3a workload (2^18 cells, /tmp prototype recorded in profiles/r2_root_direct_study.md):
// they do NOT agree everywhere -- 1.1 % of cells differ by more than 1e-12 (max 1.1e-11).  The
// cause is conditioning, not a bug: the reference iterate stops at a sign change of the
// *rounded* blend, which wanders by ~4 ulp(sie) / slope around the exact root, so wherever
// |sie| / |h(i+1) - h(i)| is large (cold, compressed states: the energy is dominated by the cold
// curve and barely moves with T) the reference's own answer is only defined to ~1e-11, and
// which value it lands on depends on every rounding of its path.  That path cannot be
// shortened; it can only be walked.  Hence two kernels:
//
//  1. rhot_T_direct_kernel, every cell: NQT logs, density-axis weights, table location exactly
//     as the reference; for on-table cells a bisection over the row pair's cells (<= 8 probes,
//     one 32-byte record each, no data-dependent trip counts beyond that) and a linear solve in
//     the bracketing cell.  The result is kept when an a-priori bound on its distance from ANY
//     value the reference iterate can return,
//         4 ulp * max|corner| / |h1 - h0| * dx * ln10   (relative error of T),
//     is below 1e-12 (measured worst case inside the bound: 3.9e-13).  Otherwise -- and for
//     non-monotone row pairs, density extrapolation and energies outside the row's range, where
//     the reference's bracket logic or its FAIL -> guess path decides the answer -- the cell's
//     index is appended to a list (warp-aggregated atomic).
//  2. rhot_T_list_kernel, the listed cells only: the reference iterate, operation for
//     operation (SpinerRhoT::T_from_rho_e, the code the per-thread eval_kernel runs), reading the
//     cached guess from lambda like the reference.  Its lanes are dense in hard cells.
//
// SGX_ROOT_MODE selects: "hybrid" (default, above), "exact" (reference iterate for every cell:
// eval_kernel<T_FROM_RHO_E, SpinerRhoT>, bit-identical lambda), "direct" (kernel 1 keeps every
// well-posed cell whatever the bound: closer to the exact root than the reference, but up to
// ~1e-11 away from the reference's value; opt-in, not a parity mode).
#include <cstdlib>
#include <cstring>

#include "sgx_internal.hpp"

namespace sgx {

namespace {

constexpr int64_t kRootChunk = int64_t(1) << 24; // cells per kernel pair (list scratch 64 MiB)

// relative-T error budget of the direct solve against the reference iterate (see header):
// keep when  kBoundCoef * max|corner| * dx <= |h1 - h0|,
// kBoundCoef = 4 * 2^-52 * ln(10) / 1e-12
constexpr double kBoundCoef = 4.0 * 2.220446049250313e-16 * 2.302585092994046 / 1.0e-12;

// One on-table cell.  Returns false when the reference iterate must decide.
template <bool BOUNDED>
SGX_D bool lT_direct(const SpinerRhoT &eos, const double *__restrict__ mono, const AxisPos &pr,
                     const double sie, double &lT) {
  // interpolation (not extrapolation) between two rows that both rise strictly with T: the
  // blend then rises strictly too and the root is unique
  if (!(pr.w1 >= 0.0 && pr.w1 <= 1.0) || ldg(mono + pr.ix) == 0.0) return false;
  int lo = 0, hi = eos.numT - 2; // candidate table cells along T
  Cell4 c;
  double h0, h1;
  int mid;
  while (true) {
    mid = (lo + hi) >> 1;
    c = ld_cell(eos.sie, eos.numT, pr.ix, mid);
    h0 = pr.w0 * c.c00 + pr.w1 * c.c10; // the blend at node mid   (w1 weight 0: exact)
    h1 = pr.w0 * c.c01 + pr.w1 * c.c11; // the blend at node mid+1
    if (sie < h0) {
      hi = mid - 1;
    } else if (sie >= h1) {
      lo = mid + 1;
    } else {
      break; // h0 <= sie < h1 (never a junction cell of a piecewise grid: there h0 == h1)
    }
    if (lo > hi) return false; // outside the row's range: the reference's bracket logic decides
  }
  const Grid1D &g = eos.gT;
  int ig = 0;
  if (g.ngrids > 1) ig = (g.ngrids > 2 && mid >= g.ptot[2]) ? 2 : (mid >= g.ptot[1] ? 1 : 0);
  const double dx = g.dx[ig];
  const double dh = h1 - h0;
  if (BOUNDED) {
    const double mag = fmax(fmax(fabs(c.c00), fabs(c.c01)), fmax(fabs(c.c10), fabs(c.c11)));
    if (!(kBoundCoef * mag * dx <= dh)) return false;
  }
  const double floor_x = (double)(mid - g.ptot[ig]) * dx + g.min[ig];
  lT = floor_x + ((sie - h0) / dh) * dx;
  return true;
}

template <bool BOUNDED>
__global__ void __launch_bounds__(256, 3)
    rhot_T_direct_kernel(const __grid_constant__ SpinerRhoT eos, const double *__restrict__ mono,
                         const double *__restrict__ rho_in, const double *__restrict__ sie_in,
                         double *__restrict__ T_out, const int64_t n, double *__restrict__ lambda,
                         const int ls, const bool vec_lambda, unsigned int *__restrict__ list,
                         unsigned int *__restrict__ count) {
  const unsigned lane = threadIdx.x & 31u;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  // warp-uniform trip count: every lane stays in the loop so the ballots below are full-warp
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31u); base < n;
       base += nthr) {
    const int64_t i = base + lane;
    const bool valid = i < n;
    bool fallback = false;
    if (valid) {
      double rho, sie;
      asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(rho) : "l"(rho_in + i));
      asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(sie) : "l"(sie_in + i));
      const double lRho = eos.lRho_(rho);
      const AxisPos pr(eos.gRho, lRho);
      const TableStatus w = eos.locRhoSie(pr, sie);
      double lT;
      if (w == OffBottom) { // eos_spiner_rho_temp.hpp:1204-1206
        lT = eos.lTMin;
      } else if (w == OffTop) { // :1207-1212
        const double Cv = interp1d(eos.dEdTMax, pr);
        const double e0 = interp1d(eos.sielTMax, pr);
        lT = eos.lT_(eos.TMax + ratio(sie - e0, Cv));
      } else {
        fallback = !lT_direct<BOUNDED>(eos, mono, pr, sie, lT);
      }
      if (!fallback) {
        asm volatile("st.global.cs.f64 [%0], %1;" ::"l"(T_out + i), "d"(eos.T_(lT)) : "memory");
        if (lambda) { // :1244-1245
          double *l = lambda + i * ls;
          if (vec_lambda) { // rows 16-byte aligned (checked at launch)
            asm volatile("st.global.cs.v2.f64 [%0], {%1,%2};" ::"l"(l), "d"(lRho), "d"(lT)
                         : "memory");
          } else {
            l[0] = lRho;
            l[1] = lT;
          }
        }
      }
    }
    const unsigned fbm = __ballot_sync(0xffffffffu, fallback);
    if (fbm) {
      unsigned pos = 0;
      if (lane == 0) pos = atomicAdd(count, (unsigned)__popc(fbm));
      pos = __shfl_sync(0xffffffffu, pos, 0);
      if (fallback) list[pos + __popc(fbm & ((1u << lane) - 1u))] = (unsigned)i;
    }
  }
}

// The listed cells through the reference iterate (SpinerRhoT::T_from_rho_e)
__global__ void __launch_bounds__(256, 4)
    rhot_T_list_kernel(const __grid_constant__ SpinerRhoT eos, const double *__restrict__ rho_in,
                       const double *__restrict__ sie_in, double *__restrict__ T_out,
                       double *__restrict__ lambda, const int ls,
                       const unsigned int *__restrict__ list,
                       const unsigned int *__restrict__ count,
                       unsigned long long *__restrict__ stats, const unsigned long long ncell) {
  const unsigned cnt = *count;
  const unsigned tid = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned nthr = gridDim.x * blockDim.x;
  if (tid == 0 && stats) {
    atomicAdd(stats + 0, ncell);
    atomicAdd(stats + 1, (unsigned long long)cnt);
  }
  for (unsigned k = tid; k < cnt; k += nthr) {
    const int64_t i = list[k];
    double lam[2];
    if (lambda) {
      lam[0] = lambda[i * ls];
      lam[1] = lambda[i * ls + 1];
    }
    T_out[i] = eos.T_from_rho_e(rho_in[i], sie_in[i], lambda ? lam : nullptr);
    if (lambda) {
      lambda[i * ls] = lam[0];
      lambda[i * ls + 1] = lam[1];
    }
  }
}

} // namespace

int root_mode() {
  const char *v = getenv("SGX_ROOT_MODE");
  if (v && strcmp(v, "exact") == 0) return 1;
  if (v && strcmp(v, "direct") == 0) return 2;
  return 0;
}

int launch_rhot_T_from_rho_e_hybrid(sgx_eos_s *e, bool bounded, const double *rho,
                                    const double *sie, double *out, int64_t n, double *lambda,
                                    int ls, cudaStream_t s) {
  const SpinerRhoT &eos = e->rhot;
  const double *mono = e->tables.d("sie#mono");
  const int64_t chunk = n < kRootChunk ? n : kRootChunk;
  unsigned int *scratch = nullptr; // [0] = count, [64...] = list
  SGX_CUDA_OK(cudaMallocAsync(&scratch, (size_t)(chunk + 64) * sizeof(unsigned int), s));
  int rc = SGX_OK;
  for (int64_t c0 = 0; c0 < n && rc == SGX_OK; c0 += chunk) {
    const int64_t m = (n - c0) < chunk ? (n - c0) : chunk;
    double *lam = lambda ? lambda + c0 * ls : nullptr;
    const bool vec_lambda = lam && (ls % 2) == 0 && (((uintptr_t)lam) & 15) == 0;
    cudaError_t err = cudaMemsetAsync(scratch, 0, sizeof(unsigned int), s);
    if (err == cudaSuccess) {
      const LaunchCfg c = stream_cfg(m, 256, 3 * 4);
      if (bounded)
        rhot_T_direct_kernel<true><<<c.blocks, c.threads, 0, s>>>(
            eos, mono, rho + c0, sie + c0, out + c0, m, lam, ls, vec_lambda, scratch + 64, scratch);
      else
        rhot_T_direct_kernel<false><<<c.blocks, c.threads, 0, s>>>(
            eos, mono, rho + c0, sie + c0, out + c0, m, lam, ls, vec_lambda, scratch + 64, scratch);
      const LaunchCfg c2 = stream_cfg(m, 256, 4);
      rhot_T_list_kernel<<<c2.blocks, c2.threads, 0, s>>>(
          eos, rho + c0, sie + c0, out + c0, lam, ls, scratch + 64, scratch,
          reinterpret_cast<unsigned long long *>(e->root_stats), (unsigned long long)m);
      count_launch(2);
      err = cudaGetLastError();
    }
    if (err != cudaSuccess)
      rc = fail(SGX_ERR_CUDA, "T(rho,sie) hybrid launch failed: %s", cudaGetErrorString(err));
  }
  cudaFreeAsync(scratch, s); // stream-ordered: released on every exit path
  return rc;
}

} // namespace sgx
