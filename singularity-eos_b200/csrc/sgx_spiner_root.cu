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

constexpr int64_t kRootChunk = int64_t(1) << 23; // cells per kernel pair
// A listed cell travels with its inputs, so the list kernel reads contiguous records instead of
// gathering 8 bytes out of every other 32-byte sector of the cell arrays.
struct RootListEntry {
  double rho, sie;
  unsigned long long idx;
};
// The cells left to the reference iterate are appended to kRootSubLists lists, chosen by block:
// one list would put every warp's atomicAdd on one address (measured: +0.9 ms per 2^26 cells).
// Each list has room for twice its fair share; a cell that finds its list full is solved on the
// spot by the same code the list kernel runs.
constexpr int kRootSubLists = 32;
constexpr int kRootCounterStride = 32; // unsigned ints between counters: one 128-byte line each
SGX_HD int64_t root_sublist_capacity(const int64_t ncell) {
  return ncell / kRootSubLists + 4096; // every cell of a fair share
}

// relative-T error budget of the direct solve against the reference iterate (see header):
// keep when  kBoundCoef * max|corner| * dx <= |h1 - h0|,
// kBoundCoef = 4 * 2^-52 * ln(10) / 1e-12
constexpr double kBoundCoef = 4.0 * 2.220446049250313e-16 * 2.302585092994046 / 1.0e-12;

// Every cell: table location as the reference, then for on-table cells the bracketing table cell
// of the (ix2, w2)-frozen blend and a linear solve in it.
//
// ncu history of this kernel (profiles/r2_root_*): a bisection with a data-dependent `break`
// ran the rest of each cell once per exit depth (5 of 32 lanes active); with a fixed trip count
// it was bound by the L1TEX wavefront rate -- 9 scattered 32-byte probes per cell, ~1.4 cycles
// per lane-probe -- at 23 Gcells/s.  So the number of scattered global loads per cell is what
// counts.  Now: the per-row metadata and the two 1-D curves of the table-location test live in
// shared memory (12 KiB), one 4-byte gather of the hint table (RootRowMeta, sgx_internal.hpp)
// starts the search 3-4 cells wide around the right cell, and the loop leaves as soon as every
// lane of the warp has its cell (1-2 probes).  The hint only narrows the search: the hit test
// h0 <= sie < h1 is always made on the table itself, and a search that comes up empty sends
// the cell to the reference iterate.
template <bool BOUNDED>
__global__ void __launch_bounds__(256, 4)
    rhot_T_direct_kernel(const __grid_constant__ SpinerRhoT eos,
                         const RootRowMeta *__restrict__ meta_g,
                         const unsigned int *__restrict__ hint,
                         const double *__restrict__ rho_in, const double *__restrict__ sie_in,
                         double *__restrict__ T_out, const int64_t n, double *__restrict__ lambda,
                         const int ls, const bool vec_lambda, RootListEntry *__restrict__ list,
                         unsigned int *__restrict__ count) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  RootRowMeta *meta = reinterpret_cast<RootRowMeta *>(smem_raw);
  double *s_sielTMax = reinterpret_cast<double *>(meta + eos.numRho);
  double *s_sieCold = s_sielTMax + eos.numRho;
  {
    const double *src = reinterpret_cast<const double *>(meta_g);
    double *dst = reinterpret_cast<double *>(smem_raw);
    for (int k = threadIdx.x; k < eos.numRho * 4; k += blockDim.x)
      dst[k] = src[k];
    for (int k = threadIdx.x; k < eos.numRho; k += blockDim.x) {
      s_sielTMax[k] = eos.sielTMax[k];
      s_sieCold[k] = eos.sieCold[k];
    }
    __syncthreads();
  }
  const unsigned lane = threadIdx.x & 31u;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  const int last_cell = eos.numT - 2;
  const unsigned sub = blockIdx.x % kRootSubLists;
  const unsigned sub_cap = (unsigned)root_sublist_capacity(n);
  unsigned int *const my_count = count + sub * kRootCounterStride;
  RootListEntry *const my_list = list + (size_t)sub * sub_cap;
  // Software pipeline: the next trip's (rho, sie) leave for HBM before this trip's chain of
  // dependent gathers (hint -> table cell -> next probe) starts; ncu had 66 % of the stall
  // cycles of this kernel on L1TEX scoreboards with < 1 eligible warp per scheduler.
  int64_t base = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31u);
  double rho_next = 0.0, sie_next = 0.0;
  if (base + lane < n) {
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];"
                 : "=d"(rho_next)
                 : "l"(rho_in + base + lane));
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];"
                 : "=d"(sie_next)
                 : "l"(sie_in + base + lane));
  }
  // warp-uniform trip count: every lane stays in the loop so the votes below are full-warp
  for (; base < n; base += nthr) {
    const int64_t i = base + lane;
    const bool valid = i < n;
    bool fallback = false, search = false, found = false;
    double rho = rho_next, sie = sie_next;
    double lRho = 0.0, lT = 0.0, w0 = 0.0, w1 = 0.0;
    int ix2 = 0, lo = 1, hi = 0;
    if (i + nthr < n) {
      asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];"
                   : "=d"(rho_next)
                   : "l"(rho_in + i + nthr));
      asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];"
                   : "=d"(sie_next)
                   : "l"(sie_in + i + nthr));
    }
    if (valid) {
      lRho = eos.lRho_(rho);
      const AxisPos pr(eos.gRho, lRho);
      // getLocDependsRhoSie_ (eos_spiner_rho_temp.hpp:1377-1392): same two 1-D interpolations,
      // same expressions as SpinerRhoT::locRhoSie, the curves read from shared memory
      const double sielTMax_ = pr.w0 * s_sielTMax[pr.ix] + pr.w1 * s_sielTMax[pr.ix + 1];
      const double sieCold_ = pr.w0 * s_sieCold[pr.ix] + pr.w1 * s_sieCold[pr.ix + 1];
      if (sie >= sielTMax_ - SpinerRhoT::SOFT_THRESH * fabs(sielTMax_)) { // OffTop :1207-1212
        const double Cv = interp1d(eos.dEdTMax, pr);
        lT = eos.lT_(eos.TMax + ratio(sie - sielTMax_, Cv));
      } else if (sie <= sieCold_ + SpinerRhoT::SOFT_THRESH * fabs(sieCold_)) { // OffBottom
        lT = eos.lTMin;                                                         // :1204-1206
      } else {
        ix2 = pr.ix;
        w0 = pr.w0;
        w1 = pr.w1;
        const RootRowMeta m0 = meta[ix2], m1 = meta[ix2 + 1];
        // interpolation (not extrapolation) between two rows that both rise strictly with T
        // (inv_du > 0 marks such pairs): the blend rises strictly too, the root is unique
        if (w1 >= 0.0 && w1 <= 1.0 && m0.inv_du > 0.0) {
          const double base_b = w0 * m0.base + w1 * m1.base, s_b = w0 * m0.s + w1 * m1.s;
          const int k = root_hint_bin(base_b, s_b, m0.umin, m0.inv_du, sie);
          const unsigned h = __ldg(hint + (size_t)ix2 * kRootHintBins + k);
          lo = (int)(h & 0xffffu);
          hi = (int)(h >> 16);
          hi = hi > last_cell ? last_cell : hi;
          search = true;
        } else {
          fallback = true;
        }
      }
    }
    // Search for the table cell with h0 <= sie < h1 (never a junction cell of a piecewise grid:
    // there h0 == h1).  Selects only; the warp leaves together.
    Cell4 c{0.0, 0.0, 0.0, 0.0};
    double h0 = 0.0, h1 = 0.0;
    int mid = 0;
    for (int it = 0; it < 20; ++it) {
      const bool live = search && !found && lo <= hi;
      if (!__any_sync(0xffffffffu, live)) break;
      if (live) {
        mid = (lo + hi) >> 1;
        c = ld_cell(eos.sie, eos.numT, ix2, mid);
        h0 = w0 * c.c00 + w1 * c.c10; // the blend at node mid   (w1 weight 0: exact)
        h1 = w0 * c.c01 + w1 * c.c11; // the blend at node mid+1
        const bool below = sie < h0, above = sie >= h1;
        found = !below && !above;
        hi = below ? mid - 1 : hi;
        lo = above ? mid + 1 : lo;
      }
    }
    if (search) {
      if (found) {
        const Grid1D &g = eos.gT;
        int ig = 0;
        if (g.ngrids > 1) ig = (g.ngrids > 2 && mid >= g.ptot[2]) ? 2 : (mid >= g.ptot[1] ? 1 : 0);
        const double dx = g.dx[ig];
        const double dh = h1 - h0;
        bool keep = true;
        if (BOUNDED) {
          const double mag = fmax(fmax(fabs(c.c00), fabs(c.c01)), fmax(fabs(c.c10), fabs(c.c11)));
          keep = kBoundCoef * mag * dx <= dh;
        }
        const double floor_x = (double)(mid - g.ptot[ig]) * dx + g.min[ig];
        lT = floor_x + ((sie - h0) / dh) * dx;
        fallback = !keep;
      } else {
        fallback = true; // outside the row pair's range: the reference's bracket logic decides
      }
    }
    if (valid && !fallback) {
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
    const unsigned fbm = __ballot_sync(0xffffffffu, fallback);
    if (fbm) {
      unsigned pos = 0;
      if (lane == 0) pos = atomicAdd(my_count, (unsigned)__popc(fbm));
      pos = __shfl_sync(0xffffffffu, pos, 0);
      if (fallback) {
        const unsigned mine = pos + __popc(fbm & ((1u << lane) - 1u));
        if (mine < sub_cap) {
          my_list[mine] = RootListEntry{rho, sie, (unsigned long long)i};
        } else { // list full (heavily skewed input): the reference iterate right here
          double lam[2];
          if (lambda) {
            lam[0] = lambda[i * ls];
            lam[1] = lambda[i * ls + 1];
          }
          T_out[i] = eos.T_from_rho_e(rho, sie, lambda ? lam : nullptr);
          if (lambda) {
            lambda[i * ls] = lam[0];
            lambda[i * ls + 1] = lam[1];
          }
        }
      }
    }
  }
}

// The listed cells through the reference iterate (SpinerRhoT::T_from_rho_e); blockIdx.y = list
__global__ void __launch_bounds__(256, 4)
    rhot_T_list_kernel(const __grid_constant__ SpinerRhoT eos, double *__restrict__ T_out,
                       double *__restrict__ lambda, const int ls,
                       const RootListEntry *__restrict__ list,
                       const unsigned int *__restrict__ count,
                       unsigned long long *__restrict__ stats, const unsigned long long ncell) {
  const unsigned sub = blockIdx.y;
  const unsigned sub_cap = (unsigned)root_sublist_capacity((int64_t)ncell);
  const unsigned raw = count[sub * kRootCounterStride];
  const unsigned cnt = raw < sub_cap ? raw : sub_cap; // the overflow was solved in place
  const RootListEntry *my_list = list + (size_t)sub * sub_cap;
  const unsigned tid = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned nthr = gridDim.x * blockDim.x;
  if (tid == 0 && stats) {
    if (sub == 0) atomicAdd(stats + 0, ncell);
    atomicAdd(stats + 1, (unsigned long long)raw);
  }
  for (unsigned k = tid; k < cnt; k += nthr) {
    const RootListEntry en = my_list[k];
    const int64_t i = (int64_t)en.idx;
    double lam[2];
    if (lambda) {
      lam[0] = lambda[i * ls];
      lam[1] = lambda[i * ls + 1];
    }
    T_out[i] = eos.T_from_rho_e(en.rho, en.sie, lambda ? lam : nullptr);
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
  const RootRowMeta *meta = reinterpret_cast<const RootRowMeta *>(e->tables.d("sie#rootmeta"));
  const unsigned int *hint = reinterpret_cast<const unsigned int *>(e->tables.d("sie#hint"));
  const size_t smem = (size_t)eos.numRho * (sizeof(RootRowMeta) + 2 * sizeof(double));
  if (smem > 48 * 1024) { // up to 96 KiB (build_spiner_rhot's `fits`): opt in once
    static const bool opted = [] {
      return cudaFuncSetAttribute(rhot_T_direct_kernel<true>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  96 * 1024) == cudaSuccess &&
             cudaFuncSetAttribute(rhot_T_direct_kernel<false>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  96 * 1024) == cudaSuccess;
    }();
    if (!opted) return fail(SGX_ERR_CUDA, "cannot reserve %zu bytes of shared memory", smem);
  }
  const int64_t chunk = n < kRootChunk ? n : kRootChunk;
  // scratch: kRootSubLists counters (one 128-byte line each), then the lists
  constexpr size_t kCounters = (size_t)kRootSubLists * kRootCounterStride;
  const size_t list_len = (size_t)kRootSubLists * (size_t)root_sublist_capacity(chunk);
  unsigned int *scratch = nullptr;
  SGX_CUDA_OK(scratch_alloc_async(
      reinterpret_cast<void **>(&scratch),
      kCounters * sizeof(unsigned int) + list_len * sizeof(RootListEntry), s));
  RootListEntry *lists = reinterpret_cast<RootListEntry *>(scratch + kCounters);
  int rc = SGX_OK;
  for (int64_t c0 = 0; c0 < n && rc == SGX_OK; c0 += chunk) {
    const int64_t m = (n - c0) < chunk ? (n - c0) : chunk;
    double *lam = lambda ? lambda + c0 * ls : nullptr;
    const bool vec_lambda = lam && (ls % 2) == 0 && (((uintptr_t)lam) & 15) == 0;
    cudaError_t err = cudaMemsetAsync(scratch, 0, kCounters * sizeof(unsigned int), s);
    if (err == cudaSuccess) {
      const LaunchCfg c = stream_cfg(m, 256, 4 * 2);
      if (bounded)
        rhot_T_direct_kernel<true><<<c.blocks, c.threads, smem, s>>>(
            eos, meta, hint, rho + c0, sie + c0, out + c0, m, lam, ls, vec_lambda, lists, scratch);
      else
        rhot_T_direct_kernel<false><<<c.blocks, c.threads, smem, s>>>(
            eos, meta, hint, rho + c0, sie + c0, out + c0, m, lam, ls, vec_lambda, lists, scratch);
      // one grid row per list; a list holds ~ m / 32 / 7 cells on the bench workload
      const LaunchCfg c2 = stream_cfg(root_sublist_capacity(m) / 4, 256, 1);
      rhot_T_list_kernel<<<dim3(c2.blocks, kRootSubLists), c2.threads, 0, s>>>(
          eos, out + c0, lam, ls, lists, scratch,
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

// ---------------------------------------------------------------------------------------
// Allocation-free variant of the hybrid T(rho, sie) launch above, for callers that cannot
// call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  It has
// no counterpart in the reference.
// The two kernels are the ones above; only the scratch (counters + sub-lists) differs: it is
// this file-scope device array, resident on every GPU that loads the library until exit.
// In host code the array's NAME is a host-side stand-in, not a device address: the address
// the kernels get is taken once per device with cudaGetSymbolAddress, that device current.
// ---------------------------------------------------------------------------------------
#include <mutex>

#define SGX_NOALLOC_MAX_CELLS_ (1ll << 20)
// counters: kRootSubLists x kRootCounterStride unsigned ints (4 KiB), then kRootSubLists lists
// of root_sublist_capacity(max) 24-byte records: 4096 + 32 * (32768 + 4096) * 24 = 28 315 648 B
#define SGX_NOALLOC_SCRATCH_BYTES_                                                             \
  (32ull * 32ull * 4ull + 32ull * (SGX_NOALLOC_MAX_CELLS_ / 32ull + 4096ull) * 24ull)
static __device__ __align__(256) unsigned char sgx_root_noalloc_scratch[SGX_NOALLOC_SCRATCH_BYTES_];

namespace sgx {
namespace {
static_assert(sizeof(RootListEntry) == 24, "scratch size above assumes 24-byte records");
// root_sublist_capacity(max) = max / kRootSubLists + 4096, written out (it is not constexpr)
static_assert(SGX_NOALLOC_SCRATCH_BYTES_ ==
                  (size_t)kRootSubLists * kRootCounterStride * sizeof(unsigned int) +
                      (size_t)kRootSubLists *
                          (size_t)(SGX_NOALLOC_MAX_CELLS_ / kRootSubLists + 4096) *
                          sizeof(RootListEntry),
              "scratch array must hold the counters and every sub-list of a maximum-size call");

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mutex;
unsigned int *g_noalloc_ptr[kNoAllocMaxDevices] = {};

// Once per device, that device current: the array's device address and the shared-memory opt-in
// of the direct kernel (function attributes are per device).  No allocation.
int noalloc_setup(const int device, const size_t smem, unsigned int **out) {
  if (device < 0 || device >= kNoAllocMaxDevices)
    return fail(SGX_ERR_UNSUPPORTED, "no-alloc T(rho,sie): device %d (this build tracks %d)",
                device, kNoAllocMaxDevices);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[device]) {
    void *p = nullptr;
    const cudaError_t err = cudaGetSymbolAddress(&p, sgx_root_noalloc_scratch);
    if (err != cudaSuccess || !p)
      return fail(SGX_ERR_CUDA, "cudaGetSymbolAddress(sgx_root_noalloc_scratch): %s",
                  cudaGetErrorString(err));
    if (cudaFuncSetAttribute(rhot_T_direct_kernel<true>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                             96 * 1024) != cudaSuccess)
      return fail(SGX_ERR_CUDA, "cannot reserve shared memory for the direct kernel");
    g_noalloc_ptr[device] = static_cast<unsigned int *>(p);
  }
  if (smem > 96 * 1024) return fail(SGX_ERR_UNSUPPORTED, "table too large: %zu B of smem", smem);
  *out = g_noalloc_ptr[device];
  return SGX_OK;
}
} // namespace
} // namespace sgx

extern "C" int64_t sgx_noalloc_max_cells(void) { return SGX_NOALLOC_MAX_CELLS_; }

extern "C" int sgx_TemperatureFromDensityInternalEnergy_noalloc(
    sgx_eos_t e, const double *rho, const double *sie, double *T, int64_t n, double *lambda,
    int lambda_stride, sgx_stream_t stream) {
  using namespace sgx;
  if (!e || e->kind != SGX_EOS_SPINER_RHOT)
    return fail(SGX_ERR_INVALID, "no-alloc T(rho,sie) needs a SpinerEOSDependsRhoT handle");
  if (n < 0) return fail(SGX_ERR_INVALID, "negative cell count");
  if (n > SGX_NOALLOC_MAX_CELLS_)
    return fail(SGX_ERR_INVALID,
                "no-alloc T(rho,sie): %lld cells exceed the static scratch's maximum of %lld; "
                "split the call or use sgx_TemperatureFromDensityInternalEnergy",
                (long long)n, (long long)SGX_NOALLOC_MAX_CELLS_);
  if (n == 0) return SGX_OK;
  if (!rho || !sie || !T) return fail(SGX_ERR_INVALID, "null array");
  if (lambda && lambda_stride < 2)
    return fail(SGX_ERR_INVALID, "lambda_stride %d < nlambda 2", lambda_stride);
  DeviceGuard guard;
  SGX_CUDA_OK(cudaSetDevice(e->device));
  const SpinerRhoT &eos = e->rhot;
  const size_t smem = (size_t)eos.numRho * (sizeof(RootRowMeta) + 2 * sizeof(double));
  unsigned int *scratch = nullptr;
  const int rc = noalloc_setup(e->device, smem, &scratch);
  if (rc != SGX_OK) return rc;
  const RootRowMeta *meta = reinterpret_cast<const RootRowMeta *>(e->tables.d("sie#rootmeta"));
  const unsigned int *hint = reinterpret_cast<const unsigned int *>(e->tables.d("sie#hint"));
  constexpr size_t kCounters = (size_t)kRootSubLists * kRootCounterStride;
  RootListEntry *lists = reinterpret_cast<RootListEntry *>(scratch + kCounters);
  cudaStream_t s = (cudaStream_t)stream;
  // the array keeps the last call's contents: the counters must start at zero; the lists need
  // no clearing (only entries below the new counts are read)
  SGX_CUDA_OK(cudaMemsetAsync(scratch, 0, kCounters * sizeof(unsigned int), s));
  const bool vec_lambda = lambda && (lambda_stride % 2) == 0 && (((uintptr_t)lambda) & 15) == 0;
  const LaunchCfg c = stream_cfg(n, 256, 4 * 2);
  rhot_T_direct_kernel<true><<<c.blocks, c.threads, smem, s>>>(
      eos, meta, hint, rho, sie, T, n, lambda, lambda_stride, vec_lambda, lists, scratch);
  const LaunchCfg c2 = stream_cfg(root_sublist_capacity(n) / 4, 256, 1);
  rhot_T_list_kernel<<<dim3(c2.blocks, kRootSubLists), c2.threads, 0, s>>>(
      eos, T, lambda, lambda_stride, lists, scratch,
      reinterpret_cast<unsigned long long *>(e->root_stats), (unsigned long long)n);
  count_launch(2);
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}
