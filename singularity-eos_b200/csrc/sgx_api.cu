// libsgx_b200 C ABI: EOS handles, vector lookup / FillEos kernels and their launchers.
// See include/sgx_b200.h for the contract and the reference interfaces each entry replaces.
//
// Kernel design (all FP64, HBM-streaming):
//  * one generic grid-stride kernel per (EOS model, method), the EOS struct passed by value
//    in constant parameter space (the reference captures the EOS by value in its
//    portableFor lambda, eos/eos_base.hpp:178-182);
//  * each thread handles two adjacent cells per trip: 16-byte ld.global.nc / st.global
//    on the cell arrays, (ρ,T)→(lRho,lT) lambda write-back as one 16-byte store when
//    lambda_stride == 2;
//  * grid = (#SMs × resident blocks) so every SM streams from the first wave;
//  * table gathers go through the read-only path (ld.global.nc); the packed tables are
//    <= 18 MiB per EOS and stay L2-resident on their own against the .cs / no_allocate cell
//    streams (measured: 99.9 % of table sectors hit L2, profiles/r1h_lookup_details.txt), so
//    no persisting access-policy window is set.
#include <cstdlib>
#include <cstring>
#include <new>
#include <type_traits>

#include "sgx_internal.hpp"

namespace sgx {

std::atomic<unsigned long long> g_launches{0};

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

LaunchCfg stream_cfg(int64_t work_items, int threads, int blocks_per_sm) {
  const int64_t need = (work_items + threads - 1) / threads;
  int64_t blocks = (int64_t)sm_count() * blocks_per_sm;
  if (need < blocks) blocks = need;
  if (blocks < 1) blocks = 1;
  return LaunchCfg{(int)blocks, threads};
}

int upload_tables(sgx_eos_s *e) {
  TableStore &ts = e->tables;
  const size_t bytes = ts.host.size() * sizeof(double);
  if (bytes == 0) return SGX_OK;
  SGX_CUDA_OK(cudaGetDevice(&e->device));
  SGX_CUDA_OK(cudaMalloc((void **)&ts.dev, bytes));
  SGX_CUDA_OK(cudaMemcpy(ts.dev, ts.host.data(), bytes, cudaMemcpyHostToDevice));
  return SGX_OK;
}

int ensure_staging(sgx_eos_s *e, size_t pinned_bytes, size_t device_bytes) {
  if (e->own_stream == nullptr) SGX_CUDA_OK(cudaStreamCreate(&e->own_stream));
  if (pinned_bytes > e->pinned_bytes) {
    if (e->pinned) cudaFreeHost(e->pinned);
    e->pinned = nullptr;
    e->pinned_bytes = 0;
    SGX_CUDA_OK(cudaMallocHost(&e->pinned, pinned_bytes));
    e->pinned_bytes = pinned_bytes;
  }
  if (device_bytes > e->dscratch_bytes) {
    if (e->dscratch) cudaFree(e->dscratch);
    e->dscratch = nullptr;
    e->dscratch_bytes = 0;
    SGX_CUDA_OK(cudaMalloc(&e->dscratch, device_bytes));
    e->dscratch_bytes = device_bytes;
  }
  return SGX_OK;
}

HostPin::HostPin(const void *p, size_t bytes) {
  static const bool enabled = [] {
    const char *v = getenv("SGX_HOST_PIN");
    return !(v && strcmp(v, "none") == 0);
  }();
  if (!enabled || !p || bytes < (size_t(1) << 20)) return; // small ranges: not worth a syscall
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
    cudaGetLastError();
    return;
  }
  if (attr.type != cudaMemoryTypeUnregistered) return; // already pinned / managed / device
  if (cudaHostRegister(const_cast<void *>(p), bytes, cudaHostRegisterDefault) == cudaSuccess)
    ptr = const_cast<void *>(p);
  else
    cudaGetLastError(); // clear; the copies still work from pageable memory
}

cudaError_t scratch_alloc_async(void **ptr, size_t bytes, cudaStream_t s) {
  static cudaMemPool_t pools[64] = {nullptr};
  static std::atomic<int> made[64];
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return err;
  if (dev < 0 || dev >= 64) return cudaMallocAsync(ptr, bytes, s);
  if (made[dev].load(std::memory_order_acquire) != 2) {
    int expect = 0;
    if (made[dev].compare_exchange_strong(expect, 1)) {
      cudaMemPoolProps props = {};
      props.allocType = cudaMemAllocationTypePinned;
      props.handleTypes = cudaMemHandleTypeNone;
      props.location.type = cudaMemLocationTypeDevice;
      props.location.id = dev;
      cudaMemPool_t pool = nullptr;
      if (cudaMemPoolCreate(&pool, &props) == cudaSuccess) {
        unsigned long long never = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &never);
        pools[dev] = pool;
      } else {
        cudaGetLastError(); // fall back to the default pool below
      }
      made[dev].store(2, std::memory_order_release);
    } else {
      while (made[dev].load(std::memory_order_acquire) != 2) {
      }
    }
  }
  if (pools[dev]) return cudaMallocFromPoolAsync(ptr, bytes, pools[dev], s);
  return cudaMallocAsync(ptr, bytes, s);
}

int ensure_host_streams(sgx_eos_s *e) {
  for (int k = 0; k < 4; ++k)
    if (e->host_streams[k] == nullptr)
      SGX_CUDA_OK(cudaStreamCreateWithFlags(&e->host_streams[k], cudaStreamNonBlocking));
  return SGX_OK;
}

// ---------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------
template <int METHOD, class EOS>
SGX_D double call_method(const EOS &eos, const double a, const double b, double *lam) {
  if constexpr (METHOD == SGX_P_FROM_RHO_T) return eos.P_from_rho_T(a, b, lam);
  if constexpr (METHOD == SGX_E_FROM_RHO_T) return eos.e_from_rho_T(a, b, lam);
  if constexpr (METHOD == SGX_T_FROM_RHO_E) return eos.T_from_rho_e(a, b, lam);
  if constexpr (METHOD == SGX_P_FROM_RHO_E) return eos.P_from_rho_e(a, b, lam);
  if constexpr (METHOD == SGX_CV_FROM_RHO_T) return eos.cv_from_rho_T(a, b, lam);
  if constexpr (METHOD == SGX_CV_FROM_RHO_E) return eos.cv_from_rho_e(a, b, lam);
  if constexpr (METHOD == SGX_BMOD_FROM_RHO_T) return eos.bmod_from_rho_T(a, b, lam);
  if constexpr (METHOD == SGX_BMOD_FROM_RHO_E) return eos.bmod_from_rho_e(a, b, lam);
  if constexpr (METHOD == SGX_GAMMA_FROM_RHO_T) return eos.gamma_from_rho_T(a, b, lam);
  if constexpr (METHOD == SGX_GAMMA_FROM_RHO_E) return eos.gamma_from_rho_e(a, b, lam);
  return 0.0;
}

// Streaming accesses on the cell arrays: read-only, no L1 allocation on loads; evict-first
// (.cs) on stores so the output stream does not push the tables out of L2.
struct D4 {
  double x, y, z, w;
};
SGX_D D4 ld_stream4(const double *p) { // 256-bit, needs 32-byte alignment
  D4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
               : "l"(p));
  return v;
}
SGX_D double2 ld_stream2(const double *p) {
  double2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];"
               : "=d"(v.x), "=d"(v.y)
               : "l"(p));
  return v;
}
SGX_D double ld_stream1(const double *p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
SGX_D void st_stream4(double *p, const double x, const double y, const double z,
                      const double w) {
  asm volatile("st.global.cs.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(x), "d"(y), "d"(z),
               "d"(w)
               : "memory");
}
SGX_D void st_stream2(double *p, const double x, const double y) {
  asm volatile("st.global.cs.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(x), "d"(y) : "memory");
}
SGX_D void st_stream1(double *p, const double x) {
  asm volatile("st.global.cs.f64 [%0], %1;" ::"l"(p), "d"(x) : "memory");
}

// Does METHOD read the cell's lambda row before (possibly) writing it?  Only then is the row
// loaded; otherwise it is produced in registers and stored once.
//  * SpinerRhoT *_FROM_RHO_E read the cached lT as the root-find guess
//    (eos_spiner_rho_temp.hpp:1218-1223);
//  * SpinerRhoSie GruneisenParamFromDensityInternalEnergy does not write lambda at all
//    (eos_spiner_rho_sie.hpp:757-761), so the row must round-trip unchanged.
template <int METHOD, class EOS>
struct LambdaTraits {
  static constexpr bool from_e = METHOD == SGX_T_FROM_RHO_E || METHOD == SGX_P_FROM_RHO_E ||
                                 METHOD == SGX_CV_FROM_RHO_E || METHOD == SGX_BMOD_FROM_RHO_E ||
                                 METHOD == SGX_GAMMA_FROM_RHO_E;
  static constexpr bool load = (std::is_same<EOS, SpinerRhoT>::value && from_e) ||
                               (std::is_same<EOS, SpinerRhoSie>::value &&
                                METHOD == SGX_GAMMA_FROM_RHO_E);
};

// Which (EOS, METHOD) pairs have the two-phase prep()/finish() form (sgx_spiner.cuh)
template <class EOS, int METHOD>
struct TwoPhase : std::false_type {};
template <int METHOD>
struct TwoPhase<SpinerRhoT, METHOD>
    : std::integral_constant<bool, SpinerRhoT::two_phase<METHOD>()> {};
template <int METHOD>
struct TwoPhase<SpinerRhoSie, METHOD>
    : std::integral_constant<bool, SpinerRhoSie::two_phase<METHOD>()> {};

// One cell with its lambda row held in registers: NL = EOS::nlambda doubles.
template <int METHOD, class EOS, int NL>
SGX_D double eval_cell(const EOS &eos, const double a, const double b, double (&lam)[NL > 0 ? NL : 1],
                       const bool has_lambda) {
  return call_method<METHOD>(eos, a, b, (NL > 0 && has_lambda) ? lam : nullptr);
}

// OUT[i] = eos.METHOD(IN1[i], IN2[i], lambda[i])   (eos_base.hpp:174-183)
// VW cells per thread per trip (4: arrays 32-byte aligned -> 256-bit accesses; 2: 16-byte;
// 1: anything).  The VW evaluations of a trip are independent, so their table gathers overlap
// (memory-level parallelism: ncu showed this kernel stalled on `long_scoreboard`).
// PACKED: lambda_stride == EOS::nlambda, rows are contiguous and stored with vector stores.
// Per-family tuning, each value measured on B200 at 2^24 cells (profiles/r1_tuning.md):
//  * SpinerRhoT direct lookups: two-phase gathers + input prefetch keep 4 cell records and the
//    next trip's inputs live -> 128 registers, 2 resident blocks (0.150 ms; capped at 80
//    registers it spills and takes 0.268 ms);
//  * SpinerRhoT root finds are issue-bound and divergent: they want warps, not registers ->
//    4 blocks (<= 64 regs), 2 cells/thread, no prefetch (2.50 ms vs 3.47 ms at 2 blocks);
//  * SpinerRhoSie lookups are branch-free, so the compiler hoists the 4 gathers by itself:
//    one-phase, no prefetch, 48 registers, 4 blocks (0.131 ms vs 0.146 ms two-phase);
//  * analytic models: light, 4 blocks, prefetch on.
template <int METHOD, class EOS>
constexpr bool is_iterative() {
  return std::is_same<EOS, SpinerRhoT>::value && LambdaTraits<METHOD, EOS>::from_e;
}
template <int METHOD, class EOS>
constexpr bool use_two_phase() {
  return std::is_same<EOS, SpinerRhoT>::value && TwoPhase<EOS, METHOD>::value;
}
template <int METHOD, class EOS>
constexpr bool use_prefetch() {
  return !is_iterative<METHOD, EOS>() && !std::is_same<EOS, SpinerRhoSie>::value;
}
template <int METHOD, class EOS>
constexpr int min_blocks() {
  return use_two_phase<METHOD, EOS>() ? 2 : 4;
}

template <int METHOD, class EOS, int VW, bool PACKED>
__global__ void __launch_bounds__(256, min_blocks<METHOD, EOS>())
    eval_kernel(const __grid_constant__ EOS eos, const double *__restrict__ in1,
                const double *__restrict__ in2, double *__restrict__ out, const int64_t n,
                double *__restrict__ lambda, const int lstride) {
  constexpr int NL = EOS::nlambda;
  constexpr int NLA = NL > 0 ? NL : 1;
  constexpr bool kLoad = LambdaTraits<METHOD, EOS>::load;
  // root-finding methods are compute-bound and register-hungry: no input prefetch for them
  constexpr bool kPrefetch = use_prefetch<METHOD, EOS>();
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  const bool has_lambda = (NL > 0) && lambda != nullptr;
  const int64_t ngroup = n / VW;
  auto load_group = [&](const int64_t g, double (&a)[VW], double (&b)[VW]) {
    const int64_t i = g * VW;
    if constexpr (VW == 4) {
      const D4 va = ld_stream4(in1 + i), vb = ld_stream4(in2 + i);
      a[0] = va.x, a[1] = va.y, a[2] = va.z, a[3] = va.w;
      b[0] = vb.x, b[1] = vb.y, b[2] = vb.z, b[3] = vb.w;
    } else if constexpr (VW == 2) {
      const double2 va = ld_stream2(in1 + i), vb = ld_stream2(in2 + i);
      a[0] = va.x, a[1] = va.y;
      b[0] = vb.x, b[1] = vb.y;
    } else {
      a[0] = ld_stream1(in1 + i);
      b[0] = ld_stream1(in2 + i);
    }
  };
  double a[VW], b[VW];
  int64_t g = tid;
  if (g < ngroup) load_group(g, a, b);
  while (g < ngroup) {
    const int64_t i = g * VW;
    // software pipeline: the next trip's rho, T leave for HBM before this trip's table gathers
    // (ncu: the first use of the streamed inputs was the top `long_scoreboard` site)
    double an[VW], bn[VW];
    const int64_t gn = g + nthr;
    if (kPrefetch && gn < ngroup) load_group(gn, an, bn);
    double r[VW], lam[VW][NLA];
    if (kLoad && has_lambda) {
#pragma unroll
      for (int c = 0; c < VW; ++c)
#pragma unroll
        for (int k = 0; k < NL; ++k)
          lam[c][k] = lambda[(i + c) * lstride + k];
    }
    if constexpr (use_two_phase<METHOD, EOS>()) {
      LookupPrep q[VW];
#pragma unroll
      for (int c = 0; c < VW; ++c)
        eos.template prep<METHOD>(a[c], b[c], q[c]);
#pragma unroll
      for (int c = 0; c < VW; ++c)
        r[c] = eos.template finish<METHOD>(a[c], b[c], q[c], has_lambda ? lam[c] : nullptr);
    } else {
#pragma unroll
      for (int c = 0; c < VW; ++c)
        r[c] = eval_cell<METHOD, EOS, NL>(eos, a[c], b[c], lam[c], has_lambda);
    }
    if constexpr (VW == 4) {
      st_stream4(out + i, r[0], r[1], r[2], r[3]);
    } else if constexpr (VW == 2) {
      st_stream2(out + i, r[0], r[1]);
    } else {
      st_stream1(out + i, r[0]);
    }
    if (has_lambda) {
      if constexpr (PACKED && NL == 2 && VW == 4) {
        double *l = lambda + i * 2; // 64 contiguous bytes, 32-byte aligned (checked at launch)
        st_stream4(l, lam[0][0], lam[0][1], lam[1][0], lam[1][1]);
        st_stream4(l + 4, lam[2][0], lam[2][1], lam[3][0], lam[3][1]);
      } else if constexpr (PACKED && NL == 1 && VW == 4) {
        st_stream4(lambda + i, lam[0][0], lam[1][0], lam[2][0], lam[3][0]);
      } else if constexpr (PACKED && NL == 2 && VW == 2) {
        st_stream4(lambda + i * 2, lam[0][0], lam[0][1], lam[1][0], lam[1][1]);
      } else {
#pragma unroll
        for (int c = 0; c < VW; ++c)
#pragma unroll
          for (int k = 0; k < NL; ++k)
            lambda[(i + c) * lstride + k] = lam[c][k];
      }
    }
    g = gn;
    if constexpr (kPrefetch) {
#pragma unroll
      for (int c = 0; c < VW; ++c) {
        a[c] = an[c];
        b[c] = bn[c];
      }
    } else {
      if (g < ngroup) load_group(g, a, b);
    }
  }
  // ragged tail: n % VW cells, one per thread
  const int64_t done = ngroup * VW;
  if (VW > 1 && tid < n - done) {
    const int64_t i = done + tid;
    double lam[NLA];
    if (kLoad && has_lambda)
      for (int k = 0; k < NL; ++k)
        lam[k] = lambda[i * lstride + k];
    out[i] = eval_cell<METHOD, EOS, NL>(eos, in1[i], in2[i], lam, has_lambda);
    if (has_lambda)
      for (int k = 0; k < NL; ++k)
        lambda[i * lstride + k] = lam[k];
  }
}

// Fused P(rho,T) + e(rho,T) (one read of rho,T; 32 B/cell)
template <class EOS>
__global__ void __launch_bounds__(256)
    eval_pe_kernel(const __grid_constant__ EOS eos, const double *__restrict__ rho,
                   const double *__restrict__ T, double *__restrict__ P,
                   double *__restrict__ E, const int64_t n, double *__restrict__ lambda,
                   const int lstride) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n; i += nthr) {
    const double r = ld_stream1(rho + i), t = ld_stream1(T + i);
    double *l = lambda ? lambda + i * lstride : nullptr;
    st_stream1(P + i, eos.P_from_rho_T(r, t, l));
    st_stream1(E + i, eos.e_from_rho_T(r, t, l));
  }
}

// FillEos vector overload (eos_base.hpp:503-518)
template <class EOS>
__global__ void __launch_bounds__(256)
    fill_kernel(const __grid_constant__ EOS eos, double *__restrict__ rho,
                double *__restrict__ T, double *__restrict__ en, double *__restrict__ P,
                double *__restrict__ cv, double *__restrict__ bmod, const int64_t n,
                const unsigned long output, double *__restrict__ lambda, const int lstride,
                int *__restrict__ err) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  const unsigned long input = ~output;
  for (int64_t i = tid; i < n; i += nthr) {
    double r = 0, t = 0, e = 0, p = 0, c = 0, b = 0;
    if ((input & TQ_DENSITY) && rho) r = rho[i];
    if ((input & TQ_TEMPERATURE) && T) t = T[i];
    if ((input & TQ_SIE) && en) e = en[i];
    if ((input & TQ_PRESSURE) && P) p = P[i];
    const bool ok = eos.fill(r, t, e, p, c, b, output, lambda ? lambda + i * lstride : nullptr);
    if (!ok) {
      if (err) *err = 1;
      continue;
    }
    if ((output & TQ_DENSITY) && rho) rho[i] = r;
    if ((output & TQ_TEMPERATURE) && T) T[i] = t;
    if ((output & TQ_SIE) && en) en[i] = e;
    if ((output & TQ_PRESSURE) && P) P[i] = p;
    if ((output & TQ_CV) && cv) cv[i] = c;
    if ((output & TQ_BMOD) && bmod) bmod[i] = b;
  }
}

__global__ void fastmath_kernel(const int which, const double *__restrict__ x,
                                double *__restrict__ y, const int64_t n) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n; i += nthr) {
    const double v = x[i];
    double r;
    switch (which) {
    case 0:
      r = fastmath::log10(v);
      break;
    case 1:
      r = fastmath::pow10(v);
      break;
    case 2:
      r = fastmath::lg(v);
      break;
    default:
      r = fastmath::pow2(v);
      break;
    }
    y[i] = r;
  }
}

__global__ void solve_kernel(const int n, double *__restrict__ a, double *__restrict__ b,
                             int *__restrict__ ok, const int64_t batch) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= batch) return;
  double la[36], lb[6];
  for (int k = 0; k < n * n; ++k)
    la[k] = a[i * n * n + k];
  for (int k = 0; k < n; ++k)
    lb[k] = b[i * n + k];
  const bool good = mix::solve_Ax_b<6>(n, la, lb);
  for (int k = 0; k < n * n; ++k)
    a[i * n * n + k] = la[k];
  for (int k = 0; k < n; ++k)
    b[i * n + k] = lb[k];
  if (ok) ok[i] = good ? 1 : 0;
}

// ---------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------
template <int METHOD, class EOS>
int launch_eval_m(const EOS &eos, const double *in1, const double *in2, double *out, int64_t n,
                  double *lambda, int lstride, cudaStream_t s, int blocks_per_sm) {
  constexpr int NL = EOS::nlambda;
  const uintptr_t bits = ((uintptr_t)in1) | ((uintptr_t)in2) | ((uintptr_t)out);
  // rows are contiguous when the caller's stride equals nlambda(); vector stores on them
  // additionally need the row array itself 32-byte aligned
  const bool packed = lambda != nullptr && NL > 0 && lstride == NL &&
                      (((uintptr_t)lambda) & 31) == 0;
#define SGX_LAUNCH(VW, PACKED)                                                             do {                                                                                       const LaunchCfg c = stream_cfg(n / VW + (VW - 1), 256, blocks_per_sm);                   eval_kernel<METHOD, EOS, VW, PACKED><<<c.blocks, c.threads, 0, s>>>(eos, in1, in2, out, n,                                                                         lambda, lstride);   } while (0)
  // Root-finding methods loop a data-dependent number of times per cell: several cells per
  // thread only stack the divergence and the registers (measured: 4/thread is 40 % slower)
  constexpr bool kIterative = is_iterative<METHOD, EOS>();
  if (!kIterative && (bits & 31) == 0 && n >= 4) {
    if (packed) SGX_LAUNCH(4, true); else SGX_LAUNCH(4, false);
  } else if ((bits & 15) == 0 && n >= 2) {
    if (packed) SGX_LAUNCH(2, true); else SGX_LAUNCH(2, false);
  } else {
    SGX_LAUNCH(1, false);
  }
#undef SGX_LAUNCH
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

template <class EOS>
int launch_eval(const EOS &eos, int method, const double *in1, const double *in2, double *out,
                int64_t n, double *lambda, int lstride, cudaStream_t s, int bps) {
  switch (method) {
#define SGX_CASE(M)                                                                      \
  case M:                                                                                \
    return launch_eval_m<M>(eos, in1, in2, out, n, lambda, lstride, s, bps);
    SGX_CASE(SGX_P_FROM_RHO_T)
    SGX_CASE(SGX_E_FROM_RHO_T)
    SGX_CASE(SGX_T_FROM_RHO_E)
    SGX_CASE(SGX_P_FROM_RHO_E)
    SGX_CASE(SGX_CV_FROM_RHO_T)
    SGX_CASE(SGX_CV_FROM_RHO_E)
    SGX_CASE(SGX_BMOD_FROM_RHO_T)
    SGX_CASE(SGX_BMOD_FROM_RHO_E)
    SGX_CASE(SGX_GAMMA_FROM_RHO_T)
    SGX_CASE(SGX_GAMMA_FROM_RHO_E)
#undef SGX_CASE
  default:
    return fail(SGX_ERR_INVALID, "unknown method %d", method);
  }
}

// Output masks the reference's scalar FillEos rejects (UNDEFINED_ERROR / EOS_ERROR aborts):
// checked on the host so the call fails loudly instead of skipping cells.
//   IdealGas   eos_ideal.hpp:236-262      Gruneisen  eos_gruneisen.hpp:527-548
//   SpinerRhoT eos_spiner_rho_temp.hpp:1043-1093   SpinerRhoSie eos_spiner_rho_sie.hpp:842-900
static const char *fill_mask_error(int kind, unsigned long output) {
  const unsigned long in = ~output;
  const bool oR = output & TQ_DENSITY, oT = output & TQ_TEMPERATURE, oE = output & TQ_SIE,
             oP = output & TQ_PRESSURE;
  switch (kind) {
  case SGX_EOS_IDEAL:
    if (oR && oE && (oP || oT)) return "IdealGas::FillEos: density and energy outputs need P and T inputs";
    if (oP && oE && (oR || oT)) return "IdealGas::FillEos: pressure and energy outputs need rho and T inputs";
    return nullptr;
  case SGX_EOS_GRUNEISEN:
    if (!((in & TQ_TEMPERATURE) && (in & TQ_PRESSURE)) && (oR || oE))
      return "Gruneisen::FillEos: density and energy must be inputs unless P and T are given";
    return nullptr;
  case SGX_EOS_SPINER_RHOT:
    if (oR && !((in & TQ_PRESSURE) && (in & TQ_TEMPERATURE)))
      return "SpinerEOSDependsRhoT::FillEos: density output needs P and T inputs";
    if (oT && !((in & TQ_DENSITY) && ((in & TQ_SIE) || (in & TQ_PRESSURE))))
      return "SpinerEOSDependsRhoT::FillEos: temperature output needs rho and (sie or P) inputs";
    return nullptr;
  case SGX_EOS_SPINER_RHOSIE:
    if (oT && oE) return "SpinerEOSDependsRhoSie::FillEos: T and sie cannot both be outputs";
    if (oR && (oP || oT)) return "SpinerEOSDependsRhoSie::FillEos: density output needs P and T inputs";
    return nullptr;
  default:
    return nullptr;
  }
}

template <class EOS>
int launch_fill(sgx_eos_s *h, const EOS &eos, double *rho, double *T, double *en, double *P,
                double *cv, double *bmod, int64_t n, unsigned long output, double *lambda,
                int lstride, cudaStream_t s) {
  if (const char *why = fill_mask_error(h->kind, output)) return fail(SGX_ERR_UNSUPPORTED, "%s", why);
  // every quantity the mask names -- input or output -- needs its array
  double *arr[6] = {rho, T, en, P, cv, bmod};
  const unsigned long bit[6] = {TQ_DENSITY, TQ_TEMPERATURE, TQ_SIE, TQ_PRESSURE, TQ_CV, TQ_BMOD};
  for (int k = 0; k < 6; ++k)
    if ((output & bit[k]) && !arr[k]) return fail(SGX_ERR_INVALID, "FillEos: null output array %d", k);
  const LaunchCfg c = stream_cfg(n, 256, 8);
  fill_kernel<EOS><<<c.blocks, c.threads, 0, s>>>(eos, rho, T, en, P, cv, bmod, n, output,
                                                  lambda, lstride, nullptr);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

template <class EOS>
int launch_pe(const EOS &eos, const double *rho, const double *T, double *P, double *E,
              int64_t n, double *lambda, int lstride, cudaStream_t s) {
  const LaunchCfg c = stream_cfg(n, 256, 8);
  eval_pe_kernel<EOS><<<c.blocks, c.threads, 0, s>>>(eos, rho, T, P, E, n, lambda, lstride);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

// Lambda slots a row must hold: nlambda(), except a Helmholtz built for type-indexed
// {Abar, Zbar} lambdas (no LogTemperature slot), which uses 2 of the 3.
static int lambda_need(const sgx_eos_s *e) {
  if (e->kind == SGX_EOS_MODIFIED) e = e->base;
  switch (e->kind) {
  case SGX_EOS_SPINER_RHOT:
    return 2;
  case SGX_EOS_SPINER_RHOSIE:
    return 1;
  case SGX_EOS_HELMHOLTZ:
    return e->helm.lambda_has_lT_slot ? 3 : 2;
  default:
    return 0;
  }
}

static int check_handle(sgx_eos_t e) {
  if (!e) return fail(SGX_ERR_INVALID, "null EOS handle");
  cudaError_t err = cudaSetDevice(e->device);
  if (err != cudaSuccess)
    return fail(SGX_ERR_CUDA, "cudaSetDevice(%d): %s", e->device, cudaGetErrorString(err));
  return SGX_OK;
}

static int new_handle(int kind, sgx_eos_s **out) {
  if (!out) return fail(SGX_ERR_INVALID, "null output handle pointer");
  *out = nullptr;
  int ndev = 0;
  cudaError_t err = cudaGetDeviceCount(&ndev);
  if (err != cudaSuccess || ndev <= 0)
    return fail(SGX_ERR_CUDA, "no CUDA device available (%s); libsgx_b200 has no CPU path",
                cudaGetErrorString(err));
  sgx_eos_s *e = new (std::nothrow) sgx_eos_s();
  if (!e) return fail(SGX_ERR_NOMEM, "out of host memory");
  e->kind = kind;
  std::memset(&e->analytic, 0, sizeof(e->analytic));
  e->analytic.kind = kind;
  err = cudaGetDevice(&e->device);
  if (err != cudaSuccess) {
    delete e;
    return fail(SGX_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(err));
  }
  *out = e;
  return SGX_OK;
}

// Gruneisen::ComputeRhoMax (eos_gruneisen.hpp:231-329), host only
static double gruneisen_rho_max(const double s1, const double s2, const double s3,
                                const double rho0) {
  auto poly = [=](double eta) { return 1 - s1 * eta - s2 * pow2i(eta) - s3 * pow3i(eta); };
  auto near_zero = [](double v, double tol) { return std::abs(v) < tol; };
  auto find_min_bounded_val = [](double v1, double v2, double mn, double mx) {
    const double minval = std::min(v1, v2), maxval = std::max(v1, v2);
    if (minval > mn && minval < mx) return minval;
    if (maxval > mn && maxval < mx) return maxval;
    return -std::numeric_limits<double>::infinity();
  };
  constexpr double EPS = kEps;
  double root_ = -1.;
  double discriminant, root1, root2;
  if (near_zero(s2, EPS) && near_zero(s3, EPS) && s1 > 1.) {
    root_ = 1. / s1;
  } else if (near_zero(s3, EPS) && !near_zero(s2, EPS)) {
    discriminant = pow2i(s1) + 4. * s2;
    if (discriminant < 0.) {
      root_ = -1.;
    } else {
      root1 = (s1 + std::sqrt(discriminant)) / (-2. * s2);
      root2 = (s1 - std::sqrt(discriminant)) / (-2. * s2);
      root_ = find_min_bounded_val(root1, root2, 0., 1.);
    }
  } else if (!near_zero(s3, EPS)) {
    discriminant = -18. * s3 * s2 * s1 - 4. * pow3i(-s2) + pow2i(s2) * pow2i(s1) -
                   4. * s3 * pow3i(s1) - 27. * pow2i(s3);
    double minbound = 0.;
    double maxbound = 1.;
    if (discriminant >= 0) {
      const double extremum1 = (s2 + std::sqrt(pow2i(s2) - 3. * s3 * s1)) / (-3. * s3);
      const double extremum2 = (s2 - std::sqrt(pow2i(s2) - 3. * s3 * s1)) / (-3. * s3);
      const double min_extremum = std::min(extremum1, extremum2);
      const double max_extremum = std::max(extremum1, extremum2);
      if (s3 < 0.) {
        minbound = std::max(min_extremum, minbound);
        maxbound = std::min(std::max(max_extremum, minbound), maxbound);
      } else {
        if (min_extremum > 0) maxbound = std::min(min_extremum, maxbound);
      }
    }
    if (poly(minbound) * poly(maxbound) < 0.) {
      constexpr double factor = 100;
      constexpr double xtol = factor * EPS;
      constexpr double ytol = factor / 10 * EPS;
      constexpr double eta_guess = 0.001;
      root::regula_falsi(poly, 0., eta_guess, minbound, maxbound, xtol, ytol, root_);
    } else if (near_zero(poly(minbound), EPS)) {
      root_ = minbound;
    } else if (near_zero(poly(maxbound), EPS)) {
      root_ = maxbound;
    } else {
      root_ = -1.;
    }
  }
  if (root_ > 0. && root_ < 1.) return rho0 / ((1 - root_) + EPS);
  return std::numeric_limits<double>::infinity();
}

} // namespace sgx

using namespace sgx;

// =======================================================================================
// extern "C"
// =======================================================================================
// the raw .sp5 contents + flags a Spiner handle was created from
static int grid_points(const sgx_grid1d &g) {
  int n = 0;
  for (int p = 0; p < g.ngrids && p < 3; ++p)
    n += g.npoints[p];
  return n;
}
static std::vector<char> sp5_recipe(int kind, const sgx_sp5_tables *raw, int reproducible,
                                    int vapor_dome) {
  Recipe r(kind);
  const int32_t flags[2] = {reproducible, vapor_dome};
  r.put(flags, sizeof(flags));
  {
    // the header record without the caller's addresses (the arrays follow as records of their
    // own): the stream depends on the contents only, so equal models give equal bytes
    sgx_sp5_tables hdr;
    std::memcpy(&hdr, raw, sizeof(hdr));
    hdr.P = hdr.sie = hdr.bMod = hdr.dPdRho = hdr.dPdE = hdr.dTdRho = hdr.dTdE = hdr.dEdRho =
        hdr.dEdT = nullptr;
    hdr.PCold = hdr.sieCold = hdr.bModCold = hdr.dPdRhoCold = nullptr;
    hdr.T_re = hdr.P_re = hdr.bMod_re = hdr.dPdRho_re = hdr.dPdE_re = hdr.dTdRho_re =
        hdr.dTdE_re = hdr.dEdRho_re = nullptr;
    r.put(&hdr, sizeof(hdr));
  }
  const size_t nr = grid_points(raw->lrho), nt = grid_points(raw->lT), ne = grid_points(raw->lE);
  const double *t2[9] = {raw->P, raw->sie, raw->bMod, raw->dPdRho, raw->dPdE, raw->dTdRho,
                         raw->dTdE, raw->dEdRho, raw->dEdT};
  for (const double *p : t2)
    r.put_array(p, nr * nt);
  const double *c1[4] = {raw->PCold, raw->sieCold, raw->bModCold, raw->dPdRhoCold};
  for (const double *p : c1)
    r.put_array(p, nr);
  const double *e2[8] = {raw->T_re, raw->P_re, raw->bMod_re, raw->dPdRho_re, raw->dPdE_re,
                         raw->dTdRho_re, raw->dTdE_re, raw->dEdRho_re};
  for (const double *p : e2)
    r.put_array(p, nr * ne);
  return std::move(r.bytes);
}

extern "C" {

int sgx_abi_version(void) { return SGX_ABI_VERSION; }
const char *sgx_last_error(void) { return last_error().c_str(); }
int sgx_device_count(void) {
  int n = 0;
  cudaError_t err = cudaGetDeviceCount(&n);
  if (err != cudaSuccess) return fail(SGX_ERR_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(err));
  return n;
}
int sgx_set_device(int device) {
  SGX_CUDA_OK(cudaSetDevice(device));
  return SGX_OK;
}
uint64_t sgx_launch_count(void) { return g_launches.load(); }

int sgx_eos_create_ideal(double gm1, double Cv, sgx_eos_t *out) {
  if (!(Cv >= 0) || !(gm1 >= 0))
    return fail(SGX_ERR_INVALID, "IdealGas::CheckParams: Cv and gm1 must be non-negative");
  sgx_eos_s *e;
  int rc = new_handle(SGX_EOS_IDEAL, &e);
  if (rc) return rc;
  e->analytic.ideal.Cv = Cv;
  e->analytic.ideal.gm1 = gm1;
  {
    Recipe r(SGX_EOS_IDEAL);
    const double a[2] = {gm1, Cv};
    r.put(a, sizeof(a));
    e->recipe = std::move(r.bytes);
  }
  *out = e;
  return SGX_OK;
}

int sgx_eos_create_gruneisen(double C0, double s1, double s2, double s3, double G0, double b,
                             double rho0, double T0, double P0, double Cv, double rho_max,
                             sgx_eos_t *out) {
  sgx_eos_s *e;
  int rc = new_handle(SGX_EOS_GRUNEISEN, &e);
  if (rc) return rc;
  Gruneisen &g = e->analytic.gr;
  g.C0 = C0;
  g.s1 = s1;
  g.s2 = s2;
  g.s3 = s3;
  g.G0 = G0;
  g.b = b;
  g.rho0 = rho0;
  g.T0 = T0;
  g.P0 = P0;
  g.Cv = Cv;
  g.rho_max = rho_max > 0 ? rho_max : 0.99 * gruneisen_rho_max(s1, s2, s3, rho0);
  {
    Recipe r(SGX_EOS_GRUNEISEN);
    const double a[11] = {C0, s1, s2, s3, G0, b, rho0, T0, P0, Cv, rho_max};
    r.put(a, sizeof(a));
    e->recipe = std::move(r.bytes);
  }
  *out = e;
  return SGX_OK;
}

int sgx_eos_create_davis_reactants(double rho0, double e0, double P0, double T0, double A,
                                   double B, double C, double G0, double Z, double alpha,
                                   double Cv0, sgx_eos_t *out) {
  sgx_eos_s *e;
  int rc = new_handle(SGX_EOS_DAVIS_REACTANTS, &e);
  if (rc) return rc;
  DavisReactants &d = e->analytic.dr;
  d.rho0 = rho0;
  d.spvol0 = 1.0 / rho0;
  d.e0 = e0;
  d.P0 = P0;
  d.T0 = T0;
  d.A = A;
  d.B = B;
  d.C = C;
  d.G0 = G0;
  d.Z = Z;
  d.alpha = alpha;
  d.i1pa = 1.0 / (1.0 + alpha);
  d.Cv0 = Cv0;
  d.A2oB = A * A / B;
  {
    Recipe r(SGX_EOS_DAVIS_REACTANTS);
    const double a[11] = {rho0, e0, P0, T0, A, B, C, G0, Z, alpha, Cv0};
    r.put(a, sizeof(a));
    e->recipe = std::move(r.bytes);
  }
  *out = e;
  return SGX_OK;
}

int sgx_eos_create_davis_products(double a, double b, double k, double n, double vc,
                                  double pc, double Cv, sgx_eos_t *out) {
  sgx_eos_s *e;
  int rc = new_handle(SGX_EOS_DAVIS_PRODUCTS, &e);
  if (rc) return rc;
  DavisProducts &d = e->analytic.dp;
  d.a = a;
  d.b = b;
  d.k = k;
  d.n = n;
  d.aon = a / n;
  d.vc = vc;
  d.pc = pc;
  d.Cv = Cv;
  d.Ts_coef = std::pow(2.0, -a * ratio(b, n)) * pc * ratio(vc, Cv * (k - 1 + a));
  {
    Recipe r(SGX_EOS_DAVIS_PRODUCTS);
    const double p7[7] = {a, b, k, n, vc, pc, Cv};
    r.put(p7, sizeof(p7));
    e->recipe = std::move(r.bytes);
  }
  *out = e;
  return SGX_OK;
}

int sgx_eos_create_spiner_rhot(const sgx_sp5_tables *raw, int reproducibility_mode,
                               int pmin_vapor_dome, sgx_eos_t *out) {
  sgx_eos_s *e;
  int rc = new_handle(SGX_EOS_SPINER_RHOT, &e);
  if (rc) return rc;
  rc = build_spiner_rhot(raw, reproducibility_mode, pmin_vapor_dome, e);
  if (rc) {
    sgx_eos_destroy(e);
    *out = nullptr;
    return rc;
  }
  e->recipe = sp5_recipe(SGX_EOS_SPINER_RHOT, raw, reproducibility_mode, pmin_vapor_dome);
  *out = e;
  return SGX_OK;
}

int sgx_eos_create_spiner_rhosie(const sgx_sp5_tables *raw, int reproducibility_mode,
                                 int pmin_vapor_dome, sgx_eos_t *out) {
  sgx_eos_s *e;
  int rc = new_handle(SGX_EOS_SPINER_RHOSIE, &e);
  if (rc) return rc;
  rc = build_spiner_rhosie(raw, reproducibility_mode, pmin_vapor_dome, e);
  if (rc) {
    sgx_eos_destroy(e);
    *out = nullptr;
    return rc;
  }
  e->recipe = sp5_recipe(SGX_EOS_SPINER_RHOSIE, raw, reproducibility_mode, pmin_vapor_dome);
  *out = e;
  return SGX_OK;
}

void sgx_helm_default_options(sgx_helm_options *opt) {
  if (!opt) return;
  opt->enable_rad = opt->enable_gas = opt->enable_coulomb = 1;
  opt->gas_ionized = opt->gas_degenerate = 1;
  opt->use_newton_raphson = 1;
  opt->lambda_has_lT_slot = 1;
}

int sgx_eos_create_helmholtz(const char *table_path, const sgx_helm_options *opt,
                             sgx_eos_t *out) {
  sgx_eos_s *e;
  int rc = new_handle(SGX_EOS_HELMHOLTZ, &e);
  if (rc) return rc;
  rc = build_helmholtz(table_path, opt, e);
  if (rc) {
    sgx_eos_destroy(e);
    *out = nullptr;
    return rc;
  }
  *out = e;
  return SGX_OK;
}

// Modifier layers (include/sgx_b200.h).  The new handle refers to the UNMODIFIED model; the
// layers accumulate in ModParams in the reference's nesting order.
static int new_modified(sgx_eos_t base, int layer, sgx_eos_s **out) {
  if (!out) return fail(SGX_ERR_INVALID, "null output handle pointer");
  *out = nullptr;
  if (!base) return fail(SGX_ERR_INVALID, "null EOS handle");
  const sgx_eos_s *root = base->kind == SGX_EOS_MODIFIED ? base->base : base;
  if (root->kind < 0 || root->kind > SGX_EOS_SPINER_RHOSIE)
    return fail(SGX_ERR_UNSUPPORTED, "modifiers are built for IdealGas, Gruneisen, Davis and "
                "Spiner models (EOS kind %d given)", root->kind);
  const int have = base->kind == SGX_EOS_MODIFIED ? base->mod.flags : 0;
  if (have >= layer) // flags are ordered: shift 1 < scale 2 < ramp 4
    return fail(SGX_ERR_UNSUPPORTED, "modifier order: the reference's default variant nests "
                "BilinearRamp(Scaled(Shifted(T))), one layer of each at most "
                "(eos/eos_type_lists.hpp:69-94)");
  DeviceGuard guard;
  SGX_CUDA_OK(cudaSetDevice(base->device));
  sgx_eos_s *e = nullptr;
  const int rc = new_handle(SGX_EOS_MODIFIED, &e);
  if (rc) return rc;
  e->base = const_cast<sgx_eos_s *>(root);
  e->base->refs.fetch_add(1, std::memory_order_relaxed);
  if (base->kind == SGX_EOS_MODIFIED) e->mod = base->mod;
  e->mod.flags |= layer;
  *out = e;
  return SGX_OK;
}
// The byte stream of a modified handle: the header with kind SGX_EOS_MODIFIED, the layer flags,
// the arguments the three sgx_eos_modify_* calls were given, then the base model's own stream
// (the reference's wrappers do the same: their own members, then the wrapped EOS's dynamic
// memory, shifted_eos.hpp:381-392).  A base without a recipe leaves this one empty.
static void set_modified_recipe(sgx_eos_s *e) {
  e->recipe.clear();
  if (e->base->recipe.empty()) return;
  Recipe r(SGX_EOS_MODIFIED);
  const int32_t flags = e->mod.flags;
  const double a[6] = {e->mod.shift, e->mod.scale, e->mod.r0, e->mod.a, e->mod.b, e->mod.c};
  r.put(&flags, sizeof(flags));
  r.put(a, sizeof(a));
  r.put(e->base->recipe.data(), e->base->recipe.size());
  e->recipe = std::move(r.bytes);
}
int sgx_eos_modify_shifted(sgx_eos_t base, double shift, sgx_eos_t *out) {
  const int rc = new_modified(base, kModShift, out);
  if (rc) return rc;
  (*out)->mod.shift = shift;
  set_modified_recipe(*out);
  return SGX_OK;
}
int sgx_eos_modify_scaled(sgx_eos_t base, double scale, sgx_eos_t *out) {
  if (!(scale != 0.0)) return fail(SGX_ERR_INVALID, "ScaledEOS: scale must be non-zero");
  const int rc = new_modified(base, kModScale, out);
  if (rc) return rc;
  (*out)->mod.scale = scale;
  (*out)->mod.inv_scale = 1. / scale; // scaled_eos.hpp:45
  set_modified_recipe(*out);
  return SGX_OK;
}
int sgx_eos_modify_bilinear_ramp(sgx_eos_t base, double r0, double a, double b, double c,
                                 sgx_eos_t *out) {
  // BilinearRampEOS::CheckParams (ramps_eos.hpp:84-92)
  if (!(r0 > 0.0) || !(a > 0.0) || !(b >= 0.0) || a == b)
    return fail(SGX_ERR_INVALID, "BilinearRampEOS: needs r0 > 0, a > 0, b >= 0, a != b");
  const double rmid = r0 * (a - b * c) / (a - b), Pmid = a * (rmid / r0 - 1.0); // :80-81
  if (rmid != rmid || Pmid != Pmid)
    return fail(SGX_ERR_INVALID, "BilinearRampEOS: mid density / pressure not defined");
  const int rc = new_modified(base, kModRamp, out);
  if (rc) return rc;
  ModParams &q = (*out)->mod;
  q.r0 = r0, q.a = a, q.b = b, q.c = c, q.rmid = rmid, q.Pmid = Pmid;
  set_modified_recipe(*out);
  return SGX_OK;
}

int sgx_eos_retain(sgx_eos_t e) {
  if (!e) return fail(SGX_ERR_INVALID, "null EOS handle");
  e->refs.fetch_add(1, std::memory_order_relaxed);
  return SGX_OK;
}

int sgx_eos_destroy(sgx_eos_t e) {
  if (!e) return SGX_OK;
  if (e->refs.fetch_sub(1, std::memory_order_acq_rel) > 1) return SGX_OK; // still shared
  if (e->base) sgx_eos_destroy(e->base); // a modified handle kept its base alive
  DeviceGuard guard;
  cudaSetDevice(e->device);
  if (e->tables.dev) cudaFree(e->tables.dev);
  if (e->pinned) cudaFreeHost(e->pinned);
  if (e->dscratch) cudaFree(e->dscratch);
  if (e->root_stats) cudaFree(e->root_stats);
  if (e->dev_struct) cudaFree(e->dev_struct);
  if (e->own_stream) cudaStreamDestroy(e->own_stream);
  for (int k = 0; k < 4; ++k)
    if (e->host_streams[k]) cudaStreamDestroy(e->host_streams[k]);
  delete e;
  return SGX_OK;
}

int sgx_eos_type(sgx_eos_t e) { return e ? e->kind : SGX_ERR_INVALID; }
int sgx_eos_nlambda(sgx_eos_t e) {
  if (!e) return SGX_ERR_INVALID;
  if (e->kind == SGX_EOS_MODIFIED) e = e->base; // shifted_eos.hpp:366, scaled_eos.hpp:224
  switch (e->kind) {
  case SGX_EOS_SPINER_RHOT:
    return 2;
  case SGX_EOS_SPINER_RHOSIE:
    return 1;
  case SGX_EOS_HELMHOLTZ:
    return 3;
  default:
    return 0;
  }
}
unsigned long sgx_eos_preferred_input(sgx_eos_t e) {
  if (!e) return 0;
  if (e->kind == SGX_EOS_MODIFIED) e = e->base; // shifted_eos.hpp:372
  switch (e->kind) {
  case SGX_EOS_IDEAL:
    return IdealGas::preferred_input;
  case SGX_EOS_GRUNEISEN:
  case SGX_EOS_DAVIS_REACTANTS:
  case SGX_EOS_DAVIS_PRODUCTS:
    return TQ_DENSITY | TQ_SIE;
  default:
    return TQ_DENSITY | TQ_TEMPERATURE;
  }
}
uint64_t sgx_eos_scratch_bytes(sgx_eos_t, int, int64_t) { return 0; }
double sgx_eos_maximum_density(sgx_eos_t e) {
  if (!e) return 0;
  if (e->kind == SGX_EOS_MODIFIED) { // scaled_eos.hpp:259-261; the other layers forward
    const double m = sgx_eos_maximum_density(e->base);
    return (e->mod.flags & kModScale) ? e->mod.inv_scale * m : m;
  }
  if (e->kind == SGX_EOS_GRUNEISEN) return e->analytic.gr.rho_max;
  if (e->kind == SGX_EOS_HELMHOLTZ) return 1e15;
  return 1e100;
}
int64_t sgx_eos_get_table(sgx_eos_t e, const char *name, double *out, int64_t cap) {
  if (!e || !name) return fail(SGX_ERR_INVALID, "null argument");
  auto it = e->tables.index.find(name);
  if (it == e->tables.index.end()) return fail(SGX_ERR_INVALID, "no table named '%s'", name);
  const int64_t n = (int64_t)it->second.second;
  if (!out) return n;
  const int64_t m = n < cap ? n : cap;
  // read back from the DEVICE copy: this is what the kernels see
  DeviceGuard guard;
  if (check_handle(e)) return SGX_ERR_CUDA;
  cudaError_t err = cudaMemcpy(out, e->tables.dev + it->second.first, m * sizeof(double),
                               cudaMemcpyDeviceToHost);
  if (err != cudaSuccess) return fail(SGX_ERR_CUDA, "cudaMemcpy: %s", cudaGetErrorString(err));
  return m;
}
int sgx_eos_get_scalar(sgx_eos_t e, const char *name, double *out) {
  if (!e || !name || !out) return fail(SGX_ERR_INVALID, "null argument");
  auto it = e->tables.scalars.find(name);
  if (it == e->tables.scalars.end())
    return fail(SGX_ERR_INVALID, "no scalar named '%s'", name);
  *out = it->second;
  return SGX_OK;
}

int sgx_eos_eval(sgx_eos_t e, int method, const double *in1, const double *in2, double *out,
                 int64_t n, double *lambda, int lambda_stride, sgx_stream_t stream) {
  DeviceGuard guard;
  int rc = check_handle(e);
  if (rc) return rc;
  if (n < 0) return fail(SGX_ERR_INVALID, "negative n");
  if (n == 0) return SGX_OK;
  if (!in1 || !in2 || !out) return fail(SGX_ERR_INVALID, "null array");
  if (lambda && lambda_stride < lambda_need(e))
    return fail(SGX_ERR_INVALID, "lambda_stride %d < %d slots this EOS uses", lambda_stride,
                lambda_need(e));
  cudaStream_t s = (cudaStream_t)stream;
  switch (e->kind) {
  case SGX_EOS_IDEAL:
    return launch_eval(e->analytic.ideal, method, in1, in2, out, n, nullptr, 0, s, 8);
  case SGX_EOS_GRUNEISEN:
    return launch_eval(e->analytic.gr, method, in1, in2, out, n, nullptr, 0, s, 8);
  case SGX_EOS_DAVIS_REACTANTS:
    return launch_eval(e->analytic.dr, method, in1, in2, out, n, nullptr, 0, s, 8);
  case SGX_EOS_DAVIS_PRODUCTS:
    return launch_eval(e->analytic.dp, method, in1, in2, out, n, nullptr, 0, s, 8);
  case SGX_EOS_SPINER_RHOT:
    // config 3a: direct inversion where it provably matches the reference iterate, the iterate
    // itself elsewhere (sgx_spiner_root.cu); SGX_ROOT_MODE=exact keeps the iterate for all cells
    if (method == SGX_T_FROM_RHO_E && root_mode() != 1)
      return launch_rhot_T_from_rho_e_hybrid(e, root_mode() == 0, in1, in2, out, n, lambda,
                                             lambda_stride, s);
    return launch_eval(e->rhot, method, in1, in2, out, n, lambda, lambda_stride, s, 8);
  case SGX_EOS_SPINER_RHOSIE:
    return launch_eval(e->rhosie, method, in1, in2, out, n, lambda, lambda_stride, s, 8);
  case SGX_EOS_HELMHOLTZ:
    return launch_helm_eval(e, method, in1, in2, out, n, lambda, lambda_stride, s);
  case SGX_EOS_MODIFIED:
    return launch_mod_eval(e, method, in1, in2, out, n, lambda, lambda_stride, s);
  default:
    return fail(SGX_ERR_INVALID, "bad EOS kind %d", e->kind);
  }
}

int sgx_PressureFromDensityTemperature(sgx_eos_t e, const double *rhos, const double *temps,
                                       double *out, int64_t n, double *lambda, int ls,
                                       sgx_stream_t s) {
  return sgx_eos_eval(e, SGX_P_FROM_RHO_T, rhos, temps, out, n, lambda, ls, s);
}
int sgx_InternalEnergyFromDensityTemperature(sgx_eos_t e, const double *rhos,
                                             const double *temps, double *out, int64_t n,
                                             double *lambda, int ls, sgx_stream_t s) {
  return sgx_eos_eval(e, SGX_E_FROM_RHO_T, rhos, temps, out, n, lambda, ls, s);
}
int sgx_TemperatureFromDensityInternalEnergy(sgx_eos_t e, const double *rhos,
                                             const double *sies, double *out, int64_t n,
                                             double *lambda, int ls, sgx_stream_t s) {
  return sgx_eos_eval(e, SGX_T_FROM_RHO_E, rhos, sies, out, n, lambda, ls, s);
}
int sgx_PressureFromDensityInternalEnergy(sgx_eos_t e, const double *rhos, const double *sies,
                                          double *out, int64_t n, double *lambda, int ls,
                                          sgx_stream_t s) {
  return sgx_eos_eval(e, SGX_P_FROM_RHO_E, rhos, sies, out, n, lambda, ls, s);
}

extern "C++" {
namespace sgx {
int launch_mod_pe(sgx_eos_s *e, const double *rho, const double *T, double *P, double *E,
                  int64_t n, double *lambda, int ls, cudaStream_t s);
}
}
int sgx_PressureEnergyFromDensityTemperature(sgx_eos_t e, const double *rhos,
                                             const double *temps, double *P, double *E,
                                             int64_t n, double *lambda, int ls,
                                             sgx_stream_t stream) {
  DeviceGuard guard;
  int rc = check_handle(e);
  if (rc) return rc;
  if (n <= 0) return n == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative n");
  if (!rhos || !temps || !P || !E) return fail(SGX_ERR_INVALID, "null array");
  if (lambda && ls < lambda_need(e))
    return fail(SGX_ERR_INVALID, "lambda_stride %d < %d slots this EOS uses", ls, lambda_need(e));
  cudaStream_t s = (cudaStream_t)stream;
  switch (e->kind) {
  case SGX_EOS_IDEAL:
    return launch_pe(e->analytic.ideal, rhos, temps, P, E, n, nullptr, 0, s);
  case SGX_EOS_GRUNEISEN:
    return launch_pe(e->analytic.gr, rhos, temps, P, E, n, nullptr, 0, s);
  case SGX_EOS_DAVIS_REACTANTS:
    return launch_pe(e->analytic.dr, rhos, temps, P, E, n, nullptr, 0, s);
  case SGX_EOS_DAVIS_PRODUCTS:
    return launch_pe(e->analytic.dp, rhos, temps, P, E, n, nullptr, 0, s);
  case SGX_EOS_SPINER_RHOT:
    return launch_pe(e->rhot, rhos, temps, P, E, n, lambda, ls, s);
  case SGX_EOS_SPINER_RHOSIE:
    return launch_pe(e->rhosie, rhos, temps, P, E, n, lambda, ls, s);
  case SGX_EOS_MODIFIED:
    return launch_mod_pe(e, rhos, temps, P, E, n, lambda, ls, s);
  default:
    return fail(SGX_ERR_UNSUPPORTED, "fused P+e not available for EOS kind %d", e->kind);
  }
}

extern "C++" {
namespace sgx {
int launch_mod_fill(sgx_eos_s *e, double *rho, double *T, double *en, double *P, double *cv,
                    double *bmod, int64_t n, unsigned long output, double *lambda, int ls,
                    cudaStream_t s);
}
}
int sgx_FillEos(sgx_eos_t e, double *rhos, double *temps, double *energies, double *presses,
                double *cvs, double *bmods, int64_t n, unsigned long output, double *lambda,
                int ls, sgx_stream_t stream) {
  DeviceGuard guard;
  int rc = check_handle(e);
  if (rc) return rc;
  if (n <= 0) return n == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative n");
  if (output == 0) return fail(SGX_ERR_UNSUPPORTED, "FillEos: output == thermalqs::none");
  // (Helmholtz checks its own row width: 2 or 3 slots, launch_helm_fill)
  if (lambda && ls < lambda_need(e))
    return fail(SGX_ERR_INVALID, "lambda_stride %d < %d slots this EOS uses", ls, lambda_need(e));
  cudaStream_t s = (cudaStream_t)stream;
  switch (e->kind) {
  case SGX_EOS_IDEAL:
    return launch_fill(e, e->analytic.ideal, rhos, temps, energies, presses, cvs, bmods, n,
                       output, nullptr, 0, s);
  case SGX_EOS_GRUNEISEN:
    return launch_fill(e, e->analytic.gr, rhos, temps, energies, presses, cvs, bmods, n,
                       output, nullptr, 0, s);
  case SGX_EOS_DAVIS_REACTANTS:
    return launch_fill(e, e->analytic.dr, rhos, temps, energies, presses, cvs, bmods, n,
                       output, nullptr, 0, s);
  case SGX_EOS_DAVIS_PRODUCTS:
    return launch_fill(e, e->analytic.dp, rhos, temps, energies, presses, cvs, bmods, n,
                       output, nullptr, 0, s);
  case SGX_EOS_SPINER_RHOT:
    return launch_fill(e, e->rhot, rhos, temps, energies, presses, cvs, bmods, n, output,
                       lambda, ls, s);
  case SGX_EOS_SPINER_RHOSIE:
    return launch_fill(e, e->rhosie, rhos, temps, energies, presses, cvs, bmods, n, output,
                       lambda, ls, s);
  case SGX_EOS_HELMHOLTZ:
    return launch_helm_fill(e, rhos, temps, energies, presses, cvs, bmods, n, output, lambda,
                            ls, s);
  case SGX_EOS_MODIFIED: { // the wrappers' FillEos (csrc/sgx_modifiers.cu)
    const unsigned long inner = (e->mod.flags & kModRamp) ? (output & ~(unsigned long)TQ_PRESSURE)
                                                           : output;
    if (inner != 0)
      if (const char *why = fill_mask_error(e->base->kind, inner))
        return fail(SGX_ERR_UNSUPPORTED, "%s", why);
    return launch_mod_fill(e, rhos, temps, energies, presses, cvs, bmods, n, output, lambda, ls,
                           s);
  }
  default:
    return fail(SGX_ERR_INVALID, "bad EOS kind %d", e->kind);
  }
}

// ---- host-pointer forms ---------------------------------------------------------------
// Which way the lambda rows travel for a vector method: `in` = the kernel reads them (or leaves
// some slots untouched, so the caller's values must survive the round trip), `out` = it writes.
static void lambda_io(const sgx_eos_s *e, int method, int ls, bool &in, bool &out) {
  const bool from_T = method == SGX_P_FROM_RHO_T || method == SGX_E_FROM_RHO_T ||
                      method == SGX_CV_FROM_RHO_T || method == SGX_BMOD_FROM_RHO_T ||
                      method == SGX_GAMMA_FROM_RHO_T;
  switch (e->kind) {
  case SGX_EOS_SPINER_RHOT:
    // (rho,T) lookups only write (lRho,lT); every *_FROM_RHO_E method reads lambda[1] as the
    // root-find guess (eos_spiner_rho_temp.hpp:1218-1223) before writing both slots.  Slots past
    // nlambda are never touched by the kernel, so a wider row must make the round trip.
    in = !from_T || ls != 2;
    out = true;
    break;
  case SGX_EOS_SPINER_RHOSIE:
    // lambda[0] = lRho is written, never read -- except GruneisenParamFromDensityInternalEnergy,
    // which does not touch lambda at all (eos_spiner_rho_sie.hpp:757-761): nothing crosses the bus
    // and the caller's rows stay as they are.
    if (method == SGX_GAMMA_FROM_RHO_E) {
      in = out = false;
    } else {
      in = ls != 1;
      out = true;
    }
    break;
  case SGX_EOS_HELMHOLTZ: // Abar, Zbar (and lT) in, lT out
    in = true;
    out = true;
    break;
  default: // analytic models: nlambda == 0
    in = false;
    out = false;
  }
}

// Host-pointer entry: the cell range is cut into chunks that flow through kHostSlots device
// slots, each on its own stream (H2D -> kernel -> D2H), so the two PCIe directions and the
// kernel overlap.  Only the lambda traffic the method needs crosses the bus.
static constexpr int kHostSlots = 4;
static constexpr int64_t kHostChunk = 1 << 21; // cells per chunk (16 MiB per array)

// On any failure inside a chunk loop the slot streams are drained before the error is returned,
// so no asynchronous copy into the caller's buffers is still pending when the call comes back.
#define SGX_PIPE_TRY(expr)                                                                 \
  do {                                                                                     \
    cudaError_t err__ = (expr);                                                            \
    if (err__ != cudaSuccess) {                                                            \
      rc = fail(SGX_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(err__));                 \
      goto drain;                                                                          \
    }                                                                                      \
  } while (0)

static int drain_host_streams(sgx_eos_s *e, int rc) {
  for (int k = 0; k < kHostSlots; ++k) {
    if (!e->host_streams[k]) continue;
    cudaError_t err = cudaStreamSynchronize(e->host_streams[k]);
    if (err != cudaSuccess && rc == SGX_OK)
      rc = fail(SGX_ERR_CUDA, "cudaStreamSynchronize: %s", cudaGetErrorString(err));
  }
  return rc;
}

int sgx_eos_eval_host(sgx_eos_t e, int method, const double *in1, const double *in2,
                      double *out, int64_t n, double *lambda, int ls) {
  DeviceGuard guard;
  int rc = check_handle(e);
  if (rc) return rc;
  if (n <= 0) return n == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative n");
  if (!in1 || !in2 || !out) return fail(SGX_ERR_INVALID, "null array");
  if (lambda && ls < lambda_need(e))
    return fail(SGX_ERR_INVALID, "lambda_stride %d < %d slots this EOS uses", ls, lambda_need(e));
  bool lam_in = false, lam_out = false;
  if (lambda) lambda_io(e, method, ls, lam_in, lam_out);
  const bool lam_dev = lam_in || lam_out; // rows that never cross the bus are not given to the kernel
  const int64_t chunk = n < kHostChunk ? n : kHostChunk;
  const size_t per_cell = (size_t)(3 + (lam_dev ? ls : 0)) * sizeof(double);
  const size_t slot_bytes = ((size_t)chunk * per_cell + 255) & ~size_t(255);
  rc = ensure_staging(e, 0, kHostSlots * slot_bytes);
  if (rc) return rc;
  rc = ensure_host_streams(e);
  if (rc) return rc;
  {
    HostPin pin_in1(in1, (size_t)n * sizeof(double)), pin_in2(in2, (size_t)n * sizeof(double)),
        pin_out(out, (size_t)n * sizeof(double)),
        pin_lam(lam_dev ? lambda : nullptr, (size_t)n * ls * sizeof(double));
    int slot = 0;
    for (int64_t i0 = 0; i0 < n; i0 += chunk, slot = (slot + 1) % kHostSlots) {
      const int64_t m = (n - i0) < chunk ? (n - i0) : chunk;
      const size_t nb = (size_t)m * sizeof(double);
      cudaStream_t s = e->host_streams[slot];
      double *d1 = (double *)((char *)e->dscratch + slot * slot_bytes);
      double *d2 = d1 + chunk, *dout = d2 + chunk;
      double *dl = lam_dev ? dout + chunk : nullptr;
      double *hl = lam_dev ? lambda + (size_t)i0 * ls : nullptr;
      SGX_PIPE_TRY(cudaMemcpyAsync(d1, in1 + i0, nb, cudaMemcpyHostToDevice, s));
      SGX_PIPE_TRY(cudaMemcpyAsync(d2, in2 + i0, nb, cudaMemcpyHostToDevice, s));
      if (lam_in) SGX_PIPE_TRY(cudaMemcpyAsync(dl, hl, nb * ls, cudaMemcpyHostToDevice, s));
      rc = sgx_eos_eval(e, method, d1, d2, dout, m, dl, ls, s);
      if (rc) goto drain;
      SGX_PIPE_TRY(cudaMemcpyAsync(out + i0, dout, nb, cudaMemcpyDeviceToHost, s));
      if (lam_out) SGX_PIPE_TRY(cudaMemcpyAsync(hl, dl, nb * ls, cudaMemcpyDeviceToHost, s));
    }
  drain:
    rc = drain_host_streams(e, rc);
  }
  return rc;
}

int sgx_FillEos_host(sgx_eos_t e, double *rhos, double *temps, double *energies,
                     double *presses, double *cvs, double *bmods, int64_t n,
                     unsigned long output, double *lambda, int ls) {
  DeviceGuard guard;
  int rc = check_handle(e);
  if (rc) return rc;
  if (n <= 0) return n == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative n");
  if (lambda && ls < lambda_need(e))
    return fail(SGX_ERR_INVALID, "lambda_stride %d < %d slots this EOS uses", ls, lambda_need(e));
  // same chunk pipeline as sgx_eos_eval_host: kHostSlots device slots, one stream each
  const int64_t chunk = n < kHostChunk ? n : kHostChunk;
  const size_t per_cell = (size_t)(6 + (lambda ? ls : 0)) * sizeof(double);
  const size_t slot_bytes = ((size_t)chunk * per_cell + 255) & ~size_t(255);
  rc = ensure_staging(e, 0, kHostSlots * slot_bytes);
  if (rc) return rc;
  rc = ensure_host_streams(e);
  if (rc) return rc;
  double *h[6] = {rhos, temps, energies, presses, cvs, bmods};
  const unsigned long bits[6] = {TQ_DENSITY, TQ_TEMPERATURE, TQ_SIE, TQ_PRESSURE, TQ_CV, TQ_BMOD};
  {
    HostPin pins[6] = {HostPin(h[0], (size_t)n * 8), HostPin(h[1], (size_t)n * 8),
                       HostPin(h[2], (size_t)n * 8), HostPin(h[3], (size_t)n * 8),
                       HostPin(h[4], (size_t)n * 8), HostPin(h[5], (size_t)n * 8)};
    HostPin pin_lam(lambda, (size_t)n * ls * sizeof(double));
    int slot = 0;
    for (int64_t i0 = 0; i0 < n; i0 += chunk, slot = (slot + 1) % kHostSlots) {
      const int64_t m = (n - i0) < chunk ? (n - i0) : chunk;
      const size_t nb = (size_t)m * sizeof(double);
      cudaStream_t s = e->host_streams[slot];
      double *base = (double *)((char *)e->dscratch + slot * slot_bytes);
      double *d[6];
      for (int k = 0; k < 6; ++k)
        d[k] = h[k] ? base + (size_t)k * chunk : nullptr;
      double *dl = lambda ? base + 6 * (size_t)chunk : nullptr;
      double *hl = lambda ? lambda + (size_t)i0 * ls : nullptr;
      for (int k = 0; k < 4; ++k) // cv / bmod are never inputs
        if (h[k] && !(output & bits[k]))
          SGX_PIPE_TRY(cudaMemcpyAsync(d[k], h[k] + i0, nb, cudaMemcpyHostToDevice, s));
      if (lambda) SGX_PIPE_TRY(cudaMemcpyAsync(dl, hl, nb * ls, cudaMemcpyHostToDevice, s));
      rc = sgx_FillEos(e, d[0], d[1], d[2], d[3], d[4], d[5], m, output, dl, ls, s);
      if (rc) goto drain;
      for (int k = 0; k < 6; ++k)
        if (h[k] && (output & bits[k]))
          SGX_PIPE_TRY(cudaMemcpyAsync(h[k] + i0, d[k], nb, cudaMemcpyDeviceToHost, s));
      if (lambda) SGX_PIPE_TRY(cudaMemcpyAsync(hl, dl, nb * ls, cudaMemcpyDeviceToHost, s));
    }
  drain:
    rc = drain_host_streams(e, rc);
  }
  return rc;
}
#undef SGX_PIPE_TRY

// ---- Serialize / DeSerialize (eos/eos_base.hpp:756-798) -----------------------------------
uint64_t sgx_eos_serialized_size(sgx_eos_t e) { return e ? (uint64_t)e->recipe.size() : 0; }
int sgx_eos_serialize(sgx_eos_t e, char *dst, uint64_t cap) {
  if (!e || !dst) return fail(SGX_ERR_INVALID, "null argument");
  if (e->recipe.empty()) return fail(SGX_ERR_UNSUPPORTED, "this handle carries no recipe");
  if (cap < e->recipe.size())
    return fail(SGX_ERR_INVALID, "buffer of %llu bytes, need %zu", (unsigned long long)cap,
                e->recipe.size());
  std::memcpy(dst, e->recipe.data(), e->recipe.size());
  return SGX_OK;
}
int sgx_eos_deserialize(const char *src, uint64_t nbytes, sgx_eos_t *out) {
  if (!src || !out) return fail(SGX_ERR_INVALID, "null argument");
  *out = nullptr;
  RecipeReader rd(src, (size_t)nbytes);
  char magic[8];
  int32_t kind = -1;
  if (!rd.get(magic, 8) || std::memcmp(magic, "SGXEOS01", 8) != 0 || !rd.get(&kind, 4))
    return fail(SGX_ERR_INVALID, "not an sgx EOS byte stream");
  double a[11];
  switch (kind) {
  case SGX_EOS_IDEAL:
    if (!rd.get(a, 2 * sizeof(double))) break;
    return sgx_eos_create_ideal(a[0], a[1], out);
  case SGX_EOS_GRUNEISEN:
    if (!rd.get(a, 11 * sizeof(double))) break;
    return sgx_eos_create_gruneisen(a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], a[8], a[9],
                                    a[10], out);
  case SGX_EOS_DAVIS_REACTANTS:
    if (!rd.get(a, 11 * sizeof(double))) break;
    return sgx_eos_create_davis_reactants(a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], a[8],
                                          a[9], a[10], out);
  case SGX_EOS_DAVIS_PRODUCTS:
    if (!rd.get(a, 7 * sizeof(double))) break;
    return sgx_eos_create_davis_products(a[0], a[1], a[2], a[3], a[4], a[5], a[6], out);
  case SGX_EOS_SPINER_RHOT:
  case SGX_EOS_SPINER_RHOSIE: {
    int32_t flags[2];
    sgx_sp5_tables raw;
    if (!rd.get(flags, sizeof(flags)) || !rd.get(&raw, sizeof(raw))) break;
    std::vector<double> store[21];
    const double **members[21] = {
        &raw.P,       &raw.sie,       &raw.bMod,       &raw.dPdRho,    &raw.dPdE,
        &raw.dTdRho,  &raw.dTdE,      &raw.dEdRho,     &raw.dEdT,      &raw.PCold,
        &raw.sieCold, &raw.bModCold,  &raw.dPdRhoCold, &raw.T_re,      &raw.P_re,
        &raw.bMod_re, &raw.dPdRho_re, &raw.dPdE_re,    &raw.dTdRho_re, &raw.dTdE_re,
        &raw.dEdRho_re};
    bool good = true;
    for (int k = 0; k < 21 && good; ++k)
      good = rd.get_array(store[k], *members[k]);
    if (!good) break;
    return kind == SGX_EOS_SPINER_RHOT
               ? sgx_eos_create_spiner_rhot(&raw, flags[0], flags[1], out)
               : sgx_eos_create_spiner_rhosie(&raw, flags[0], flags[1], out);
  }
  case SGX_EOS_HELMHOLTZ: {
    sgx_helm_options o;
    std::vector<double> store;
    const double *raw = nullptr;
    if (!rd.get(&o, sizeof(o)) || !rd.get_array(store, raw)) break;
    sgx_eos_s *e;
    int rc = new_handle(SGX_EOS_HELMHOLTZ, &e);
    if (rc) return rc;
    rc = build_helmholtz_raw(raw, store.size(), &o, e);
    if (rc) {
      sgx_eos_destroy(e);
      return rc;
    }
    *out = e;
    return SGX_OK;
  }
  case SGX_EOS_MODIFIED: { // the base model from the rest of the stream, then the layers again
    int32_t flags = 0;
    if (!rd.get(&flags, sizeof(flags)) || !rd.get(a, 6 * sizeof(double))) break;
    if (flags <= 0 || (flags & ~(kModShift | kModScale | kModRamp)))
      return fail(SGX_ERR_INVALID, "modifier flags %d in the byte stream", flags);
    sgx_eos_t cur = nullptr;
    int rc = sgx_eos_deserialize(rd.p, (uint64_t)(rd.end - rd.p), &cur);
    if (rc) return rc;
    for (int layer = kModShift; layer <= kModRamp && !rc; layer <<= 1) {
      if (!(flags & layer)) continue;
      sgx_eos_t next = nullptr;
      rc = layer == kModShift   ? sgx_eos_modify_shifted(cur, a[0], &next)
           : layer == kModScale ? sgx_eos_modify_scaled(cur, a[1], &next)
                                : sgx_eos_modify_bilinear_ramp(cur, a[2], a[3], a[4], a[5], &next);
      sgx_eos_destroy(cur); // the new handle keeps the base model alive
      cur = next;
    }
    if (rc) return rc; // `cur` is null: the failed call stored nothing
    *out = cur;
    return SGX_OK;
  }
  default:
    return fail(SGX_ERR_INVALID, "unknown EOS kind %d in the byte stream", kind);
  }
  return fail(SGX_ERR_INVALID, "truncated EOS byte stream");
}

int sgx_eos_root_stats(sgx_eos_t e, uint64_t out[2], int reset) {
  if (!e || !out) return fail(SGX_ERR_INVALID, "null argument");
  out[0] = out[1] = 0;
  if (!e->root_stats) return SGX_OK;
  DeviceGuard guard;
  int rc = check_handle(e);
  if (rc) return rc;
  SGX_CUDA_OK(cudaMemcpy(out, e->root_stats, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  if (reset) SGX_CUDA_OK(cudaMemset(e->root_stats, 0, 2 * sizeof(uint64_t)));
  return SGX_OK;
}

// ---- building blocks ------------------------------------------------------------------
int sgx_fastmath_eval(int which, const double *x, double *y, int64_t n, sgx_stream_t stream) {
  if (n <= 0) return n == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative n");
  if (!x || !y) return fail(SGX_ERR_INVALID, "null array");
  const LaunchCfg c = stream_cfg(n, 256, 8);
  fastmath_kernel<<<c.blocks, c.threads, 0, (cudaStream_t)stream>>>(which, x, y, n);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}
int sgx_solve_Ax_b(int n, double *a, double *b, int *ok, int64_t batch, sgx_stream_t stream) {
  if (n < 1 || n > 6 || n == 5) return fail(SGX_ERR_INVALID, "n must be 1..4 or 6");
  if (batch <= 0) return batch == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative batch");
  const int threads = 128;
  const int blocks = (int)((batch + threads - 1) / threads);
  solve_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(n, a, b, ok, batch);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

} // extern "C"
