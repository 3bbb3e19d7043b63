// sgx_tune_lookup.cu -- occupancy / vector-width study of the headline lookup
// (SpinerEOSDependsRhoT::PressureFromDensityTemperature + lambda write-back).
// Not a product entry point: `sgx_tune_lookup` runs ONE of the kernel shapes below on device
// arrays so that tools/studies/lookup_shapes.py can time them side by side on a B200.  The
// arithmetic is prep()/finish() of sgx_spiner.cuh in every shape, so all of them return the
// same bits as sgx_eos_eval (checked by the study script before it times anything).
#include "sgx_internal.hpp"

namespace sgx {
namespace {

struct D4 {
  double x, y, z, w;
};
SGX_D D4 ld4(const double *p) {
  D4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
               : "l"(p));
  return v;
}
SGX_D double2 ld2(const double *p) {
  double2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];"
               : "=d"(v.x), "=d"(v.y)
               : "l"(p));
  return v;
}
SGX_D double ld1(const double *p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
SGX_D void st4(double *p, double x, double y, double z, double w) {
  asm volatile("st.global.cs.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(x), "d"(y), "d"(z),
               "d"(w)
               : "memory");
}
SGX_D void st2(double *p, double x, double y) {
  asm volatile("st.global.cs.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(x), "d"(y) : "memory");
}
SGX_D void st1(double *p, double x) {
  asm volatile("st.global.cs.f64 [%0], %1;" ::"l"(p), "d"(x) : "memory");
}

// VW cells per thread per trip, MINB resident blocks of 256 threads, PF = software prefetch of
// the next trip's inputs, SINGLE = both axes are single-patch grids (told to the compiler).
template <int VW, int MINB, bool PF, bool SINGLE, bool LAMBDA>
__global__ void __launch_bounds__(256, MINB)
    lookup_shape_kernel(const __grid_constant__ SpinerRhoT eos, const double *__restrict__ rho,
                        const double *__restrict__ T, double *__restrict__ out,
                        const int64_t n, double *__restrict__ lambda) {
  if (SINGLE) {
    __builtin_assume(eos.gRho.ngrids == 1);
    __builtin_assume(eos.gT.ngrids == 1);
  }
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  const int64_t ngroup = n / VW; // the study passes n % VW == 0
  auto load_group = [&](const int64_t g, double (&a)[VW], double (&b)[VW]) {
    const int64_t i = g * VW;
    if constexpr (VW == 4) {
      const D4 va = ld4(rho + i), vb = ld4(T + i);
      a[0] = va.x, a[1] = va.y, a[2] = va.z, a[3] = va.w;
      b[0] = vb.x, b[1] = vb.y, b[2] = vb.z, b[3] = vb.w;
    } else if constexpr (VW == 2) {
      const double2 va = ld2(rho + i), vb = ld2(T + i);
      a[0] = va.x, a[1] = va.y;
      b[0] = vb.x, b[1] = vb.y;
    } else {
      a[0] = ld1(rho + i);
      b[0] = ld1(T + i);
    }
  };
  double a[VW], b[VW];
  int64_t g = tid;
  if (g < ngroup) load_group(g, a, b);
  while (g < ngroup) {
    const int64_t i = g * VW;
    double an[VW], bn[VW];
    const int64_t gn = g + nthr;
    if (PF && gn < ngroup) load_group(gn, an, bn);
    double r[VW], lam[VW][2];
    LookupPrep q[VW];
#pragma unroll
    for (int c = 0; c < VW; ++c)
      eos.template prep<SGX_P_FROM_RHO_T>(a[c], b[c], q[c]);
#pragma unroll
    for (int c = 0; c < VW; ++c)
      r[c] = eos.template finish<SGX_P_FROM_RHO_T>(a[c], b[c], q[c], LAMBDA ? lam[c] : nullptr);
    if constexpr (VW == 4) {
      st4(out + i, r[0], r[1], r[2], r[3]);
      if (LAMBDA) {
        st4(lambda + i * 2, lam[0][0], lam[0][1], lam[1][0], lam[1][1]);
        st4(lambda + i * 2 + 4, lam[2][0], lam[2][1], lam[3][0], lam[3][1]);
      }
    } else if constexpr (VW == 2) {
      st2(out + i, r[0], r[1]);
      if (LAMBDA) st4(lambda + i * 2, lam[0][0], lam[0][1], lam[1][0], lam[1][1]);
    } else {
      st1(out + i, r[0]);
      if (LAMBDA) st2(lambda + i * 2, lam[0][0], lam[0][1]);
    }
    g = gn;
    if constexpr (PF) {
#pragma unroll
      for (int c = 0; c < VW; ++c) {
        a[c] = an[c];
        b[c] = bn[c];
      }
    } else {
      if (g < ngroup) load_group(g, a, b);
    }
  }
}

template <int VW, int MINB, bool PF, bool SINGLE>
void launch_shape(const SpinerRhoT &eos, const double *rho, const double *T, double *out,
                  int64_t n, double *lambda, int bps, cudaStream_t s) {
  const LaunchCfg c = stream_cfg(n / VW, 256, bps > 0 ? bps : MINB);
  if (lambda)
    lookup_shape_kernel<VW, MINB, PF, SINGLE, true><<<c.blocks, c.threads, 0, s>>>(eos, rho, T,
                                                                                  out, n, lambda);
  else
    lookup_shape_kernel<VW, MINB, PF, SINGLE, false><<<c.blocks, c.threads, 0, s>>>(
        eos, rho, T, out, n, lambda);
}

} // namespace
} // namespace sgx

// variant: 0 = 4 cells/thread, 2 blocks, prefetch (the product kernel's shape)
//          1 = 2 cells, 4 blocks, prefetch      2 = 2 cells, 4 blocks, no prefetch
//          3 = 4 cells, 3 blocks, no prefetch   4 = 2 cells, 3 blocks, prefetch
//          5 = 1 cell, 4 blocks, prefetch       6 = 1 cell, 6 blocks, no prefetch
//          +10: the same with single-patch grids assumed at compile time
// blocks_per_sm <= 0: grid = SMs x resident blocks.
extern "C" int sgx_tune_lookup(sgx_eos_t e, int variant, const double *rho, const double *T,
                               double *out, int64_t n, double *lambda, int blocks_per_sm,
                               sgx_stream_t stream) {
  using namespace sgx;
  if (!e || e->kind != SGX_EOS_SPINER_RHOT) return fail(SGX_ERR_INVALID, "needs a SpinerRhoT");
  if (n % 4) return fail(SGX_ERR_INVALID, "study sizes are multiples of 4");
  DeviceGuard guard;
  cudaSetDevice(e->device);
  cudaStream_t s = (cudaStream_t)stream;
  const bool single = variant >= 10;
  if (single && (e->rhot.gRho.ngrids != 1 || e->rhot.gT.ngrids != 1))
    return fail(SGX_ERR_INVALID, "variant >= 10 needs single-patch grids");
#define SGX_SHAPE(V, VW, MINB, PF)                                                           \
  case V:                                                                                    \
    launch_shape<VW, MINB, PF, false>(e->rhot, rho, T, out, n, lambda, blocks_per_sm, s);    \
    break;                                                                                   \
  case V + 10:                                                                               \
    launch_shape<VW, MINB, PF, true>(e->rhot, rho, T, out, n, lambda, blocks_per_sm, s);     \
    break;
  switch (variant) {
    SGX_SHAPE(0, 4, 2, true)
    SGX_SHAPE(1, 2, 4, true)
    SGX_SHAPE(2, 2, 4, false)
    SGX_SHAPE(3, 4, 3, false)
    SGX_SHAPE(4, 2, 3, true)
    SGX_SHAPE(5, 1, 4, true)
    SGX_SHAPE(6, 1, 6, false)
  default:
    return fail(SGX_ERR_INVALID, "unknown lookup shape %d", variant);
  }
#undef SGX_SHAPE
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}
