// The launcher of the lane-refill state machine (sgx_pte_sm.cuh), in a header so that the
// analytic accessors (sgx_pte_rhot.cu) and the tabulated / modified ones (sgx_pte_tab_sm*.cu)
// instantiate it in translation units of their own.
#ifndef SGX_PTE_RHOT_LAUNCH_CUH_
#define SGX_PTE_RHOT_LAUNCH_CUH_
#include "sgx_internal.hpp"
#include "sgx_pte_kernel.cuh"
#include "sgx_pte_sm.cuh"

namespace sgx {

// once per device: cudaLimitStackSize -> 2048 B if lower (sgx_pte_rhot.cu)
void raise_stack_limit_once();

// Cells go through in chunks: init kernel -> scratch record -> state-machine kernel, both on
// the caller's stream; the scratch is stream-ordered (scratch_alloc_async, the library's own pool): (2*NM+1)*8 B/cell for the
// hand-over record plus one 64-bit cell cursor per chunk.
constexpr int64_t kPteChunk = int64_t(1) << 23;

template <int NM, class Acc, int THREADS = kSmThreads, int MINB = 1>
inline int launch_sm(const EosSet &set, const MixParamsDev &prm, double *rho, double *vfrac,
                     double *sie, double *temp, double *press, const double *sie_tot,
                     const double *vfrac_tot, const double *Tguess, uint64_t *status,
                     double *residual, uint64_t *stats, int64_t ncell, cudaStream_t s,
                     double *lambda = nullptr, const unsigned char *skip = nullptr) {
  const int64_t chunk = ncell < kPteChunk ? ncell : kPteChunk;
  const int64_t nchunk = (ncell + chunk - 1) / chunk;
  const size_t cursor_bytes = ((size_t)nchunk * sizeof(unsigned long long) + 255) & ~size_t(255);
  char *scratch = nullptr;
  SGX_CUDA_OK(scratch_alloc_async(
      reinterpret_cast<void **>(&scratch),
      cursor_bytes + (size_t)chunk * (2 * NM + 1) * sizeof(double), s));
  unsigned long long *cursors = reinterpret_cast<unsigned long long *>(scratch);
  double *init = reinterpret_cast<double *>(scratch + cursor_bytes);
  cudaError_t err = cudaMemsetAsync(cursors, 0, cursor_bytes, s);
  int64_t k = 0;
  for (int64_t c0 = 0; c0 < ncell && err == cudaSuccess; c0 += chunk, ++k) {
    const int64_t n = (ncell - c0) < chunk ? (ncell - c0) : chunk;
    const int64_t o = c0 * NM;
    const LaunchCfg ci = stream_cfg(n, 256, 4);
    pte_rhot_init_kernel<NM, Acc><<<ci.blocks, ci.threads, 0, s>>>(
        set, prm, rho + o, vfrac + o, sie + o, temp + o, press + o,
        sie_tot ? sie_tot + c0 : nullptr, vfrac_tot ? vfrac_tot + c0 : nullptr,
        Tguess ? Tguess + c0 : nullptr, init, n);
    // one block of THREADS per SM (fewer when the chunk has fewer cells than lanes)
    const int threads = THREADS;
    int64_t blocks = (int64_t)sm_count() * MINB;
    const int64_t need = (n + threads - 1) / threads;
    if (blocks > need) blocks = need;
    pte_rhot_sm_kernel<NM, Acc, THREADS, MINB><<<(unsigned)blocks, threads, 0, s>>>(
        set, prm, rho + o, vfrac + o, sie + o, temp + o, press + o,
        sie_tot ? sie_tot + c0 : nullptr, vfrac_tot ? vfrac_tot + c0 : nullptr, init,
        status ? reinterpret_cast<unsigned long long *>(status) + c0 : nullptr,
        residual ? residual + c0 : nullptr, reinterpret_cast<unsigned long long *>(stats), n,
        cursors + k, lambda ? lambda + c0 * NM * (Acc::kCacheSlots > 0 ? Acc::kCacheSlots : 1) : nullptr,
        skip ? skip + c0 : nullptr);
    count_launch(2);
    err = cudaGetLastError();
  }
  cudaFreeAsync(scratch, s); // stream-ordered; released on every exit path
  if (err != cudaSuccess)
    return fail(SGX_ERR_CUDA, "PTE RhoT launch failed: %s", cudaGetErrorString(err));
  return SGX_OK;
}


} // namespace sgx
#endif // SGX_PTE_RHOT_LAUNCH_CUH_
