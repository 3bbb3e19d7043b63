// PTESolverRhoT launchers (closure/mixed_cell_models.hpp:883-1193 + PTESolver :2188-2291).
// Default: the lane-refill state-machine kernel of sgx_pte_sm.cuh.  The straight
// one-thread-walks-one-cell transcription (pte_kernel in sgx_pte_kernel.cuh) is kept for the
// benchmark mixture as the comparison variant (SGX_PTE_KERNEL=percell).
#include <cstdlib>
#include <cstring>

#include "sgx_internal.hpp"
#include "sgx_pte_kernel.cuh"
#include "sgx_pte_sm.cuh"

namespace sgx {

// read per call so a test can run both variants in one process
static bool percell_variant() {
  const char *s = getenv("SGX_PTE_KERNEL");
  return s && strcmp(s, "percell") == 0;
}

// Cells go through in chunks: init kernel -> scratch record -> state-machine kernel, both on
// the caller's stream; the scratch is stream-ordered (cudaMallocAsync), (2*NM+1)*8 B/cell.
static constexpr int64_t kPteChunk = int64_t(1) << 22;

template <int NM, class Acc, int THREADS = kSmThreads, int MINB = 1>
static int launch_sm(const EosSet &set, const MixParamsDev &prm, double *rho, double *vfrac,
                     double *sie, double *temp, double *press, const double *sie_tot,
                     const double *vfrac_tot, const double *Tguess, uint64_t *status,
                     double *residual, uint64_t *stats, int64_t ncell, cudaStream_t s) {
  const int64_t chunk = ncell < kPteChunk ? ncell : kPteChunk;
  double *init = nullptr;
  SGX_CUDA_OK(cudaMallocAsync(&init, (size_t)chunk * (2 * NM + 1) * sizeof(double), s));
  for (int64_t c0 = 0; c0 < ncell; c0 += chunk) {
    const int64_t n = (ncell - c0) < chunk ? (ncell - c0) : chunk;
    const int64_t o = c0 * NM;
    const LaunchCfg ci = stream_cfg(n, 256, 4);
    pte_rhot_init_kernel<NM, Acc><<<ci.blocks, ci.threads, 0, s>>>(
        set, prm, rho + o, vfrac + o, sie + o, temp + o, press + o,
        sie_tot ? sie_tot + c0 : nullptr, vfrac_tot ? vfrac_tot + c0 : nullptr,
        Tguess ? Tguess + c0 : nullptr, init, n);
    // one block of kSmThreads per SM; each warp owns a contiguous range of >= 32 cells
    const int threads = THREADS;
    int64_t warps = (int64_t)sm_count() * MINB * (threads / 32);
    const int64_t min_warps = (n + 31) / 32;
    if (warps > min_warps) warps = min_warps;
    const int64_t blocks = (warps + threads / 32 - 1) / (threads / 32);
    warps = blocks * (threads / 32);
    const int64_t cpw = (n + warps - 1) / warps;
    pte_rhot_sm_kernel<NM, Acc, THREADS, MINB><<<(unsigned)blocks, threads, 0, s>>>(
        set, prm, rho + o, vfrac + o, sie + o, temp + o, press + o,
        sie_tot ? sie_tot + c0 : nullptr, vfrac_tot ? vfrac_tot + c0 : nullptr, init,
        status ? reinterpret_cast<unsigned long long *>(status) + c0 : nullptr,
        residual ? residual + c0 : nullptr, reinterpret_cast<unsigned long long *>(stats), n,
        cpw);
    count_launch(2);
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) { // the stream-ordered scratch is released on every exit path
      cudaFreeAsync(init, s);
      return fail(SGX_ERR_CUDA, "PTE RhoT launch failed: %s", cudaGetErrorString(err));
    }
  }
  SGX_CUDA_OK(cudaFreeAsync(init, s));
  return SGX_OK;
}

int launch_pte_rhot(const EosSet &set, bool bench_mix, const MixParamsDev &prm, double *rho,
                    double *vfrac, double *sie, double *temp, double *press,
                    const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                    uint64_t *status, double *residual, uint64_t *stats, int64_t ncell,
                    cudaStream_t s) {
#define SGX_PTE_ARGS                                                                     \
  set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, status, residual,   \
      stats, ncell, s
  if (bench_mix) {
    if (percell_variant()) return launch_pte_percell<3, Mix3GrDrDp, SGX_PTE_RHOT>(SGX_PTE_ARGS);
    return launch_sm<3, Mix3GrDrDp>(SGX_PTE_ARGS);
  }
  switch (set.nmat) {
  case 1:
    return launch_sm<1, DynEos>(SGX_PTE_ARGS);
  case 2:
    return launch_sm<2, DynEos>(SGX_PTE_ARGS);
  case 3:
    return launch_sm<3, DynEos>(SGX_PTE_ARGS);
  default:
    return launch_sm<4, DynEos>(SGX_PTE_ARGS);
  }
#undef SGX_PTE_ARGS
}

} // namespace sgx
