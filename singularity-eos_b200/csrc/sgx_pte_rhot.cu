// PTESolverRhoT launchers (closure/mixed_cell_models.hpp:883-1193 + PTESolver :2188-2291).
// Default: the lane-refill state-machine kernel of sgx_pte_sm.cuh.  The straight
// one-thread-walks-one-cell transcription (pte_kernel in sgx_pte_kernel.cuh) is kept for the
// benchmark mixture as the comparison variant (SGX_PTE_KERNEL=percell).
#include <atomic>
#include <cstdlib>
#include <cstring>

#include "sgx_internal.hpp"
#include "sgx_pte_kernel.cuh"
#include "sgx_pte_sm.cuh"
#include "sgx_pte_rhot_launch.cuh"

namespace sgx {

// read per call so a test can run both variants in one process
static bool percell_variant() {
  const char *s = getenv("SGX_PTE_KERNEL");
  return s && strcmp(s, "percell") == 0;
}

static bool warp_variant() {
  const char *s = getenv("SGX_PTE_KERNEL");
  return s && strcmp(s, "warp") == 0;
}

// The warp-per-cell comparison variant (SGX_PTE_KERNEL=warp), benchmark mixture only.
template <int NM, class Acc>
static int launch_warp(const EosSet &set, const MixParamsDev &prm, double *rho, double *vfrac,
                       double *sie, double *temp, double *press, const double *sie_tot,
                       const double *vfrac_tot, const double *Tguess, uint64_t *status,
                       double *residual, int64_t ncell, cudaStream_t s) {
  const int64_t chunk = ncell < kPteChunk ? ncell : kPteChunk;
  double *init = nullptr;
  SGX_CUDA_OK(scratch_alloc_async(reinterpret_cast<void **>(&init),
                                  (size_t)chunk * (2 * NM + 1) * sizeof(double), s));
  cudaError_t err = cudaSuccess;
  for (int64_t c0 = 0; c0 < ncell && err == cudaSuccess; c0 += chunk) {
    const int64_t n = (ncell - c0) < chunk ? (ncell - c0) : chunk;
    const int64_t o = c0 * NM;
    const LaunchCfg ci = stream_cfg(n, 256, 4);
    pte_rhot_init_kernel<NM, Acc><<<ci.blocks, ci.threads, 0, s>>>(
        set, prm, rho + o, vfrac + o, sie + o, temp + o, press + o,
        sie_tot ? sie_tot + c0 : nullptr, vfrac_tot ? vfrac_tot + c0 : nullptr,
        Tguess ? Tguess + c0 : nullptr, init, n);
    const LaunchCfg cw = stream_cfg(n * 32, 256, 2);
    pte_rhot_warp_kernel<NM, Acc><<<cw.blocks, cw.threads, 0, s>>>(
        set, prm, rho + o, vfrac + o, sie + o, temp + o, press + o,
        sie_tot ? sie_tot + c0 : nullptr, vfrac_tot ? vfrac_tot + c0 : nullptr, init,
        status ? reinterpret_cast<unsigned long long *>(status) + c0 : nullptr,
        residual ? residual + c0 : nullptr, n);
    count_launch(2);
    err = cudaGetLastError();
  }
  cudaFreeAsync(init, s);
  if (err != cudaSuccess)
    return fail(SGX_ERR_CUDA, "PTE RhoT warp-per-cell launch failed: %s", cudaGetErrorString(err));
  return SGX_OK;
}

// The state-machine kernels keep a spill frame and call the out-of-line libm cores, so their
// local memory is laid out by the DEVICE stack-size limit.  At CUDA's default of 1024 bytes per
// thread the layout costs 4-13 % (3-material benchmark mixture 112.6 -> 119.5 M solves/s at
// 2048; profiles/r2ab_pte_stack_limit.txt) -- which is why the same kernel ran faster in any
// process that had created a multi-rank NCCL communicator: NCCL raises the limit to 1872 for
// its own kernels.  Once per device, if the limit is below 2048 it is raised to 2048 (never
// lowered; about 0.3 GB more local-memory backing on a B200).  SGX_STACK_LIMIT=<bytes> picks
// another value, SGX_STACK_LIMIT=0 leaves the device alone.
void raise_stack_limit_once() {
  static std::atomic<unsigned long long> done{0};
  int device = 0;
  if (cudaGetDevice(&device) != cudaSuccess || device < 0 || device >= 64) return;
  const unsigned long long bit = 1ull << device;
  if (done.fetch_or(bit) & bit) return;
  size_t want = 2048;
  if (const char *v = getenv("SGX_STACK_LIMIT")) want = (size_t)strtoull(v, nullptr, 10);
  if (want == 0) return;
  size_t now = 0;
  if (cudaDeviceGetLimit(&now, cudaLimitStackSize) != cudaSuccess || now >= want) return;
  if (cudaDeviceSetLimit(cudaLimitStackSize, want) != cudaSuccess) cudaGetLastError();
}

int launch_pte_rhot(const EosSet &set, bool bench_mix, const MixParamsDev &prm, double *rho,
                    double *vfrac, double *sie, double *temp, double *press,
                    const double *sie_tot, const double *vfrac_tot, const double *Tguess,
                    uint64_t *status, double *residual, uint64_t *stats, int64_t ncell,
                    cudaStream_t s) {
#define SGX_PTE_ARGS                                                                     \
  set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot, Tguess, status, residual,   \
      stats, ncell, s
  raise_stack_limit_once();
  if (bench_mix) {
    if (percell_variant()) return launch_pte_percell<3, Mix3GrDrDp, SGX_PTE_RHOT>(SGX_PTE_ARGS);
    if (warp_variant() && !stats)
      return launch_warp<3, Mix3GrDrDp>(set, prm, rho, vfrac, sie, temp, press, sie_tot, vfrac_tot,
                                        Tguess, status, residual, ncell, s);
    // SGX_PTE_SM_SHAPE: block shape study of the state machine: threads per
    // block x resident blocks per SM -> register budget.  Default 512 x 1 (128 registers).
    static const int shape = [] {
      const char *v = getenv("SGX_PTE_SM_SHAPE");
      return v ? atoi(v) : 0;
    }();
    if (shape == 384) return launch_sm<3, Mix3GrDrDp, 384, 1>(SGX_PTE_ARGS); // 168 registers
    if (shape == 256) return launch_sm<3, Mix3GrDrDp, 256, 1>(SGX_PTE_ARGS); // 255 registers
    if (shape == 2562) return launch_sm<3, Mix3GrDrDp, 256, 2>(SGX_PTE_ARGS); // 2 blocks, 128
    if (shape == 640) return launch_sm<3, Mix3GrDrDp, 640, 1>(SGX_PTE_ARGS); // 96 registers
    if (shape == 768) return launch_sm<3, Mix3GrDrDp, 768, 1>(SGX_PTE_ARGS); // 80 registers
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
