// Vector lookups on MODIFIED handles (include/sgx_b200.h sgx_eos_modify_*): the reference's
// ShiftedEOS / ScaledEOS / BilinearRampEOS (eos/modifiers/shifted_eos.hpp, scaled_eos.hpp,
// ramps_eos.hpp) around any of the six non-Helmholtz models, nested as in its default variant
// (eos/eos_type_lists.hpp:69-94).  The layers themselves live in TabEos (sgx_pte_tab.cuh), the
// run-time accessor the mixed-cell kernels use, so that lookups, PTE solves and get_sg_eos see
// one implementation; this file is the cell-parallel kernel over it for sgx_eos_eval.
//
//   OUT[i] = eos.METHOD(IN1[i], IN2[i], lambda_i)      (eos_base.hpp:174-183)
//
// lambda_i is handed to the unmodified model exactly as the reference's wrappers do
// (`t_.Method(..., lambda)`): its nlambda slots are loaded, offered to the model and stored back.
// The model is picked by a run-time switch per cell: these are not the 80 %-of-HBM kernels of
// sgx_api.cu (DESIGN.md 7 has the measured rate).
#include "sgx_internal.hpp"
#include "sgx_pte_tab.cuh"

namespace sgx {
namespace {

__global__ void __launch_bounds__(256)
    mod_eval_kernel(const __grid_constant__ EosSet set, const int method,
                    const double *__restrict__ in1, const double *__restrict__ in2,
                    double *__restrict__ out, const int64_t n, double *__restrict__ lambda,
                    const int ls, const int nl) {
  TabEos eos(set);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n; i += nthr) {
    double lam[kMaxNumLambdas];
    const bool has_lambda = lambda != nullptr && nl > 0;
    if (has_lambda)
      for (int k = 0; k < nl; ++k)
        lam[k] = lambda[i * ls + k];
    eos.lam = has_lambda ? lam : nullptr;
    const double a = in1[i], b = in2[i];
    double r = 0.0;
    switch (method) {
    case SGX_P_FROM_RHO_T:
      r = eos.P_from_rho_T(0, a, b);
      break;
    case SGX_E_FROM_RHO_T:
      r = eos.e_from_rho_T(0, a, b);
      break;
    case SGX_T_FROM_RHO_E:
      r = eos.T_from_rho_e(0, a, b);
      break;
    case SGX_P_FROM_RHO_E:
      r = eos.P_from_rho_e(0, a, b);
      break;
    case SGX_CV_FROM_RHO_T:
      r = eos.cv_from_rho_T(0, a, b);
      break;
    case SGX_CV_FROM_RHO_E:
      r = eos.cv_from_rho_e(0, a, b);
      break;
    case SGX_BMOD_FROM_RHO_T:
      r = eos.bmod_from_rho_T(0, a, b);
      break;
    case SGX_BMOD_FROM_RHO_E:
      r = eos.bmod_from_rho_e(0, a, b);
      break;
    case SGX_GAMMA_FROM_RHO_T:
      r = eos.gamma_from_rho_T(0, a, b);
      break;
    default:
      r = eos.gamma_from_rho_e(0, a, b);
      break;
    }
    out[i] = r;
    if (has_lambda)
      for (int k = 0; k < nl; ++k)
        lambda[i * ls + k] = lam[k];
  }
}

} // namespace

int launch_mod_eval(sgx_eos_s *e, int method, const double *in1, const double *in2, double *out,
                    int64_t n, double *lambda, int ls, cudaStream_t s) {
  if (method < 0 || method >= SGX_METHOD_COUNT)
    return fail(SGX_ERR_INVALID, "unknown method %d", method);
  EosSet set;
  set.nmat = 1;
  if (!e->base || !set_material(e, set, 0))
    return fail(SGX_ERR_INVALID, "modified handle without a usable base model");
  const int nl = e->base->kind == SGX_EOS_SPINER_RHOT ? 2
                 : e->base->kind == SGX_EOS_SPINER_RHOSIE ? 1 : 0;
  const LaunchCfg c = stream_cfg(n, 256, 8);
  mod_eval_kernel<<<c.blocks, c.threads, 0, s>>>(set, method, in1, in2, out, n, lambda, ls, nl);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

} // namespace sgx
