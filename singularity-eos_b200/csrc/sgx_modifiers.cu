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

// P and e of one (rho, T) pair in one pass (sgx_PressureEnergyFromDensityTemperature): the two
// wrapper calls back to back on the same lambda row, i.e. what a host code's P-then-e sequence
// computes, with one read of the inputs and one round trip of the row.
__global__ void __launch_bounds__(256)
    mod_pe_kernel(const __grid_constant__ EosSet set, const double *__restrict__ rho,
                  const double *__restrict__ T, double *__restrict__ P, double *__restrict__ E,
                  const int64_t n, double *__restrict__ lambda, const int ls, const int nl) {
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
    const double a = rho[i], b = T[i];
    const double p = eos.P_from_rho_T(0, a, b);
    const double e = eos.e_from_rho_T(0, a, b);
    P[i] = p;
    E[i] = e;
    if (has_lambda)
      for (int k = 0; k < nl; ++k)
        lambda[i * ls + k] = lam[k];
  }
}

// ---- FillEos of the wrappers (shifted_eos.hpp:148-170, scaled_eos.hpp:160-180,
// ramps_eos.hpp:217-239), quirks included: with a (rho, sie)-preferring base the shifted / scaled
// energy is a LOCAL of the wrapper, so an energy the base computes is not handed back; ScaledEOS
// hands a local scaled density down, so a density the base computes is not handed back either;
// BilinearRampEOS always recomputes bmod from (rho, energy) at the end.
__device__ inline bool base_fill(const TabEos &eos, double &rho, double &temp, double &energy,
                                 double &press, double &cv, double &bmod,
                                 const unsigned long output) {
  const AnyEos &b = eos.mat(0);
  switch (b.kind) {
  case 0:
    return b.ideal.fill(rho, temp, energy, press, cv, bmod, output, nullptr);
  case 1:
    return b.gr.fill(rho, temp, energy, press, cv, bmod, output, nullptr);
  case 2:
    return b.dr.fill(rho, temp, energy, press, cv, bmod, output, nullptr);
  case 3:
    return b.dp.fill(rho, temp, energy, press, cv, bmod, output, nullptr);
  case 4:
    return b.rhot->fill(rho, temp, energy, press, cv, bmod, output, eos.cache(0));
  default:
    return b.rhosie->fill(rho, temp, energy, press, cv, bmod, output, eos.cache(0));
  }
}
__device__ inline bool shifted_fill(const TabEos &eos, const ModParams &q, double &rho,
                                    double &temp, double &energy, double &press, double &cv,
                                    double &bmod, const unsigned long output) {
  if (!(q.flags & kModShift)) return base_fill(eos, rho, temp, energy, press, cv, bmod, output);
  const unsigned long pref = eos.preferred(0);
  if (pref == (TQ_DENSITY | TQ_TEMPERATURE)) {
    const bool ok = base_fill(eos, rho, temp, energy, press, cv, bmod, output);
    energy = energy + q.shift;
    return ok;
  }
  if (pref == (TQ_DENSITY | TQ_SIE)) {
    double senergy = energy - q.shift;
    return base_fill(eos, rho, temp, senergy, press, cv, bmod, output);
  }
  return false; // EOS_ERROR("Didn't find a valid input for ShiftedEOS::FillEOS")
}
__device__ inline bool scaled_fill(const TabEos &eos, const ModParams &q, double &rho,
                                   double &temp, double &energy, double &press, double &cv,
                                   double &bmod, const unsigned long output) {
  if (!(q.flags & kModScale))
    return shifted_fill(eos, q, rho, temp, energy, press, cv, bmod, output);
  const unsigned long pref = eos.preferred(0);
  double srho = q.scale * rho;
  if (pref == (TQ_DENSITY | TQ_TEMPERATURE)) {
    const bool ok = shifted_fill(eos, q, srho, temp, energy, press, cv, bmod, output);
    energy = q.scale * energy;
    return ok;
  }
  if (pref == (TQ_DENSITY | TQ_SIE)) {
    double senergy = q.inv_scale * energy;
    return shifted_fill(eos, q, srho, temp, senergy, press, cv, bmod, output);
  }
  return false;
}
__device__ inline bool mod_fill(const TabEos &eos, double &rho, double &temp, double &energy,
                                double &press, double &cv, double &bmod,
                                const unsigned long output) {
  const ModParams &q = eos.mod(0);
  if (!(q.flags & kModRamp)) return scaled_fill(eos, q, rho, temp, energy, press, cv, bmod, output);
  const unsigned long ramp_out = output & ~(unsigned long)TQ_PRESSURE;
  if (output & TQ_PRESSURE) {
    if (!(output & TQ_SIE)) {
      press = eos.P_from_rho_e(0, rho, energy);
    } else if (!(output & TQ_TEMPERATURE)) {
      press = eos.P_from_rho_T(0, rho, temp);
    }
  }
  const bool ok = scaled_fill(eos, q, rho, temp, energy, press, cv, bmod, ramp_out);
  double rho_ramp = rho;
  if (output & TQ_DENSITY) // get_ramp_density (ramps_eos.hpp:113-118)
    rho_ramp = press < q.Pmid ? q.r0 * (ratio(press, q.a) + 1.0) : q.r0 * (ratio(press, q.b) + q.c);
  rho = rho_ramp < rho ? rho_ramp : rho;
  bmod = eos.bmod_from_rho_e(0, rho, energy);
  return ok;
}

__global__ void __launch_bounds__(256)
    mod_fill_kernel(const __grid_constant__ EosSet set, double *__restrict__ rho,
                    double *__restrict__ T, double *__restrict__ en, double *__restrict__ P,
                    double *__restrict__ cv, double *__restrict__ bmod, const int64_t n,
                    const unsigned long output, double *__restrict__ lambda, const int ls,
                    const int nl, int *__restrict__ err) {
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
    // the reference's vector overload hands the array ELEMENTS to the scalar method
    // (eos_base.hpp:503-518), so what a wrapper does not assign keeps the array's value
    double r = 0, t = 0, e = 0, p = 0, c = 0, b = 0;
    if (rho) r = rho[i];
    if (T) t = T[i];
    if (en) e = en[i];
    if (P) p = P[i];
    if (cv) c = cv[i];
    if (bmod) b = bmod[i];
    const bool ok = mod_fill(eos, r, t, e, p, c, b, output);
    if (!ok) {
      if (err) *err = 1;
      continue;
    }
    // ... and whatever it does assign lands in the caller's arrays, inputs included: e.g.
    // ShiftedEOS over a (rho, T)-preferring base adds the shift to an INPUT energy
    // (shifted_eos.hpp:157-158), BilinearRampEOS over Gruneisen turns the pressure into an input
    // of the inner call, which then rewrites density and energy (eos_gruneisen.hpp:532-533)
    if (rho) rho[i] = r;
    if (T) T[i] = t;
    if (en) en[i] = e;
    if (P) P[i] = p;
    if (cv) cv[i] = c;
    if (bmod) bmod[i] = b;
    if (has_lambda)
      for (int k = 0; k < nl; ++k)
        lambda[i * ls + k] = lam[k];
  }
}

} // namespace

int launch_mod_fill(sgx_eos_s *e, double *rho, double *T, double *en, double *P, double *cv,
                    double *bmod, int64_t n, unsigned long output, double *lambda, int ls,
                    cudaStream_t s) {
  EosSet set;
  set.nmat = 1;
  if (!e->base || !set_material(e, set, 0))
    return fail(SGX_ERR_INVALID, "modified handle without a usable base model");
  if (e->base->kind == SGX_EOS_IDEAL && (e->mod.flags & (kModShift | kModScale)))
    return fail(SGX_ERR_UNSUPPORTED,
                "FillEos: ShiftedEOS / ScaledEOS switch on the base's PreferredInput(), and "
                "IdealGas's (rho | sie | T | P) matches neither case: the reference raises "
                "EOS_ERROR (shifted_eos.hpp:167, scaled_eos.hpp:177)");
  const int nl = e->base->kind == SGX_EOS_SPINER_RHOT ? 2
                 : e->base->kind == SGX_EOS_SPINER_RHOSIE ? 1 : 0;
  const LaunchCfg c = stream_cfg(n, 256, 8);
  mod_fill_kernel<<<c.blocks, c.threads, 0, s>>>(set, rho, T, en, P, cv, bmod, n, output, lambda,
                                                  ls, nl, nullptr);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

int launch_mod_pe(sgx_eos_s *e, const double *rho, const double *T, double *P, double *E,
                  int64_t n, double *lambda, int ls, cudaStream_t s) {
  EosSet set;
  set.nmat = 1;
  if (!e->base || !set_material(e, set, 0))
    return fail(SGX_ERR_INVALID, "modified handle without a usable base model");
  const int nl = e->base->kind == SGX_EOS_SPINER_RHOT ? 2
                 : e->base->kind == SGX_EOS_SPINER_RHOSIE ? 1 : 0;
  const LaunchCfg c = stream_cfg(n, 256, 8);
  mod_pe_kernel<<<c.blocks, c.threads, 0, s>>>(set, rho, T, P, E, n, lambda, ls, nl);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

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
