// The rest of the reference's vector surface, one translation unit:
//   MinInternalEnergyFromDensity          eos_base.hpp:432-470 (vector), :604-611 (default),
//                                         eos_gruneisen.hpp:372-377, eos_davis.hpp:100-107,
//                                         eos_spiner_rho_temp.hpp:823-827, eos_spiner_rho_sie.hpp:688-692
//   InternalEnergyFromDensityPressure     eos_base.hpp:680-722 (regula falsi on P(rho, sie)),
//                                         eos_ideal.hpp:208-214 (closed form)
//   Entropy / GibbsFreeEnergy From Density{Temperature,InternalEnergy}
//                                         eos_ideal.hpp:101-114, eos_base.hpp:384-402; every other
//                                         model here aborts with EntropyIsNotEnabled
//   DensityEnergyFromPressureTemperature  per cell; the models' own overloads
//                                         (eos_ideal.hpp:199-206, eos_gruneisen.hpp:478-526,
//                                         eos_davis.hpp:476-496,564-582,
//                                         eos_spiner_rho_temp.hpp:1007-1016 -> lRhoFromPlT_ :1135-1190,
//                                         eos_spiner_rho_sie.hpp:836-848 -> lRhoFromPlT_ :941-986)
// Plain grid-stride kernels: these are the rarely-called members of the surface; the streaming
// design work lives in sgx_api.cu.
#include "sgx_internal.hpp"

namespace sgx {
namespace {

enum XOp { X_MIN_E = 0, X_E_FROM_RHO_P, X_S_RT, X_S_RE, X_G_RT, X_G_RE, X_RHO_E_FROM_PT };

// ---- per-model pieces --------------------------------------------------------------------
SGX_D double min_e(const IdealGas &, const double, double *) { return 0.0; }
SGX_D double min_e(const Gruneisen &g, const double rho_in, double *) {
  const double rho = stdmin(rho_in, g.rho_max);
  return g.e_from_rho_T(rho, 0.); // EosBase default (:604-611) at the clamped density
}
SGX_D double min_e(const DavisReactants &d, const double rho, double *) {
  const double es = d.Es(rho);
  const double ts = d.Ts(rho);
  return es - (d.Cv0 * ts) * d.i1pa;
}
SGX_D double min_e(const DavisProducts &, const double, double *) { return 0.0; } // rejected on host
SGX_D double min_e(const SpinerRhoT &s, const double rho, double *) {
  return interp1d(s.sieCold, s.gRho, s.lRho_(rho));
}
SGX_D double min_e(const SpinerRhoSie &s, const double rho, double *) {
  return interp1d(s.sieCold, s.gRho, to_log(rho, s.lRhoOffset));
}

// IdealGas(gm1, Cv): _EntropyT0 = _T0 = ROOM_TEMPERATURE, _EntropyRho0 = _rho0 = P0/(gm1 Cv T0)
// (eos_ideal.hpp:41-47, base/constants.hpp:41-42)
SGX_D double ideal_entropy(const IdealGas &g, const double rho, const double T) {
  constexpr double T0 = 293, P0 = 1e6;
  const double rho0 = P0 / (g.gm1 * g.Cv * T0);
  return g.Cv * xlog(ratio(T, T0)) + g.gm1 * g.Cv * xlog(ratio(rho0, rho));
}

template <class EOS>
SGX_D double min_temperature(const EOS &eos) {
  return eos.MinimumTemperature();
}
// SpinerEOSDependsRhoT::MinimumTemperature() = T_(lTMin_) (eos_spiner_rho_temp.hpp:252-253)
SGX_D double min_temperature(const SpinerRhoT &s) { return s.T_(s.lTMin); }

// EosBase::InternalEnergyFromDensityPressure (:680-722), the no-guess form
template <class EOS>
SGX_D double e_from_rho_P(const EOS &eos, const double rho, const double P, double *lambda) {
  const double sie_min = eos.e_from_rho_T(rho, min_temperature(eos), nullptr);
  const double sie_max = eos.e_from_rho_T(rho, 1e20, nullptr);
  double sie = 0.5 * (sie_min + sie_max);
  auto f = [&](const double s) { return eos.P_from_rho_e(rho, s, lambda); };
  const double sie_guess = (((sie_min < sie) && (sie < sie_max)) ? 0.5 * (sie_min + sie_max) : sie);
  root::regula_falsi(f, P, sie_guess, sie_min, sie_max, kEps, kEps, sie);
  return sie;
}
SGX_D double e_from_rho_P(const IdealGas &g, const double rho, const double P, double *) {
  return P / (g.gm1 * rho);
}

template <class EOS>
SGX_D void rho_e_from_PT(const EOS &eos, const double P, const double T, double *lambda,
                         double &rho, double &sie) {
  eos.rho_e_from_P_T(P, T, lambda, rho, sie);
}
// eos_spiner_rho_sie.hpp:836-848
SGX_D void rho_e_from_PT(const SpinerRhoSie &s, const double P, const double T, double *lambda,
                         double &rho, double &sie) {
  const double lT = to_log(T, s.lTOffset);
  const double lRho = s.lRhoFromPlT(P, lT, lambda);
  rho = from_log(lRho, s.lRhoOffset);
  sie = interp2c(s.sie, s.gRho, s.gT, s.numT, lRho, lT);
}

template <class EOS>
__global__ void __launch_bounds__(256)
    extra_kernel(const __grid_constant__ EOS eos, const int op, const double *__restrict__ in1,
                 const double *__restrict__ in2, double *__restrict__ out1,
                 double *__restrict__ out2, const int64_t n, double *__restrict__ lambda,
                 const int ls) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid; i < n; i += nthr) {
    double *lam = lambda ? lambda + i * ls : nullptr;
    switch (op) {
    case X_MIN_E:
      out1[i] = min_e(eos, in1[i], lam);
      break;
    case X_E_FROM_RHO_P:
      if constexpr (!std::is_same<EOS, SpinerRhoSie>::value)
        out1[i] = e_from_rho_P(eos, in1[i], in2[i], lam);
      break;
    case X_RHO_E_FROM_PT: {
      double rho = out1[i], sie = 0.0; // the incoming density is the guess where a model uses one
      rho_e_from_PT(eos, in1[i], in2[i], lam, rho, sie);
      out1[i] = rho;
      out2[i] = sie;
    } break;
    default:
      if constexpr (std::is_same<EOS, IdealGas>::value) {
        const double rho = in1[i];
        const double T = (op == X_S_RT || op == X_G_RT) ? in2[i] : eos.T_from_rho_e(rho, in2[i]);
        const double S = ideal_entropy(eos, rho, T);
        if (op == X_S_RT || op == X_S_RE) {
          out1[i] = S;
        } else {
          const double sie = (op == X_G_RT) ? eos.e_from_rho_T(rho, T) : in2[i];
          const double P = eos.P_from_rho_T(rho, T);
          out1[i] = sie + (P / rho) - T * S;
        }
      }
      break;
    }
  }
}

template <class EOS>
int launch_extra(const EOS &eos, int op, const double *in1, const double *in2, double *out1,
                 double *out2, int64_t n, double *lambda, int ls, cudaStream_t s) {
  const LaunchCfg c = stream_cfg(n, 256, 8);
  extra_kernel<EOS><<<c.blocks, c.threads, 0, s>>>(eos, op, in1, in2, out1, out2, n, lambda, ls);
  count_launch();
  SGX_CUDA_OK(cudaGetLastError());
  return SGX_OK;
}

int run_extra(sgx_eos_t e, int op, const double *in1, const double *in2, double *out1,
              double *out2, int64_t n, double *lambda, int ls, sgx_stream_t stream,
              const char *name) {
  DeviceGuard guard;
  if (!e) return fail(SGX_ERR_INVALID, "null EOS handle");
  SGX_CUDA_OK(cudaSetDevice(e->device));
  if (n <= 0) return n == 0 ? SGX_OK : fail(SGX_ERR_INVALID, "negative n");
  if (!in1 || !out1 || (op != X_MIN_E && !in2) || (op == X_RHO_E_FROM_PT && !out2))
    return fail(SGX_ERR_INVALID, "%s: null array", name);
  const bool entropy = op == X_S_RT || op == X_S_RE || op == X_G_RT || op == X_G_RE;
  if (entropy && e->kind != SGX_EOS_IDEAL)
    return fail(SGX_ERR_UNSUPPORTED, "%s: entropy is not enabled for this EOS "
                "(EntropyIsNotEnabled in the reference)", name);
  if (op == X_MIN_E && e->kind == SGX_EOS_DAVIS_PRODUCTS)
    return fail(SGX_ERR_UNSUPPORTED, "MinInternalEnergyFromDensity is not enabled for "
                "DavisProducts (eos_davis.hpp:255-259)");
  if (op == X_E_FROM_RHO_P && e->kind == SGX_EOS_SPINER_RHOSIE)
    return fail(SGX_ERR_UNSUPPORTED, "%s: not built for SpinerEOSDependsRhoSie", name);
  cudaStream_t s = (cudaStream_t)stream;
  switch (e->kind) {
  case SGX_EOS_IDEAL:
    return launch_extra(e->analytic.ideal, op, in1, in2, out1, out2, n, nullptr, 0, s);
  case SGX_EOS_GRUNEISEN:
    return launch_extra(e->analytic.gr, op, in1, in2, out1, out2, n, nullptr, 0, s);
  case SGX_EOS_DAVIS_REACTANTS:
    return launch_extra(e->analytic.dr, op, in1, in2, out1, out2, n, nullptr, 0, s);
  case SGX_EOS_DAVIS_PRODUCTS:
    return launch_extra(e->analytic.dp, op, in1, in2, out1, out2, n, nullptr, 0, s);
  case SGX_EOS_SPINER_RHOT:
    if (lambda && ls < 2) return fail(SGX_ERR_INVALID, "lambda_stride %d < 2", ls);
    return launch_extra(e->rhot, op, in1, in2, out1, out2, n, lambda, ls, s);
  case SGX_EOS_SPINER_RHOSIE:
    if (lambda && ls < 1) return fail(SGX_ERR_INVALID, "lambda_stride %d < 1", ls);
    return launch_extra(e->rhosie, op, in1, in2, out1, out2, n, lambda, ls, s);
  default:
    return fail(SGX_ERR_UNSUPPORTED, "%s: not built for EOS kind %d", name, e->kind);
  }
}

} // namespace
} // namespace sgx

using namespace sgx;

extern "C" {

int sgx_MinInternalEnergyFromDensity(sgx_eos_t e, const double *rhos, double *sies, int64_t n,
                                     double *lambda, int ls, sgx_stream_t s) {
  return run_extra(e, X_MIN_E, rhos, nullptr, sies, nullptr, n, lambda, ls, s,
                   "MinInternalEnergyFromDensity");
}
int sgx_InternalEnergyFromDensityPressure(sgx_eos_t e, const double *rhos, const double *Ps,
                                          double *sies, int64_t n, double *lambda, int ls,
                                          sgx_stream_t s) {
  return run_extra(e, X_E_FROM_RHO_P, rhos, Ps, sies, nullptr, n, lambda, ls, s,
                   "InternalEnergyFromDensityPressure");
}
int sgx_EntropyFromDensityTemperature(sgx_eos_t e, const double *rhos, const double *temps,
                                      double *entropies, int64_t n, double *lambda, int ls,
                                      sgx_stream_t s) {
  return run_extra(e, X_S_RT, rhos, temps, entropies, nullptr, n, lambda, ls, s,
                   "EntropyFromDensityTemperature");
}
int sgx_EntropyFromDensityInternalEnergy(sgx_eos_t e, const double *rhos, const double *sies,
                                         double *entropies, int64_t n, double *lambda, int ls,
                                         sgx_stream_t s) {
  return run_extra(e, X_S_RE, rhos, sies, entropies, nullptr, n, lambda, ls, s,
                   "EntropyFromDensityInternalEnergy");
}
int sgx_GibbsFreeEnergyFromDensityTemperature(sgx_eos_t e, const double *rhos,
                                              const double *temps, double *gibbs, int64_t n,
                                              double *lambda, int ls, sgx_stream_t s) {
  return run_extra(e, X_G_RT, rhos, temps, gibbs, nullptr, n, lambda, ls, s,
                   "GibbsFreeEnergyFromDensityTemperature");
}
int sgx_GibbsFreeEnergyFromDensityInternalEnergy(sgx_eos_t e, const double *rhos,
                                                 const double *sies, double *gibbs, int64_t n,
                                                 double *lambda, int ls, sgx_stream_t s) {
  return run_extra(e, X_G_RE, rhos, sies, gibbs, nullptr, n, lambda, ls, s,
                   "GibbsFreeEnergyFromDensityInternalEnergy");
}
int sgx_DensityEnergyFromPressureTemperature(sgx_eos_t e, const double *presses,
                                             const double *temps, double *rhos, double *sies,
                                             int64_t n, double *lambda, int ls, sgx_stream_t s) {
  return run_extra(e, X_RHO_E_FROM_PT, presses, temps, rhos, sies, n, lambda, ls, s,
                   "DensityEnergyFromPressureTemperature");
}

} // extern "C"
