// TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
//
// oracle/_ref/libsgref.so: the UNMODIFIED reference (lanl/singularity-eos headers where they
// lie under /root/reference) compiled for the host with PORTABILITY_STRATEGY_NONE, wrapped in
// a small C API so tests/ and bench.py's cpu_baseline / --impl reference legs can run the
// reference's own code on the same inputs as the CUDA path.
//
// What runs here is reference code:
//   vector lookups  -> Variant vector overloads, eos/eos_variant.hpp:56-143 -> eos_base.hpp:169-230
//   FillEos         -> eos/eos_variant.hpp:652-708 -> eos_base.hpp:503-531
//   Spiner EOS      -> the real SpinerEOSDependsRhoT / SpinerEOSDependsRhoSie constructors
//                      (eos_spiner_rho_temp.hpp:373-424, eos_spiner_rho_sie.hpp:980-1040) reading
//                      an in-memory .sp5 tree through oracle/hdf5_shim (libhdf5 is absent)
//   PTE             -> the cell loop of test/test_pte.cpp:177-265 around PTESolverRhoT/PT +
//                      PTESolver (closure/mixed_cell_models.hpp:883-1446,2188-2291), Eigen 3x3
// The only additions are the C wrapper, the .sp5 tree builder, and an OpenMP `parallel for`
// over cells / chunks (the reference's portableFor is a serial loop without Kokkos,
// utils/ports-of-call/ports-of-call/portability.hpp:224-236).
//
// `private`/`protected` are opened up so the harness can read the post-fix-up DataBoxes for
// table-level parity checks; no reference behaviour changes.
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include <unistd.h>

#define private public
#define protected public
#include <singularity-eos/closure/mixed_cell_models.hpp>
#include <singularity-eos/eos/eos.hpp>
#undef private
#undef protected

#include <singularity-eos/base/spiner_table_utils.hpp>

#ifdef _OPENMP
#include <omp.h>
#endif

// Second build of this file: oracle/_ref/libsgref_kokkos.so, -O3 -DNDEBUG with
// PORTABILITY_STRATEGY_KOKKOS against the vendored Kokkos (utils/kokkos, OpenMP backend).  There
// portableFor IS Kokkos::parallel_for (utils/ports-of-call/ports-of-call/portability.hpp:228-236),
// so the timed baselines below hand the whole array to ONE reference vector call / ONE
// portableFor over cells: the reference's Kokkos-OpenMP CPU path, not a loop of ours.
#ifdef SGREF_KOKKOS
#include <Kokkos_Core.hpp>
#endif

#include "../include/sgx_b200.h" // POD layouts only (sgx_sp5_tables, sgx_mix_params, enums)
#include "sp5_tree.hpp"

using namespace singularity;
using DataBox = spiner_common::DataBox;

namespace {

struct StridedLambda {
  Real *base;
  int stride;
  PORTABLE_FORCEINLINE_FUNCTION Real *operator[](const int i) const {
    return base ? base + static_cast<std::ptrdiff_t>(stride) * i : nullptr;
  }
};

// Type-indexed lambda with NO LogTemperature slot (finding 6):
// {MeanAtomicMass, MeanAtomicNumber} at lambda[stride*i + {0,1}].
using AZIndexer = IndexerUtils::VariadicPointerIndexer<IndexableTypes::MeanAtomicMass,
                                                       IndexableTypes::MeanAtomicNumber>;
struct StridedAZLambda {
  Real *base;
  int stride;
  PORTABLE_FORCEINLINE_FUNCTION AZIndexer operator[](const int i) const {
    return AZIndexer(base + static_cast<std::ptrdiff_t>(stride) * i);
  }
};

#ifdef SGREF_KOKKOS
void kokkos_init_once(int nthreads) {
  if (Kokkos::is_initialized()) return;
  Kokkos::InitializationSettings st;
  if (nthreads > 0) st.set_num_threads(nthreads);
  st.set_disable_warnings(true);
  Kokkos::initialize(st);
  std::atexit([] {
    if (Kokkos::is_initialized() && !Kokkos::is_finalized()) Kokkos::finalize();
  });
}
#endif

int nthreads_for(int requested) {
#ifdef _OPENMP
  return requested > 0 ? requested : omp_get_max_threads();
#else
  (void)requested;
  return 1;
#endif
}

using sgx_oracle::build_sp5;

template <typename F>
int guarded(F &&f) {
  try {
    f();
    return 0;
  } catch (const std::exception &e) {
    std::fprintf(stderr, "[sgref] reference threw: %s\n", e.what());
    return -1;
  }
}

int64_t copy_db(const DataBox &db, double *out, int64_t cap) {
  const int64_t n = db.size();
  if (out == nullptr) return n;
  const int64_t m = std::min(n, cap);
  for (int64_t i = 0; i < m; ++i)
    out[i] = db.dataView_.data()[i]; // data() returns the base allocation even for slices
  return m;
}

} // namespace

extern "C" {

int ref_abi_version() { return SGX_ABI_VERSION; }
int ref_num_threads() { return nthreads_for(0); }
// 1 when this library is the Kokkos-OpenMP build (timed baselines), 0 for the serial-portableFor
// build used as the parity oracle
int ref_is_kokkos() {
#ifdef SGREF_KOKKOS
  return 1;
#else
  return 0;
#endif
}

// ---- FastMath (base/fast-math/logs.hpp) -----------------------------------------
void ref_fastmath_eval(int which, const double *x, double *y, int64_t n) {
  for (int64_t i = 0; i < n; ++i) {
    switch (which) {
    case 0:
      y[i] = FastMath::log10(x[i]);
      break;
    case 1:
      y[i] = FastMath::pow10(x[i]);
      break;
    case 2:
      y[i] = FastMath::lg(x[i]);
      break;
    default:
      y[i] = FastMath::pow2(x[i]);
      break;
    }
  }
}

// ---- regula falsi on a 1-D piecewise-linear table (root_finding.hpp:154-240) ------
// f(x) = DataBox(y over RegularGrid1D(xmin,xmax,ny)).interpToReal(x)
int ref_regula_falsi_table(const double *y, int ny, double xmin, double xmax, double ytarget,
                           double guess, double a, double b, double xtol, double ytol,
                           double *xroot) {
  DataBox db(Spiner::AllocationTarget::Host, ny);
  db.setRange(0, spiner_common::Grid_t(
                     std::vector<Spiner::RegularGrid1D<Real>>{{xmin, xmax, (size_t)ny}}));
  for (int i = 0; i < ny; ++i)
    db(i) = y[i];
  const callable_interp::interp f(db);
  auto st = RootFinding1D::regula_falsi(f, ytarget, guess, a, b, xtol, ytol, *xroot);
  free(db);
  return st == RootFinding1D::Status::SUCCESS ? 0 : 1;
}

// ---- solver utils (closure/mixed_cell_models.hpp:176-305) -------------------------
int ref_solve_Ax_b(int n, double *a, double *b) {
  double scr[16];
  return mix_impl::solve_Ax_b_wscr(n, a, b, scr) ? 1 : 0;
}
// batch form for the bit-equality test of the device linear solver (a, b as in the C ABI's
// sgx_solve_Ax_b: [batch][n*n] row-major and [batch][n], both overwritten)
void ref_solve_Ax_b_batch(int n, double *a, double *b, int *ok, long batch) {
  double scr[16];
  for (long i = 0; i < batch; ++i)
    ok[i] = mix_impl::solve_Ax_b_wscr(n, a + i * n * n, b + i * n, scr) ? 1 : 0;
}
double ref_sum_neumaier(const double *x, int n) { return mix_impl::sum_neumaier(x, n); }

// ---- EOS handles ----------------------------------------------------------------
void *ref_eos_ideal(double gm1, double Cv) { return new EOS(IdealGas(gm1, Cv)); }
void *ref_eos_gruneisen(double C0, double s1, double s2, double s3, double G0, double b,
                        double rho0, double T0, double P0, double Cv, double rho_max) {
  if (rho_max > 0) return new EOS(Gruneisen(C0, s1, s2, s3, G0, b, rho0, T0, P0, Cv, rho_max));
  return new EOS(Gruneisen(C0, s1, s2, s3, G0, b, rho0, T0, P0, Cv));
}
void *ref_eos_davis_reactants(double rho0, double e0, double P0, double T0, double A,
                              double B, double C, double G0, double Z, double alpha,
                              double Cv0) {
  return new EOS(DavisReactants(rho0, e0, P0, T0, A, B, C, G0, Z, alpha, Cv0));
}
void *ref_eos_davis_products(double a, double b, double k, double n, double vc, double pc,
                             double Cv) {
  return new EOS(DavisProducts(a, b, k, n, vc, pc, Cv));
}
// Modifier layers through the reference's own Variant::Modify (eos_variant.hpp:213-223): a NEW
// handle holding Mod<T>(copy of T, args).  which: 0 ShiftedEOS(shift = p[0]), 1 ScaledEOS(scale
// = p[0]), 2 BilinearRampEOS(r0, a, b, c = p[0..3]).  Null when the modified type is not in the
// oracle's variant.  Tabulated models are shallow copies: free the modified handle with
// ref_eos_free_shallow and keep the base alive.
void *ref_eos_modify(void *h, int which, const double *p) {
  EOS *e = static_cast<EOS *>(h);
  void *out = nullptr;
  guarded([&]() {
    if (which == 0 && e->ModifiedInVariant<ShiftedEOS>())
      out = new EOS(e->Modify<ShiftedEOS>(p[0]));
    else if (which == 1 && e->ModifiedInVariant<ScaledEOS>())
      out = new EOS(e->Modify<ScaledEOS>(p[0]));
    else if (which == 2 && e->ModifiedInVariant<BilinearRampEOS>())
      out = new EOS(e->Modify<BilinearRampEOS>(p[0], p[1], p[2], p[3]));
  });
  return out;
}
void ref_eos_free_shallow(void *h) { delete static_cast<EOS *>(h); }
void *ref_eos_vinet(double rho0, double T0, double B0, double BP0, double A0, double Cv0,
                    double E0, double S0, const double *d2to40) {
  return new EOS(Vinet(rho0, T0, B0, BP0, A0, Cv0, E0, S0, d2to40));
}
void *ref_eos_spiner_rhot(const sgx_sp5_tables *raw, int reproducible, int pmin_vapor_dome) {
  EOS *out = nullptr;
  int rc = guarded([&]() {
    const int matid = 1;
    std::string f = build_sp5(raw, matid);
    out = new EOS(SpinerEOSDependsRhoT(f, matid, reproducible != 0, pmin_vapor_dome != 0));
    h5shim::drop_file(f);
    std::remove(f.c_str());
  });
  return rc == 0 ? out : nullptr;
}
void *ref_eos_spiner_rhosie(const sgx_sp5_tables *raw, int reproducible,
                            int pmin_vapor_dome) {
  EOS *out = nullptr;
  int rc = guarded([&]() {
    const int matid = 1;
    std::string f = build_sp5(raw, matid);
    out = new EOS(SpinerEOSDependsRhoSie(f, matid, reproducible != 0, pmin_vapor_dome != 0));
    h5shim::drop_file(f);
    std::remove(f.c_str());
  });
  return rc == 0 ? out : nullptr;
}
void *ref_eos_helmholtz(const char *path, const sgx_helm_options *o) {
  EOS *out = nullptr;
  int rc = guarded([&]() {
    out = new EOS(Helmholtz(std::string(path), o->enable_rad != 0, o->enable_gas != 0,
                            o->enable_coulomb != 0, o->gas_ionized != 0,
                            o->gas_degenerate != 0, false, o->use_newton_raphson != 0));
  });
  return rc == 0 ? out : nullptr;
}
void ref_eos_free(void *h) {
  EOS *e = static_cast<EOS *>(h);
  if (!e) return;
  e->Finalize();
  delete e;
}
int ref_eos_nlambda(void *h) { return static_cast<EOS *>(h)->nlambda(); }
unsigned long ref_eos_preferred_input(void *h) {
  return static_cast<EOS *>(h)->PreferredInput();
}
double ref_eos_maximum_density(void *h) { return static_cast<EOS *>(h)->MaximumDensity(); }

// Post-fix-up tables of the real Spiner classes, by the reference's member names.
int64_t ref_eos_get_table(void *h, const char *name, double *out, int64_t cap) {
  EOS *e = static_cast<EOS *>(h);
  const std::string nm(name);
  if (e->IsType<SpinerEOSDependsRhoT>()) {
    const auto s = e->get<SpinerEOSDependsRhoT>();
    const std::map<std::string, const DataBox *> m = {
        {"P", &s.P_},
        {"sie", &s.sie_},
        {"bMod", &s.bMod_},
        {"dPdRho", &s.dPdRho_},
        {"dPdE", &s.dPdE_},
        {"dTdRho", &s.dTdRho_},
        {"dTdE", &s.dTdE_},
        {"dEdRho", &s.dEdRho_},
        {"dEdT", &s.dEdT_},
        {"PMax", &s.PMax_},
        {"sielTMax", &s.sielTMax_},
        {"dEdTMax", &s.dEdTMax_},
        {"gm1Max", &s.gm1Max_},
        {"lTColdCrit", &s.lTColdCrit_},
        {"PCold", &s.PCold_},
        {"sieCold", &s.sieCold_},
        {"bModCold", &s.bModCold_},
        {"dPdRhoCold", &s.dPdRhoCold_},
        {"dPdECold", &s.dPdECold_},
        {"dTdRhoCold", &s.dTdRhoCold_},
        {"dTdECold", &s.dTdECold_},
        {"dEdTCold", &s.dEdTCold_},
        {"rho_at_pmin", &s.rho_at_pmin_}};
    auto it = m.find(nm);
    return it == m.end() ? -1 : copy_db(*it->second, out, cap);
  }
  if (e->IsType<SpinerEOSDependsRhoSie>()) {
    const auto s = e->get<SpinerEOSDependsRhoSie>();
    const std::map<std::string, const DataBox *> m = {
        {"sie", &s.sie_},
        {"T_re", &s.T_},
        {"rho_at_pmin", &s.rho_at_pmin_},
        {"P", &s.dependsRhoT_.P},
        {"bMod", &s.dependsRhoT_.bMod},
        {"dPdRho", &s.dependsRhoT_.dPdRho},
        {"dPdE", &s.dependsRhoT_.dPdE},
        {"dTdRho", &s.dependsRhoT_.dTdRho},
        {"dTdE", &s.dependsRhoT_.dTdE},
        {"dEdRho", &s.dependsRhoT_.dEdRho},
        {"P_re", &s.dependsRhoSie_.P},
        {"bMod_re", &s.dependsRhoSie_.bMod},
        {"dPdRho_re", &s.dependsRhoSie_.dPdRho},
        {"dPdE_re", &s.dependsRhoSie_.dPdE},
        {"dTdRho_re", &s.dependsRhoSie_.dTdRho},
        {"dTdE_re", &s.dependsRhoSie_.dTdE},
        {"dEdRho_re", &s.dependsRhoSie_.dEdRho},
        {"PlRhoMax", &s.PlRhoMax_},
        {"dPdRhoMax", &s.dPdRhoMax_},
        {"PCold", &s.PCold_},
        {"sieCold", &s.sieCold_},
        {"bModCold", &s.bModCold_},
        {"dPdRhoCold", &s.dPdRhoCold_}};
    auto it = m.find(nm);
    return it == m.end() ? -1 : copy_db(*it->second, out, cap);
  }
  return -1;
}
int ref_eos_get_scalar(void *h, const char *name, double *out) {
  EOS *e = static_cast<EOS *>(h);
  const std::string nm(name);
  if (e->IsType<SpinerEOSDependsRhoT>()) {
    const auto s = e->get<SpinerEOSDependsRhoT>();
    const std::map<std::string, double> m = {
        {"lRhoMin", s.lRhoMin_},       {"lRhoMax", s.lRhoMax_},   {"rhoMax", s.rhoMax_},
        {"lTMin", s.lTMin_},           {"lTMax", s.lTMax_},       {"TMax", s.TMax_},
        {"PMin", s.PMin_},             {"rhoNormal", s.rhoNormal_}, {"TNormal", s.TNormal_},
        {"sieNormal", s.sieNormal_},   {"PNormal", s.PNormal_},   {"CvNormal", s.CvNormal_},
        {"bModNormal", s.bModNormal_}, {"dPdENormal", s.dPdENormal_},
        {"dVdTNormal", s.dVdTNormal_}, {"lRhoOffset", s.lRhoOffset_},
        {"lTOffset", s.lTOffset_}};
    auto it = m.find(nm);
    if (it == m.end()) return -1;
    *out = it->second;
    return 0;
  }
  if (e->IsType<SpinerEOSDependsRhoSie>()) {
    const auto s = e->get<SpinerEOSDependsRhoSie>();
    const std::map<std::string, double> m = {
        {"lRhoMin", s.lRhoMin_},       {"lRhoMax", s.lRhoMax_},   {"rhoMax", s.rhoMax_},
        {"PMin", s.PMin_},             {"rhoNormal", s.rhoNormal_}, {"TNormal", s.TNormal_},
        {"sieNormal", s.sieNormal_},   {"PNormal", s.PNormal_},   {"CvNormal", s.CvNormal_},
        {"bModNormal", s.bModNormal_}, {"dPdENormal", s.dPdENormal_},
        {"dVdTNormal", s.dVdTNormal_}, {"lRhoOffset", s.lRhoOffset_},
        {"lTOffset", s.lTOffset_},     {"lEOffset", s.lEOffset_}};
    auto it = m.find(nm);
    if (it == m.end()) return -1;
    *out = it->second;
    return 0;
  }
  return -1;
}

// ---- vector lookups through the reference's Variant vector API -------------------
// method: enum sgx_method.  The array is cut into one chunk per OpenMP thread; each chunk is
// ONE call of the reference vector overload with Exec::Host() on a thread-private copy of the
// (shallow, trivially copyable) EOS object.
extern "C++" {
namespace {
template <class Lam, class... Lead>
void dispatch_method(EOS &local, int method, const double *a, const double *b, double *o,
                     int cnt, Lam &lam, Lead &&...lead) {
  switch (method) {
  case SGX_P_FROM_RHO_T:
    local.PressureFromDensityTemperature(lead..., a, b, o, cnt, lam);
    break;
  case SGX_E_FROM_RHO_T:
    local.InternalEnergyFromDensityTemperature(lead..., a, b, o, cnt, lam);
    break;
  case SGX_T_FROM_RHO_E:
    local.TemperatureFromDensityInternalEnergy(lead..., a, b, o, cnt, lam);
    break;
  case SGX_P_FROM_RHO_E:
    local.PressureFromDensityInternalEnergy(lead..., a, b, o, cnt, lam);
    break;
  case SGX_CV_FROM_RHO_T:
    local.SpecificHeatFromDensityTemperature(lead..., a, b, o, cnt, lam);
    break;
  case SGX_CV_FROM_RHO_E:
    local.SpecificHeatFromDensityInternalEnergy(lead..., a, b, o, cnt, lam);
    break;
  case SGX_BMOD_FROM_RHO_T:
    local.BulkModulusFromDensityTemperature(lead..., a, b, o, cnt, lam);
    break;
  case SGX_BMOD_FROM_RHO_E:
    local.BulkModulusFromDensityInternalEnergy(lead..., a, b, o, cnt, lam);
    break;
  case SGX_GAMMA_FROM_RHO_T:
    local.GruneisenParamFromDensityTemperature(lead..., a, b, o, cnt, lam);
    break;
  case SGX_GAMMA_FROM_RHO_E:
    local.GruneisenParamFromDensityInternalEnergy(lead..., a, b, o, cnt, lam);
    break;
  default:
    throw std::runtime_error("bad method");
  }
}
} // namespace
} // extern "C++"

int ref_eos_eval(void *h, int method, const double *in1, const double *in2, double *out,
                 int64_t n, double *lambda, int lambda_stride, int nthreads) {
  const EOS *eos = static_cast<const EOS *>(h);
#ifdef SGREF_KOKKOS
  { // the whole array in one call of the reference's vector overload (Kokkos::parallel_for)
    kokkos_init_once(nthreads);
    EOS local = *eos;
    StridedLambda lam{lambda, lambda_stride};
    return guarded([&]() {
      dispatch_method(local, method, in1, in2, out, static_cast<int>(n), lam);
      Kokkos::fence();
    });
  }
#endif
  const int nt = nthreads_for(nthreads);
  int rc = 0;
#pragma omp parallel for num_threads(nt) schedule(static)
  for (int t = 0; t < nt; ++t) {
    const int64_t lo = n * t / nt, hi = n * (t + 1) / nt;
    const int cnt = static_cast<int>(hi - lo);
    if (cnt <= 0) continue;
    EOS local = *eos;
    const double *a = in1 + lo, *b = in2 + lo;
    double *o = out + lo;
    StridedLambda lam{lambda ? lambda + lo * lambda_stride : nullptr, lambda_stride};
    auto host = PortsOfCall::Exec::Host();
    int lrc = guarded([&]() { dispatch_method(local, method, a, b, o, cnt, lam, host); });
    if (lrc != 0) {
#pragma omp atomic write
      rc = -1;
    }
  }
  return rc;
}

// The remaining vector-surface members, applied per cell through the Variant's scalar forwards:
// op 0 MinInternalEnergyFromDensity, 1 InternalEnergyFromDensityPressure (no-guess form,
// eos_base.hpp:709-721), 2/3 EntropyFromDensity{Temperature,InternalEnergy}, 4/5
// GibbsFreeEnergyFromDensity{Temperature,InternalEnergy}, 6 DensityEnergyFromPressureTemperature
// (out1 = rho in/out, out2 = sie).
int ref_eos_eval_extra(void *h, int op, const double *in1, const double *in2, double *out1,
                       double *out2, int64_t n, double *lambda, int lambda_stride) {
  const EOS *eos = static_cast<const EOS *>(h);
  return guarded([&]() {
    if (op == 1) { // only the vector overload reaches the no-guess form through the Variant
      StridedLambda lam{lambda, lambda_stride};
      eos->InternalEnergyFromDensityPressure(PortsOfCall::Exec::Host(), in1, in2, out1,
                                             static_cast<int>(n), lam);
      return;
    }
    for (int64_t i = 0; i < n; ++i) {
      Real *lam = lambda ? lambda + i * lambda_stride : nullptr;
      switch (op) {
      case 0:
        out1[i] = eos->MinInternalEnergyFromDensity(in1[i], lam);
        break;
      case 2:
        out1[i] = eos->EntropyFromDensityTemperature(in1[i], in2[i], lam);
        break;
      case 3:
        out1[i] = eos->EntropyFromDensityInternalEnergy(in1[i], in2[i], lam);
        break;
      case 4:
        out1[i] = eos->GibbsFreeEnergyFromDensityTemperature(in1[i], in2[i], lam);
        break;
      case 5:
        out1[i] = eos->GibbsFreeEnergyFromDensityInternalEnergy(in1[i], in2[i], lam);
        break;
      case 6: {
        Real rho = out1[i], sie = 0.0;
        eos->DensityEnergyFromPressureTemperature(in1[i], in2[i], lam, rho, sie);
        out1[i] = rho;
        out2[i] = sie;
      } break;
      default:
        throw std::runtime_error("bad op");
      }
    }
  });
}

// FillEos vector overload.  az_only != 0 => type-indexed {Abar,Zbar} lambda (no lT slot).
int ref_eos_fill(void *h, double *rho, double *T, double *e, double *P, double *cv,
                 double *bmod, int64_t n, unsigned long output, double *lambda,
                 int lambda_stride, int az_only, int nthreads) {
  const EOS *eos = static_cast<const EOS *>(h);
#ifdef SGREF_KOKKOS
  { // one call of the reference's FillEos vector overload over the whole array
    kokkos_init_once(nthreads);
    EOS local = *eos;
    return guarded([&]() {
      if (az_only) {
        StridedAZLambda lam{lambda, lambda_stride};
        local.FillEos(rho, T, e, P, cv, bmod, static_cast<int>(n), output, lam);
      } else {
        StridedLambda lam{lambda, lambda_stride};
        local.FillEos(rho, T, e, P, cv, bmod, static_cast<int>(n), output, lam);
      }
      Kokkos::fence();
    });
  }
#endif
  const int nt = nthreads_for(nthreads);
  int rc = 0;
#pragma omp parallel for num_threads(nt) schedule(static)
  for (int t = 0; t < nt; ++t) {
    const int64_t lo = n * t / nt, hi = n * (t + 1) / nt;
    const int cnt = static_cast<int>(hi - lo);
    if (cnt <= 0) continue;
    EOS local = *eos;
    auto host = PortsOfCall::Exec::Host();
    int lrc = guarded([&]() {
      if (az_only) {
        StridedAZLambda lam{lambda + lo * lambda_stride, lambda_stride};
        local.FillEos(host, rho + lo, T + lo, e + lo, P + lo, cv + lo, bmod + lo, cnt, output,
                      lam);
      } else {
        StridedLambda lam{lambda ? lambda + lo * lambda_stride : nullptr, lambda_stride};
        local.FillEos(host, rho + lo, T + lo, e + lo, P + lo, cv + lo, bmod + lo, cnt, output,
                      lam);
      }
    });
    if (lrc != 0) {
#pragma omp atomic write
      rc = -1;
    }
  }
  return rc;
}

// ---- PTE: the cell loop of test/test_pte.cpp:177-265 ------------------------------
namespace {
struct RowIndexer { // == Indexer2D over a [ncell][nmat] row-major array
  double *row;
  PORTABLE_FORCEINLINE_FUNCTION double &operator[](const int i) const { return row[i]; }
};
struct EosIndexer {
  EOS *e;
  PORTABLE_FORCEINLINE_FUNCTION EOS &operator[](const int i) const { return e[i]; }
};
MixParams to_params(const sgx_mix_params *p) {
  MixParams m;
  if (!p) return m;
  m.verbose = p->verbose != 0;
  m.iterate_t_guess = p->iterate_t_guess != 0;
  m.derivative_eps = p->derivative_eps;
  m.pte_rel_tolerance_p = p->pte_rel_tolerance_p;
  m.pte_rel_tolerance_t = p->pte_rel_tolerance_t;
  m.pte_rel_tolerance_e = p->pte_rel_tolerance_e;
  m.pte_rel_tolerance_v = p->pte_rel_tolerance_v;
  m.pte_abs_tolerance_p = p->pte_abs_tolerance_p;
  m.pte_abs_tolerance_t = p->pte_abs_tolerance_t;
  m.pte_slow_convergence_thresh = p->pte_slow_convergence_thresh;
  m.pte_rel_tolerance_p_sufficient = p->pte_rel_tolerance_p_sufficient;
  m.pte_rel_tolerance_t_sufficient = p->pte_rel_tolerance_t_sufficient;
  m.pte_rel_tolerance_e_sufficient = p->pte_rel_tolerance_e_sufficient;
  m.pte_rel_tolerance_v_sufficient = p->pte_rel_tolerance_v_sufficient;
  m.pte_abs_tolerance_p_sufficient = p->pte_abs_tolerance_p_sufficient;
  m.pte_abs_tolerance_t_sufficient = p->pte_abs_tolerance_t_sufficient;
  m.pte_max_iter_per_mat = p->pte_max_iter_per_mat;
  m.line_search_alpha = p->line_search_alpha;
  m.line_search_max_iter = p->line_search_max_iter;
  m.line_search_fac = p->line_search_fac;
  m.vfrac_safety_fac = p->vfrac_safety_fac;
  m.temperature_limit = p->temperature_limit;
  m.default_tguess = p->default_tguess;
  m.min_dtde = p->min_dtde;
  m.pte_small_step_tries = p->pte_small_step_tries;
  m.pte_small_step_thresh = p->pte_small_step_thresh;
  m.pte_max_dpdv = p->pte_max_dpdv;
  m.pte_reduced_system_exclude_idx = p->pte_reduced_system_exclude_idx;
  return m;
}
uint64_t pack_status(const SolverStatus &s, std::size_t niter) {
  auto clamp = [](std::size_t v, std::size_t mx) { return v > mx ? mx : v; };
  uint64_t w = 0;
  w |= s.converged ? 1ull : 0ull;
  w |= s.slow_convergence_detected ? 2ull : 0ull;
  w |= static_cast<uint64_t>(clamp(niter, 0xffff)) << 8;
  w |= static_cast<uint64_t>(clamp(s.max_line_niter, 0xff)) << 24;
  w |= static_cast<uint64_t>(clamp(s.small_step_iters, 0xffff)) << 32;
  w |= static_cast<uint64_t>(clamp(s.max_niter, 0xffff)) << 48;
  return w;
}
} // namespace

void ref_mix_params_default(sgx_mix_params *p) {
  MixParams m;
  p->verbose = m.verbose;
  p->iterate_t_guess = m.iterate_t_guess;
  p->derivative_eps = m.derivative_eps;
  p->pte_rel_tolerance_p = m.pte_rel_tolerance_p;
  p->pte_rel_tolerance_t = m.pte_rel_tolerance_t;
  p->pte_rel_tolerance_e = m.pte_rel_tolerance_e;
  p->pte_rel_tolerance_v = m.pte_rel_tolerance_v;
  p->pte_abs_tolerance_p = m.pte_abs_tolerance_p;
  p->pte_abs_tolerance_t = m.pte_abs_tolerance_t;
  p->pte_slow_convergence_thresh = m.pte_slow_convergence_thresh;
  p->pte_rel_tolerance_p_sufficient = m.pte_rel_tolerance_p_sufficient;
  p->pte_rel_tolerance_t_sufficient = m.pte_rel_tolerance_t_sufficient;
  p->pte_rel_tolerance_e_sufficient = m.pte_rel_tolerance_e_sufficient;
  p->pte_rel_tolerance_v_sufficient = m.pte_rel_tolerance_v_sufficient;
  p->pte_abs_tolerance_p_sufficient = m.pte_abs_tolerance_p_sufficient;
  p->pte_abs_tolerance_t_sufficient = m.pte_abs_tolerance_t_sufficient;
  p->pte_max_iter_per_mat = m.pte_max_iter_per_mat;
  p->line_search_alpha = m.line_search_alpha;
  p->line_search_max_iter = m.line_search_max_iter;
  p->line_search_fac = m.line_search_fac;
  p->vfrac_safety_fac = m.vfrac_safety_fac;
  p->temperature_limit = m.temperature_limit;
  p->default_tguess = m.default_tguess;
  p->min_dtde = m.min_dtde;
  p->pte_small_step_tries = m.pte_small_step_tries;
  p->pte_small_step_thresh = m.pte_small_step_thresh;
  p->pte_max_dpdv = m.pte_max_dpdv;
  p->pte_reduced_system_exclude_idx = m.pte_reduced_system_exclude_idx;
}

extern "C++" {
namespace {
using singularity::mix_impl::CacheAccessor;
// One cell of the loop of test/test_pte.cpp:177-265.
struct PteArgs {
  int solver, nmat;
  double *rho, *vfrac, *sie, *temp, *press;
  const double *sie_tot_in, *vfrac_tot_in, *Tguess_in;
  uint64_t *status;
  double *residual, *Tguess_out;
  double *lambda; // [ncell][nmat][MAX_NUM_LAMBDAS] caches (CacheAccessor layout) or null
};
template <typename LambdaIndexer>
inline int pte_one_cell_l(const PteArgs &A, const MixParams &mp, const EosIndexer &eos,
                          double *scratch, const int64_t t, LambdaIndexer lambda) {
  const int nmat = A.nmat;
  RowIndexer r{A.rho + t * nmat}, vf{A.vfrac + t * nmat}, e{A.sie + t * nmat},
      tp{A.temp + t * nmat}, pr{A.press + t * nmat};
  Real sie_tot = 0.0, rho_tot = 0.0;
  for (int i = 0; i < nmat; i++) { // test/test_pte.cpp:188-195
    const Real rhobar = r[i] * vf[i];
    rho_tot += rhobar;
    sie_tot += rhobar * e[i];
  }
  sie_tot /= rho_tot;
  if (A.sie_tot_in) sie_tot = A.sie_tot_in[t];
  const Real vfrac_tot = A.vfrac_tot_in ? A.vfrac_tot_in[t] : 1.0;
  Real Tguess;
  if (A.Tguess_in) {
    Tguess = A.Tguess_in[t];
  } else {
    Tguess = ApproxTemperatureFromRhoMatU(nmat, eos, rho_tot * sie_tot, r, vf, 0.0, lambda);
  }
  if (A.Tguess_out) A.Tguess_out[t] = Tguess;
  SolverStatus st;
  std::size_t niter = 0;
  const int solver = A.solver;
  int lrc = guarded([&]() {
    if (solver == SGX_PTE_RHOT) {
      PTESolverRhoT<EosIndexer, RowIndexer, LambdaIndexer> method(
          nmat, eos, vfrac_tot, sie_tot, r, vf, e, tp, pr, lambda, scratch, Tguess, mp);
      st = PTESolver(method);
      niter = method.Niter();
    } else if (solver == SGX_PTE_RHOU) {
      PTESolverRhoU<EosIndexer, RowIndexer, LambdaIndexer> method(
          nmat, eos, vfrac_tot, sie_tot, r, vf, e, tp, pr, lambda, scratch, Tguess, mp);
      st = PTESolver(method);
      niter = method.Niter();
    } else if (solver == SGX_PTE_FIXED_P) {
      PTESolverFixedP<EosIndexer, RowIndexer, LambdaIndexer> method(
          nmat, eos, vfrac_tot, Tguess /* P */, r, vf, e, tp, pr, lambda, scratch, mp);
      st = PTESolver(method);
      niter = method.Niter();
    } else if (solver == SGX_PTE_FIXED_T) {
      PTESolverFixedT<EosIndexer, RowIndexer, LambdaIndexer> method(
          nmat, eos, vfrac_tot, Tguess /* T_true */, r, vf, e, tp, pr, lambda, scratch, mp);
      st = PTESolver(method);
      niter = method.Niter();
    } else {
      PTESolverPT<EosIndexer, RowIndexer, LambdaIndexer> method(
          nmat, eos, vfrac_tot, sie_tot, r, vf, e, tp, pr, lambda, scratch, Tguess, mp);
      st = PTESolver(method);
      niter = method.Niter();
    }
  });
  if (A.status) A.status[t] = pack_status(st, niter);
  if (A.residual) A.residual[t] = st.residual;
  return lrc;
}
inline int pte_one_cell(const PteArgs &A, const MixParams &mp, const EosIndexer &eos,
                        double *scratch, const int64_t t) {
  if (A.lambda)
    return pte_one_cell_l(A, mp, eos, scratch, t,
                          CacheAccessor(A.lambda + t * A.nmat * MAX_NUM_LAMBDAS));
  return pte_one_cell_l(A, mp, eos, scratch, t, NullIndexer());
}
} // namespace
} // extern "C++"

int ref_pte_solve_cached(int solver, void **eos_handles, int nmat, int64_t ncell, double *rho,
                         double *vfrac, double *sie, double *temp, double *press,
                         const double *sie_tot_in, const double *vfrac_tot_in,
                         const double *Tguess_in, const sgx_mix_params *params,
                         uint64_t *status, double *residual, double *Tguess_out, double *lambda,
                         int nthreads);
int ref_pte_solve(int solver, void **eos_handles, int nmat, int64_t ncell, double *rho,
                  double *vfrac, double *sie, double *temp, double *press,
                  const double *sie_tot_in, const double *vfrac_tot_in,
                  const double *Tguess_in, const sgx_mix_params *params, uint64_t *status,
                  double *residual, double *Tguess_out, int nthreads) {
  return ref_pte_solve_cached(solver, eos_handles, nmat, ncell, rho, vfrac, sie, temp, press,
                              sie_tot_in, vfrac_tot_in, Tguess_in, params, status, residual,
                              Tguess_out, nullptr, nthreads);
}
// lambda: per-cell CacheAccessor storage [ncell][nmat][MAX_NUM_LAMBDAS], or null (NullIndexer)
int ref_pte_solve_cached(int solver, void **eos_handles, int nmat, int64_t ncell, double *rho,
                         double *vfrac, double *sie, double *temp, double *press,
                         const double *sie_tot_in, const double *vfrac_tot_in,
                         const double *Tguess_in, const sgx_mix_params *params,
                         uint64_t *status, double *residual, double *Tguess_out, double *lambda,
                         int nthreads) {
  const MixParams mp = to_params(params);
  std::vector<EOS> eos_vec(nmat);
  for (int m = 0; m < nmat; ++m)
    eos_vec[m] = *static_cast<EOS *>(eos_handles[m]);
  const PteArgs A{solver, nmat, rho, vfrac, sie, temp, press, sie_tot_in, vfrac_tot_in,
                  Tguess_in, status, residual, Tguess_out, lambda};
  const int nscr = (int)std::max<std::size_t>(
      std::max<std::size_t>(PTESolverRhoTRequiredScratch(nmat), PTESolverPTRequiredScratch(nmat)),
      std::max<std::size_t>(std::max<std::size_t>(PTESolverFixedTRequiredScratch(nmat),
                                                  PTESolverFixedPRequiredScratch(nmat)),
                            PTESolverRhoURequiredScratch(nmat)));
#ifdef SGREF_KOKKOS
  { // the reference test's own shape: one portableFor over cells, per-cell scratch slices
    // (test/test_pte.cpp:146-150,177: `scratch + nscratch_vars * t`)
    kokkos_init_once(nthreads);
    std::vector<double> scratch_all(static_cast<std::size_t>(ncell) * nscr);
    double *scr = scratch_all.data();
    EOS *eos_p = eos_vec.data();
    int nfail = 0;
    portableReduce(
        "PTE", 0, static_cast<int>(ncell),
        PORTABLE_LAMBDA(const int t, int &nf) {
          EosIndexer eos{eos_p};
          nf += pte_one_cell(A, mp, eos, scr + static_cast<std::size_t>(t) * nscr, t) != 0;
        },
        nfail);
    return nfail ? -1 : 0;
  }
#endif
  const int nt = nthreads_for(nthreads);
  int rc = 0;
#pragma omp parallel num_threads(nt)
  {
    std::vector<EOS> eos_local(eos_vec);
    EosIndexer eos{eos_local.data()};
    std::vector<double> scratch(nscr);
#pragma omp for schedule(dynamic, 256)
    for (int64_t t = 0; t < ncell; ++t) {
      if (pte_one_cell(A, mp, eos, scratch.data(), t) != 0) {
#pragma omp atomic write
        rc = -1;
      }
    }
  }
  return rc;
}

// Demonstration for DESIGN.md 7: PTESolverFixedP::Init() with iterate_t_guess = true reads the
// base-class members uscale / utotal_scale (mixed_cell_models.hpp:537) before anything has
// written them (:1747-1755 -- the constructor chain :371-387 leaves them uninitialised).  The
// solver object of ONE cell is placement-constructed over a buffer pre-filled with the bit
// pattern of `fill`, Init() runs, and the caller gets back what Init() left in sie[] (=
// e(rho, Tguess), i.e. the temperature guess it settled on).  If the answer depends on `fill`,
// the reference has no defined result to match.
int ref_pte_fixedp_init_probe(void **eos_handles, int nmat, double *rho, double *vfrac,
                              double *sie, double *temp, double *press, double P,
                              const sgx_mix_params *params, double fill) {
  const MixParams mp = to_params(params);
  std::vector<EOS> eos_vec(nmat);
  for (int m = 0; m < nmat; ++m)
    eos_vec[m] = *static_cast<EOS *>(eos_handles[m]);
  EosIndexer eos{eos_vec.data()};
  std::vector<double> scratch(PTESolverFixedPRequiredScratch(nmat));
  using Solver = PTESolverFixedP<EosIndexer, RowIndexer, NullIndexer>;
  constexpr std::size_t nd = (sizeof(Solver) + sizeof(double) - 1) / sizeof(double);
  alignas(alignof(Solver) > alignof(double) ? alignof(Solver) : alignof(double)) double buf[nd];
  for (std::size_t k = 0; k < nd; ++k)
    buf[k] = fill;
  RowIndexer r{rho}, vf{vfrac}, e{sie}, tp{temp}, pr{press};
  NullIndexer no_lambda;
  return guarded([&]() {
    Solver *method = new (buf) Solver(nmat, eos, 1.0, P, r, vf, e, tp, pr, no_lambda,
                                      scratch.data(), mp);
    method->Init();
    method->~Solver();
  });
}

// The reference's own get_sg_eos (eos/get_sg_eos.cpp:36-280; Kokkos-only code, so only the
// Kokkos build of this library has it).  eos_handles[k] are this harness's EOS objects.
#ifdef SGREF_KOKKOS
// the reference's C entry point (eos/singularity_eos.hpp:114-137, defined in get_sg_eos.cpp)
int get_sg_eos(int nmat, int ncell, int cell_dim, int input_int, int *eos_offsets,
               singularity::EOS *eos, int *offsets, double *press, double *pmax, double *vol,
               double *spvol, double *sie, double *temp, double *bmod, double *dpde, double *cv,
               double *frac_mass, double *frac_vol, double *frac_ie, double *frac_bmod,
               double *frac_dpde, double *frac_cv, double mass_frac_cutoff);
#endif
int ref_get_sg_eos(int nmat, int ncell, int cell_dim, int input_int, int *eos_offsets,
                   void **eos_handles, int neos, int *offsets, double *press, double *pmax,
                   double *vol, double *spvol, double *sie, double *temp, double *bmod,
                   double *dpde, double *cv, double *frac_mass, double *frac_vol,
                   double *frac_ie, double *frac_bmod, double *frac_dpde, double *frac_cv,
                   double mass_frac_cutoff) {
#ifdef SGREF_KOKKOS
  kokkos_init_once(0);
  std::vector<EOS> eos_vec(neos);
  for (int k = 0; k < neos; ++k)
    eos_vec[k] = *static_cast<EOS *>(eos_handles[k]);
  return guarded([&]() {
    get_sg_eos(nmat, ncell, cell_dim, input_int, eos_offsets, eos_vec.data(), offsets, press,
               pmax, vol, spvol, sie, temp, bmod, dpde, cv, frac_mass, frac_vol, frac_ie,
               frac_bmod, frac_dpde, frac_cv, mass_frac_cutoff);
  });
#else
  return -99; // PORTABILITY_STRATEGY_NONE: the reference's get_sg_eos is an empty function
#endif
}

// ApproxTemperatureFromRhoMatU alone (closure/mixed_cell_models.hpp:790-847)
void ref_pte_approx_T(void **eos_handles, int nmat, int64_t ncell, const double *rho,
                      const double *vfrac, const double *u_tot, double *Tout) {
  std::vector<EOS> eos_vec(nmat);
  for (int m = 0; m < nmat; ++m)
    eos_vec[m] = *static_cast<EOS *>(eos_handles[m]);
  EosIndexer eos{eos_vec.data()};
  for (int64_t t = 0; t < ncell; ++t) {
    RowIndexer r{const_cast<double *>(rho) + t * nmat},
        vf{const_cast<double *>(vfrac) + t * nmat};
    Tout[t] = ApproxTemperatureFromRhoMatU(nmat, eos, u_tot[t], r, vf);
  }
}

// ---- synthetic-table helpers: reference Bounds grids (base/spiner_table_utils.hpp:56-70) --
// Single-patch log grid exactly as sesame2spiner builds it; also returns the linear-space
// node positions i2lin(i) (:137-146).
int ref_bounds_log_grid(double xmin, double xmax, int n, sgx_grid1d *grid, double *offset,
                        double *nodes_lin) {
  table_utils::Bounds<3> b(xmin, xmax, n, true);
  grid->ngrids = 1;
  grid->npoints[0] = n;
  grid->min[0] = b.grid.min();
  grid->max[0] = b.grid.max();
  grid->dx[0] = b.grid.dx(0);
  *offset = b.offset;
  for (int i = 0; i < n; ++i)
    nodes_lin[i] = b.i2lin(i);
  return 0;
}

} // extern "C"
