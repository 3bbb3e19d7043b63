// The reference-side plugin (plugins/sgx/*.hpp) compiled against the reference headers where they
// lie and run through singularity::Variant: the scenarios of test/test_eos_vector.cpp (IdealGas
// gm1 = 0.4, Cv = 5; density {1,2,5}, energy {5,10,15}, temperature {50,100,150}; gold arrays
// :53-57,139-143,309-313; tolerance 1e-12 as array_compare) with
//   * the Space-first raw-pointer overloads on HOST arrays (Exec::Host: CheckHostExecVectorApi,
//     test_eos_vector.cpp:36-118) and on DEVICE arrays (Exec::Device),
//   * the scratch-taking overloads and the no-Space overloads (eos_variant.hpp:56-143),
//   * a LambdaIndexer&& (strided rows, eos_base.hpp:196-203) on the tabulated EOS,
//   * FillEos, and a non-pointer indexer (falls back to the reference's own per-element loop),
// then SgxSpinerRhoT against the reference's own SpinerEOSDependsRhoT in the same Variant.
// TEST INFRASTRUCTURE: built by __graft_entry__.build() where /root/reference is mounted (include
// paths of oracle/Makefile); the binary travels to the GPU box.
// usage: test_plugin [--link-only] [sp5_raw_file]
#include <cuda_runtime_api.h>

#include <array>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include <ports-of-call/portability.hpp>
#include <singularity-eos/eos/eos.hpp>

#include "../../oracle/sp5_tree.hpp"
#include "../../plugins/sgx/sgx_variant.hpp"

using namespace singularity;
using PortsOfCall::Exec::Device;
using PortsOfCall::Exec::Host;

static int failures = 0;
#define CHECK(cond)                                                                      \
  do {                                                                                   \
    if (!(cond)) {                                                                       \
      std::printf("CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #cond);                \
      ++failures;                                                                        \
    }                                                                                    \
  } while (0)

// isClose / array_compare of test/eos_unit_test_helpers.hpp:60-81
static bool isClose(double a, double b, double eps = 1e-12) {
  return std::abs(b - a) / (std::abs(a + b) + 1e-20) <= eps;
}
template <class A, class B>
static void compare(int n, const A &got, const B &gold, const char *what, double eps = 1e-12) {
  for (int i = 0; i < n; ++i)
    if (!isClose(got[i], gold[i], eps)) {
      std::printf("%s[%d]: got %.17g want %.17g\n", what, i, (double)got[i], (double)gold[i]);
      ++failures;
    }
}
struct DevArray {
  double *p = nullptr;
  size_t n;
  explicit DevArray(size_t n_) : n(n_) { cudaMalloc((void **)&p, n * sizeof(double)); }
  DevArray(const std::vector<double> &h) : DevArray(h.size()) {
    cudaMemcpy(p, h.data(), n * sizeof(double), cudaMemcpyHostToDevice);
  }
  ~DevArray() { cudaFree(p); }
  std::vector<double> host() const {
    cudaDeviceSynchronize(); // the caller fences (test/test_eos_vector.cpp:200-203)
    std::vector<double> h(n);
    cudaMemcpy(h.data(), p, n * sizeof(double), cudaMemcpyDeviceToHost);
    return h;
  }
};
// lambda rows as data + stride * i (what lambdaIndexer2D of singularity_eos.cpp:38-49 is)
struct StridedLambda {
  Real *base;
  int stride;
  Real *operator[](const int i) const { return base + (std::ptrdiff_t)stride * i; }
};
// an indexer that is not a raw pointer (test_eos_vector.cpp uses Kokkos views / std::vector)
struct VecIndexer {
  double *p;
  double &operator[](const int i) const { return p[i]; }
};

struct RawFile {
  sgx_sp5_tables t;
  std::vector<std::vector<double>> store;
};
static bool read_raw(const char *path, RawFile &rf) {
  std::ifstream f(path, std::ios::binary);
  char magic[8];
  if (!f.read(magic, 8) || std::memcmp(magic, "SGXSP5R1", 8) != 0) return false;
  if (!f.read(reinterpret_cast<char *>(&rf.t), sizeof(rf.t))) return false;
  const double **members[] = {
      &rf.t.P,       &rf.t.sie,       &rf.t.bMod,      &rf.t.dPdRho,    &rf.t.dPdE,
      &rf.t.dTdRho,  &rf.t.dTdE,      &rf.t.dEdRho,    &rf.t.dEdT,      &rf.t.PCold,
      &rf.t.sieCold, &rf.t.bModCold,  &rf.t.dPdRhoCold, &rf.t.T_re,     &rf.t.P_re,
      &rf.t.bMod_re, &rf.t.dPdRho_re, &rf.t.dPdE_re,   &rf.t.dTdRho_re, &rf.t.dTdE_re,
      &rf.t.dEdRho_re};
  rf.store.resize(sizeof(members) / sizeof(members[0]));
  for (size_t k = 0; k < rf.store.size(); ++k) {
    int64_t n = 0;
    if (!f.read(reinterpret_cast<char *>(&n), 8)) return false;
    rf.store[k].resize(n);
    if (n && !f.read(reinterpret_cast<char *>(rf.store[k].data()), n * 8)) return false;
    *members[k] = n ? rf.store[k].data() : nullptr;
  }
  return true;
}

static void ideal_gas_scenarios() {
  constexpr Real Cv = 5.0, gm1 = 0.4;
  constexpr int num = 3;
  SgxEOS host_eos = MakeSgxIdealGas(gm1, Cv);
  SgxEOS eos = host_eos.GetOnDevice();
  CHECK(eos.EosType() == "SgxIdealGas");
  CHECK(eos.IsType<SgxIdealGas>());
  const std::array<Real, num> temperature_true{1., 2., 3.}, pressure_true{2., 8., 30.},
      bulkmodulus_true{2.8, 11.2, 42.}, cv_true{Cv, Cv, Cv}, gruneisen_true{gm1, gm1, gm1};
  const std::array<Real, num> energy_true{250., 500., 750.}, pressure_true2{100., 400., 1500.},
      bulkmodulus_true2{140., 560., 2100.};
  NullIndexer lambdas{};
  { // ---- Exec::Host, raw pointers to std::vector storage (test_eos_vector.cpp:36-118) --------
    std::vector<Real> density{1.0, 2.0, 5.0}, energy{5.0, 10.0, 15.0};
    std::vector<Real> temperature(num, -1.0), pressure(num, -1.0), scratch(num, -1.0);
    eos.TemperatureFromDensityInternalEnergy(Host(), density.data(), energy.data(),
                                             temperature.data(), num, lambdas);
    compare(num, temperature, temperature_true, "Host T(rho,e)");
    eos.PressureFromDensityInternalEnergy(Host(), density.data(), energy.data(), pressure.data(),
                                          scratch.data(), num, lambdas);
    compare(num, pressure, pressure_true, "Host P(rho,e) [scratch overload]");
    std::vector<Real> fT(num, -1.0), fP(num, -1.0), fcv(num, -1.0), fb(num, -1.0);
    eos.FillEos(Host(), density.data(), fT.data(), energy.data(), fP.data(), fcv.data(),
                fb.data(), num,
                thermalqs::temperature | thermalqs::pressure | thermalqs::specific_heat |
                    thermalqs::bulk_modulus,
                lambdas);
    compare(num, fT, temperature_true, "Host FillEos T");
    compare(num, fP, pressure_true, "Host FillEos P");
    compare(num, fcv, cv_true, "Host FillEos cv");
    compare(num, fb, bulkmodulus_true, "Host FillEos bmod");
  }
  { // ---- Exec::Device, device arrays ("Vector EOS" scenario, :120-330) ----------------------
    DevArray density({1.0, 2.0, 5.0}), energy({5.0, 10.0, 15.0}), Tin({50.0, 100.0, 150.0});
    DevArray out(num), scratch(num);
    eos.TemperatureFromDensityInternalEnergy(Device(), density.p, energy.p, out.p, num, lambdas);
    compare(num, out.host(), temperature_true, "Device T(rho,e)");
    eos.PressureFromDensityInternalEnergy(Device(), density.p, energy.p, out.p, num);
    compare(num, out.host(), pressure_true, "Device P(rho,e) [no lambda]");
    eos.BulkModulusFromDensityInternalEnergy(Device(), density.p, energy.p, out.p, scratch.p, num,
                                             lambdas);
    compare(num, out.host(), bulkmodulus_true, "Device bmod(rho,e) [scratch]");
    eos.SpecificHeatFromDensityInternalEnergy(density.p, energy.p, out.p, num); // no Space
    compare(num, out.host(), cv_true, "Device cv(rho,e) [no Space]");
    eos.GruneisenParamFromDensityInternalEnergy(density.p, energy.p, out.p, scratch.p, num,
                                                lambdas);
    compare(num, out.host(), gruneisen_true, "Device gamma(rho,e) [no Space, scratch]");
    eos.InternalEnergyFromDensityTemperature(Device(), density.p, Tin.p, out.p, num, lambdas);
    compare(num, out.host(), energy_true, "Device e(rho,T)");
    eos.PressureFromDensityTemperature(Device(), density.p, Tin.p, out.p, num, lambdas);
    compare(num, out.host(), pressure_true2, "Device P(rho,T)");
    eos.BulkModulusFromDensityTemperature(Device(), density.p, Tin.p, out.p, num, lambdas);
    compare(num, out.host(), bulkmodulus_true2, "Device bmod(rho,T)");
    DevArray fe(num), fP(num), fcv(num), fb(num);
    eos.FillEos(Device(), density.p, Tin.p, fe.p, fP.p, fcv.p, fb.p, num,
                thermalqs::specific_internal_energy | thermalqs::pressure |
                    thermalqs::specific_heat | thermalqs::bulk_modulus,
                lambdas);
    compare(num, fe.host(), energy_true, "Device FillEos e");
    compare(num, fP.host(), pressure_true2, "Device FillEos P");
    compare(num, fb.host(), bulkmodulus_true2, "Device FillEos bmod");
  }
  { // ---- a non-pointer indexer keeps the reference's own per-element loop (host) -------------
    std::vector<Real> density{1.0, 2.0, 5.0}, energy{5.0, 10.0, 15.0}, T(num, -1.0);
    VecIndexer d{density.data()}, e{energy.data()}, t{T.data()};
    eos.TemperatureFromDensityInternalEnergy(Host(), d, e, t, num, lambdas);
    compare(num, T, temperature_true, "indexer T(rho,e)");
  }
  // scalar calls still answer (the reference class inside the plugin)
  CHECK(isClose(eos.PressureFromDensityTemperature(5.0, 150.0), 1500.0));
  eos.Finalize();
  host_eos.Finalize();
}

static void spiner_scenarios(const char *raw_path) {
  RawFile rf;
  if (!read_raw(raw_path, rf)) {
    std::printf("cannot read %s\n", raw_path);
    ++failures;
    return;
  }
  const int matid = 1;
  const std::string fname = sgx_oracle::build_sp5(&rf.t, matid); // in-memory .sp5 (hdf5 shim)
  SgxEOS ours = MakeSgxSpinerRhoT(fname, matid, rf.t);
  SgxEOS theirs = SpinerEOSDependsRhoT(fname, matid);
  CHECK(ours.EosType() == "SgxSpinerEOSDependsRhoT");
  CHECK(ours.nlambda() == 2);
  const int n = 4099; // ragged
  std::vector<Real> rho(n), T(n), P_ref(n), T_ref(n), sie(n);
  unsigned long long s = 0x9E3779B97F4A7C15ull;
  auto u01 = [&]() {
    s ^= s << 13;
    s ^= s >> 7;
    s ^= s << 17;
    return (double)(s >> 11) * (1.0 / 9007199254740992.0);
  };
  for (int i = 0; i < n; ++i) {
    rho[i] = std::pow(10.0, 0.02 + 1.43 * u01());
    T[i] = std::pow(10.0, 0.05 + 4.9 * u01());
  }
  const int stride = 3; // wider than nlambda on purpose
  std::vector<Real> lam_ref(n * stride, -7.0), lam_h(n * stride, -7.0);
  StridedLambda L_ref{lam_ref.data(), stride};
  theirs.PressureFromDensityTemperature(Host(), rho.data(), T.data(), P_ref.data(), n, L_ref);
  theirs.InternalEnergyFromDensityTemperature(Host(), rho.data(), T.data(), sie.data(), n, L_ref);
  // device arrays + a strided LambdaIndexer&& through the Variant
  DevArray d_rho(rho), d_T(T), d_P(n), d_lam(lam_h), d_sie(sie), d_Tout(n);
  ours.PressureFromDensityTemperature(Device(), d_rho.p, d_T.p, d_P.p, n,
                                      StridedLambda{d_lam.p, stride});
  compare(n, d_P.host(), P_ref, "SgxSpinerRhoT P(rho,T) vs reference class");
  const auto lam_got = d_lam.host();
  for (int i = 0; i < n; ++i) {
    if (lam_got[i * stride] != lam_ref[i * stride] ||
        lam_got[i * stride + 1] != lam_ref[i * stride + 1] || lam_got[i * stride + 2] != -7.0) {
      std::printf("lambda row %d differs\n", i);
      ++failures;
      break;
    }
  }
  // T(rho, sie): the regula-falsi path, cold start from NaN rows on both sides
  std::vector<Real> nanrows(n * stride, std::nan(""));
  lam_ref = nanrows;
  theirs.TemperatureFromDensityInternalEnergy(Host(), rho.data(), sie.data(), T_ref.data(), n,
                                              StridedLambda{lam_ref.data(), stride});
  DevArray d_lam2(nanrows);
  ours.TemperatureFromDensityInternalEnergy(Device(), d_rho.p, d_sie.p, d_Tout.p, n,
                                            StridedLambda{d_lam2.p, stride});
  compare(n, d_Tout.host(), T_ref, "SgxSpinerRhoT T(rho,sie) vs reference class");
  // host arrays through the same object (Exec::Host -> the library's staged entry point)
  std::vector<Real> P_h(n, -1.0);
  ours.PressureFromDensityTemperature(Host(), rho.data(), T.data(), P_h.data(), n,
                                      StridedLambda{lam_h.data(), stride});
  compare(n, P_h, P_ref, "SgxSpinerRhoT P(rho,T) host arrays");
  // a lambda layout that is not data + stride * i is rejected, not mis-read
  struct Scattered {
    Real *a, *b;
    Real *operator[](const int i) const { return (i % 2) ? b + i : a + 5 * i; }
  };
  bool threw = false;
  try {
    ours.PressureFromDensityTemperature(Host(), rho.data(), T.data(), P_h.data(), n,
                                        Scattered{lam_h.data(), lam_ref.data()});
  } catch (const std::exception &) {
    threw = true;
  }
  CHECK(threw);
  ours.Finalize();
  theirs.Finalize();
}

int main(int argc, char **argv) {
  if (argc > 1 && std::strcmp(argv[1], "--link-only") == 0) {
    std::printf("abi %d\n", sgx_abi_version());
    return sgx_abi_version() == SGX_ABI_VERSION ? 0 : 1;
  }
  ideal_gas_scenarios();
  if (argc > 1) spiner_scenarios(argv[1]);
  std::printf(failures ? "FAILED (%d)\n" : "ok\n", failures);
  return failures ? 1 : 0;
}
