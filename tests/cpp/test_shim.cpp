// C++ host-side test of the sgx:: shim classes (include/sgx/eos.hpp, closure.hpp) over the C
// ABI.  Mirrors test/test_eos_vector.cpp:139-143,309-313 (IdealGas golden values through the
// raw-pointer vector overloads) and the acceptance checks of test/test_pte.cpp:204-262.
// Build: g++ -std=c++17 -Iinclude -I/usr/local/cuda/include tests/cpp/test_shim.cpp
//            -Lsingularity-eos_b200/lib -lsgx_b200 -L/usr/local/cuda/lib64 -lcudart
// `--link-only` exits 0 before touching the device (used by the CPU suite).
#include <cuda_runtime_api.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>

#include "sgx/closure.hpp"
#include "sgx/eos.hpp"

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
struct DevArray {
  double *p = nullptr;
  size_t n;
  explicit DevArray(size_t n_) : n(n_) { cudaMalloc((void **)&p, n * sizeof(double)); }
  DevArray(const std::vector<double> &h) : DevArray(h.size()) {
    cudaMemcpy(p, h.data(), n * sizeof(double), cudaMemcpyHostToDevice);
  }
  ~DevArray() { cudaFree(p); }
  std::vector<double> host() const {
    std::vector<double> h(n);
    cudaMemcpy(h.data(), p, n * sizeof(double), cudaMemcpyDeviceToHost);
    return h;
  }
};
static void expect(const DevArray &a, const std::vector<double> &gold, const char *what) {
  cudaDeviceSynchronize(); // the caller fences (test/test_eos_vector.cpp:200-203)
  const auto h = a.host();
  for (size_t i = 0; i < gold.size(); ++i)
    if (!isClose(h[i], gold[i])) {
      std::printf("%s[%zu]: got %.17g want %.17g\n", what, i, h[i], gold[i]);
      ++failures;
    }
}

int main(int argc, char **argv) {
  if (argc > 1 && std::strcmp(argv[1], "--link-only") == 0) {
    std::printf("abi %d\n", sgx_abi_version());
    return sgx_abi_version() == SGX_ABI_VERSION ? 0 : 1;
  }
  using namespace sgx;
  // --- "Vector EOS", GIVEN an ideal gas with gm1 = 0.4, Cv = 5 ------------------------
  {
    IdealGas host_eos(0.4, 5.0);
    auto eos = host_eos.GetOnDevice();
    constexpr int num = 3;
    DevArray density({1.0, 2.0, 5.0}), energy({5.0, 10.0, 15.0});
    DevArray temperature(num), pressure(num), bulkmodulus(num), heatcapacity(num), gruneisen(num);
    eos.TemperatureFromDensityInternalEnergy(density.p, energy.p, temperature.p, num);
    expect(temperature, {1.0, 2.0, 3.0}, "T(rho,e)");
    eos.PressureFromDensityInternalEnergy(density.p, energy.p, pressure.p, num);
    expect(pressure, {2.0, 8.0, 30.0}, "P(rho,e)");
    eos.SpecificHeatFromDensityInternalEnergy(density.p, energy.p, heatcapacity.p, num);
    expect(heatcapacity, {5.0, 5.0, 5.0}, "cv(rho,e)");
    eos.BulkModulusFromDensityInternalEnergy(density.p, energy.p, bulkmodulus.p, num);
    expect(bulkmodulus, {2.8, 11.2, 42.0}, "bmod(rho,e)");
    eos.GruneisenParamFromDensityInternalEnergy(density.p, energy.p, gruneisen.p, num);
    expect(gruneisen, {0.4, 0.4, 0.4}, "gamma(rho,e)");
    DevArray T2({50.0, 100.0, 150.0}), sie(num);
    eos.InternalEnergyFromDensityTemperature(density.p, T2.p, sie.p, num);
    expect(sie, {250.0, 500.0, 750.0}, "e(rho,T)");
    // scratch-taking overload (ignored scratch, eos_base.hpp:196-203)
    eos.PressureFromDensityTemperature(density.p, T2.p, pressure.p, nullptr, num);
    expect(pressure, {100.0, 400.0, 1500.0}, "P(rho,T)");
    eos.BulkModulusFromDensityTemperature(density.p, T2.p, bulkmodulus.p, num);
    expect(bulkmodulus, {140.0, 560.0, 2100.0}, "bmod(rho,T)");
    // FillEos: (rho,T) -> e, P, cv, bmod
    DevArray e2(num), P2(num), cv2(num), b2(num);
    eos.FillEos(density.p, T2.p, e2.p, P2.p, cv2.p, b2.p, num,
                thermalqs::specific_internal_energy | thermalqs::pressure |
                    thermalqs::specific_heat | thermalqs::bulk_modulus);
    expect(e2, {250.0, 500.0, 750.0}, "FillEos e");
    expect(P2, {100.0, 400.0, 1500.0}, "FillEos P");
    eos.Finalize();
    // errors are exceptions, never silent: a finalized EOS cannot be evaluated
    bool threw = false;
    try {
      eos.PressureFromDensityTemperature(density.p, T2.p, pressure.p, num);
    } catch (const Error &) {
      threw = true;
    }
    CHECK(threw);
  }
  // --- PTE: 3 analytic materials, acceptance checks of test/test_pte.cpp:204-262 --------
  {
    auto eos = set_eos_3mat_analytic();
    constexpr int nmat = 3, ncell = 2000;
    const double rho0[nmat] = {8.93, 1.89, 2.5};
    std::vector<double> rho(ncell * nmat), vf(ncell * nmat), sie(ncell * nmat),
        temp(ncell * nmat, 600.0), press(ncell * nmat, 0.0), sie0(nmat);
    for (int m = 0; m < nmat; ++m) { // sie = e(rho_m, 600) (pte_test_3mat_analytic.hpp:66-68)
      DevArray r(std::vector<double>{rho0[m]}), t(std::vector<double>{600.0}), o(1);
      eos[m].InternalEnergyFromDensityTemperature(r.p, t.p, o.p, 1);
      cudaDeviceSynchronize();
      sie0[m] = o.host()[0];
    }
    unsigned long long s = 88172645463325252ull;
    auto u = [&]() { // xorshift64 uniform in (0,1)
      s ^= s << 13; s ^= s >> 7; s ^= s << 17;
      return ((s >> 11) + 0.5) * (1.0 / 9007199254740992.0);
    };
    for (int c = 0; c < ncell; ++c) {
      double v[nmat], sum = 0;
      for (int m = 0; m < nmat; ++m) sum += (v[m] = -std::log(1.0 - u()) + 2.2e-15);
      for (int m = 0; m < nmat; ++m) {
        rho[c * nmat + m] = rho0[m];
        vf[c * nmat + m] = v[m] / sum;
        sie[c * nmat + m] = sie0[m];
      }
    }
    std::vector<double> rhobar0(ncell * nmat);
    for (size_t i = 0; i < rhobar0.size(); ++i) rhobar0[i] = rho[i] * vf[i];
    DevArray d_rho(rho), d_vf(vf), d_sie(sie), d_T(temp), d_P(press);
    unsigned long long *d_status = nullptr;
    cudaMalloc((void **)&d_status, ncell * sizeof(unsigned long long));
    MixParams params; // library defaults
    PTESolverRhoT(eos, ncell, d_rho.p, d_vf.p, d_sie.p, d_T.p, d_P.p, params,
                  reinterpret_cast<std::uint64_t *>(d_status));
    cudaDeviceSynchronize();
    std::vector<unsigned long long> status(ncell);
    cudaMemcpy(status.data(), d_status, ncell * sizeof(unsigned long long),
               cudaMemcpyDeviceToHost);
    cudaFree(d_status);
    const auto r = d_rho.host(), v = d_vf.host(), T = d_T.host(), P = d_P.host();
    int nconv = 0;
    for (int c = 0; c < ncell; ++c) {
      const SolverStatus st = SolverStatus::unpack(status[c]);
      if (!st.converged) continue;
      ++nconv;
      double vsum = 0;
      for (int m = 0; m < nmat; ++m) {
        vsum += v[c * nmat + m];
        CHECK(isClose(r[c * nmat + m] * v[c * nmat + m], rhobar0[c * nmat + m], 1e-12));
      }
      CHECK(isClose(vsum, 1.0, 1e-12));
      for (int m = 0; m + 1 < nmat; ++m) {
        CHECK(isClose(T[c * nmat + m], T[c * nmat + m + 1], 1e-5));
        CHECK(isClose(P[c * nmat + m], P[c * nmat + m + 1], 1e-5));
      }
    }
    std::printf("PTE RhoT: %d/%d cells converged\n", nconv, ncell);
    CHECK(nconv >= ncell / 2); // test/test_pte.cpp:368
  }
  std::printf(failures ? "FAILED (%d)\n" : "ok\n", failures);
  return failures ? 1 : 0;
}
