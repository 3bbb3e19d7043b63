// CPU check of csrc/sgx_libm.cuh (compiled as plain C++) against the live host libm:
// bit equality of exp / log / pow / log10 / cbrt over seeded input families that cover the
// operand ranges of the Davis EOS, Helmholtz, and every special-case branch.
// Usage: libm_check <n per family> ; prints one line per function, exits 1 on any mismatch.
// Test infrastructure (tests/test_libm_parity.py builds and runs it).
#include <cinttypes>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <gnu/libc-version.h>

#include "../../singularity-eos_b200/csrc/sgx_libm.cuh"

static uint64_t splitmix(uint64_t &s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
static double u01(uint64_t &s) { return (splitmix(s) >> 11) * 0x1p-53; }
static double bitsd(uint64_t u) { double d; std::memcpy(&d, &u, 8); return d; }
static uint64_t dbits(double d) { uint64_t u; std::memcpy(&u, &d, 8); return u; }
static bool same(double a, double b) {
  if (a != a && b != b) return true;  // any NaN equals any NaN
  return dbits(a) == dbits(b);
}

static long bad_total = 0;
template <class F, class G, class Gen>
static void run1(const char *name, const char *family, long n, uint64_t seed, F mine, G ref, Gen gen) {
  long bad = 0;
  uint64_t s = seed;
  for (long i = 0; i < n; ++i) {
    const double x = gen(s);
    volatile double xv = x;
    const double a = mine(x), b = ref(xv);
    if (!same(a, b)) {
      if (bad < 5) std::printf("  MISMATCH %s[%s](%a) = %a, libm %a\n", name, family, x, a, b);
      ++bad;
    }
  }
  std::printf("%-6s %-28s n=%ld mismatches=%ld\n", name, family, n, bad);
  bad_total += bad;
}
template <class Gen>
static void run2(const char *family, long n, uint64_t seed, Gen gen) {
  long bad = 0;
  uint64_t s = seed;
  for (long i = 0; i < n; ++i) {
    double x, y;
    gen(s, x, y);
    volatile double xv = x, yv = y;
    const double a = sgx::libm::pow(x, y), b = std::pow(xv, yv);
    if (!same(a, b)) {
      if (bad < 5) std::printf("  MISMATCH pow[%s](%a, %a) = %a, libm %a\n", family, x, y, a, b);
      ++bad;
    }
  }
  std::printf("%-6s %-28s n=%ld mismatches=%ld\n", "pow", family, n, bad);
  bad_total += bad;
}

int main(int argc, char **argv) {
  const long n = argc > 1 ? std::atol(argv[1]) : 1000000;
  std::printf("glibc %s\n", gnu_get_libc_version());
  auto e = [](double x) { return sgx::libm::exp(x); };
  auto er = [](double x) { return std::exp(x); };
  run1("exp", "uniform[-750,715]", n, 1, e, er, [](uint64_t &s) { return -750.0 + 1465.0 * u01(s); });
  run1("exp", "uniform[-40,40]", n, 2, e, er, [](uint64_t &s) { return -40.0 + 80.0 * u01(s); });
  run1("exp", "ln10*[-320,310]", n, 3, e, er, [](uint64_t &s) { return 2.30258509299405e+00 * (-320.0 + 630.0 * u01(s)); });
  run1("exp", "tiny+-2^[-70,-1]", n, 4, e, er, [](uint64_t &s) { double m = std::ldexp(1.0 + u01(s), -70 + int(u01(s) * 69)); return u01(s) < 0.5 ? m : -m; });
  run1("exp", "random bits", n, 5, e, er, [](uint64_t &s) { return bitsd(splitmix(s)); });
  auto l = [](double x) { return sgx::libm::log(x); };
  auto lr = [](double x) { return std::log(x); };
  run1("log", "log-uniform all normals", n, 11, l, lr, [](uint64_t &s) { return std::ldexp(1.0 + u01(s), -1022 + int(u01(s) * 2045)); });
  run1("log", "[0.9,1.1]", n, 12, l, lr, [](uint64_t &s) { return 0.9 + 0.2 * u01(s); });
  run1("log", "[1e-3,1e3]", n, 13, l, lr, [](uint64_t &s) { return std::exp(-6.9 + 13.8 * u01(s)); });
  run1("log", "subnormal", n / 4, 14, l, lr, [](uint64_t &s) { return bitsd(splitmix(s) >> 12); });
  run1("log", "random bits", n, 15, l, lr, [](uint64_t &s) { return bitsd(splitmix(s)); });
  auto l10 = [](double x) { return sgx::libm::log10(x); };
  auto l10r = [](double x) { return std::log10(x); };
  run1("log10", "log-uniform all normals", n, 21, l10, l10r, [](uint64_t &s) { return std::ldexp(1.0 + u01(s), -1022 + int(u01(s) * 2045)); });
  run1("log10", "T in 10^[3,13]", n, 22, l10, l10r, [](uint64_t &s) { return std::exp(2.302585092994046 * (3.0 + 10.0 * u01(s))); });
  run1("log10", "[0.5,2]", n, 23, l10, l10r, [](uint64_t &s) { return 0.5 + 1.5 * u01(s); });
  run1("log10", "subnormal", n / 4, 24, l10, l10r, [](uint64_t &s) { return bitsd(splitmix(s) >> 12); });
  run1("log10", "random bits", n, 25, l10, l10r, [](uint64_t &s) { return bitsd(splitmix(s)); });
  auto c = [](double x) { return sgx::libm::cbrt(x); };
  auto cr = [](double x) { return std::cbrt(x); };
  run1("cbrt", "log-uniform all normals +-", n, 31, c, cr, [](uint64_t &s) { double m = std::ldexp(1.0 + u01(s), -1022 + int(u01(s) * 2045)); return u01(s) < 0.5 ? m : -m; });
  run1("cbrt", "[1e-12,1e15]", n, 32, c, cr, [](uint64_t &s) { return std::exp(-27.6 + 62.2 * u01(s)); });
  run1("cbrt", "subnormal", n / 4, 33, c, cr, [](uint64_t &s) { return bitsd(splitmix(s) >> 12); });
  run1("cbrt", "random bits", n, 34, c, cr, [](uint64_t &s) { return bitsd(splitmix(s)); });
  // pow: Davis exponents (eos/eos_davis.hpp:353-394: pow(vvc, +-n), pow(.., a/n), ...) on
  // compression ratios, general positive bases, negative bases with integer y, specials
  run2("x in [0.05,20], y in [-8,8]", n, 41, [](uint64_t &s, double &x, double &y) { x = std::exp(-3.0 + 6.0 * u01(s)); y = -8.0 + 16.0 * u01(s); });
  run2("x log-uniform, y in [-3,3]", n, 42, [](uint64_t &s, double &x, double &y) { x = std::ldexp(1.0 + u01(s), -1022 + int(u01(s) * 2045)); y = -3.0 + 6.0 * u01(s); });
  run2("x near 1, y huge", n, 43, [](uint64_t &s, double &x, double &y) { x = 1.0 + (u01(s) - 0.5) * 1e-6; y = (u01(s) - 0.5) * 2e9; });
  run2("x<0, y integer", n, 44, [](uint64_t &s, double &x, double &y) { x = -std::exp(-5.0 + 10.0 * u01(s)); y = std::floor(-40.0 + 80.0 * u01(s)); });
  run2("x<0, y any", n / 4, 45, [](uint64_t &s, double &x, double &y) { x = -std::exp(-5.0 + 10.0 * u01(s)); y = -40.0 + 80.0 * u01(s); });
  run2("over/underflow edge", n, 46, [](uint64_t &s, double &x, double &y) { x = std::exp(-5.0 + 10.0 * u01(s)); y = (u01(s) < 0.5 ? 1.0 : -1.0) * (700.0 + 60.0 * u01(s)) / std::log(x); });
  run2("x subnormal", n / 4, 47, [](uint64_t &s, double &x, double &y) { x = bitsd(splitmix(s) >> 12); y = -2.0 + 4.0 * u01(s); });
  run2("y tiny / huge", n / 4, 48, [](uint64_t &s, double &x, double &y) { x = std::exp(-5.0 + 10.0 * u01(s)); y = std::ldexp(1.0 + u01(s), u01(s) < 0.5 ? -80 + int(u01(s) * 20) : 60 + int(u01(s) * 10)); if (u01(s) < 0.5) y = -y; });
  run2("random bits both", n, 49, [](uint64_t &s, double &x, double &y) { x = bitsd(splitmix(s)); y = bitsd(splitmix(s)); });
  run2("specials grid", 11 * 11, 50, [](uint64_t &s, double &x, double &y) {
    static const double v[11] = {0.0, -0.0, 1.0, -1.0, 2.0, -2.0, 0.5, HUGE_VAL, -HUGE_VAL, NAN, 3.0};
    static int idx = 0; x = v[idx / 11]; y = v[idx % 11]; ++idx; (void)s; });
  std::printf("total mismatches %ld\n", bad_total);
  return bad_total ? 1 : 0;
}
