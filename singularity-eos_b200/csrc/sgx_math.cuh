// Device-side numerics shared by every kernel of libsgx_b200: the reference's
// "not-quite-transcendental" logs, robust division, Spiner grid weights and multilinear
// interpolation, and the 1-D root finders.  FP64 throughout, compiled with -fmad=false so
// every expression rounds exactly like the reference's x86-64 build (no FMA contraction).
//
// Reference (paths relative to the lanl/singularity-eos checkout):
//   base/fast-math/logs.hpp:188-292              lg / pow2 / log10 / pow10 (NQT order 2, aliased)
//   base/robust_utils.hpp:27-70                  SMALL, EPS, sgn, ratio
//   base/math_utils.hpp:32-60                    pow<N>, pow10
//   eos/eos_spiner_common.hpp:52-57              to_log / from_log
//   utils/spiner/spiner/regular_grid_1d.hpp:66-87     bound / index / weights
//   utils/spiner/spiner/piecewise_grid_1d.hpp:73-126  findGrid / weights
//   utils/spiner/spiner/databox.hpp:451-475      interpToReal 1-D / 2-D
//   base/root-finding-1d/root_finding.hpp:114-310     check_bracket, set_bracket,
//                                                regula_falsi, newton_raphson
#ifndef SGX_MATH_CUH_
#define SGX_MATH_CUH_

#include <cfloat>
#include <climits>
#include <cstdint>

#ifdef __CUDACC__
#define SGX_HD __host__ __device__ __forceinline__
#define SGX_HDI __host__ __device__ inline
#define SGX_D __device__ __forceinline__
#else
#define SGX_HD inline
#define SGX_HDI inline
#define SGX_D inline
#endif
#include <cmath>
#include <cstring>

#include "sgx_libm.cuh"

namespace sgx {

// ---------------------------------------------------------------------------------------
// The reference's std::exp / log / pow / log10 / cbrt.  Host code calls the host libm (the
// one the reference links); device code runs the same glibc algorithm bit for bit
// (sgx_libm.cuh), so both sides of the parity contract see identical transcendental values.
// ---------------------------------------------------------------------------------------
#ifdef __CUDA_ARCH__
SGX_HD double xexp(const double x) { return libm::exp(x); }
SGX_HD double xlog(const double x) { return libm::log(x); }
SGX_HD double xpow(const double x, const double y) { return libm::pow(x, y); }
SGX_HD double xlog10(const double x) { return libm::log10(x); }
SGX_HD double xcbrt(const double x) { return libm::cbrt(x); }
#else
SGX_HD double xexp(const double x) { return std::exp(x); }
SGX_HD double xlog(const double x) { return std::log(x); }
SGX_HD double xpow(const double x, const double y) { return std::pow(x, y); }
SGX_HD double xlog10(const double x) { return std::log10(x); }
SGX_HD double xcbrt(const double x) { return std::cbrt(x); }
#endif

// ---------------------------------------------------------------------------------------
// robust utils
// ---------------------------------------------------------------------------------------
constexpr double kSmall = 10.0 * DBL_MIN;    // robust::SMALL()
constexpr double kEps = 10.0 * DBL_EPSILON;  // robust::EPS()

SGX_HD int sgn(const double v) { return (0.0 <= v) - (v < 0.0); }
// a / (b + sgn(b) * SMALL)
SGX_HD double ratio(const double a, const double b) { return a / (b + sgn(b) * kSmall); }
// std::max / std::min argument-order semantics (matter only for NaN / signed zero)
SGX_HD double stdmax(const double a, const double b) { return (a < b) ? b : a; }
SGX_HD double stdmin(const double a, const double b) { return (b < a) ? b : a; }

// math_utils::pow<N>: same association as the recursive template
SGX_HD double pow2i(const double x) { return x * x; }
SGX_HD double pow3i(const double x) { return x * (x * x); }
SGX_HD double pow4i(const double x) { return (x * x) * (x * x); }
SGX_HD double pow5i(const double x) { return x * ((x * x) * (x * x)); }

// static_cast<int>(double) as the x86-64 host does it (cvttsd2si: "integer indefinite" =
// INT_MIN when out of range or NaN); CUDA's cvt.rzi saturates instead.
SGX_HD int cvt_int_x86(const double v) {
#ifdef __CUDA_ARCH__
  return (v >= 2147483648.0 || v <= -2147483649.0 || v != v) ? INT_MIN : __double2int_rz(v);
#else
  return static_cast<int>(v);
#endif
}
SGX_HD long long cvt_ll_x86(const double v) {
#ifdef __CUDA_ARCH__
  return (v >= 9223372036854775808.0 || v < -9223372036854775808.0 || v != v)
             ? LLONG_MIN
             : __double2ll_rz(v);
#else
  return static_cast<long long>(v);
#endif
}
SGX_HD long long as_int(const double x) {
#ifdef __CUDA_ARCH__
  return __double_as_longlong(x);
#else
  long long i;
  std::memcpy(&i, &x, sizeof(i));
  return i;
#endif
}
SGX_HD double as_double(const long long i) {
#ifdef __CUDA_ARCH__
  return __longlong_as_double(i);
#else
  double x;
  std::memcpy(&x, &i, sizeof(x));
  return x;
#endif
}
SGX_HD double sqrt_rn(const double x) {
#ifdef __CUDA_ARCH__
  return __dsqrt_rn(x);
#else
  return std::sqrt(x);
#endif
}

// ---------------------------------------------------------------------------------------
// FastMath NQT order-2, integer aliased (the reference default build)
// ---------------------------------------------------------------------------------------
namespace fastmath {
constexpr long long one = 1;
constexpr long long one_as_int = (one << 62) - (one << 52);
constexpr double scale_down = 2.22044604925031e-16; // literal as in logs.hpp:112
constexpr double scale_up = (double)(one << 52);
constexpr long long mantissa_mask = (one << 52) - one;
constexpr long long low_mask = (one << 26) - 1;

// floor(n / 3) for 0 <= n < 2^53 with 32-bit pieces.  With n = a 2^32 + b and
// 2^32 = 3 * 0x55555555 + 1:  n = 3 (a * 0x55555555) + (a + b), so
// floor(n/3) = a * 0x55555555 + floor((a + b) / 3); when a + b carries out of 32 bits,
// 2^32 + d = 3 * 0x55555555 + (d + 1).  floor(s/3) for a 32-bit s is umulhi(s, 0xAAAAAAAB) >> 1.
// (The generic signed 64-bit division the compiler emits for `/ 3` costs ~16 instructions
// with its sign fix-ups; the numerator in lg() is provably in [0, 2^52).)
SGX_HD unsigned long long div3_u53(const unsigned long long n) {
  const unsigned int a = (unsigned int)(n >> 32), b = (unsigned int)n;
  const unsigned int s = a + b;
  const bool carry = s < a;
  const unsigned int t = carry ? s + 1u : s; // cannot wrap: a < 2^21 so s < 2^21 when carrying
#ifdef __CUDA_ARCH__
  unsigned int q = __umulhi(t, 0xAAAAAAABu) >> 1;
#else
  unsigned int q = (unsigned int)(((unsigned long long)t * 0xAAAAAAABull) >> 33);
#endif
  return (unsigned long long)a * 0x55555555ull + (unsigned long long)q +
         (carry ? 0x55555555ull : 0ull);
}

// lg_o2_aliased (logs.hpp:188-199), restated on explicit 32-bit pieces.  With
// f = x_as_int & mantissa_mask in [0, 2^52): fh = f >> 26 and fl = f & low_mask are < 2^26, so
// fh*fh and fh*fl are single 32x32->64 products, and f - frac_squared is in [0, 2^52)
// (frac_squared ~ f^2 / 2^52 <= f), so C++'s truncating `/ 3` is floor division.
// Bit-identical to the literal formula: tests/test_cpu_oracle.py::test_lg_restatement_bits.
SGX_HD double lg(const double x) {
#ifdef __CUDA_ARCH__
  // Device form: the same integers on explicit 32-bit registers.  nvcc does not see that
  // (f >> 26) has a zero upper word and emulates full 64x64 products (3 IMAD each); here every
  // product is one IMAD.WIDE.U32 and every shift one funnel shift.  f = {fhi20 : flo32}.
  const long long xi = as_int(x) - one_as_int;
  const unsigned int lo = (unsigned int)xi;
  const unsigned int fhi20 = (unsigned int)((unsigned long long)xi >> 32) & 0xfffffu;
  const unsigned int fh = __funnelshift_l(lo, fhi20, 6); // f >> 26, < 2^26
  const unsigned int fl = lo & (unsigned int)low_mask;
  const unsigned long long cross = (unsigned long long)fh * fl; // < 2^52
  const unsigned int cross_sh =
      __funnelshift_r((unsigned int)cross, (unsigned int)(cross >> 32), 25); // < 2^27
  const unsigned long long sq = (unsigned long long)fh * fh + cross_sh;
  const unsigned long long f = ((unsigned long long)fhi20 << 32) | lo;
  const long long third_d = (long long)div3_u53(f - sq);
  return __ll2double_rn(xi + third_d) * scale_down;
#else
  const long long x_as_int = as_int(x) - one_as_int;
  const unsigned long long frac_as_int = (unsigned long long)(x_as_int & mantissa_mask);
  const unsigned int frac_high = (unsigned int)(frac_as_int >> 26);
  const unsigned int frac_low = (unsigned int)frac_as_int & (unsigned int)low_mask;
  const unsigned long long frac_squared =
      (unsigned long long)frac_high * frac_high +
      (((unsigned long long)frac_high * frac_low) >> 25);
  const long long third = (long long)div3_u53(frac_as_int - frac_squared);
  return static_cast<double>(x_as_int + third) * scale_down;
#endif
}
// the literal formula of logs.hpp:188-199, kept for the bit-equality test
SGX_HD double lg_literal(const double x) {
  const long long x_as_int = as_int(x) - one_as_int;
  const long long frac_as_int = x_as_int & mantissa_mask;
  const long long frac_high = frac_as_int >> 26;
  const long long frac_low = frac_as_int & low_mask;
  const long long frac_squared = frac_high * frac_high + ((frac_high * frac_low) >> 25);
  return static_cast<double>(x_as_int + ((frac_as_int - frac_squared) / 3)) * scale_down;
}
SGX_HD double pow2(const double x) {
  constexpr long long a = 9007199254740992ll;  // 2 * 2^52
  constexpr double b = 67108864.0;             // 2^26
  constexpr long long c = 18014398509481984ll; // 4 * 2^52
  const long long x_as_int = cvt_ll_x86(x * scale_up);
  const long long frac_as_int = x_as_int & mantissa_mask;
  const long long frac_sqrt =
      cvt_ll_x86(b * sqrt_rn(static_cast<double>(c - 3 * frac_as_int)));
  return as_double(x_as_int + a - frac_sqrt - frac_as_int + one_as_int);
}
SGX_HD double log10(const double x) {
  constexpr double LOG2OLOG10 = 0.301029995663981195;
  return LOG2OLOG10 * lg(x);
}
SGX_HD double pow10(const double x) {
  constexpr double LOG10OLOG2 = 3.321928094887362626;
  return pow2(LOG10OLOG2 * x);
}
} // namespace fastmath

// spiner_common::to_log / from_log
SGX_HD double to_log(const double x, const double offset) {
  return fastmath::log10(fabs(stdmax(x, -offset) + offset) + kEps);
}
SGX_HD double from_log(const double lx, const double offset) {
  return fastmath::pow10(lx) - offset;
}

// ---------------------------------------------------------------------------------------
// Spiner grids + interpolation
// ---------------------------------------------------------------------------------------
struct Grid1D {
  int ngrids;
  int npoints[3];
  int ptot[3]; // pointTotals_
  double min[3], max[3], dx[3], idx[3];

  SGX_HD int n_total() const { return ptot[ngrids - 1] + npoints[ngrids - 1]; }
  SGX_HD double gmin() const { return min[0]; }
  SGX_HD double gmax() const { return max[ngrids - 1]; }
  // PiecewiseGrid1D::findGrid + findGridFromPosition
  SGX_HD int find_grid(const double x) const {
    if (ngrids == 1) return 0;
    int l = 0, r = ngrids - 1;
    for (int iter = 0; iter < ngrids; iter++) {
      const int m = (l + r) / 2;
      const int d = (x < min[m]) ? -1 : ((x > max[m]) ? 1 : 0);
      if (d < 0) {
        if (m == 0) return 0;
        r = m - 1;
      } else if (d > 0) {
        if (m >= ngrids - 1) return ngrids - 1;
        l = m + 1;
      } else {
        return m;
      }
    }
    return 0; // unreachable for well-formed grids (reference aborts)
  }
  // PiecewiseGrid1D::x(i)
  SGX_HD double x(const int i) const {
    int ig = 0;
    for (int g = ngrids - 1; g > 0; --g) {
      if (i >= ptot[g]) {
        ig = g;
        break;
      }
    }
    return (i - ptot[ig]) * dx[ig] + min[ig];
  }
  // bound(int(v)) of RegularGrid1D::index (regular_grid_1d.hpp:66-74) for a patch of n points.
  // On x86-64 the conversion yields INT_MIN when v is NaN or out of int range, which the
  // clamp then maps to 0.  CUDA's cvt.rzi saturates instead: NaN -> 0 and v <= -2^31 ->
  // INT_MIN both clamp to 0 as well, so only v >= 2^31 (saturates to INT_MAX, would clamp to
  // n-2) needs an explicit test.
  SGX_HD static int index_clamped(const double v, const int n) {
#ifdef __CUDA_ARCH__
    int i = __double2int_rz(v);
    if (v >= 2147483648.0) i = 0;
#else
    int i = static_cast<int>(v);
#endif
    if (i < 0) i = 0;
    if (i >= n - 1) i = n - 2;
    return i;
  }
  SGX_HD void weights(const double x, int &ix, double &w0, double &w1) const {
    if (ngrids == 1) {
      // single patch (the common case): constant member indices, no patch search, so the
      // grid metadata stays in uniform registers instead of register-indexed constant loads
      const int i = index_clamped(idx[0] * (x - min[0]), npoints[0]);
      const double floor_x = (double)i * dx[0] + min[0];
      w1 = idx[0] * (x - floor_x);
      w0 = (1. - w1);
      ix = i; // pointTotals_[0] == 0
      return;
    }
    const int ig = find_grid(x);
    const int i = index_clamped(idx[ig] * (x - min[ig]), npoints[ig]);
    const double floor_x = (double)i * dx[ig] + min[ig];
    w1 = idx[ig] * (x - floor_x);
    w0 = (1. - w1);
    ix = i + ptot[ig];
  }
};

SGX_HD double ldg(const double *p) {
#ifdef __CUDA_ARCH__
  return __ldg(p);
#else
  return *p;
#endif
}

// DataBox::interpToReal(x), rank 1
SGX_HD double interp1d(const double *__restrict__ a, const Grid1D &g, const double x) {
  int ix;
  double w0, w1;
  g.weights(x, ix, w0, w1);
  return w0 * ldg(a + ix) + w1 * ldg(a + ix + 1);
}
// The four-corner blend of DataBox::interpToReal(x2, x1), databox.hpp:471-474
SGX_HD double blend2d(const double *__restrict__ a, const int n1, const int ix2,
                     const double w20, const double w21, const int ix1, const double w10,
                     const double w11) {
  const double *r0 = a + (size_t)ix2 * n1 + ix1;
  const double *r1 = r0 + n1;
  return (w20 * (w10 * ldg(r0) + w11 * ldg(r0 + 1)) + w21 * (w10 * ldg(r1) + w11 * ldg(r1 + 1)));
}
// DataBox::interpToReal(x2, x1), rank 2: g1 = fastest axis (range(0)), g2 = range(1)
SGX_HD double interp2d(const double *__restrict__ a, const Grid1D &g2, const Grid1D &g1,
                      const int n1, const double x2, const double x1) {
  int ix1, ix2;
  double w10, w11, w20, w21;
  g1.weights(x1, ix1, w10, w11);
  g2.weights(x2, ix2, w20, w21);
  return blend2d(a, n1, ix2, w20, w21, ix1, w10, w11);
}
// Position along one axis, computed once and shared by several tables
struct AxisPos {
  int ix;
  double w0, w1;
  SGX_HD AxisPos() {}
  SGX_HD AxisPos(const Grid1D &g, const double x) { g.weights(x, ix, w0, w1); }
};
SGX_HD double interp1d(const double *__restrict__ a, const AxisPos &p) {
  return p.w0 * ldg(a + p.ix) + p.w1 * ldg(a + p.ix + 1);
}
SGX_HD double interp2d(const double *__restrict__ a, const int n1, const AxisPos &p2,
                      const AxisPos &p1) {
  return blend2d(a, n1, p2.ix, p2.w0, p2.w1, p1.ix, p1.w0, p1.w1);
}

// ---------------------------------------------------------------------------------------
// Cell-major ("packed") 2-D tables for the device hot path.  At create time every 2-D table
// A[n2][n1] is repacked into one 32-byte record per table cell (j, i), j < n2-1, i < n1-1:
//   {A(j,i), A(j,i+1), A(j+1,i), A(j+1,i+1)}
// so the four corners of DataBox::interpToReal(x2, x1) are ONE aligned sector fetched by ONE
// 256-bit load (LDG.E.256) instead of four gathers over two cache lines.  ncu on the
// node-major layout showed the L1TEX wavefront path, not HBM, bounding the lookup kernels
// (profiles/r1a_lookup_details.txt).  Values and evaluation order are unchanged
// (databox.hpp:471-474), so results are bit-identical to the node-major form.
// The grid's weights() clamps ix to [0, n-2] within a patch, so (ix2, ix1) is always a cell.
// ---------------------------------------------------------------------------------------
struct Cell4 {
  double c00, c01, c10, c11;
};
SGX_HD Cell4 ld_cell(const double *__restrict__ cells, const int n1, const int ix2,
                     const int ix1) {
  const double *p = cells + ((size_t)ix2 * (size_t)(n1 - 1) + (size_t)ix1) * 4;
  Cell4 c;
#ifdef __CUDA_ARCH__
  asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
      : "=d"(c.c00), "=d"(c.c01), "=d"(c.c10), "=d"(c.c11)
      : "l"(p));
#else
  c.c00 = p[0];
  c.c01 = p[1];
  c.c10 = p[2];
  c.c11 = p[3];
#endif
  return c;
}
SGX_HD double blend_cell(const Cell4 &c, const double w20, const double w21, const double w10,
                         const double w11) {
  return (w20 * (w10 * c.c00 + w11 * c.c01) + w21 * (w10 * c.c10 + w11 * c.c11));
}
SGX_HD double blend2c(const double *__restrict__ cells, const int n1, const int ix2,
                      const double w20, const double w21, const int ix1, const double w10,
                      const double w11) {
  return blend_cell(ld_cell(cells, n1, ix2, ix1), w20, w21, w10, w11);
}
SGX_HD double interp2c(const double *__restrict__ cells, const int n1, const AxisPos &p2,
                       const AxisPos &p1) {
  return blend2c(cells, n1, p2.ix, p2.w0, p2.w1, p1.ix, p1.w0, p1.w1);
}
SGX_HD double interp2c(const double *__restrict__ cells, const Grid1D &g2, const Grid1D &g1,
                       const int n1, const double x2, const double x1) {
  return interp2c(cells, n1, AxisPos(g2, x2), AxisPos(g1, x1));
}
// host-side repack: node-major A[n2][n1] -> cells[(n2-1)*(n1-1)][4]
inline void pack_cells(const double *a, const int n2, const int n1, double *cells) {
  for (int j = 0; j < n2 - 1; ++j)
    for (int i = 0; i < n1 - 1; ++i) {
      double *c = cells + ((size_t)j * (n1 - 1) + i) * 4;
      const double *r0 = a + (size_t)j * n1 + i;
      c[0] = r0[0];
      c[1] = r0[1];
      c[2] = r0[n1];
      c[3] = r0[n1 + 1];
    }
}

// ---------------------------------------------------------------------------------------
// Root finding
// ---------------------------------------------------------------------------------------
namespace root {
enum Status { SUCCESS = 0, FAIL = 1 };

SGX_HD bool check_bracket(const double ya, const double yb) { return (ya * yb <= 0.0); }

// set_bracket, root_finding.hpp:118-151.  f already has ytarget subtracted.
template <typename F>
SGX_HDI bool set_bracket(const F &f, double &a, const double guess, double &b,
                                   double &ya, const double yg, double &yb) {
  constexpr int max_search_depth = 6;
  const double dx = b - a;
  for (int level = 0; level < max_search_depth; level++) {
    const int nlev = (1 << level);
    for (int i = 0; i < nlev; i++) {
      const double x = a + (i + 0.5) * dx;
      const double yx = f(x);
      if (check_bracket(yx, yg)) {
        if (x < guess) {
          a = x;
          ya = yx;
          b = guess;
          yb = yg;
        } else {
          a = guess;
          ya = yg;
          b = x;
          yb = yx;
        }
        return true;
      }
    }
  }
  return false;
}

// regula_falsi (Illinois variant), root_finding.hpp:154-240.  `niter` receives the loop count.
template <typename F>
SGX_HDI Status regula_falsi(const F &f, const double ytarget, const double guess,
                                      double a, double b, const double xtol,
                                      const double ytol, double &xroot, int *niter = nullptr) {
  constexpr int max_iter = 1000; // SECANT_NITER_MAX
  auto func = [&](const double x) { return f(x) - ytarget; };
  double ya = func(a);
  double yg = func(guess);
  double yb;

  if (check_bracket(ya, yg)) {
    b = guess;
    yb = yg;
  } else {
    yb = func(b);
    if (check_bracket(yg, yb)) {
      a = guess;
      ya = yg;
    } else {
      if (!set_bracket(func, a, guess, b, ya, yg, yb)) {
        if (niter) *niter = 0;
        return FAIL;
      }
    }
  }

  const double sign = (ya < 0 ? 1.0 : -1.0);
  ya *= sign;
  yb *= sign;

  int b1 = 0;
  int b2 = 0;
  int iteration_count = 0;
  while (b - a > 2.0 * xtol && (fabs(ya) > ytol || fabs(yb) > ytol) &&
         iteration_count < max_iter) {
    const double c = (a * yb - b * ya) / (yb - ya);
    if (c == a) {
      b = a;
      continue;
    } else if (c == b) {
      a = b;
      continue;
    }
    const double yc = sign * func(c);
    if (yc > 0.0) {
      b = c;
      yb = yc;
      b1++;
      ya *= (b1 > 1 ? 0.5 : 1.0);
      b2 = 0;
    } else if (yc < 0.0) {
      a = c;
      ya = yc;
      b2++;
      yb *= (b2 > 1 ? 0.5 : 1.0);
      b1 = 0;
    } else {
      a = c;
      b = c;
    }
    iteration_count++;
  }
  if (niter) *niter = iteration_count;
  xroot = 0.5 * (a + b);
  return (iteration_count == max_iter) ? FAIL : SUCCESS;
}

// newton_raphson, root_finding.hpp:246-310.  f(x, y, dy) returns value and derivative.
template <typename F>
SGX_HDI Status newton_raphson(const F &f, const double ytarget, const double guess,
                                        const double a, const double b, const double ytol,
                                        double &xroot, const bool fail_on_bound_root = true) {
  constexpr int max_iter = 100; // NEWTON_RAPHSON_NITER_MAX
  double _x = guess;
  double _xold = 0.0;
  Status status = SUCCESS;
  double yg, dfunc;
  int iter;
  for (iter = 0; iter < max_iter; iter++) {
    f(_x, yg, dfunc);
    if (fabs(yg - ytarget) < fabs(ytol * ytarget)) break;
    _xold = _x;
    _x = _x - (yg - ytarget) / dfunc;
    if ((_x <= a && _xold <= a) || (_x >= b && _xold >= b)) {
      if (fail_on_bound_root) status = FAIL;
      break;
    }
    _x = stdmax(stdmin(_x, b), a);
  }
  if (iter >= max_iter) status = FAIL;
  xroot = _x;
  return status;
}
} // namespace root

} // namespace sgx

#endif // SGX_MATH_CUH_
