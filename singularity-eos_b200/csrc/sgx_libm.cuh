// exp / log / pow / log10 / cbrt for device code, bit-identical to the host libm the
// reference links (glibc 2.39, x86-64, the FMA ifunc variants every AVX2 host selects).
//
// Why this file exists: the reference calls std::pow/exp/log/log10/cbrt in the Davis EOS
// (eos/eos_davis.hpp:353-394,415,439), in Helmholtz (eos/eos_helmholtz.hpp:750-752,769,
// 1404-1477, base/math_utils.hpp:56-59) and nowhere else on the hot path.  None of those is
// correctly rounded, so CUDA's own libdevice answers differ from glibc's in the last bit, and
// the PTE Newton iteration (finite-difference Jacobian, tolerances below round-off in
// test/test_pte.cpp:164-175) turns one ulp into a different iteration count.  Running the SAME
// published algorithm with the SAME tables and the SAME fused/unfused operations on both sides
// removes the only arithmetic that was outside the parity contract (hard part 2).
//
// The algorithms are Szabolcs Nagy's table-driven double exp/log/pow (ARM optimized-routines,
// in glibc since 2.28: sysdeps/ieee754/dbl-64/e_exp.c, e_log.c, e_pow.c), fdlibm's log10
// (e_log10.c, calling the new log) and glibc's s_cbrt.c.  glibc builds the first three twice;
// on a host with FMA+AVX2 the ifunc resolver picks the copy compiled with -mfma, in which gcc
// contracted some a*b+c into fused operations.  Which ones is a property of that binary, so
// the fused/unfused choice below was read operation by operation from the disassembly of
// __exp_fma / __log_fma / __pow_fma in this image's libm.so.6 (build-id in
// sgx_libm_tables.inc); every `fma()` here is a vfmadd there and every plain `a*b + c` is a
// vmulsd + vaddsd.  tests/test_libm_parity.py checks bit equality against the live libm on the
// CPU (this header compiles as plain C++) and on the GPU box (device vs host, >= 2^22 inputs
// per function).
//
// The library is compiled with -fmad=false, so nothing here is contracted by nvcc.
#ifndef SGX_LIBM_CUH_
#define SGX_LIBM_CUH_

#include <cstdint>
#include <cstring>

#ifdef __CUDACC__
#define SGX_LM __device__ __forceinline__
#define SGX_LM_NOINLINE static __device__ __noinline__
#define SGX_LIBM_TABLE static __device__ const
#else
#define SGX_LM inline
#define SGX_LM_NOINLINE inline
#define SGX_LIBM_TABLE static const
#endif

namespace sgx {
namespace libm {

#include "sgx_libm_tables.inc"

SGX_LM uint64_t asu(const double x) {
#ifdef __CUDA_ARCH__
  return static_cast<uint64_t>(__double_as_longlong(x));
#else
  uint64_t u;
  std::memcpy(&u, &x, sizeof(u));
  return u;
#endif
}
SGX_LM double asd(const uint64_t u) {
#ifdef __CUDA_ARCH__
  return __longlong_as_double(static_cast<long long>(u));
#else
  double x;
  std::memcpy(&x, &u, sizeof(x));
  return x;
#endif
}
SGX_LM double fma(const double a, const double b, const double c) {
#ifdef __CUDA_ARCH__
  return __fma_rn(a, b, c);
#else
  return __builtin_fma(a, b, c);
#endif
}
// two 8-byte table words in one 16-byte load
SGX_LM void ld2(const unsigned long long *p, uint64_t &a, uint64_t &b) {
#ifdef __CUDA_ARCH__
  const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2 *>(p));
  a = v.x;
  b = v.y;
#else
  a = p[0];
  b = p[1];
#endif
}
SGX_LM uint32_t top12(const double x) { return static_cast<uint32_t>(asu(x) >> 52); }

constexpr double kInf = __builtin_huge_val();

// ---------------------------------------------------------------------------------------
// exp: e_exp.c (__exp_fma).  x = k ln2/128 + r; exp(x) = 2^(k/128) (1 + tail + expm1(r)).
// ---------------------------------------------------------------------------------------
// tmp = tail + r + r2 (C2 + r C3) + r2 r2 (C4 + r C5) as the FMA build evaluates it.
SGX_LM double exp_poly(const double r, const double tail) {
  const double p23 = fma(r, kExpC3, kExpC2);
  const double tr = tail + r;
  const double r2 = r * r;
  const double p45 = fma(r, kExpC5, kExpC4);
  const double lo = fma(p23, r2, tr);
  const double r4 = r2 * r2;
  return fma(r4, p45, lo);
}

// |x| >= 512 (or pow's |y log x| >= 512): the scale 2^(k/128) is not representable.
// `signed_scale`: pow's variant (the result may be negative).
template <bool signed_scale>
SGX_LM_NOINLINE double exp_specialcase(const double tmp, uint64_t sbits, const uint64_t ki) {
  if ((ki & 0x80000000u) == 0) {
    // k > 0: the exponent of scale might have overflowed by <= 460
    sbits -= 1009ull << 52;
    const double scale = asd(sbits);
    return 0x1p1009 * fma(scale, tmp, scale);
  }
  // k < 0: care in the subnormal range
  sbits += 1022ull << 52;
  const double scale = asd(sbits);
  const double st = scale * tmp;  // not fused in this branch
  double y = scale + st;
  const double ay = signed_scale ? (y < 0.0 ? -y : y) : y;
  if (ay < 1.0) {
    double one = 1.0;
    if (signed_scale && y < 0.0) one = -1.0;
    double lo = (scale - y) + st;
    const double hi = one + y;
    lo = ((one - hi) + y) + lo;
    y = (hi + lo) - one;
    if (y == 0.0) y = signed_scale ? asd(sbits & 0x8000000000000000ull) : 0.0;
  }
  return 0x1p-1022 * y;
}

SGX_LM double exp(const double x) {
  uint32_t abstop = top12(x) & 0x7ff;
  if (abstop - 0x3c9u >= 0x3fu) {  // |x| < 2^-54 or |x| >= 512 or nan
    if (abstop - 0x3c9u >= 0x80000000u) return 1.0 + x;  // tiny
    if (abstop >= 0x409u) {                              // |x| >= 1024
      if (asu(x) == asu(-kInf)) return 0.0;
      if (abstop >= 0x7ffu) return 1.0 + x;
      return (asu(x) >> 63) ? 0.0 : kInf;  // __math_uflow / __math_oflow
    }
    abstop = 0;  // large x: special-cased below
  }
  double kd = fma(x, kExpInvLn2N, kExpShift);
  const uint64_t ki = asu(kd);
  kd -= kExpShift;
  double r = fma(kd, kExpNegLn2hiN, x);
  r = fma(kd, kExpNegLn2loN, r);
  uint64_t tbits, sbits;
  ld2(kExpTab + 2 * (ki & 127), tbits, sbits);
  sbits += ki << 45;
  const double tmp = exp_poly(r, asd(tbits));
  if (abstop == 0) return exp_specialcase<false>(tmp, sbits, ki);
  const double scale = asd(sbits);
  return fma(scale, tmp, scale);
}

// ---------------------------------------------------------------------------------------
// log: e_log.c (__log_fma)
// ---------------------------------------------------------------------------------------
SGX_LM double log_near1(const double x) {
  const double r = x - 1.0;
  const double p12 = fma(r, kLogB2, kLogB1);
  const double p45 = fma(r, kLogB5, kLogB4);
  const double r2 = r * r;
  const double p78 = fma(r, kLogB8, kLogB7);
  const double p123 = fma(r2, kLogB3, p12);
  const double p456 = fma(r2, kLogB6, p45);
  const double r3 = r * r2;
  double q = fma(r2, kLogB9, p78);
  q = fma(r3, kLogB10, q);
  q = fma(q, r3, p456);
  q = fma(q, r3, p123);  // B1 + r B2 + r2 B3 + r3 (B4 + ... )
  // rhi = r + w - w with w = r 2^27, both roundings fused in the FMA build
  const double rw = fma(r, 0x1p27, r);
  const double rhi = fma(-0x1p27, r, rw);
  const double rhi2 = rhi * rhi;
  const double rlo = r - rhi;
  const double hi = fma(rhi2, kLogB0, r);
  const double rmh = r - hi;
  const double rsum = r + rhi;
  double lo = fma(rhi2, kLogB0, rmh);
  const double b0rlo = kLogB0 * rlo;
  lo = fma(b0rlo, rsum, lo);
  const double y = fma(q, r3, lo);
  return hi + y;
}

SGX_LM double log(const double x) {
  uint64_t ix = asu(x);
  const uint32_t top = static_cast<uint32_t>(ix >> 48);
  // 1 - 2^-4 <= x < 1 + 0x1.09p-4
  if (ix - 0x3fee000000000000ull < 0x3090000000000ull) {
    if (ix == 0x3ff0000000000000ull) return 0.0;
    return log_near1(x);
  }
  if (top - 0x0010u >= 0x7ff0u - 0x0010u) {
    // x < 2^-1022 or inf or nan
    if (ix * 2 == 0) return -kInf;
    if (ix == asu(kInf)) return x;
    if ((top & 0x8000u) || (top & 0x7ff0u) == 0x7ff0u) return (x - x) / (x - x);
    ix = asu(x * 0x1p52);  // subnormal: normalize
    ix -= 52ull << 52;
  }
  const uint64_t tmp = ix - 0x3fe6000000000000ull;
  const int i = static_cast<int>((tmp >> 45) & 127);
  const int k = static_cast<int>(static_cast<int64_t>(tmp) >> 52);
  const uint64_t iz = ix - (tmp & (0xfffull << 52));
  uint64_t uinvc, ulogc;
  ld2(kLogTab + 2 * i, uinvc, ulogc);
  const double z = asd(iz);
  const double kd = static_cast<double>(k);
  const double w = fma(kd, kLogLn2hi, asd(ulogc));
  const double r = fma(z, asd(uinvc), -1.0);
  const double p12 = fma(r, kLogA2, kLogA1);
  const double hi = r + w;
  const double r2 = r * r;
  double lo = (w - hi) + r;
  lo = fma(kd, kLogLn2lo, lo);
  const double r3 = r * r2;
  const double p34 = fma(r, kLogA4, kLogA3);
  lo = fma(r2, kLogA0, lo);
  const double p = fma(p34, r2, p12);
  const double y = fma(r3, p, lo);
  return y + hi;
}

// ---------------------------------------------------------------------------------------
// pow: e_pow.c (__pow_fma)
// ---------------------------------------------------------------------------------------
// hi + *tail = log(x) in double-double; ix = bits of a positive normal(ised) x
SGX_LM double pow_log_inline(const uint64_t ix, double *tail) {
  const uint64_t tmp = ix - 0x3fe6955500000000ull;
  const int i = static_cast<int>((tmp >> 45) & 127);
  const int k = static_cast<int>(static_cast<int64_t>(tmp) >> 52);
  const uint64_t iz = ix - (tmp & (0xfffull << 52));
  const double z = asd(iz);
  const double kd = static_cast<double>(k);
#ifdef __CUDA_ARCH__
  const ulonglong4 rec = *reinterpret_cast<const ulonglong4 *>(kPowLogTab + 4 * i);
  const double invc = asd(rec.x), logc = asd(rec.z), logctail = asd(rec.w);
#else
  const double invc = asd(kPowLogTab[4 * i]), logc = asd(kPowLogTab[4 * i + 2]),
               logctail = asd(kPowLogTab[4 * i + 3]);
#endif
  const double t1 = fma(kd, kPowLogLn2hi, logc);
  const double lo1 = fma(kd, kPowLogLn2lo, logctail);
  const double r = fma(z, invc, -1.0);
  const double ar = r * kPowLogA0;
  const double p12 = fma(r, kPowLogA2, kPowLogA1);
  const double p34 = fma(r, kPowLogA4, kPowLogA3);
  const double t2 = r + t1;
  const double lo2 = (t1 - t2) + r;
  const double ar2 = r * ar;
  const double ar3 = r * ar2;
  const double lo3 = fma(ar, r, -ar2);
  const double hi = t2 + ar2;
  const double p56 = fma(r, kPowLogA6, kPowLogA5);
  const double lo4 = (t2 - hi) + ar2;
  const double q = fma(p56, ar2, p34);
  const double p = fma(ar2, q, p12);
  double lo = lo1 + lo2;
  lo = lo + lo3;
  lo = lo + lo4;
  lo = fma(ar3, p, lo);
  const double y = hi + lo;
  *tail = (hi - y) + lo;
  return y;
}

// exp(x + xtail) with the sign of the result in sign_bias (0 or 0x800 << 7)
SGX_LM double pow_exp_inline(const double x, const double xtail, const uint32_t sign_bias) {
  uint32_t abstop = top12(x) & 0x7ff;
  if (abstop - 0x3c9u >= 0x3fu) {
    if (abstop - 0x3c9u >= 0x80000000u) {
      const double one = 1.0 + x;
      return sign_bias ? -one : one;
    }
    if (abstop >= 0x409u) {
      // inf and nan are already handled: under- or overflow
      if (asu(x) >> 63) return sign_bias ? -0.0 : 0.0;
      return sign_bias ? -kInf : kInf;
    }
    abstop = 0;
  }
  double kd = fma(x, kExpInvLn2N, kExpShift);
  const uint64_t ki = asu(kd);
  kd -= kExpShift;
  double r = fma(kd, kExpNegLn2hiN, x);
  r = fma(kd, kExpNegLn2loN, r);
  r = xtail + r;
  uint64_t tbits, sbits;
  ld2(kExpTab + 2 * (ki & 127), tbits, sbits);
  sbits += (ki + sign_bias) << 45;
  const double tmp = exp_poly(r, asd(tbits));
  if (abstop == 0) return exp_specialcase<true>(tmp, sbits, ki);
  const double scale = asd(sbits);
  return fma(tmp, scale, scale);
}

// 0 if y is not an integer, 1 if odd, 2 if even
SGX_LM int pow_checkint(const uint64_t iy) {
  const int e = static_cast<int>((iy >> 52) & 0x7ff);
  if (e < 0x3ff) return 0;
  if (e > 0x3ff + 52) return 2;
  if (iy & ((1ull << (0x3ff + 52 - e)) - 1)) return 0;
  if (iy & (1ull << (0x3ff + 52 - e))) return 1;
  return 2;
}
SGX_LM bool pow_zeroinfnan(const uint64_t i) {
  return 2 * i - 1 >= 2 * 0x7ff0000000000000ull - 1;
}

// The rare operands: x <= 0, subnormal, inf, nan; |y| tiny, huge, inf, nan.  Returns true
// when *res is the answer, false when ix / sign_bias were normalised for the main path.
SGX_LM_NOINLINE bool pow_special(const double x, const double y, uint64_t &ix,
                                 uint32_t &sign_bias, double *res) {
  const uint64_t iy = asu(y);
  uint32_t topx = top12(x);
  const uint32_t topy = top12(y);
  constexpr uint64_t one = 0x3ff0000000000000ull, inf = 0x7ff0000000000000ull;
  if (pow_zeroinfnan(iy)) {
    // (signalling NaNs are not distinguished: no device code produces one)
    if (2 * iy == 0) { *res = 1.0; return true; }
    if (ix == one) { *res = 1.0; return true; }
    if (2 * ix > 2 * inf || 2 * iy > 2 * inf) { *res = x + y; return true; }
    if (2 * ix == 2 * one) { *res = 1.0; return true; }
    if ((2 * ix < 2 * one) == !(iy >> 63)) { *res = 0.0; return true; }
    *res = y * y;
    return true;
  }
  if (pow_zeroinfnan(ix)) {
    double x2 = x * x;
    if ((ix >> 63) && pow_checkint(iy) == 1) x2 = -x2;
    *res = (iy >> 63) ? 1.0 / x2 : x2;
    return true;
  }
  // x and y are non-zero finite
  if (ix >> 63) {
    const int yint = pow_checkint(iy);
    if (yint == 0) { *res = (x - x) / (x - x); return true; }
    if (yint == 1) sign_bias = 0x800u << 7;
    ix &= 0x7fffffffffffffffull;
    topx &= 0x7ff;
  }
  if ((topy & 0x7ff) - 0x3beu >= 0x80u) {
    // sign_bias == 0 here because y is not odd
    if (ix == one) { *res = 1.0; return true; }
    if ((topy & 0x7ff) < 0x3beu) { *res = ix > one ? 1.0 + y : 1.0 - y; return true; }
    *res = ((ix > one) == (topy < 0x800u)) ? kInf : 0.0;
    return true;
  }
  if (topx == 0) {
    // subnormal x: normalise so the exponent becomes negative
    ix = asu(x * 0x1p52);
    ix &= 0x7fffffffffffffffull;
    ix -= 52ull << 52;
  }
  return false;
}

SGX_LM double pow(const double x, const double y) {
  uint32_t sign_bias = 0;
  uint64_t ix = asu(x);
  const uint32_t topx = top12(x), topy = top12(y);
  if (topx - 0x001u >= 0x7ffu - 0x001u || (topy & 0x7ff) - 0x3beu >= 0x43eu - 0x3beu) {
    double res;
    if (pow_special(x, y, ix, sign_bias, &res)) return res;
  }
  double lo;
  const double hi = pow_log_inline(ix, &lo);
  const double ehi = y * hi;
  const double c = fma(hi, y, -ehi);
  const double elo = fma(y, lo, c);
  return pow_exp_inline(ehi, elo, sign_bias);
}

// ---------------------------------------------------------------------------------------
// log10: e_log10.c (fdlibm form around the log above; compiled without FMA)
// ---------------------------------------------------------------------------------------
SGX_LM double log10(double x) {
  int64_t hx = static_cast<int64_t>(asu(x));
  int32_t k = 0;
  if (hx < 0x0010000000000000ll) {  // x < 2^-1022
    if ((hx & 0x7fffffffffffffffll) == 0) return -kInf;  // log(+-0) = -inf
    if (hx < 0) return (x - x) / (x - x);                // log(-#) = NaN
    k -= 54;
    x *= 0x1p54;  // subnormal, scale up
    hx = static_cast<int64_t>(asu(x));
  }
  if (static_cast<uint64_t>(hx) >= 0x7ff0000000000000ull) return x + x;
  k += static_cast<int32_t>(hx >> 52) - 1023;
  const int64_t i = static_cast<int64_t>(static_cast<uint64_t>(static_cast<int64_t>(k)) >> 63);
  hx = (hx & 0x000fffffffffffffll) | ((0x3ff - i) << 52);
  const double yk = static_cast<double>(k + i);
  const double lx = log(asd(static_cast<uint64_t>(hx)));
  const double z = yk * kLog10_2lo + kLog10IvLn10 * lx;
  return z + yk * kLog10_2hi;
}

// ---------------------------------------------------------------------------------------
// cbrt: s_cbrt.c (frexp, degree-6 initial approximation, one Halley step; no FMA)
// ---------------------------------------------------------------------------------------
SGX_LM double cbrt(const double x) {
  const uint64_t ux = asu(x);
  uint64_t ax = ux & 0x7fffffffffffffffull;
  if (ax == 0 || ax >= 0x7ff0000000000000ull) return x + x;  // 0, inf, nan
  int xe = static_cast<int>(ax >> 52);
  if (xe == 0) {  // subnormal: frexp scales by 2^54 first
    ax = asu(asd(ax) * 0x1p54);
    xe = static_cast<int>(ax >> 52) - 54;
  }
  xe -= 1022;  // frexp exponent: |x| = xm 2^xe, xm in [0.5, 1)
  const double xm = asd((ax & 0x000fffffffffffffull) | 0x3fe0000000000000ull);
  double u = kCbrtP5 - kCbrtP6 * xm;
  u = u * xm - kCbrtP4;
  u = u * xm + kCbrtP3;
  u = u * xm - kCbrtP2;
  u = u * xm + kCbrtP1;
  u = u * xm + kCbrtP0;
  const double t2 = (u * u) * u;
  const int q = xe / 3;  // truncating, like C
  const int rem = xe - 3 * q;
#ifdef __CUDA_ARCH__
  const double fac = asd(__ldg(kCbrtFactor + 2 + rem));
#else
  const double fac = asd(kCbrtFactor[2 + rem]);
#endif
  double ym = (((xm + xm) + t2) * u) / ((t2 + t2) + xm) * fac;
  if (!(x > 0.0)) ym = -ym;
  // ldexp(ym, q): ym in [0.31, 1.6), |q| <= 358, so the product is exact
  return ym * asd(static_cast<uint64_t>(1023 + q) << 52);
}

}  // namespace libm
}  // namespace sgx

#endif  // SGX_LIBM_CUH_
