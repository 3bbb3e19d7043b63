// The small dense linear algebra of the PTE solvers, usable from host and device code:
// error_utils::bad_value, mix_impl::sum_neumaier (closure/mixed_cell_models.hpp:287-305),
// mix_impl::solve_Ax_b_wscr (:176-278) with Eigen's colPivHouseholderQr().solve() restated
// operation for operation for n >= 3 (utils/eigen/Eigen/src/QR/ColPivHouseholderQR.h).
// Compiles as plain C++ (tests/cpp/linsolve_check.cpp compares it bit for bit with the
// reference on the CPU) and as device code (sgx_pte.cuh).
#ifndef SGX_LINSOLVE_CUH_
#define SGX_LINSOLVE_CUH_

#include "sgx_math.cuh"

namespace sgx {
namespace mix {

// error_utils::bad_value == !is_normal_or_zero (base/error_utils.hpp:31-76)
SGX_HD bool is_normal(const double v) {
  const double a = fabs(v);
  return (a >= DBL_MIN) && (a <= DBL_MAX); // excludes 0, subnormal, inf, NaN
}
SGX_HD bool bad_value(const double v) {
  const bool good = (v == 0.0) || (is_normal(1.0e10 * v) && is_normal(v));
  return !good;
}
SGX_HD bool isfinite_ref(const double a) { return (a == a) && ((a == 0) || (a != 2 * a)); }

// sum_neumaier (:287-305)
template <typename Op>
SGX_HD double sum_neumaier(const double *data, const int n, const int offset, const int iskip,
                          const Op &op) {
  double sum = 0;
  double c = 0;
#pragma unroll
  for (int i = 0; i < n; ++i) {
    if (iskip >= 0 && i == iskip) continue;
    const double x = op(data[i + offset]);
    const double t = sum + x;
    // (sum - t) + x  or  (x - t) + sum, as selects: the data-dependent branch ran each side
    // with a handful of lanes (ncu r2j: 4.5 % of the PTE kernel's instructions at 5-9 lanes)
    const bool sum_is_big = fabs(sum) >= fabs(x);
    const double big = sum_is_big ? sum : x, small = sum_is_big ? x : sum;
    c += (big - t) + small;
    sum = t;
  }
  return sum + c;
}
SGX_HD double sum_neumaier(const double *data, const int n, const int offset = 0,
                          const int iskip = -1) {
  return sum_neumaier(data, n, offset, iskip, [](const double x) { return x; });
}

// Sum of `count` (<= CNT, compile-time after unrolling) terms in the order Eigen's
// linear-vectorised redux uses with 2-wide packets (Core/Redux.h, SSE2, alignedStart = 0):
// two packet accumulators over groups of four, a third packet if count % 4 >= 2, the
// horizontal add, then the scalar tail.  count <= 3 degenerates to the plain order.
template <int CNT>
SGX_HD double eigen_redux_sum(const double *p, const int count) {
  double res = 0.0;
#pragma unroll
  for (int c = 1; c <= CNT; ++c) {
    if (c == count) {
      constexpr int dummy = 0;
      (void)dummy;
      const int aligned = (c / 2) * 2, aligned2 = (c / 4) * 4;
      if (aligned == 0) {
        res = p[0];
      } else {
        double r0x = p[0], r0y = p[1];
        if (aligned > 2) {
          double r1x = p[2], r1y = p[3];
#pragma unroll
          for (int idx = 4; idx + 3 < CNT + 0 && idx < aligned2; idx += 4) {
            r0x = r0x + p[idx];
            r0y = r0y + p[idx + 1];
            r1x = r1x + p[idx + 2];
            r1y = r1y + p[idx + 3];
          }
          r0x = r0x + r1x;
          r0y = r0y + r1y;
          if (aligned > aligned2) {
            r0x = r0x + p[aligned2];
            r0y = r0y + p[aligned2 + 1];
          }
        }
        res = r0x + r0y;
#pragma unroll
        for (int idx = 0; idx < CNT; ++idx)
          if (idx >= aligned && idx < c) res = res + p[idx];
      }
    }
  }
  return res;
}

// Column-pivoted Householder QR solve for n >= 3.  The reference delegates to Eigen
// (`A.colPivHouseholderQr().solve(B)` on a dynamic row-major Map, :266-273), so the parity
// contract for the Newton step is Eigen's arithmetic.  This is that algorithm operation for
// operation (utils/eigen/Eigen/src/QR/ColPivHouseholderQR.h computeInPlace + _solve_impl,
// Householder/Householder.h makeHouseholder + applyHouseholderOnTheLeft, the scalar paths of
// the gemv / outer-product / triangular-solve kernels they reach for n <= 4 without FMA):
// column norms with LAPACK's down-dating (lawn176), first-maximum pivot, essential = tail /
// (c0 - beta), tau = (beta - c0) / beta, tmp = essential^T bottom + row0, row0 -= tau tmp,
// bottom -= (tau essential) tmp, rank detection by threshold_helper, back substitution that
// skips zero right-hand sides.  tests: test_solve_Ax_b_bit_identical_to_eigen.
// a is row-major N x N (destroyed), b -> x.  N is a compile-time constant and every loop is
// unrolled with static indices: the pivot column is brought to position k by predicated
// swaps, never by a[i*n + piv], so the matrix stays in registers.
template <int N>
SGX_HD void qr_solve_static(double *a, double *b) {
  int perm[N];
  double nu[N], nd[N]; // m_colNormsUpdated, m_colNormsDirect
#pragma unroll
  for (int j = 0; j < N; ++j) {
    perm[j] = j;
    double s = a[j] * a[j];
#pragma unroll
    for (int i = 1; i < N; ++i)
      s = s + a[i * N + j] * a[i * N + j];
    nd[j] = sqrt(s);
    nu[j] = nd[j];
  }
  double maxnorm = nu[0];
#pragma unroll
  for (int j = 1; j < N; ++j)
    if (nu[j] > maxnorm) maxnorm = nu[j];
  const double me = maxnorm * DBL_EPSILON;
  const double threshold_helper = (me * me) / static_cast<double>(N);
  constexpr double norm_downdate_threshold = 1.4901161193847656e-08; // sqrt(epsilon) = 2^-26
  int nonzero_pivots = N;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    int piv = k;
    double best = nu[k];
#pragma unroll
    for (int j = k + 1; j < N; ++j) {
      if (nu[j] > best) {
        best = nu[j];
        piv = j;
      }
    }
    if (nonzero_pivots == N && best * best < threshold_helper * static_cast<double>(N - k))
      nonzero_pivots = k;
#pragma unroll
    for (int j = k + 1; j < N; ++j) {
      if (piv == j) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
          const double t = a[i * N + k];
          a[i * N + k] = a[i * N + j];
          a[i * N + j] = t;
        }
        const int tp = perm[k];
        perm[k] = perm[j];
        perm[j] = tp;
        double t = nu[k];
        nu[k] = nu[j];
        nu[j] = t;
        t = nd[k];
        nd[k] = nd[j];
        nd[j] = t;
      }
    }
    // makeHouseholderInPlace on column k, rows k..N-1
    const double c0 = a[k * N + k];
    double tail = 0;
    if constexpr (true) {
      bool first = true;
#pragma unroll
      for (int i = k + 1; i < N; ++i) {
        const double sq = a[i * N + k] * a[i * N + k];
        tail = first ? sq : tail + sq;
        first = false;
      }
    }
    double tau, beta;
    if (tail <= DBL_MIN) {
      tau = 0.0;
      beta = c0;
#pragma unroll
      for (int i = k + 1; i < N; ++i)
        a[i * N + k] = 0.0;
    } else {
      beta = sqrt(c0 * c0 + tail);
      if (c0 >= 0.0) beta = -beta;
      const double den = c0 - beta;
#pragma unroll
      for (int i = k + 1; i < N; ++i)
        a[i * N + k] = a[i * N + k] / den;
      tau = (beta - c0) / beta;
    }
    a[k * N + k] = beta;
    if (k < N - 1 && tau != 0.0) {
      // trailing columns: tmp = essential^T bottom; tmp += row0; row0 -= tau tmp;
      // bottom -= (tau essential) tmp
#pragma unroll
      for (int j = k + 1; j < N; ++j) {
        double tmp = 0.0;
#pragma unroll
        for (int i = k + 1; i < N; ++i)
          tmp = a[i * N + k] * a[i * N + j] + tmp;
        tmp = tmp + a[k * N + j];
        a[k * N + j] = a[k * N + j] - tau * tmp;
#pragma unroll
        for (int i = k + 1; i < N; ++i)
          a[i * N + j] = a[i * N + j] - (tau * a[i * N + k]) * tmp;
      }
      // the same reflector on the right-hand side (Q^T b applies H_0 first; only the first
      // nonzero_pivots reflectors are used)
      if (k < nonzero_pivots) {
        double tmp = 0.0;
#pragma unroll
        for (int i = k + 1; i < N; ++i)
          tmp = tmp + a[i * N + k] * b[i];
        tmp = tmp + b[k];
        b[k] = b[k] - tau * tmp;
#pragma unroll
        for (int i = k + 1; i < N; ++i)
          b[i] = b[i] - (tau * a[i * N + k]) * tmp;
      }
    }
    // norm down-date of the remaining columns
#pragma unroll
    for (int j = k + 1; j < N; ++j) {
      if (nu[j] != 0.0) {
        double temp = fabs(a[k * N + j]) / nu[j];
        temp = (1.0 + temp) * (1.0 - temp);
        temp = temp < 0.0 ? 0.0 : temp;
        const double r = nu[j] / nd[j];
        const double temp2 = temp * (r * r);
        if (temp2 <= norm_downdate_threshold) {
          double s = 0.0;
          bool first = true;
#pragma unroll
          for (int i = k + 1; i < N; ++i) {
            const double sq = a[i * N + j] * a[i * N + j];
            s = first ? sq : s + sq;
            first = false;
          }
          nd[j] = sqrt(s);
          nu[j] = nd[j];
        } else {
          nu[j] = nu[j] * sqrt(temp);
        }
      }
    }
  }
  // R(0:np,0:np) y = c(0:np): rows from the bottom, zero right-hand sides are not divided
  double y[N];
#pragma unroll
  for (int i = N - 1; i >= 0; --i) {
    double v = 0.0;
    if (i < nonzero_pivots) {
      v = b[i];
      if constexpr (N > 1) {
        // row i of the np x np triangle: np-1-i products summed in Eigen's redux order
        double pr[N - 1];
#pragma unroll
        for (int j = i + 1; j < N; ++j)
          pr[j - i - 1] = a[i * N + j] * y[j];
        const int cnt = nonzero_pivots - 1 - i;
        if (cnt > 0) v = v - eigen_redux_sum<N - 1>(pr, cnt);
      }
      if (v != 0.0) v = v / a[i * N + i];
    }
    y[i] = v;
  }
  // x[perm[j]] = y[j]
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double x = 0.0;
#pragma unroll
    for (int j = 0; j < N; ++j)
      if (perm[j] == i) x = y[j];
    b[i] = x;
  }
}

// solve_Ax_b_wscr (:176-278), N a compile-time constant
template <int N>
SGX_HD bool solve_Ax_b_static(double *a, double *b) {
  bool nonzero_rows = true;
#pragma unroll
  for (int row = 0; row < N; ++row) {
    if (nonzero_rows) { // the reference returns at the first all-zero row
      double maxabs = 0;
#pragma unroll
      for (int column = 0; column < N; ++column) {
        maxabs = stdmax(fabs(a[row * N + column]), maxabs);
      }
      if (maxabs == 0) {
        nonzero_rows = false;
      } else {
#pragma unroll
        for (int column = 0; column < N; ++column) {
          a[row * N + column] /= maxabs;
        }
        b[row] /= maxabs;
      }
    }
  }
  if (!nonzero_rows) return false;
  if constexpr (N == 1) {
    b[0] = ratio(b[0], a[0]);
  } else if constexpr (N == 2) {
    const double det = a[0] * a[3] - a[1] * a[2];
    const double x0 = ratio(a[3] * b[0] - a[1] * b[1], det);
    const double x1 = ratio(a[0] * b[1] - a[2] * b[0], det);
    b[0] = x0;
    b[1] = x1;
  } else {
    qr_solve_static<N>(a, b);
  }
  bool retval = true;
#pragma unroll
  for (int i = 0; i < N; ++i)
    if (!isfinite_ref(b[i])) retval = false;
  return retval;
}

// runtime-n front end (building-block entry sgx_solve_Ax_b)
template <int NMAX>
SGX_HDI bool solve_Ax_b(const int n, double *a, double *b) {
  if (n == 1) return solve_Ax_b_static<1>(a, b);
  if (n == 2) return solve_Ax_b_static<2>(a, b);
  if constexpr (NMAX >= 4) {
    if (n == 4) return solve_Ax_b_static<4>(a, b);
  }
  if constexpr (NMAX >= 6) {
    if (n == 6) return solve_Ax_b_static<6>(a, b);
  }
  return solve_Ax_b_static<3>(a, b);
}

} // namespace mix
} // namespace sgx

#endif // SGX_LINSOLVE_CUH_
