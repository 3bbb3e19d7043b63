// Analytic EOS models on the hot path, as trivially-copyable device structs passed to
// kernels by value (the role GetOnDevice() plays in the reference).
//
// Reference:
//   IdealGas        eos/eos_ideal.hpp:38-262
//   Gruneisen       eos/eos_gruneisen.hpp:36-570
//   DavisReactants  eos/eos_davis.hpp:40-209, 397-530
//   DavisProducts   eos/eos_davis.hpp:211-395, 531-612
//   EosBase bounds  eos/eos_base.hpp:533-557
// Expression association is kept literally (the parity contract is 1e-12 relative, and
// bit-exact for IdealGas); libm calls map to CUDA's pow/exp (<= 2 ulp).
#ifndef SGX_ANALYTIC_CUH_
#define SGX_ANALYTIC_CUH_

#include "sgx_math.cuh"

namespace sgx {

constexpr unsigned long TQ_DENSITY = 1ul << 0;
constexpr unsigned long TQ_SIE = 1ul << 1;
constexpr unsigned long TQ_PRESSURE = 1ul << 2;
constexpr unsigned long TQ_TEMPERATURE = 1ul << 3;
constexpr unsigned long TQ_CV = 1ul << 4;
constexpr unsigned long TQ_BMOD = 1ul << 5;

#define SGX_MYMAX(a, b) ((a) > (b) ? (a) : (b)) // eos_ideal.hpp:32

// The density-only terms of a Mie-Grueneisen-form model, e(rho,T) = Es + f(T, Ts) and
// P(rho,e) = Ps + rho*Gamma*(e - Es).  The PTE kernels evaluate several temperatures at one
// density (Jacobian dT column, the T-guess bisection); staging the evaluation as
// rho_part() + *_from_part() lets them reuse these terms.  The staged form performs exactly
// the operations of the one-shot functions in the same order, so results
// are bit-identical whichever way a value is obtained.
struct RhoPart {
  double Ts, Es, Ps, rG; // reference temperature / energy / pressure curves, rho*Gamma
};

// ---------------------------------------------------------------------------------------
struct IdealGas {
  double Cv, gm1;
  static constexpr unsigned long preferred_input =
      TQ_DENSITY | TQ_SIE | TQ_TEMPERATURE | TQ_PRESSURE;
  static constexpr int nlambda = 0;

  SGX_HD double T_from_rho_e(double rho, double sie, double * = nullptr) const {
    return SGX_MYMAX(0.0, sie / Cv);
  }
  SGX_HD double e_from_rho_T(double rho, double T, double * = nullptr) const {
    return SGX_MYMAX(0.0, Cv * T);
  }
  SGX_HD double P_from_rho_T(double rho, double T, double * = nullptr) const {
    return SGX_MYMAX(0.0, gm1 * rho * Cv * T);
  }
  SGX_HD double P_from_rho_e(double rho, double sie, double * = nullptr) const {
    return SGX_MYMAX(0.0, gm1 * rho * sie);
  }
  SGX_HD double cv_from_rho_T(double, double, double * = nullptr) const { return Cv; }
  SGX_HD double cv_from_rho_e(double, double, double * = nullptr) const { return Cv; }
  SGX_HD double bmod_from_rho_T(double rho, double T, double * = nullptr) const {
    return SGX_MYMAX(0.0, (gm1 + 1) * gm1 * rho * Cv * T);
  }
  SGX_HD double bmod_from_rho_e(double rho, double sie, double * = nullptr) const {
    return SGX_MYMAX(0.0, (gm1 + 1) * gm1 * rho * sie);
  }
  SGX_HD double gamma_from_rho_T(double, double, double * = nullptr) const { return gm1; }
  SGX_HD double gamma_from_rho_e(double, double, double * = nullptr) const { return gm1; }
  SGX_HD void rho_e_from_P_T(double press, double temp, double *, double &rho,
                             double &sie) const {
    sie = SGX_MYMAX(0.0, Cv * temp);
    rho = SGX_MYMAX(0.0, press / (gm1 * sie));
  }
  // eos_ideal.hpp:236-262.  Returns false where the reference hits UNDEFINED_ERROR.
  SGX_HD bool fill(double &rho, double &temp, double &sie, double &press, double &cv,
                   double &bmod, const unsigned long output, double *lambda) const {
    if ((output & TQ_DENSITY) && (output & TQ_SIE)) {
      if ((output & TQ_PRESSURE) || (output & TQ_TEMPERATURE)) return false;
      rho_e_from_P_T(press, temp, lambda, rho, sie);
    }
    if ((output & TQ_PRESSURE) && (output & TQ_SIE)) {
      if ((output & TQ_DENSITY) || (output & TQ_TEMPERATURE)) return false;
      sie = e_from_rho_T(rho, temp);
    }
    if ((output & TQ_TEMPERATURE) && (output & TQ_SIE)) {
      sie = press / (gm1 * rho);
    }
    if (output & TQ_PRESSURE) press = P_from_rho_e(rho, sie);
    if (output & TQ_TEMPERATURE) temp = T_from_rho_e(rho, sie);
    if (output & TQ_BMOD) bmod = bmod_from_rho_e(rho, sie);
    if (output & TQ_CV) cv = cv_from_rho_e(rho, sie);
    return true;
  }
  // staged form (no density-only terms)
  SGX_HD void rho_part(double, RhoPart &p) const { p.Ts = p.Es = p.Ps = p.rG = 0.0; }
  SGX_HD double e_from_part(const RhoPart &, double rho, double T) const {
    return e_from_rho_T(rho, T);
  }
  SGX_HD double P_from_part(const RhoPart &, double rho, double sie) const {
    return P_from_rho_e(rho, sie);
  }
  SGX_HD double cv_from_part(const RhoPart &, double rho, double sie) const {
    return cv_from_rho_e(rho, sie);
  }
  // bounds used by the PTE solvers (EosBase defaults)
  SGX_HD double MinimumDensity() const { return 0; }
  SGX_HD double MinimumTemperature() const { return 0; }
  SGX_HD double MaximumDensity() const { return 1e100; }
  SGX_HD double MinimumPressure() const { return 0; }
  SGX_HD double MaximumPressureAtTemperature(double) const { return 1e100; }
  SGX_HD double RhoPmin(double) const { return 0.0; }
};

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------------
struct Gruneisen {
  double C0, s1, s2, s3, G0, b, rho0, T0, P0, Cv, rho_max;
  static constexpr unsigned long preferred_input = TQ_DENSITY | TQ_SIE;
  static constexpr int nlambda = 0;

  SGX_D double Gamma(const double rho_in) const {
    const double rho = stdmin(rho_in, rho_max);
    return rho < rho0 ? G0 : G0 * rho0 / rho + b * (1 - rho0 / rho);
  }
  SGX_D double T_from_rho_e(double, double sie, double * = nullptr) const {
    return T0 + sie / Cv;
  }
  SGX_D double e_from_rho_T(double, double temp, double * = nullptr) const {
    return Cv * (temp - T0);
  }
  // eos_gruneisen.hpp:353-371, staged: Hugoniot reference curves, then the thermal term
  SGX_D void rho_part(const double rho_in, RhoPart &p) const {
    const double rho = stdmin(rho_in, rho_max);
    double P_H, E_H;
    if (rho >= rho0) {
      const double eta = 1 - rho0 / rho;
      const double s = s1 * eta + s2 * pow2i(eta) + s3 * pow3i(eta);
      P_H = P0 + pow2i(C0) * rho0 * eta / pow2i(1 - s);
      E_H = (P_H + P0) * eta / rho0 / 2.;
    } else {
      P_H = P0 + pow2i(C0) * (rho - rho0);
      E_H = 0.;
    }
    p.Ps = P_H;
    p.Es = E_H;
    p.rG = Gamma(rho) * rho;
    p.Ts = T0;
  }
  SGX_D double e_from_part(const RhoPart &, double, const double temp) const {
    return Cv * (temp - T0);
  }
  SGX_D double P_from_part(const RhoPart &p, double, const double sie) const {
    return p.Ps + p.rG * (sie - p.Es);
  }
  SGX_D double cv_from_part(const RhoPart &, double, double) const { return Cv; }
  SGX_D double P_from_rho_e(const double rho_in, const double sie, double * = nullptr) const {
    RhoPart p;
    rho_part(rho_in, p);
    return P_from_part(p, rho_in, sie);
  }
  SGX_D double P_from_rho_T(const double rho_in, const double temp, double * = nullptr) const {
    const double rho = stdmin(rho_in, rho_max);
    return P_from_rho_e(rho, e_from_rho_T(rho, temp));
  }
  SGX_D double cv_from_rho_T(double, double, double * = nullptr) const { return Cv; }
  SGX_D double cv_from_rho_e(double, double, double * = nullptr) const { return Cv; }
  // eos_gruneisen.hpp:331-351
  SGX_D double dPres_drho_e(const double rho_in, const double sie) const {
    const double rho = stdmin(rho_in, rho_max);
    if (rho < rho0) {
      return pow2i(C0) + Gamma(rho) * sie;
    } else {
      const double eta = 1 - rho0 / rho;
      const double s = s1 * eta + s2 * pow2i(eta) + s3 * pow3i(eta);
      const double ds = s1 + 2 * s2 * eta + 3 * s3 * pow2i(eta);
      const double deta = rho0 / pow2i(rho);
      const double dGam = (b - G0) * deta;
      const double P_H = P0 + pow2i(C0) * rho0 * eta / pow2i(1 - s);
      const double dP_H =
          pow2i(C0) * rho0 / pow2i(1 - s) * deta * (1 + 2 * eta * ds / (1 - s));
      const double E_H = (P_H + P0) * eta / rho0 / 2.;
      const double dE_H = deta * (P_H + P0) / rho0 / 2. + eta / rho0 / 2 * dP_H;
      return dP_H + Gamma(rho) * (sie - E_H) + rho * dGam * (sie - E_H) -
             rho * Gamma(rho) * dE_H;
    }
  }
  // eos_gruneisen.hpp:382-397
  SGX_D double bmod_from_rho_e(const double rho_in, const double sie,
                               double * = nullptr) const {
    const double rho = stdmin(rho_in, rho_max);
    if (rho < rho0) {
      return rho * pow2i(C0) + G0 * (rho * sie + P_from_rho_e(rho, sie));
    } else {
      const double dPdr_e = dPres_drho_e(rho, sie);
      const double dPde_r = rho * Gamma(rho);
      return rho * dPdr_e + P_from_rho_e(rho, sie) / rho * dPde_r;
    }
  }
  SGX_D double bmod_from_rho_T(const double rho_in, const double temp,
                               double * = nullptr) const {
    const double rho = stdmin(rho_in, rho_max);
    return bmod_from_rho_e(rho, e_from_rho_T(rho, temp));
  }
  SGX_D double gamma_from_rho_T(double rho, double, double * = nullptr) const {
    return Gamma(stdmin(rho, rho_max));
  }
  SGX_D double gamma_from_rho_e(double rho, double, double * = nullptr) const {
    return Gamma(stdmin(rho, rho_max));
  }
  // eos_gruneisen.hpp:423-477.  `ok` is cleared where the reference aborts / EOS_ERRORs.
  __device__ inline double MaxStableDensityAtTemperature(const double temperature) const {
    const double slope_at_max_density =
        dPres_drho_e(rho_max, e_from_rho_T(rho_max, temperature));
    const double slope_at_ref_density = dPres_drho_e(rho0, e_from_rho_T(rho0, temperature));
    if (slope_at_max_density >= 0) return rho_max;
    (void)slope_at_ref_density; // reference aborts when < 0; we continue (never trap)
    auto dPdrho_T = [this, temperature](const double r) {
      return dPres_drho_e(r, e_from_rho_T(r, temperature));
    };
    const double rho_lower = rho0;
    const double rho_upper = stdmin(rho_max, 1.0e4);
    const double rho_guess = (rho_lower + rho_upper) / 2.;
    double rho_at_max_P = rho_guess;
    root::regula_falsi(dPdrho_T, 0., rho_guess, rho_lower, rho_upper, 1.0e-8, 1.0e-8,
                       rho_at_max_P);
    return rho_at_max_P;
  }
  // eos_gruneisen.hpp:478-526
  __device__ inline void rho_e_from_P_T(const double press, const double temp, double *,
                                        double &rho, double &sie) const {
    const double Pref = P_from_rho_T(rho0, temp);
    double rho_lower, rho_upper, rho_guess;
    double p_used = press;
    if (press < Pref) {
      rho_lower = 0.;
      rho_upper = 1.1 * rho0;
      rho_guess = rho0;
    } else {
      rho_lower = 0.9 * rho0;
      rho_upper = MaxStableDensityAtTemperature(temp);
      const double pres_max = P_from_rho_T(rho_upper, temp);
      if (press > pres_max) p_used = pres_max;
      const double slope = (rho_upper - rho0) / (pres_max - Pref);
      rho_guess = rho0 + slope * (p_used - Pref);
    }
    auto PofRatT = [this, temp](const double r) { return P_from_rho_T(r, temp); };
    root::regula_falsi(PofRatT, p_used, rho_guess, rho_lower, rho_upper, 1.0e-8, 1.0e-8, rho);
    sie = e_from_rho_T(rho, temp);
  }
  // eos_gruneisen.hpp:527-548
  __device__ inline bool fill(double &rho_in, double &temp, double &sie, double &press,
                              double &cv, double &bmod, const unsigned long output,
                              double *lambda) const {
    const unsigned long input = ~output;
    if ((TQ_TEMPERATURE & input) && (TQ_PRESSURE & input)) {
      rho_e_from_P_T(press, temp, lambda, rho_in, sie);
    } else if ((TQ_DENSITY & output) || (TQ_SIE & output)) {
      return false;
    }
    const double rho = stdmin(rho_in, rho_max);
    if (output & TQ_PRESSURE) press = P_from_rho_e(rho, sie);
    if (output & TQ_TEMPERATURE) temp = T_from_rho_e(rho, sie);
    if (output & TQ_BMOD) bmod = bmod_from_rho_e(rho, sie);
    if (output & TQ_CV) cv = cv_from_rho_e(rho, sie);
    return true;
  }
  SGX_D double MinimumDensity() const { return 0; }
  SGX_D double MinimumTemperature() const { return 0; }
  SGX_D double MaximumDensity() const { return rho_max; }
  SGX_D double MinimumPressure() const { return P_from_rho_e(0, 0); }
  // returns a *density* -- reference quirk, eos_gruneisen.hpp:174-177
  SGX_D double MaximumPressureAtTemperature(const double T) const {
    return MaxStableDensityAtTemperature(T);
  }
  SGX_D double RhoPmin(double) const { return 0.0; }
};

// ---------------------------------------------------------------------------------------
struct DavisReactants {
  double rho0, spvol0, e0, P0, T0, A, B, C, G0, Z, alpha, i1pa, Cv0, A2oB;
  static constexpr unsigned long preferred_input = TQ_DENSITY | TQ_SIE;
  static constexpr int nlambda = 0;
  static constexpr double onethird = 1.0 / 3.0;

  // eos_davis.hpp:402-417
  SGX_D double Ps(const double rho) const {
    if (rho <= 0) return 0;
    const double y = 1.0 - ratio(rho0, stdmax(rho, 0.));
    const double phat = 0.25 * A2oB * rho0;
    const double b4y = 4.0 * B * y;
    if (rho >= rho0) {
      return phat *
             (b4y + 0.5 * (pow2i(b4y) + onethird * (pow3i(b4y) + C * pow4i(b4y) * 0.25)) +
              ratio(pow2i(y), pow4i(1 - y)));
    } else {
      return phat * (xexp(b4y) - 1.0);
    }
  }
  // eos_davis.hpp:418-434
  SGX_D double Es(const double rho) const {
    if (rho <= 0) return 0;
    const double y = 1 - ratio(rho0, stdmax(rho, 0.));
    const double phat = 0.25 * A2oB * rho0;
    const double b4y = 4 * B * y;
    double e_s;
    if (y > 0.0) {
      const double z = rho * spvol0 - 1;
      e_s = 0.5 * y * b4y *
                (1.0 + onethird * b4y * (1.0 + 0.25 * b4y * (1.0 + C * 0.2 * b4y))) +
            onethird * pow3i(z);
    } else {
      e_s = -y - ratio(1.0 - xexp(b4y), 4.0 * B);
    }
    return e0 + P0 * (spvol0 - ratio(1.0, stdmax(rho, 0.))) + phat * spvol0 * e_s;
  }
  // eos_davis.hpp:435-440
  SGX_D double Ts(const double rho) const {
    if (rho <= 0) return T0;
    const double rho0overrho = ratio(rho0, stdmax(rho, 0.));
    const double y = 1 - rho0overrho;
    return T0 * xexp(-Z * y) * xpow(rho0overrho, -G0 - Z);
  }
  // eos_davis.hpp:441-448
  SGX_D double Gamma(const double rho) const {
    if (rho >= rho0) {
      const double y = 1 - ratio(rho0, stdmax(rho, 0.));
      return G0 + Z * y;
    } else {
      return G0;
    }
  }
  SGX_D double DimlessEdiff(const double rho, const double sie) const {
    return ratio(1.0 + alpha, Ts(rho) * Cv0) * (sie - Es(rho)) + 1.0;
  }
  SGX_D double T_from_rho_e(const double rho, const double sie, double * = nullptr) const {
    const double power_base = DimlessEdiff(rho, sie);
    if (power_base <= 0) return 0.;
    const double tmp = xpow(power_base, i1pa);
    return Ts(rho) * tmp;
  }
  // staged form of :76-84 (e), :92-97 (P), :129-139 (cv)
  // Ps, Es, Ts, Gamma share y = 1 - rho0/rho, 4By and exp(4By); each is evaluated once here
  // (same operands => same bits as the one-shot forms above, expressions kept literally)
  SGX_D void rho_part(const double rho, RhoPart &p) const {
    if (rho <= 0) { // the one-shot forms' guards (:402-440); Gamma(rho < rho0) = G0
      p.Ps = 0;
      p.Es = 0;
      p.Ts = T0;
      p.rG = G0 * rho;
      return;
    }
    const double rho0overrho = ratio(rho0, stdmax(rho, 0.));
    const double y = 1.0 - rho0overrho;
    const double phat = 0.25 * A2oB * rho0;
    const double b4y = 4.0 * B * y;
    const bool compressed = rho >= rho0;
    const bool y_pos = y > 0.0;
    double eb4y = 0.0;
    if (!compressed || !y_pos) eb4y = xexp(b4y);
    if (compressed) {
      p.Ps = phat *
             (b4y + 0.5 * (pow2i(b4y) + onethird * (pow3i(b4y) + C * pow4i(b4y) * 0.25)) +
              ratio(pow2i(y), pow4i(1 - y)));
    } else {
      p.Ps = phat * (eb4y - 1.0);
    }
    double e_s;
    if (y_pos) {
      const double z = rho * spvol0 - 1;
      e_s = 0.5 * y * b4y *
                (1.0 + onethird * b4y * (1.0 + 0.25 * b4y * (1.0 + C * 0.2 * b4y))) +
            onethird * pow3i(z);
    } else {
      e_s = -y - ratio(1.0 - eb4y, 4.0 * B);
    }
    p.Es = e0 + P0 * (spvol0 - ratio(1.0, stdmax(rho, 0.))) + phat * spvol0 * e_s;
    p.Ts = T0 * xexp(-Z * y) * xpow(rho0overrho, -G0 - Z);
    p.rG = (compressed ? G0 + Z * y : G0) * rho;
  }
  SGX_D double e_from_part(const RhoPart &p, double, const double temp) const {
    return p.Es + Cv0 * p.Ts * i1pa * (xpow(ratio(temp, p.Ts), 1.0 + alpha) - 1.0);
  }
  SGX_D double P_from_part(const RhoPart &p, double, const double sie) const {
    return p.Ps + p.rG * (sie - p.Es);
  }
  SGX_D double cv_from_part(const RhoPart &p, double, const double sie) const {
    const double power_base = ratio(1.0 + alpha, p.Ts * Cv0) * (sie - p.Es) + 1.0;
    if (power_base <= 0) return 0.;
    return ratio(Cv0, xpow(power_base, -alpha * i1pa));
  }
  SGX_D double e_from_rho_T(const double rho, const double temp, double * = nullptr) const {
    const double t_s = Ts(rho);
    return Es(rho) + Cv0 * t_s * i1pa * (xpow(ratio(temp, t_s), 1.0 + alpha) - 1.0);
  }
  SGX_D double P_from_rho_e(const double rho, const double sie, double * = nullptr) const {
    return Ps(rho) + Gamma(rho) * rho * (sie - Es(rho));
  }
  SGX_D double P_from_rho_T(const double rho, const double temp, double * = nullptr) const {
    return P_from_rho_e(rho, e_from_rho_T(rho, temp));
  }
  // eos_davis.hpp:129-139
  SGX_D double cv_from_rho_e(const double rho, const double sie, double * = nullptr) const {
    const double power_base = DimlessEdiff(rho, sie);
    if (power_base <= 0) return 0.;
    return ratio(Cv0, xpow(power_base, -alpha * i1pa));
  }
  SGX_D double cv_from_rho_T(const double rho, const double temp, double * = nullptr) const {
    return cv_from_rho_e(rho, e_from_rho_T(rho, temp));
  }
  // eos_davis.hpp:450-474
  SGX_D double bmod_from_rho_e(const double rho, const double sie, double * = nullptr) const {
    const double y = 1 - ratio(rho0, stdmax(rho, 0.));
    const double phat = 0.25 * A2oB * rho0;
    const double b4y = 4 * B * y;
    const double gamma = Gamma(stdmax(rho, 0.));
    const double esv = -Ps(stdmax(rho, 0.));
    const double psv =
        (rho >= rho0)
            ? -phat * rho0 *
                  (4 * B * (1 + b4y + 0.5 * (pow2i(b4y) + onethird * C * pow3i(b4y))) +
                   3 * ratio(y, pow4i(1 - y)) + 4 * ratio(pow2i(y), pow5i(1 - y)))
            : -phat * 4 * B * rho0 * xexp(b4y);
    const double gammav = (rho >= rho0) ? -Z * rho0 : 0.0;
    const double es = Es(rho);
    const double numerator =
        -(psv + (sie - es) * stdmax(rho, 0.) * (gammav - gamma * stdmax(rho, 0.)) -
          gamma * stdmax(rho, 0.) * esv);
    const double ke = ratio(numerator, stdmax(rho, 0.));
    const double p = -esv + rho * gamma * (sie - es);
    return ke + gamma * p;
  }
  SGX_D double bmod_from_rho_T(const double rho, const double temp,
                               double * = nullptr) const {
    return bmod_from_rho_e(rho, e_from_rho_T(rho, temp));
  }
  SGX_D double gamma_from_rho_T(double rho, double, double * = nullptr) const {
    return Gamma(rho);
  }
  SGX_D double gamma_from_rho_e(double rho, double, double * = nullptr) const {
    return Gamma(rho);
  }
  // eos_davis.hpp:476-496
  __device__ inline void rho_e_from_P_T(const double press, const double temp, double *,
                                        double &rho, double &sie) const {
    auto PofRatT = [&](const double r) {
      return (Ps(r) + Gamma(r) * r * Cv0 * Ts(r) * i1pa *
                          (xpow(ratio(temp, Ts(r)), 1 + alpha) - 1.0));
    };
    root::regula_falsi(PofRatT, press, rho0, 1.0e-5, 1.0e3, 1.0e-8, 1.0e-8, rho);
    rho = stdmax(0., rho);
    sie = e_from_rho_T(rho, temp);
  }
  __device__ inline bool fill(double &rho, double &temp, double &sie, double &press,
                              double &cv, double &bmod, const unsigned long output,
                              double *) const {
    if (output & TQ_PRESSURE) press = P_from_rho_e(rho, sie);
    if (output & TQ_TEMPERATURE) temp = T_from_rho_e(rho, sie);
    if (output & TQ_BMOD) bmod = bmod_from_rho_e(rho, sie);
    if (output & TQ_CV) cv = cv_from_rho_e(rho, sie);
    return true;
  }
  SGX_D double MinimumDensity() const { return kEps; }
  SGX_D double MinimumTemperature() const { return 0; }
  SGX_D double MaximumDensity() const { return 1e100; }
  SGX_D double MinimumPressure() const { return 0; }
  SGX_D double MaximumPressureAtTemperature(double) const { return 1e100; }
  SGX_D double RhoPmin(double) const { return 0.0; }
};

// ---------------------------------------------------------------------------------------
struct DavisProducts {
  double a, b, k, n, aon, vc, pc, Cv;
  // pow(2, -a*b/n) * pc * vc / (Cv*(k-1+a)): the density-independent factor of Ts (:381-391),
  // evaluated once on the host at create time (the reference evaluates it with the host libm
  // on every call)
  double Ts_coef;
  static constexpr unsigned long preferred_input = TQ_DENSITY | TQ_SIE;
  static constexpr int nlambda = 0;

  // eos_davis.hpp:353-394
  SGX_D double F(const double rho) const {
    if (rho <= 0) return 0.;
    const double vvc = ratio(1.0, rho * vc);
    return 2.0 * ratio(a, xpow(vvc, 2 * n) + 1.0);
  }
  SGX_D double Ps(const double rho) const {
    if (rho <= 0) return 0.;
    const double vvc = ratio(1, rho * vc);
    return pc * ratio(xpow(0.5 * (xpow(vvc, n) + xpow(vvc, -n)), aon), xpow(vvc, k + a)) *
           ratio(k - 1.0 + F(rho), k - 1.0 + a);
  }
  SGX_D double Es(const double rho) const {
    if (rho <= 0) return 0.;
    const double vvc = ratio(1.0, rho * vc);
    const double ec = pc * ratio(vc, k - 1.0 + a);
    return ratio(ec * xpow(0.5 * (xpow(vvc, n) + xpow(vvc, -n)), aon), xpow(vvc, k - 1.0 + a));
  }
  SGX_D double Ts(const double rho) const {
    if (rho <= 0) return 0.;
    const double vvc = ratio(1.0, rho * vc);
    return Ts_coef *
           ratio(xpow(0.5 * (xpow(vvc, n) + xpow(vvc, -n)), aon * (1 - b)),
                 xpow(vvc, k - 1.0 + a * (1 - b)));
  }
  SGX_D double Gamma(const double rho) const { return k - 1.0 + (1.0 - b) * F(rho); }
  SGX_D double T_from_rho_e(const double rho, const double sie, double * = nullptr) const {
    return Ts(rho) + ratio(sie - Es(rho), Cv);
  }
  // staged form of :235-240 (e), :248-253 (P)
  // Ts, Es, Ps and F share pow(vvc, +-n) and the cosh-like mixture; each distinct pow is
  // evaluated once here (same function, same operands => same bits as the one-shot forms
  // above, whose expressions and order are kept literally).
  SGX_D void rho_part(const double rho, RhoPart &p) const {
    if (rho <= 0) { // the one-shot forms' guards
      p.Ts = 0.;
      p.Es = 0.;
      p.Ps = 0.;
      p.rG = rho * (k - 1.0 + (1.0 - b) * 0.);
      return;
    }
    const double vvc = ratio(1.0, rho * vc);
    const double mix = 0.5 * (xpow(vvc, n) + xpow(vvc, -n));
    const double mix_aon = xpow(mix, aon);
    const double Fv = 2.0 * ratio(a, xpow(vvc, 2 * n) + 1.0);
    p.Ts = Ts_coef * ratio(xpow(mix, aon * (1 - b)), xpow(vvc, k - 1.0 + a * (1 - b)));
    const double ec = pc * ratio(vc, k - 1.0 + a);
    p.Es = ratio(ec * mix_aon, xpow(vvc, k - 1.0 + a));
    p.Ps = pc * ratio(mix_aon, xpow(vvc, k + a)) * ratio(k - 1.0 + Fv, k - 1.0 + a);
    p.rG = rho * (k - 1.0 + (1.0 - b) * Fv);
  }
  SGX_D double e_from_part(const RhoPart &p, double, const double temp) const {
    return Cv * (temp - p.Ts) + p.Es;
  }
  SGX_D double P_from_part(const RhoPart &p, double, const double sie) const {
    return p.Ps + p.rG * (sie - p.Es);
  }
  SGX_D double cv_from_part(const RhoPart &, double, double) const { return Cv; }
  SGX_D double e_from_rho_T(const double rho, const double temp, double * = nullptr) const {
    return Cv * (temp - Ts(rho)) + Es(rho);
  }
  SGX_D double P_from_rho_e(const double rho, const double sie, double * = nullptr) const {
    return Ps(rho) + rho * Gamma(rho) * (sie - Es(rho));
  }
  SGX_D double P_from_rho_T(const double rho, const double temp, double * = nullptr) const {
    return P_from_rho_e(rho, e_from_rho_T(rho, temp));
  }
  SGX_D double cv_from_rho_T(double, double, double * = nullptr) const { return Cv; }
  SGX_D double cv_from_rho_e(double, double, double * = nullptr) const { return Cv; }
  // eos_davis.hpp:531-562
  SGX_D double bmod_from_rho_e(const double rho, const double sie, double * = nullptr) const {
    if (rho <= 0) return 0.;
    const double vvc = ratio(1.0, rho * vc);
    const double Fx = -4 * a * n * ratio(xpow(vvc, 2 * n - 1), pow2i(1 + xpow(vvc, 2 * n)));
    const double tmp = ratio(xpow(0.5 * (xpow(vvc, n) + xpow(vvc, -n)), aon), xpow(vvc, k + a));
    const double tmp_x =
        0.5 * a * (xpow(vvc, n - 1) - xpow(vvc, -n - 1)) *
            ratio(xpow(0.5 * (xpow(vvc, n) + xpow(vvc, -n)), aon - 1), xpow(vvc, k + a)) -
        (k + a) * ratio(tmp, vvc);
    const double psv =
        ratio(pc, k - 1 + a) * ratio(tmp * Fx + (k - 1 + F(rho)) * tmp_x, vc);
    const double esv = ratio(pc, k - 1 + a) * (tmp + vvc * tmp_x);
    const double gamma = Gamma(rho);
    const double es = Es(rho);
    const double gammav = (1 - b) * Fx / vc;
    const double ke =
        -ratio(psv + (sie - es) * rho * (gammav - gamma * rho) - gamma * rho * esv, rho);
    const double p = Ps(rho) + rho * gamma * (sie - es);
    return ke + gamma * p;
  }
  SGX_D double bmod_from_rho_T(const double rho, const double temp,
                               double * = nullptr) const {
    return bmod_from_rho_e(rho, e_from_rho_T(rho, temp));
  }
  SGX_D double gamma_from_rho_T(double rho, double, double * = nullptr) const {
    return Gamma(rho);
  }
  SGX_D double gamma_from_rho_e(double rho, double, double * = nullptr) const {
    return Gamma(rho);
  }
  // eos_davis.hpp:564-582
  __device__ inline void rho_e_from_P_T(const double press, const double temp, double *,
                                        double &rho, double &sie) const {
    auto PofRatT = [&](const double r) {
      return (Ps(r) + Gamma(r) * r * Cv * (temp - Ts(r)));
    };
    const double rho0 = ratio(1.0, vc);
    root::regula_falsi(PofRatT, press, rho0, 1.0e-5, 1.0e3, 1.0e-8, 1.0e-8, rho);
    rho = stdmax(rho, 0.);
    sie = e_from_rho_T(rho, temp);
  }
  __device__ inline bool fill(double &rho, double &temp, double &sie, double &press,
                              double &cv, double &bmod, const unsigned long output,
                              double *) const {
    if (output & TQ_PRESSURE) press = P_from_rho_e(rho, sie);
    if (output & TQ_TEMPERATURE) temp = T_from_rho_e(rho, sie);
    if (output & TQ_BMOD) bmod = bmod_from_rho_e(rho, sie);
    if (output & TQ_CV) cv = cv_from_rho_e(rho, sie);
    return true;
  }
  SGX_D double MinimumDensity() const { return kEps; }
  SGX_D double MinimumTemperature() const { return 0; }
  SGX_D double MaximumDensity() const { return 1e100; }
  SGX_D double MinimumPressure() const { return 0; }
  SGX_D double MaximumPressureAtTemperature(double) const { return 1e100; }
  SGX_D double RhoPmin(double) const { return 0.0; }
};
#endif // __CUDACC__

} // namespace sgx
#endif // SGX_ANALYTIC_CUH_
