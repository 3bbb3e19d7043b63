// Tabulated Spiner EOS on device: SpinerEOSDependsRhoT and SpinerEOSDependsRhoSie
// (NullTransform).  One contiguous device allocation per EOS holds every table; the structs
// below carry raw pointers into it plus grid metadata and are passed to kernels by value.
//
// Reference:
//   eos/eos_spiner_rho_temp.hpp:777-1123  public scalar methods
//   eos/eos_spiner_rho_temp.hpp:1125-1404 getLogsRhoT_, lRhoFromPlT_, lTFromlRhoSie_,
//                                         lTFromlRhoP_, *FromlRholT_, getLocDepends*
//   eos/eos_spiner_rho_sie.hpp:654-939    scalar methods, FillEos, interpRhoT_/interpRhoSie_
//   eos/eos_spiner_rho_sie.hpp:941-986    lRhoFromPlT_
#ifndef SGX_SPINER_CUH_
#define SGX_SPINER_CUH_

#include "../../include/sgx_b200.h" // enum sgx_method
#include "sgx_analytic.cuh"
#include "sgx_math.cuh"

namespace sgx {

enum TableStatus { OnTable = 0, OffBottom = 1, OffTop = 2 };

// Two-phase form of the direct (non-root-finding) lookups.  prep() does the NQT logs and grid
// weights and ISSUES the 256-bit load of the table cell -- unconditionally: weights() clamps the
// indices, so the address is always a valid cell even for off-table states.  finish() blends and
// handles the rare off-table branches.  A kernel that runs prep() for several cells before any
// finish() has all of their gathers in flight together (memory-level parallelism); with the
// branches inside one scalar method the compiler cannot hoist the loads across cells.
// Arithmetic and evaluation order are those of the one-phase methods, so results are identical.
struct LookupPrep {
  double x2, x1; // lRho and lT (or lE)
  AxisPos p2, p1;
  Cell4 cell;
};

struct SpinerRhoT {
  Grid1D gRho, gT; // gT = range(0) (fastest), gRho = range(1)
  int numRho, numT;
  // (lRho, lT) tables, CELL-MAJOR packed (sgx_math.cuh: Cell4 / interp2c)
  const double *P, *sie, *bMod, *dPdRho, *dPdE, *dTdRho, *dTdE, *dEdRho, *dEdT;
  // 1-D over lRho
  const double *PMax, *sielTMax, *dEdTMax, *gm1Max, *lTColdCrit;
  const double *PCold, *sieCold, *bModCold, *dPdRhoCold, *dPdECold, *dTdRhoCold, *dTdECold,
      *dEdTCold;
  // 1-D over lT
  const double *rho_at_pmin;
  double lRhoMin, lRhoMax, rhoMax, lTMin, lTMax, TMax, PMin;
  double lRhoOffset, lTOffset;
  int reproducible;

  static constexpr unsigned long preferred_input = TQ_DENSITY | TQ_TEMPERATURE;
  static constexpr int nlambda = 2;
  static constexpr double ROOT_THRESH = 1e-14;
  static constexpr double SOFT_THRESH = 1e-8;

#ifdef __CUDACC__
  SGX_D double lRho_(const double rho) const { return to_log(rho, lRhoOffset); }
  SGX_D double lT_(const double T) const { return to_log(T, lTOffset); }
  SGX_D double rho_(const double lRho) const {
    const double rho = from_log(lRho, lRhoOffset);
    return rho < 0 ? 0 : rho;
  }
  SGX_D double T_(const double lT) const { return from_log(lT, lTOffset); }

  SGX_D void getLogsRhoT(const double rho, const double T, double &lRho, double &lT,
                         double *lambda) const {
    lRho = lRho_(rho);
    lT = lT_(T);
    if (lambda) {
      lambda[0] = lRho;
      lambda[1] = lT;
    }
  }
  SGX_D TableStatus locRhoT(const double lRho, const double lT) const {
    if (lT <= (1 + SOFT_THRESH) * lTMin) return OffBottom;
    if (lT >= (1 - SOFT_THRESH) * lTMax) return OffTop;
    return OnTable;
  }
  SGX_D TableStatus locRhoSie(const AxisPos &pr, const double sie) const {
    const double sielTMax_ = interp1d(sielTMax, pr);
    const double sieCold_ = interp1d(sieCold, pr);
    if (sie >= sielTMax_ - SOFT_THRESH * fabs(sielTMax_)) return OffTop;
    if (sie <= sieCold_ + SOFT_THRESH * fabs(sieCold_)) return OffBottom;
    return OnTable;
  }

  SGX_D double sieFromlRhoTlT(const AxisPos &pr, const double T, const double lT,
                              const TableStatus w) const {
    if (w == OffBottom) return interp1d(sieCold, pr);
    if (w == OffTop) {
      const double Cv = interp1d(dEdTMax, pr);
      const double e0 = interp1d(sielTMax, pr);
      return e0 + Cv * (T - TMax);
    }
    return interp2c(sie, numT, pr, AxisPos(gT, lT));
  }
  SGX_D double PFromRholRhoTlT(const double rho, const AxisPos &pr, const double T,
                               const double lT, const TableStatus w) const {
    if (w == OffBottom) return interp1d(PCold, pr);
    if (w == OffTop) {
      const double Cv = interp1d(dEdTMax, pr);
      const double gm1 = interp1d(gm1Max, pr);
      const double e0 = interp1d(sielTMax, pr);
      const double e = e0 + Cv * (T - TMax);
      return gm1 * rho * e;
    }
    return interp2c(P, numT, pr, AxisPos(gT, lT));
  }
  SGX_D double CvFromlRholT(const AxisPos &pr, const double lT, const TableStatus w) const {
    double Cv;
    if (w == OffBottom) {
      Cv = interp1d(dEdTCold, pr);
    } else if (w == OffTop) {
      Cv = interp1d(dEdTMax, pr);
    } else {
      Cv = interp2c(dEdT, numT, pr, AxisPos(gT, lT));
    }
    return Cv > kEps ? Cv : kEps;
  }
  SGX_D double bModFromRholRhoTlT(const double rho, const AxisPos &pr, const double T,
                                  const double lT, const TableStatus w) const {
    double bm;
    if (w == OffBottom) {
      bm = interp1d(bModCold, pr);
    } else if (w == OffTop) {
      const double Cv = interp1d(dEdTMax, pr);
      const double gm1 = interp1d(gm1Max, pr);
      const double e0 = interp1d(sielTMax, pr);
      const double e = e0 + Cv * (T - TMax);
      bm = (gm1 + 1) * gm1 * rho * e;
    } else {
      bm = interp2c(bMod, numT, pr, AxisPos(gT, lT));
    }
    return bm > kEps ? bm : kEps;
  }

  // lTFromlRhoSie_ (eos_spiner_rho_temp.hpp:1192-1250).  The density-axis position `pr` is
  // computed once; every regula-falsi evaluation is then the reference's four-corner blend
  // with fixed (ix2, w2) -- identical arithmetic, fewer instructions.
  __device__ inline double lTFromlRhoSie(const double lRho, const AxisPos &pr,
                                         const double sie_in, TableStatus &w, double *lambda,
                                         int *root_status = nullptr,
                                         int *niter = nullptr) const {
    double lT;
    int status = root::SUCCESS;
    if (niter) *niter = 0;
    w = locRhoSie(pr, sie_in);
    if (w == OffBottom) {
      lT = lTMin;
    } else if (w == OffTop) {
      const double Cv = interp1d(dEdTMax, pr);
      const double e0 = interp1d(sielTMax, pr);
      const double T = TMax + ratio(sie_in - e0, Cv);
      lT = lT_(T);
    } else {
      double lTGuess = reproducible ? lTMin : 0.5 * (lTMin + lTMax);
      if (lambda) {
        const double lT_cache = lambda[1];
        if ((lTMin <= lT_cache) && (lT_cache <= lTMax)) lTGuess = lT_cache;
      }
      const double *tab = sie;
      const int n1 = numT;
      const Grid1D &g1 = gT;
      auto sieFunc = [&](const double x) {
        int ix1;
        double w10, w11;
        g1.weights(x, ix1, w10, w11);
        return blend2c(tab, n1, pr.ix, pr.w0, pr.w1, ix1, w10, w11);
      };
      status = root::regula_falsi(sieFunc, sie_in, lTGuess, lTMin, lTMax, ROOT_THRESH,
                                  ROOT_THRESH, lT, niter);
      if (status != root::SUCCESS) lT = reproducible ? lTMin : lTGuess;
    }
    if (lambda) {
      lambda[0] = lRho;
      lambda[1] = lT;
    }
    if (root_status) *root_status = status;
    return lT;
  }

  // lTFromlRhoP_ (eos_spiner_rho_temp.hpp:1252-1305).  NB the reference passes the
  // (uninitialised) local `lT` as the lambda *index* (:1278); with a Real* lambda that reads
  // an arbitrary slot, so the cache is never trusted here: guess = midpoint.
  __device__ inline double lTFromlRhoP(const double lRho, const AxisPos &pr,
                                       const double press, TableStatus &w,
                                       double *lambda) const {
    double lT;
    const double PCold_ = interp1d(PCold, pr);
    const double PMax_ = interp1d(PMax, pr);
    if (press <= PCold_) {
      w = OffBottom;
      lT = lTMin;
    } else if (press >= PMax_) {
      w = OffTop;
      lT = lTMax;
    } else {
      w = OnTable;
      const double lTGuess = 0.5 * (lTMin + lTMax);
      const double *tab = P;
      const int n1 = numT;
      const Grid1D &g1 = gT;
      auto PFunc = [&](const double x) {
        int ix1;
        double w10, w11;
        g1.weights(x, ix1, w10, w11);
        return blend2c(tab, n1, pr.ix, pr.w0, pr.w1, ix1, w10, w11);
      };
      const int status = root::regula_falsi(PFunc, press, lTGuess, lTMin, lTMax, ROOT_THRESH,
                                            ROOT_THRESH, lT);
      if (status != root::SUCCESS) lT = reproducible ? lTMin : lTGuess;
    }
    if (lambda) {
      lambda[0] = lRho;
      lambda[1] = lT;
    }
    return lT;
  }

  // lRhoFromPlT_ (eos_spiner_rho_temp.hpp:1135-1190)
  __device__ inline double lRhoFromPlT(const double Pin, const double lT, TableStatus &w,
                                       double *lambda) const {
    int status = root::SUCCESS;
    const double rhopmin = lRho_(interp1d(rho_at_pmin, gT, lT));
    double lRho;
    double lRhoGuess = reproducible ? lRhoMax : 0.5 * (rhopmin + lRhoMax);
    if (lambda) {
      const double lRho_cache = lambda[0];
      if ((rhopmin <= lRho_cache) && (lRho_cache <= lRhoMax)) lRhoGuess = lRho_cache;
    }
    const Grid1D &g2 = gRho;
    if (lT <= lTMin) {
      w = OffBottom;
      const double *tab = PCold;
      auto PFunc = [&](const double x) { return interp1d(tab, g2, x); };
      status = root::regula_falsi(PFunc, Pin, lRhoGuess, rhopmin, lRhoMax, ROOT_THRESH,
                                  ROOT_THRESH, lRho);
    } else if (lT >= lTMax) {
      w = OffTop;
      const double *t1 = gm1Max, *t2 = dEdTMax;
      const double r = lT;
      auto PFunc = [&](const double x) {
        return interp1d(t1, g2, x) * interp1d(t2, g2, x) * r * x;
      };
      status = root::regula_falsi(PFunc, Pin, lRhoGuess, rhopmin, lRhoMax, ROOT_THRESH,
                                  ROOT_THRESH, lRho);
    } else {
      w = OnTable;
      const AxisPos pt(gT, lT);
      const double *tab = P;
      const int n1 = numT;
      auto PFunc = [&](const double x) {
        return interp2c(tab, n1, AxisPos(g2, x), pt);
      };
      status = root::regula_falsi(PFunc, Pin, lRhoGuess, rhopmin, lRhoMax, ROOT_THRESH,
                                  ROOT_THRESH, lRho);
    }
    if (status != root::SUCCESS) lRho = reproducible ? lRhoMax : lRhoGuess;
    if (lambda) {
      lambda[0] = lRho;
      lambda[1] = lT;
    }
    return lRho;
  }

  // ---- public scalar methods ---------------------------------------------------------
  SGX_D double T_from_rho_e(const double rho, const double sie_in, double *lambda = nullptr) const {
    TableStatus w;
    const double lRho = lRho_(rho);
    const AxisPos pr(gRho, lRho);
    const double lT = lTFromlRhoSie(lRho, pr, sie_in, w, lambda);
    return T_(lT);
  }
  SGX_D double e_from_rho_T(const double rho, const double T, double *lambda = nullptr) const {
    double lRho, lT;
    getLogsRhoT(rho, T, lRho, lT, lambda);
    const TableStatus w = locRhoT(lRho, lT);
    return sieFromlRhoTlT(AxisPos(gRho, lRho), T, lT, w);
  }
  SGX_D double P_from_rho_T(const double rho, const double T, double *lambda = nullptr) const {
    double lRho, lT;
    getLogsRhoT(rho, T, lRho, lT, lambda);
    const TableStatus w = locRhoT(lRho, lT);
    return PFromRholRhoTlT(rho, AxisPos(gRho, lRho), T, lT, w);
  }
  SGX_D double P_from_rho_e(const double rho, const double sie_in, double *lambda = nullptr) const {
    TableStatus w;
    const double lRho = lRho_(rho);
    const AxisPos pr(gRho, lRho);
    const double lT = lTFromlRhoSie(lRho, pr, sie_in, w, lambda);
    if (w == OffBottom) return interp1d(PCold, pr);
    if (w == OffTop) {
      const double gm1 = interp1d(gm1Max, pr);
      return gm1 * rho * sie_in;
    }
    return interp2c(P, numT, pr, AxisPos(gT, lT));
  }
  SGX_D double cv_from_rho_T(const double rho, const double T, double *lambda = nullptr) const {
    double lRho, lT;
    getLogsRhoT(rho, T, lRho, lT, lambda);
    const TableStatus w = locRhoT(lRho, lT);
    return CvFromlRholT(AxisPos(gRho, lRho), lT, w);
  }
  SGX_D double cv_from_rho_e(const double rho, const double sie_in, double *lambda = nullptr) const {
    TableStatus w;
    double Cv;
    const double lRho = lRho_(rho);
    const AxisPos pr(gRho, lRho);
    const double lT = lTFromlRhoSie(lRho, pr, sie_in, w, lambda);
    if (w == OffBottom) {
      Cv = interp1d(dEdTCold, pr);
    } else if (w == OffTop) {
      Cv = interp1d(dEdTMax, pr);
    } else {
      Cv = interp2c(dEdT, numT, pr, AxisPos(gT, lT));
    }
    return Cv > kEps ? Cv : kEps;
  }
  SGX_D double bmod_from_rho_T(const double rho, const double T, double *lambda = nullptr) const {
    double lRho, lT;
    getLogsRhoT(rho, T, lRho, lT, lambda);
    const TableStatus w = locRhoT(lRho, lT);
    return bModFromRholRhoTlT(rho, AxisPos(gRho, lRho), T, lT, w);
  }
  SGX_D double bmod_from_rho_e(const double rho, const double sie_in, double *lambda = nullptr) const {
    TableStatus w;
    const double lRho = lRho_(rho);
    const AxisPos pr(gRho, lRho);
    const double lT = lTFromlRhoSie(lRho, pr, sie_in, w, lambda);
    if (w == OffBottom) return interp1d(bModCold, pr);
    if (w == OffTop) {
      const double gm1 = interp1d(gm1Max, pr);
      return (gm1 + 1) * gm1 * rho * sie_in;
    }
    return interp2c(bMod, numT, pr, AxisPos(gT, lT));
  }
  SGX_D double gamma_from_rho_T(const double rho, const double T, double *lambda = nullptr) const {
    double lRho, lT;
    getLogsRhoT(rho, T, lRho, lT, lambda);
    const AxisPos pr(gRho, lRho);
    const TableStatus w = locRhoT(lRho, lT);
    if (w == OffBottom) {
      const double dpde = interp1d(dPdECold, pr);
      return ratio(fabs(dpde), fabs(rho));
    }
    if (w == OffTop) return interp1d(gm1Max, pr);
    const double dpde = interp2c(dPdE, numT, pr, AxisPos(gT, lT));
    return ratio(fabs(dpde), fabs(rho));
  }
  SGX_D double gamma_from_rho_e(const double rho, const double sie_in, double *lambda = nullptr) const {
    TableStatus w;
    const double lRho = lRho_(rho);
    const AxisPos pr(gRho, lRho);
    const double lT = lTFromlRhoSie(lRho, pr, sie_in, w, lambda);
    if (w == OffBottom) {
      const double dpde = interp1d(dPdECold, pr);
      return ratio(fabs(dpde), fabs(rho));
    }
    if (w == OffTop) return interp1d(gm1Max, pr);
    const double dpde = interp2c(dPdE, numT, pr, AxisPos(gT, lT));
    return ratio(fabs(dpde), fabs(rho));
  }
  // ---- two-phase (rho,T) lookups: P, sie, Cv, bMod, Gruneisen ---------------------------
  template <int METHOD>
  static constexpr bool two_phase() {
    return METHOD == SGX_P_FROM_RHO_T || METHOD == SGX_E_FROM_RHO_T ||
           METHOD == SGX_CV_FROM_RHO_T || METHOD == SGX_BMOD_FROM_RHO_T ||
           METHOD == SGX_GAMMA_FROM_RHO_T;
  }
  template <int METHOD>
  SGX_D const double *table_of() const {
    if constexpr (METHOD == SGX_P_FROM_RHO_T) return P;
    if constexpr (METHOD == SGX_E_FROM_RHO_T) return sie;
    if constexpr (METHOD == SGX_CV_FROM_RHO_T) return dEdT;
    if constexpr (METHOD == SGX_BMOD_FROM_RHO_T) return bMod;
    return dPdE;
  }
  template <int METHOD>
  SGX_D void prep(const double rho, const double T, LookupPrep &q) const {
    q.x2 = lRho_(rho);
    q.x1 = lT_(T);
    q.p2 = AxisPos(gRho, q.x2);
    q.p1 = AxisPos(gT, q.x1);
    q.cell = ld_cell(table_of<METHOD>(), numT, q.p2.ix, q.p1.ix);
  }
  template <int METHOD>
  SGX_D double finish(const double rho, const double T, const LookupPrep &q,
                      double *lambda) const {
    if (lambda) { // getLogsRhoT_ (:1125-1133)
      lambda[0] = q.x2;
      lambda[1] = q.x1;
    }
    const TableStatus w = locRhoT(q.x2, q.x1);
    const double tab = blend_cell(q.cell, q.p2.w0, q.p2.w1, q.p1.w0, q.p1.w1);
    if constexpr (METHOD == SGX_P_FROM_RHO_T) {
      if (w == OnTable) return tab;
      return PFromRholRhoTlT(rho, q.p2, T, q.x1, w);
    } else if constexpr (METHOD == SGX_E_FROM_RHO_T) {
      if (w == OnTable) return tab;
      return sieFromlRhoTlT(q.p2, T, q.x1, w);
    } else if constexpr (METHOD == SGX_CV_FROM_RHO_T) {
      if (w == OnTable) return tab > kEps ? tab : kEps;
      return CvFromlRholT(q.p2, q.x1, w);
    } else if constexpr (METHOD == SGX_BMOD_FROM_RHO_T) {
      if (w == OnTable) return tab > kEps ? tab : kEps;
      return bModFromRholRhoTlT(rho, q.p2, T, q.x1, w);
    } else {
      if (w == OnTable) return ratio(fabs(tab), fabs(rho));
      if (w == OffBottom) return ratio(fabs(interp1d(dPdECold, q.p2)), fabs(rho));
      return interp1d(gm1Max, q.p2);
    }
  }
  // FillEos (eos_spiner_rho_temp.hpp:1043-1093)
  __device__ inline bool fill(double &rho, double &temp, double &energy, double &press,
                              double &cv, double &bmod, const unsigned long output,
                              double *lambda) const {
    double lRho, lT;
    TableStatus w;
    const unsigned long input = ~output;
    if (output == 0) return false;
    if (output & TQ_DENSITY) {
      if ((input & TQ_PRESSURE) && (input & TQ_TEMPERATURE)) {
        lT = lT_(temp);
        lRho = lRhoFromPlT(press, lT, w, lambda);
        rho = rho_(lRho);
      } else {
        return false;
      }
    } else {
      lRho = lRho_(rho);
    }
    const AxisPos pr(gRho, lRho);
    if (output & TQ_TEMPERATURE) {
      if ((input & TQ_DENSITY) && (input & TQ_SIE)) {
        lT = lTFromlRhoSie(lRho, pr, energy, w, lambda);
      } else if ((input & TQ_DENSITY) && (input & TQ_PRESSURE)) {
        lT = lTFromlRhoP(lRho, pr, press, w, lambda);
      } else {
        return false;
      }
      temp = T_(lT);
    } else {
      lT = lT_(temp);
    }
    w = locRhoT(lRho, lT);
    if (output & TQ_SIE) energy = sieFromlRhoTlT(pr, temp, lT, w);
    if (output & TQ_PRESSURE) press = PFromRholRhoTlT(rho, pr, temp, lT, w);
    if (output & TQ_CV) cv = CvFromlRholT(pr, lT, w);
    if (output & TQ_BMOD) bmod = bModFromRholRhoTlT(rho, pr, temp, lT, w);
    if (lambda) {
      lambda[0] = lRho;
      lambda[1] = lT;
    }
    return true;
  }
  SGX_D void rho_e_from_P_T(const double press, const double temp, double *lambda,
                            double &rho, double &sie_out) const {
    TableStatus w;
    const double lT = lT_(temp);
    const double lRho = lRhoFromPlT(press, lT, w, lambda);
    rho = rho_(lRho);
    sie_out = e_from_rho_T(rho, temp, lambda);
  }
  SGX_D double RhoPmin(const double temp) const {
    const double lT = lT_(temp);
    if (lT <= lTMin) return ldg(rho_at_pmin);
    if (lT >= lTMax) return 0.0;
    return interp1d(rho_at_pmin, gT, lT);
  }
  // bounds the PTE solvers ask for (eos_spiner_rho_temp.hpp:226-246; EosBase defaults :533-557)
  SGX_D double MinimumDensity() const { return rho_(lRhoMin); }
  SGX_D double MinimumTemperature() const { return T_(lTMin); }
  SGX_D double MaximumDensity() const { return rhoMax; }
  SGX_D double MinimumPressure() const { return PMin; }
  SGX_D double MaximumPressureAtTemperature(double) const { return 1e100; }
#endif
};

struct SpinerRhoSie {
  Grid1D gRho, gT, gE;
  int numRho, numT, numE;
  // (lRho, lT), CELL-MAJOR packed (sgx_math.cuh: Cell4 / interp2c)
  const double *sie, *P, *bMod, *dPdRho, *dPdE, *dTdRho, *dTdE, *dEdRho;
  // (lRho, lE), cell-major packed
  const double *T_re, *P_re, *bMod_re, *dPdRho_re, *dPdE_re, *dTdRho_re, *dTdE_re, *dEdRho_re;
  // 1-D over lRho
  const double *PCold, *sieCold, *bModCold, *dPdRhoCold;
  // 1-D over lT
  const double *rho_at_pmin, *PlRhoMax, *dPdRhoMax;
  double lRhoMin, lRhoMax, rhoMax, PMin;
  double lRhoOffset, lTOffset, lEOffset;
  int reproducible;

  static constexpr unsigned long preferred_input = TQ_DENSITY | TQ_TEMPERATURE;
  static constexpr int nlambda = 1;

#ifdef __CUDACC__
  // interpRhoT_ / interpRhoSie_ (eos_spiner_rho_sie.hpp:915-939), NullTransform
  SGX_D double interpRhoT(const double rho, const double T, const double *db,
                          double *lambda) const {
    const double lRho = to_log(rho, lRhoOffset);
    const double lT = to_log(T, lTOffset);
    if (lambda) lambda[0] = lRho;
    return interp2c(db, gRho, gT, numT, lRho, lT);
  }
  SGX_D double interpRhoSie(const double rho, const double sie_in, const double *db,
                            double *lambda) const {
    const double lRho = to_log(rho, lRhoOffset);
    const double lE = to_log(sie_in, lEOffset);
    if (lambda) lambda[0] = lRho;
    return interp2c(db, gRho, gE, numE, lRho, lE);
  }
  SGX_D double T_from_rho_e(double rho, double e, double *lambda = nullptr) const {
    return interpRhoSie(rho, e, T_re, lambda);
  }
  SGX_D double e_from_rho_T(double rho, double T, double *lambda = nullptr) const {
    return interpRhoT(rho, T, sie, lambda);
  }
  SGX_D double P_from_rho_T(double rho, double T, double *lambda = nullptr) const {
    return interpRhoT(rho, T, P, lambda);
  }
  SGX_D double P_from_rho_e(double rho, double e, double *lambda = nullptr) const {
    return interpRhoSie(rho, e, P_re, lambda);
  }
  SGX_D double cv_from_rho_T(double rho, double T, double *lambda = nullptr) const {
    return 1. / interpRhoT(rho, T, dTdE, lambda);
  }
  SGX_D double cv_from_rho_e(double rho, double e, double *lambda = nullptr) const {
    return 1. / interpRhoSie(rho, e, dTdE_re, lambda);
  }
  SGX_D double bmod_from_rho_T(double rho, double T, double *lambda = nullptr) const {
    return interpRhoT(rho, T, bMod, lambda);
  }
  SGX_D double bmod_from_rho_e(double rho, double e, double *lambda = nullptr) const {
    return interpRhoSie(rho, e, bMod_re, lambda);
  }
  SGX_D double gamma_from_rho_T(double rho, double T, double *lambda = nullptr) const {
    const double dpde = interpRhoT(rho, T, dPdE, lambda);
    return dpde / rho;
  }
  // NB: no lambda write on this one in the reference (eos_spiner_rho_sie.hpp:757-761)
  SGX_D double gamma_from_rho_e(double rho, double e, double *) const {
    const double lRho = to_log(rho, lRhoOffset);
    const double lE = to_log(e, lEOffset);
    const double dpde = interp2c(dPdE_re, gRho, gE, numE, lRho, lE);
    return dpde / rho;
  }
  // ---- two-phase lookups: every vector method is one bilinear interpolation ------------
  template <int METHOD>
  static constexpr bool two_phase() {
    return METHOD >= 0 && METHOD < SGX_METHOD_COUNT;
  }
  template <int METHOD>
  static constexpr bool from_e() {
    return METHOD == SGX_T_FROM_RHO_E || METHOD == SGX_P_FROM_RHO_E ||
           METHOD == SGX_CV_FROM_RHO_E || METHOD == SGX_BMOD_FROM_RHO_E ||
           METHOD == SGX_GAMMA_FROM_RHO_E;
  }
  template <int METHOD>
  SGX_D const double *table_of() const {
    if constexpr (METHOD == SGX_P_FROM_RHO_T) return P;
    if constexpr (METHOD == SGX_E_FROM_RHO_T) return sie;
    if constexpr (METHOD == SGX_CV_FROM_RHO_T) return dTdE;
    if constexpr (METHOD == SGX_BMOD_FROM_RHO_T) return bMod;
    if constexpr (METHOD == SGX_GAMMA_FROM_RHO_T) return dPdE;
    if constexpr (METHOD == SGX_T_FROM_RHO_E) return T_re;
    if constexpr (METHOD == SGX_P_FROM_RHO_E) return P_re;
    if constexpr (METHOD == SGX_CV_FROM_RHO_E) return dTdE_re;
    if constexpr (METHOD == SGX_BMOD_FROM_RHO_E) return bMod_re;
    return dPdE_re;
  }
  template <int METHOD>
  SGX_D void prep(const double rho, const double x, LookupPrep &q) const {
    q.x2 = to_log(rho, lRhoOffset);
    q.p2 = AxisPos(gRho, q.x2);
    if constexpr (from_e<METHOD>()) {
      q.x1 = to_log(x, lEOffset);
      q.p1 = AxisPos(gE, q.x1);
      q.cell = ld_cell(table_of<METHOD>(), numE, q.p2.ix, q.p1.ix);
    } else {
      q.x1 = to_log(x, lTOffset);
      q.p1 = AxisPos(gT, q.x1);
      q.cell = ld_cell(table_of<METHOD>(), numT, q.p2.ix, q.p1.ix);
    }
  }
  template <int METHOD>
  SGX_D double finish(const double rho, const double, const LookupPrep &q,
                      double *lambda) const {
    // every method but GruneisenParamFromDensityInternalEnergy caches lRho (:915-939,757-761)
    if (METHOD != SGX_GAMMA_FROM_RHO_E && lambda) lambda[0] = q.x2;
    const double tab = blend_cell(q.cell, q.p2.w0, q.p2.w1, q.p1.w0, q.p1.w1);
    if constexpr (METHOD == SGX_CV_FROM_RHO_T || METHOD == SGX_CV_FROM_RHO_E) return 1. / tab;
    if constexpr (METHOD == SGX_GAMMA_FROM_RHO_T || METHOD == SGX_GAMMA_FROM_RHO_E)
      return tab / rho;
    return tab;
  }
  // lRhoFromPlT_ (eos_spiner_rho_sie.hpp:941-986)
  __device__ inline double lRhoFromPlT(const double Pin, const double lT,
                                       double *lambda) const {
    const double rhopmin = to_log(interp1d(rho_at_pmin, gT, lT), lRhoOffset);
    double lRho;
    const double dPdRhoMax_ = interp1d(dPdRhoMax, gT, lT);
    const double PMax_ = interp1d(PlRhoMax, gT, lT);
    if (dPdRhoMax_ > 0 && Pin > PMax_) {
      const double rho = (Pin - PMax_) / dPdRhoMax_ + rhoMax;
      lRho = to_log(rho, lRhoOffset);
    } else {
      double lRhoGuess = reproducible ? lRhoMax : 0.5 * (rhopmin + lRhoMax);
      if (lambda) {
        const double lRho_cache = lambda[0];
        if ((rhopmin <= lRho_cache) && (lRho_cache <= lRhoMax)) lRhoGuess = lRho_cache;
      }
      const AxisPos pt(gT, lT);
      const double *tab = P;
      const int n1 = numT;
      const Grid1D &g2 = gRho;
      auto PFunc = [&](const double x) { return interp2c(tab, n1, AxisPos(g2, x), pt); };
      const int status =
          root::regula_falsi(PFunc, Pin, lRhoGuess, rhopmin, lRhoMax, kEps, kEps, lRho);
      if (status != root::SUCCESS) lRho = reproducible ? lRhoMax : lRhoGuess;
    }
    if (lambda) lambda[0] = lRho;
    return lRho;
  }
  // DensityEnergyFromPressureTemperature (eos_spiner_rho_sie.hpp:836-848)
  __device__ inline void rho_e_from_P_T(const double press, const double temp, double *lambda,
                                        double &rho, double &sie_out) const {
    const double lT = to_log(temp, lTOffset);
    const double lRho = lRhoFromPlT(press, lT, lambda);
    rho = from_log(lRho, lRhoOffset);
    sie_out = interp2c(sie, gRho, gT, numT, lRho, lT);
  }
  // bounds the PTE solvers ask for (eos_spiner_rho_sie.hpp:242-270)
  SGX_D double RhoPmin(const double temp) const {
    return interp1d(rho_at_pmin, gT, to_log(temp, lTOffset));
  }
  SGX_D double MinimumDensity() const { return from_log(lRhoMin, lRhoOffset); }
  SGX_D double MinimumTemperature() const { return from_log(gT.gmin(), lTOffset); }
  SGX_D double MaximumDensity() const { return rhoMax; }
  SGX_D double MinimumPressure() const { return PMin; }
  SGX_D double MaximumPressureAtTemperature(double) const { return 1e100; }
  // FillEos (eos_spiner_rho_sie.hpp:842-900)
  __device__ inline bool fill(double &rho, double &temp, double &energy, double &press,
                              double &cv, double &bmod, const unsigned long output,
                              double *lambda) const {
    double lRho, lT, lE;
    if (output == 0) return false;
    if ((output & TQ_TEMPERATURE) && (output & TQ_SIE)) return false;
    if (output & TQ_DENSITY) {
      if (!((output & TQ_PRESSURE) || (output & TQ_TEMPERATURE))) {
        lT = to_log(temp, lTOffset);
        lRho = lRhoFromPlT(press, lT, lambda);
        rho = from_log(lRho, lRhoOffset);
      } else {
        return false;
      }
    } else {
      lRho = to_log(rho, lRhoOffset);
      if (lambda) lambda[0] = lRho;
    }
    const AxisPos pr(gRho, lRho);
    if (output & TQ_TEMPERATURE) {
      lE = to_log(energy, lEOffset);
      const AxisPos pe(gE, lE);
      temp = interp2c(T_re, numE, pr, pe);
      if (output & TQ_PRESSURE) press = interp2c(P_re, numE, pr, pe);
      if (output & TQ_CV) cv = 1. / interp2c(dTdE_re, numE, pr, pe);
      if (output & TQ_BMOD) bmod = interp2c(bMod_re, numE, pr, pe);
    }
    if (output & TQ_SIE) {
      lT = to_log(temp, lTOffset);
      const AxisPos pt(gT, lT);
      energy = interp2c(sie, numT, pr, pt);
      if (output & TQ_PRESSURE) press = interp2c(P, numT, pr, pt);
      if (output & TQ_CV) cv = 1. / interp2c(dTdE, numT, pr, pt);
      if (output & TQ_BMOD) bmod = interp2c(bMod, numT, pr, pt);
    }
    return true;
  }
#endif
};

} // namespace sgx
#endif // SGX_SPINER_CUH_
