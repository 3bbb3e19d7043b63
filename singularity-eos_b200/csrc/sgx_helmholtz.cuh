// Helmholtz EOS on device (Timmes & Swesty free-energy table + radiation + ideal ions +
// Coulomb corrections).  One thread evaluates one cell; the 21 tables are repacked at create
// time into one CELL-MAJOR record per (jT, iRho) table cell -- the 36 biquintic values
// followed by the 3 x 16 bicubic values, each group in the reference's own `fi[]` order -- so
// a lookup reads 21 contiguous, 32-byte-aligned sectors instead of 84 scattered doubles.
//
// Reference: singularity-eos/eos/eos_helmholtz.hpp
//   :744-758   GetElectronDensities_, lTAnalytic_
//   :723-742   ComputeGamma1_/ComputeGamma3_
//   :818-865   FillEos                      :867-960   lTFromRhoSie_
//   :962-989   InitDataFile_                :997-1278  HelmElectrons::GetFromDensityTemperature
//   :1280-1320 HelmRad                      :1322-1367 HelmIon
//   :1369-1504 HelmCoulomb (+ butterworth :395-401)
//   :1506-1572 GetFromDensityLogTemperature_
// base/hermite.hpp:24-92 (psi*/h5/h3), base/math_utils.hpp:57-60 (pow10 = exp(ln10 x)),
// base/root-finding-1d/root_finding.hpp:246-310 (newton_raphson).
// Every expression keeps the reference's association; the library is compiled with
// -fmad=false.  libm calls (log, log10, exp, sqrt, cbrt, pow) map to CUDA's (<= 2 ulp).
#ifndef SGX_HELMHOLTZ_CUH_
#define SGX_HELMHOLTZ_CUH_
#include "sgx_analytic.cuh"
#include "sgx_math.cuh"

namespace sgx {

namespace helm {
constexpr int NTEMP = 101, NRHO = 271;
constexpr int NCELL_T = NTEMP - 1, NCELL_R = NRHO - 1;
constexpr int REC = 84;  // doubles per packed cell record: 36 (f) + 16 (dpdf) + 16 (ef) + 16 (xf)
constexpr double lTMin = 3, lTMax = 13, lRhoMin = -12, lRhoMax = 15;
constexpr double TMin = 1e3, TMax = 1e13, rhoMin = 1.0 / 1e12, rhoMax = 1e15;
enum { VAL = 0, DDR = 1, DDT = 2, DDA = 3, DDZ = 4, NDERIV = 5 };
// HelmRad / HelmIon / HelmCoulomb constants (:291-296, :309-318, :336-342)
constexpr double STEFAN_BOLTZMANN_CONSTANT = 5.67040047374e-5;
constexpr double CL = 2.99792458e10;
constexpr double SIG_O_C = STEFAN_BOLTZMANN_CONSTANT / CL;
constexpr double PREFACTOR = (4. / 3.) * SIG_O_C;
constexpr double KB = 1.3806504e-16;
constexpr double NA = 6.02214199e23;
constexpr double KBNA = KB * NA;
constexpr double KBi = 1.0 / KB;
constexpr double LSWOT15 = 46.682612633059801;
constexpr double LN10 = 2.30258509299405e+00;
constexpr double ELECTRON_CHARGE_ESU = 4.80320427e-10;
constexpr double PI = 3.14159265358979323846; // M_PI
constexpr double E_ = 2.7182818284590452354;  // M_E
constexpr double ROOT_THRESH = 1e-14;
constexpr double HELM_EOS_EPS = 1e-10;
} // namespace helm

struct HelmDev {
  const double *cells = nullptr; // [NCELL_T * NCELL_R][REC]
  const double *rho_ = nullptr;  // [NRHO]  pow10(lRho grid)
  const double *T_ = nullptr;    // [NTEMP] pow10(lT grid)
  // RegularGrid1D(min, max, N) of the two table axes (regular_grid_1d.hpp:60-63)
  double lR_min, lR_dx, lR_idx, lT_min, lT_dx, lT_idx;
  double log10e; // std::log10(M_E), evaluated once on the host with the same libm
  int enable_rad, enable_gas, enable_coulomb, gas_ionized, gas_degenerate;
  int use_newton_raphson, lambda_has_lT_slot;

#ifdef __CUDACC__
  SGX_D static double pow10(const double x) { return xexp(helm::LN10 * x); }
  // hermite basis (base/hermite.hpp:24-56, 74-82)
  SGX_D static double psi0(double z) { return z * z * z * (z * (-6.0 * z + 15.0) - 10.0) + 1.0; }
  SGX_D static double dpsi0(double z) { return z * z * (z * (-30.0 * z + 60.0) - 30.0); }
  SGX_D static double ddpsi0(double z) { return z * (z * (-120.0 * z + 180.0) - 60.0); }
  SGX_D static double psi1(double z) { return z * (z * z * (z * (-3.0 * z + 8.0) - 6.0) + 1.0); }
  SGX_D static double dpsi1(double z) { return z * z * (z * (-15.0 * z + 32.0) - 18.0) + 1.0; }
  SGX_D static double ddpsi1(double z) { return z * (z * (-60.0 * z + 96.0) - 36.0); }
  SGX_D static double psi2(double z) { return 0.5 * z * z * (z * (z * (-z + 3.0) - 3.0) + 1.0); }
  SGX_D static double dpsi2(double z) {
    return 0.5 * z * (z * (z * (-5.0 * z + 12.0) - 9.0) + 2.0);
  }
  SGX_D static double ddpsi2(double z) {
    return 0.5 * (z * (z * (-20.0 * z + 36.0) - 18.0) + 2.0);
  }
  SGX_D static double xpsi0(double z) { return z * z * (2.0 * z - 3.0) + 1.0; }
  SGX_D static double xdpsi0(double z) { return z * (6.0 * z - 6.0); }
  SGX_D static double xpsi1(double z) { return z * (z * (z - 2.0) + 1.0); }
  SGX_D static double xdpsi1(double z) { return z * (3.0 * z - 4.0) + 1.0; }

  // biquintic / bicubic sums, left-to-right like hermite::h5 / h3 (base/hermite.hpp:59-71,84-91)
  SGX_D static double h5(const double *fi, double w0t, double w1t, double w2t, double w0mt,
                         double w1mt, double w2mt, double w0d, double w1d, double w2d,
                         double w0md, double w1md, double w2md) {
    return fi[0] * w0d * w0t + fi[1] * w0md * w0t + fi[2] * w0d * w0mt + fi[3] * w0md * w0mt +
           fi[4] * w0d * w1t + fi[5] * w0md * w1t + fi[6] * w0d * w1mt + fi[7] * w0md * w1mt +
           fi[8] * w0d * w2t + fi[9] * w0md * w2t + fi[10] * w0d * w2mt + fi[11] * w0md * w2mt +
           fi[12] * w1d * w0t + fi[13] * w1md * w0t + fi[14] * w1d * w0mt +
           fi[15] * w1md * w0mt + fi[16] * w2d * w0t + fi[17] * w2md * w0t +
           fi[18] * w2d * w0mt + fi[19] * w2md * w0mt + fi[20] * w1d * w1t +
           fi[21] * w1md * w1t + fi[22] * w1d * w1mt + fi[23] * w1md * w1mt +
           fi[24] * w2d * w1t + fi[25] * w2md * w1t + fi[26] * w2d * w1mt +
           fi[27] * w2md * w1mt + fi[28] * w1d * w2t + fi[29] * w1md * w2t +
           fi[30] * w1d * w2mt + fi[31] * w1md * w2mt + fi[32] * w2d * w2t +
           fi[33] * w2md * w2t + fi[34] * w2d * w2mt + fi[35] * w2md * w2mt;
  }
  SGX_D static double h3(const double *fi, double w0t, double w1t, double w0mt, double w1mt,
                         double w0d, double w1d, double w0md, double w1md) {
    return fi[0] * w0d * w0t + fi[1] * w0md * w0t + fi[2] * w0d * w0mt + fi[3] * w0md * w0mt +
           fi[4] * w0d * w1t + fi[5] * w0md * w1t + fi[6] * w0d * w1mt + fi[7] * w0md * w1mt +
           fi[8] * w1d * w0t + fi[9] * w1md * w0t + fi[10] * w1d * w0mt + fi[11] * w1md * w0mt +
           fi[12] * w1d * w1t + fi[13] * w1md * w1t + fi[14] * w1d * w1mt + fi[15] * w1md * w1mt;
  }

  // 16-byte read-only loads of `n` doubles (n even) from a packed cell record
  template <int N>
  SGX_D static void load_rec(const double *__restrict__ p, double *fi) {
#pragma unroll
    for (int k = 0; k < N; k += 2) {
      const double2 v = __ldg(reinterpret_cast<const double2 *>(p + k));
      fi[k] = v.x;
      fi[k + 1] = v.y;
    }
  }

  // RegularGrid1D::index (regular_grid_1d.hpp:66-74)
  SGX_D static int grid_index(const double x, const double mn, const double idx, const int n) {
    int ix = cvt_int_x86(idx * (x - mn));
    if (ix < 0) ix = 0;
    if (ix >= n - 1) ix = n - 2;
    return ix;
  }

  // HelmElectrons::GetFromDensityTemperature (:997-1278)
  __device__ inline void electrons(double rho, const double lT, const double Ye,
                                   const double Ytot, double De, double lDe, double *pele,
                                   double *eele, double *sele, double *etaele, double *xne,
                                   const bool only_e) const {
    using namespace helm;
    rho = stdmin(rhoMax, stdmax(rhoMin, rho));
    De = stdmin(rhoMax, stdmax(rhoMin, De));
    lDe = stdmin(lRhoMax, stdmax(lRhoMin, lDe));
    const double T = pow10(lT);
    const int iat = grid_index(lDe, lR_min, lR_idx, NRHO);
    const int jat = grid_index(lT, lT_min, lT_idx, NTEMP);

    const double Tj = ldg(T_ + jat);
    const double dth = ldg(T_ + jat + 1) - Tj;
    const double dt2 = dth * dth;
    const double dti = ratio(1.0, dth);
    const double dt2i = dti * dti;
    const double ri = ldg(rho_ + iat);
    const double dd = ldg(rho_ + iat + 1) - ri;
    const double dd2 = dd * dd;
    const double ddi = ratio(1.0, dd);

    const double *rec = cells + (size_t)(jat * NCELL_R + iat) * REC;
    double fi[36];
    load_rec<36>(rec, fi);

    const double xt = stdmax(0.0, (T - Tj) * dti);
    const double xd = stdmax(0.0, (De - ri) * ddi);
    const double mxt = 1. - xt;
    const double mxd = 1. - xd;

    double si0t = psi0(xt);
    double si1t = psi1(xt) * dth;
    const double si2t = psi2(xt) * dt2;
    double si0mt = psi0(mxt);
    double si1mt = -psi1(mxt) * dth;
    const double si2mt = psi2(mxt) * dt2;
    double si0d = psi0(xd);
    double si1d = psi1(xd) * dd;
    const double si2d = psi2(xd) * dd2;
    double si0md = psi0(mxd);
    double si1md = -psi1(mxd) * dd;
    const double si2md = psi2(mxd) * dd2;

    double dsi0t = dpsi0(xt) * dti;
    double dsi1t = dpsi1(xt);
    const double dsi2t = dpsi2(xt) * dth;
    double dsi0mt = -dpsi0(mxt) * dti;
    double dsi1mt = dpsi1(mxt);
    const double dsi2mt = -dpsi2(mxt) * dth;
    double dsi0d = dpsi0(xd) * ddi;
    double dsi1d = dpsi1(xd);
    const double dsi2d = dpsi2(xd) * dd;
    double dsi0md = -dpsi0(mxd) * ddi;
    double dsi1md = dpsi1(mxd);
    const double dsi2md = -dpsi2(mxd) * dd;

    const double ddsi0t = ddpsi0(xt) * dt2i;
    const double ddsi1t = ddpsi1(xt) * dti;
    const double ddsi2t = ddpsi2(xt);
    const double ddsi0mt = ddpsi0(mxt) * dt2i;
    const double ddsi1mt = -ddpsi1(mxt) * dti;
    const double ddsi2mt = ddpsi2(mxt);

    const double free_en =
        h5(fi, si0t, si1t, si2t, si0mt, si1mt, si2mt, si0d, si1d, si2d, si0md, si1md, si2md);
    const double df_t = h5(fi, dsi0t, dsi1t, dsi2t, dsi0mt, dsi1mt, dsi2mt, si0d, si1d, si2d,
                           si0md, si1md, si2md);
    const double df_tt = h5(fi, ddsi0t, ddsi1t, ddsi2t, ddsi0mt, ddsi1mt, ddsi2mt, si0d, si1d,
                            si2d, si0md, si1md, si2md);
    if (only_e) {
      eele[0] = Ye * free_en + T * (-df_t) * Ye;
      eele[2] = T * (-df_tt) * Ye;
      return;
    }
    const double df_d = h5(fi, si0t, si1t, si2t, si0mt, si1mt, si2mt, dsi0d, dsi1d, dsi2d,
                           dsi0md, dsi1md, dsi2md);
    const double df_dt = h5(fi, dsi0t, dsi1t, dsi2t, dsi0mt, dsi1mt, dsi2mt, dsi0d, dsi1d,
                            dsi2d, dsi0md, dsi1md, dsi2md);

    si0t = xpsi0(xt);
    si1t = xpsi1(xt) * dth;
    si0mt = xpsi0(mxt);
    si1mt = -xpsi1(mxt) * dth;
    si0d = xpsi0(xd);
    si1d = xpsi1(xd) * dd;
    si0md = xpsi0(mxd);
    si1md = -xpsi1(mxd) * dd;
    dsi0t = xdpsi0(xt) * dti;
    dsi1t = xdpsi1(xt);
    dsi0mt = -xdpsi0(mxt) * dti;
    dsi1mt = xdpsi1(mxt);
    dsi0d = xdpsi0(xd) * ddi;
    dsi1d = xdpsi1(xd);
    dsi0md = -xdpsi0(mxd) * ddi;
    dsi1md = xdpsi1(mxd);

    // dpdf block
    load_rec<16>(rec + 36, fi);
    pele[1] = stdmax(0.0, Ye * h3(fi, si0t, si1t, si0mt, si1mt, si0d, si1d, si0md, si1md));
    // ef block
    load_rec<16>(rec + 52, fi);
    etaele[0] = h3(fi, si0t, si1t, si0mt, si1mt, si0d, si1d, si0md, si1md);
    double x = h3(fi, si0t, si1t, si0mt, si1mt, dsi0d, dsi1d, dsi0md, dsi1md);
    etaele[1] = Ye * x;
    etaele[2] = h3(fi, dsi0t, dsi1t, dsi0mt, dsi1mt, si0d, si1d, si0md, si1md);
    etaele[3] = -x * De * Ytot;
    etaele[4] = x * rho * Ytot;
    // xf block
    load_rec<16>(rec + 68, fi);
    xne[0] = h3(fi, si0t, si1t, si0mt, si1mt, si0d, si1d, si0md, si1md);
    x = stdmax(0.0, h3(fi, si0t, si1t, si0mt, si1mt, dsi0d, dsi1d, dsi0md, dsi1md));
    xne[1] = Ye * x;
    xne[2] = h3(fi, dsi0t, dsi1t, dsi0mt, dsi1mt, si0d, si1d, si0md, si1md);
    xne[3] = -x * De * Ytot;
    xne[4] = x * rho * Ytot;

    x = De * De;
    pele[0] = x * df_d;
    pele[2] = x * df_dt;
    const double s = pele[1] / Ye - 2.0 * De * df_d;
    pele[3] = -Ytot * (2.0 * pele[0] + s * De);
    pele[4] = rho * Ytot * (2.0 * De * df_d + s);

    x = Ye * Ye;
    sele[0] = -df_t * Ye;
    sele[2] = -df_tt * Ye;
    sele[1] = -df_dt * x;
    sele[3] = Ytot * (Ye * df_dt * De - sele[0]);
    sele[4] = -Ytot * (Ye * df_dt * rho + df_t);

    eele[0] = Ye * free_en + T * sele[0];
    eele[2] = T * sele[2];
    eele[1] = x * df_d + T * sele[1];
    eele[3] = -Ye * Ytot * (free_en + df_d * De) + T * sele[3];
    eele[4] = Ytot * (free_en + Ye * df_d * rho) + T * sele[4];
  }

  // HelmRad::GetFromDensityTemperature (:1280-1320)
  SGX_D static void rad(const double rho, const double temp, double *prad, double *erad,
                        double *srad) {
    using namespace helm;
    const double rhoi = ratio(1.0, rho);
    const double tempi = ratio(1.0, temp);
    const double T3 = pow3i(temp);
    const double T4 = T3 * temp;
    const double P = PREFACTOR * T4;
    const double dPdT = 4.0 * PREFACTOR * T3;
    prad[VAL] = P;
    prad[DDR] = 0;
    prad[DDT] = dPdT;
    prad[DDA] = 0;
    prad[DDZ] = 0;
    const double PoR = P * rhoi;
    const double e = 3.0 * PoR;
    const double dedR = -e * rhoi;
    const double dedT = 3.0 * dPdT * rhoi;
    erad[VAL] = e;
    erad[DDR] = dedR;
    erad[DDT] = dedT;
    erad[DDA] = 0;
    erad[DDZ] = 0;
    constexpr double dPdR = 0;
    const double dEdR = erad[DDR];
    const double s = (PoR + e) * tempi;
    const double dsdR = ((dPdR - PoR) * rhoi + dEdR) * tempi;
    const double dsdT = (dPdT * rhoi + dedT - s) * tempi;
    srad[VAL] = s;
    srad[DDR] = dsdR;
    srad[DDT] = dsdT;
    srad[DDA] = 0;
    srad[DDZ] = 0;
  }

  // HelmIon::GetFromDensityTemperature (:1322-1367)
  SGX_D static void ion(const double rho, const double temp, const double y, const double ytot,
                        const double xni, const double dxnidd, const double dxnida,
                        double *pion, double *eion, double *sion) {
    using namespace helm;
    const double kbT = KB * temp;
    const double P = xni * kbT;
    const double dPdR = dxnidd * kbT;
    const double dPdT = xni * KB;
    const double dPdA = dxnida * kbT;
    pion[VAL] = P;
    pion[DDR] = dPdR;
    pion[DDT] = dPdT;
    pion[DDA] = dPdA;
    pion[DDZ] = 0;
    const double rhoi = ratio(1.0, rho);
    const double PoR = P * rhoi;
    const double e = 1.5 * PoR;
    const double dedR = (1.5 * dPdR - e) * rhoi;
    const double dedT = 1.5 * dPdT * rhoi;
    const double dedA = 1.5 * dPdA * rhoi;
    eion[VAL] = e;
    eion[DDR] = dedR;
    eion[DDT] = dedT;
    eion[DDA] = dedA;
    eion[DDZ] = 0;
    const double tempi = ratio(1.0, temp);
    const double KBNAY = KBNA * ytot;
    const double s = (PoR + e) * tempi + KBNAY * y;
    const double dsdR = (dPdR * rhoi - P * rhoi * rhoi + dedR) * tempi - KBNAY * rhoi;
    const double dsdT =
        (dPdT * rhoi + dedT) * tempi - (PoR + e) * tempi * tempi + 1.5 * KBNAY * tempi;
    const double dsdA = (dPdA * rhoi + dedA) * tempi + KBNAY * ytot * (2.5 - y);
    sion[VAL] = s;
    sion[DDR] = dsdR;
    sion[DDT] = dsdT;
    sion[DDA] = dsdA;
    sion[DDZ] = 0;
  }

  // butterworth<12> (:395-401): pow<24> = (pow<12>)^2, pow<12> = (pow<6>)^2, pow<6> =
  // (pow<3>)^2, pow<3> = x * pow<2>; pow<23> = x * pow<22>, pow<22> = (pow<11>)^2,
  // pow<11> = x * pow<10>, pow<10> = (pow<5>)^2, pow<5> = x * pow<4> (math_utils.hpp:32-49)
  SGX_D static double pow24(const double x) {
    const double p3 = x * (x * x);
    const double p6 = p3 * p3;
    const double p12 = p6 * p6;
    return p12 * p12;
  }
  SGX_D static double pow23(const double x) {
    const double p2 = x * x;
    const double p4 = p2 * p2;
    const double p5 = x * p4;
    const double p10 = p5 * p5;
    const double p11 = x * p10;
    const double p22 = p11 * p11;
    return x * p22;
  }
  SGX_D static void butterworth12(const double freq, const double cfreq, double &g,
                                  double &dgdf) {
    g = ratio(1.0, 1.0 + pow24(ratio(freq, cfreq)));
    dgdf = ratio(-pow2i(g) * 2 * 12, pow24(cfreq)) * pow23(freq);
  }

  // HelmCoulomb::GetFromDensityTemperature (:1369-1504)
  __device__ inline static void coulomb(const double rho, const double temp, const double ytot,
                                        const double abar, const double zbar, const double xni,
                                        const double dxnidd, const double dxnida,
                                        double *pcoul, double *ecoul, double *scoul) {
    using namespace helm;
    constexpr double a1 = -0.898004, b1 = 0.96786, c1 = 0.220703, d1 = -0.86097, e1 = 2.5269;
    constexpr double a2 = 0.29561, b2 = 1.9885, c2 = 0.288675;
    const double kbT = KB * temp;
    const double pion = xni * kbT;
    const double dpiondd = dxnidd * kbT;
    const double dpiondt = xni * KB;
    const double dpionda = dxnida * kbT;
    constexpr double dpiondz = 0.0;
    constexpr double four_thirds = 4.0 / 3.0;
    constexpr double ftpi = four_thirds * PI;
    double s = ftpi * xni;
    const double dsdd = ftpi * dxnidd;
    const double dsda = ftpi * dxnida;
    constexpr double one_third = 1.0 / 3.0;
    const double si = ratio(1.0, s);
    const double lami = ratio(1.0, xcbrt(s));
    const double lamidd = -lami * one_third * dsdd * si;
    const double lamida = -lami * one_third * dsda * si;
    const double plasg = pow2i(ELECTRON_CHARGE_ESU * zbar) / (kbT * lami);
    const double plasg_o_lami = ratio(plasg, lami);
    const double plasgdd = -plasg_o_lami * lamidd;
    const double plasgda = -plasg_o_lami * lamida;
    const double plasgdt = -ratio(plasg, temp);
    const double plasgdz = 2.0 * ratio(plasg, zbar);
    const double abari = ratio(1, abar);
    if (plasg >= 0) {
      const double x = sqrt(sqrt(plasg));
      double y = KBNA * ytot;
      const double c1_o_x = ratio(c1, x);
      ecoul[VAL] = y * temp * (a1 * plasg + b1 * x + c1_o_x + d1);
      pcoul[VAL] = one_third * rho * ecoul[0];
      scoul[VAL] = -y * (3.0 * b1 * x - 5.0 * c1_o_x + d1 * (xlog(plasg) - 1.0) - e1);
      y = KBNA * temp * ytot * (a1 + ratio(0.25, plasg) * (b1 * x - c1_o_x));
      ecoul[DDR] = y * plasgdd;
      ecoul[DDT] = y * plasgdt + ratio(ecoul[0], temp);
      ecoul[DDA] = y * plasgda - ecoul[0] * abari;
      ecoul[DDZ] = y * plasgdz;
      y = one_third * rho;
      pcoul[DDR] = one_third * ecoul[0] + y * ecoul[1];
      pcoul[DDT] = y * ecoul[2];
      pcoul[DDA] = y * ecoul[3];
      pcoul[DDZ] = y * ecoul[4];
      y = -KBNA * plasg * abari * (0.75 * b1 * x + 1.25 * c1_o_x + d1);
      scoul[DDR] = y * plasgdd;
      scoul[DDT] = y * plasgdt;
      scoul[DDA] = y * plasgda - scoul[0] * abari;
      scoul[DDZ] = y * plasgdz;
    } else {
      const double x = plasg * sqrt(fabs(plasg));
      const double y = xpow(plasg, b2);
      const double z = c2 * x - one_third * a2 * y;
      const double plasgi = ratio(1., plasg);
      const double rhoi = ratio(1.0, rho);
      pcoul[VAL] = -pion * z;
      ecoul[VAL] = 3.0 * pcoul[0] * rhoi;
      scoul[VAL] = -KBNA * abari * (c2 * x - a2 * ratio(b2 - 1.0, b2) * y);
      s = 1.5 * c2 * x * plasgi - one_third * a2 * b2 * y * plasgi;
      pcoul[DDR] = -dpiondd * z - pion * s * plasgdd;
      pcoul[DDT] = -dpiondt * z - pion * s * plasgdt;
      pcoul[DDA] = -dpionda * z - pion * s * plasgda;
      pcoul[DDZ] = -dpiondz * z - pion * s * plasgdz;
      s = 3.0 * rhoi;
      ecoul[DDR] = s * pcoul[1] - ecoul[0] * rhoi;
      ecoul[DDT] = s * pcoul[2];
      ecoul[DDA] = s * pcoul[3];
      ecoul[DDZ] = s * pcoul[4];
      s = -KBNA * abari * plasgi * (1.5 * c2 * x - a2 * (b2 - 1.0) * y);
      scoul[DDR] = s * plasgdd;
      scoul[DDT] = s * plasgdt;
      scoul[DDA] = s * plasgda - scoul[0] * abari;
      scoul[DDZ] = s * plasgdz;
    }
    double g_r, dgdf_r, g_t, dgdf_t;
    butterworth12(xlog10(temp) - 4.5, 3.0, g_r, dgdf_r);
    butterworth12(xlog10(rho) + 1.0, 6.0, g_t, dgdf_t);
    const double gain = (1.0 - g_t * g_r);
    const double dgaindt = -ratio(g_r * dgdf_t, temp * LN10);
    const double dgaindd = -ratio(g_t * dgdf_r, rho * LN10);
    pcoul[VAL] = pcoul[VAL] * gain;
    ecoul[VAL] = ecoul[VAL] * gain;
    scoul[VAL] = scoul[VAL] * gain;
    pcoul[DDR] = gain * pcoul[DDR] + ratio(pcoul[VAL] * dgaindd, gain);
    pcoul[DDT] = gain * pcoul[DDT] + ratio(pcoul[VAL] * dgaindt, gain);
    pcoul[DDA] = gain * pcoul[DDA];
    pcoul[DDZ] = gain * pcoul[DDZ];
    ecoul[DDR] = gain * ecoul[DDR] + ratio(ecoul[VAL] * dgaindd, gain);
    ecoul[DDT] = gain * ecoul[DDT] + ratio(ecoul[VAL] * dgaindt, gain);
    ecoul[DDA] = gain * ecoul[DDA];
    ecoul[DDZ] = gain * ecoul[DDZ];
    scoul[DDR] = gain * scoul[DDR] + ratio(scoul[VAL] * dgaindd, gain);
    scoul[DDT] = gain * scoul[DDT] + ratio(scoul[VAL] * dgaindt, gain);
    scoul[DDA] = gain * scoul[DDA];
    scoul[DDZ] = gain * scoul[DDZ];
  }

  struct Comp { // per-cell composition-derived scalars (GetElectronDensities_, :744-753)
    double abar, zbar, ytot, ye, ywot, De, lDe;
  };
  SGX_D static Comp electron_densities(const double rho, const double abar,
                                       const double zbar) {
    Comp c;
    c.abar = abar;
    c.zbar = zbar;
    c.ytot = ratio(1.0, abar);
    c.ye = zbar * c.ytot;
    c.ywot = xlog(ratio(abar * abar * sqrt(abar), rho * helm::NA));
    c.De = rho * c.ye;
    c.lDe = xlog10(c.De);
    return c;
  }
  SGX_D static void mass_fractions(const double rho, const double ytot, double &xni,
                                   double &dxnidd, double &dxnida) {
    xni = helm::NA * ytot * rho;
    dxnidd = helm::NA * ytot;
    dxnida = -xni * ytot;
  }

  // GetFromDensityLogTemperature_ (:1506-1572).  p/s may be null when only_e.
  __device__ inline void get(const double rho, const double T, const Comp &c, double *p,
                             double *e, double *s, double *etaele, double *nep,
                             const bool only_e) const {
    using namespace helm;
    double prad[NDERIV] = {0}, pion[NDERIV] = {0}, pele[NDERIV] = {0}, pcoul[NDERIV] = {0};
    double erad[NDERIV] = {0}, eion[NDERIV] = {0}, eele[NDERIV] = {0}, ecoul[NDERIV] = {0};
    double srad[NDERIV] = {0}, sion[NDERIV] = {0}, sele[NDERIV] = {0}, scoul[NDERIV] = {0};
#pragma unroll
    for (int i = 0; i < 5; ++i)
      etaele[i] = nep[i] = 0;
    const double lT = xlog10(T);
    const double lnT = lT / log10e;
    if (enable_rad) rad(rho, T, prad, erad, srad);
    if (enable_gas) {
      const bool do_unmodified_ions = (gas_ionized && gas_degenerate) || !gas_ionized;
      if (do_unmodified_ions) {
        double xni, dxnidd, dxnida;
        mass_fractions(rho, c.ytot, xni, dxnidd, dxnida);
        const double y = c.ywot + LSWOT15 * lnT;
        ion(rho, T, y, c.ytot, xni, dxnidd, dxnida, pion, eion, sion);
      } else {
        const double abar_ion = ratio(c.abar, c.zbar + 1);
        const double ytot_ion = ratio(1.0, abar_ion);
        const double ywot_ion = ratio(xlog(abar_ion * abar_ion * sqrt(abar_ion)), rho * NA);
        const double y_ion = ywot_ion + LSWOT15 * lnT;
        double xni, dxnidd, dxnida;
        mass_fractions(rho, ytot_ion, xni, dxnidd, dxnida);
        ion(rho, T, y_ion, ytot_ion, xni, dxnidd, dxnida, pion, eion, sion);
      }
      if (gas_degenerate) {
        electrons(rho, lT, c.ye, c.ytot, c.De, c.lDe, pele, eele, sele, etaele, nep, only_e);
        if (enable_coulomb) {
          double xni, dxnidd, dxnida;
          mass_fractions(rho, c.ytot, xni, dxnidd, dxnida);
          coulomb(rho, T, c.ytot, c.abar, c.zbar, xni, dxnidd, dxnida, pcoul, ecoul, scoul);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 5; ++i)
      e[i] = erad[i] + eion[i] + eele[i] + ecoul[i];
    if (!only_e) {
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        p[i] = prad[i] + pion[i] + pele[i] + pcoul[i];
        s[i] = srad[i] + sion[i] + sele[i] + scoul[i];
      }
    }
  }

  SGX_D static double lTAnalytic(const double rho, const double e, const double ni,
                                 const double ne) {
    return xlog10((2.0 / 3.0) * ratio(e * rho, ni + ne) * helm::KBi);
  }

  // lTFromRhoSie_ (:867-960).  `lambda` is the cell's lambda row (Abar, Zbar[, lT]).
  __device__ inline double lTFromRhoSie(const double rho, const double e, const Comp &c,
                                        const double *lambda, int *root_its = nullptr) const {
    using namespace helm;
    const double abari = ratio(1.0, c.abar);
    const double ni = abari * rho * NA;
    const double ne = c.zbar * ni;
    double lT = 0.0;
    double T = 0.0;
    if (enable_rad || gas_degenerate || enable_coulomb) {
      // SafeGet<LogTemperature>: reads lambda[2] when the slot exists; with a type-indexed
      // {Abar,Zbar} lambda the reference leaves lTguess uninitialised -- we take the
      // out-of-range branch (analytic guess), see DESIGN.md.
      double lTguess = lambda_has_lT_slot ? lambda[2] : -1.0;
      if (!((lTMin <= lTguess) && (lTguess <= lTMax))) {
        lTguess = lTAnalytic(rho, e, ni, ne);
        if (!((lTMin <= lTguess) && (lTguess <= lTMax))) lTguess = 8;
      }
      const double Tguess = pow10(lTguess);
      const double Ta = pow10(lTMin), Tb = pow10(lTMax);
      auto e_only = [&](const double Tx, double &val, double &der) {
        double ee[NDERIV], eta[NDERIV], nep[NDERIV];
        get(rho, Tx, c, nullptr, ee, nullptr, eta, nep, true);
        val = ee[VAL];
        der = ee[DDT];
      };
      auto e_val = [&](const double Tx) {
        double v, d;
        e_only(Tx, v, d);
        return v;
      };
      int status;
      if (use_newton_raphson) {
        status = root::newton_raphson(e_only, e, Tguess, Ta, Tb, HELM_EOS_EPS, T, false);
        if (status != root::SUCCESS) {
          status = root::regula_falsi(e_val, e, Tguess, Ta, Tb, ROOT_THRESH, ROOT_THRESH, T);
          if (status != root::SUCCESS) {
            lT = lTAnalytic(rho, e, ni, gas_ionized * ne);
            T = pow10(lT);
          }
        }
      } else {
        status = root::regula_falsi(e_val, e, Tguess, Ta, Tb, ROOT_THRESH, ROOT_THRESH, T);
        if (status != root::SUCCESS) {
          lT = lTAnalytic(rho, e, ni, gas_ionized * ne);
          T = pow10(lT);
        }
      }
    } else {
      lT = lTAnalytic(rho, e, ni, gas_ionized * ne);
      T = pow10(lT);
    }
    lT = xlog10(T);
    if (lT < lTMin) lT = lTMin;
    if (lT > lTMax) lT = lTMax;
    // :958 SafeGet (sic): with an lT slot the computed value is overwritten by the cache
    if (lambda_has_lT_slot) lT = lambda[2];
    return lT;
  }

  SGX_D static double gamma1(const double rho, const double T, const double *p,
                             const double *e) {
    const double chit = ratio(T, p[0]) * p[helm::DDT];
    const double chid = ratio(p[helm::DDR] * rho, p[0]);
    const double x = ratio(p[0] * chit, rho * T * e[helm::DDT]);
    return chit * x + chid;
  }
  SGX_D static double gamma3(const double rho, const double T, const double *p,
                             const double *e) {
    const double chit = ratio(T, p[0]) * p[helm::DDT];
    const double x = ratio(p[0] * chit, rho * T * e[helm::DDT]);
    return x + 1.0;
  }

  // FillEos (:818-865).  Returns false where the reference aborts.
  __device__ inline bool fill(double &rho, double &temp, double &energy, double &press,
                              double &cv, double &bmod, const unsigned long output,
                              double *lambda) const {
    using namespace helm;
    const bool need_temp = (output & TQ_TEMPERATURE);
    const bool need_sie = (output & TQ_SIE);
    const bool need_rho = (output & TQ_DENSITY);
    if (need_rho || (need_temp && need_sie) || lambda == nullptr) return false;
    const Comp c = electron_densities(rho, lambda[0], lambda[1]);
    double lT;
    if (need_temp) {
      lT = lTFromRhoSie(rho, energy, c, lambda);
      temp = pow10(lT);
    } else {
      lT = xlog10(temp);
      if (lambda_has_lT_slot) lambda[2] = lT;
    }
    double p[NDERIV], e[NDERIV], s[NDERIV], etaele[NDERIV], nep[NDERIV];
    get(rho, temp, c, p, e, s, etaele, nep, false);
    if (output & TQ_SIE) energy = e[0];
    if (output & TQ_PRESSURE) press = p[0];
    if (output & TQ_CV) cv = e[DDT];
    if (output & TQ_BMOD) bmod = stdmax(kEps, p[0] * gamma1(rho, temp, p, e));
    return true;
  }
  // GruneisenParamFromDensity{Temperature,InternalEnergy} (:632-661)
  __device__ inline double gruneisen(const double rho, const double in2, const bool from_e,
                                     double *lambda) const {
    using namespace helm;
    const Comp c = electron_densities(rho, lambda[0], lambda[1]);
    double T = in2;
    if (from_e) {
      const double lT = lTFromRhoSie(rho, in2, c, lambda);
      T = pow10(lT);
    }
    double p[NDERIV], e[NDERIV], s[NDERIV], etaele[NDERIV], nep[NDERIV];
    get(rho, T, c, p, e, s, etaele, nep, false);
    return gamma3(rho, T, p, e) - 1.0;
  }
#endif
};

} // namespace sgx
#endif
