// Host-side construction of the tabulated Spiner EOS: everything the reference does once at
// load time after reading the .sp5 databoxes, then one upload to the device.
//
// Reference:
//   SpinerEOSDependsRhoT::loadDataboxes_   eos/eos_spiner_rho_temp.hpp:545-687
//   fixBulkModulus_                        eos/eos_spiner_rho_temp.hpp:689-714
//   setlTColdCrit_                         eos/eos_spiner_rho_temp.hpp:716-775
//   SetRhoPMin                             eos/eos_spiner_common.hpp:59-109
//   SpinerEOSDependsRhoSie::loadDataboxes_ eos/eos_spiner_rho_sie.hpp:434-577
//   calcBMod_                              eos/eos_spiner_rho_sie.hpp:579-601
// The arithmetic reuses the same host/device inline functions the kernels use
// (sgx_math.cuh), so derived tables equal the reference's to the last bit when both run on
// x86-64 without FMA contraction.
#include <algorithm>
#include <cmath>
#include <limits>

#include "sgx_internal.hpp"

namespace sgx {

namespace {
// Append the cell-major copy "<name>#cells" of the (already fixed-up) node-major table
// `name` [n2][n1]; kernels read the packed copy, sgx_eos_get_table the node-major one.
void add_packed(TableStore &ts, const std::string &name, const int n2, const int n1) {
  const size_t ncell = (size_t)(n2 - 1) * (size_t)(n1 - 1);
  const size_t off = ts.add(name + "#cells", nullptr, ncell * 4);
  pack_cells(ts.h(name), n2, n1, ts.host.data() + off);
}

constexpr double ROOM_TEMPERATURE = 293; // base/constants.hpp:42

Grid1D make_grid(const sgx_grid1d &g) {
  Grid1D o{};
  o.ngrids = g.ngrids;
  int tot = 0;
  for (int p = 0; p < 3; ++p) {
    if (p < g.ngrids) {
      o.npoints[p] = g.npoints[p];
      o.ptot[p] = tot;
      tot += g.npoints[p];
      o.min[p] = g.min[p];
      o.max[p] = g.max[p];
      o.dx[p] = g.dx[p];
      o.idx[p] = 1. / g.dx[p]; // RegularGrid1D::loadHDF: idx_ = 1./dx_
    } else {
      o.npoints[p] = 0;
      o.ptot[p] = tot;
      o.min[p] = o.max[p] = o.dx[p] = o.idx[p] = std::numeric_limits<double>::quiet_NaN();
    }
  }
  return o;
}

bool grid_ok(const sgx_grid1d &g) {
  if (g.ngrids < 1 || g.ngrids > 3) return false;
  for (int p = 0; p < g.ngrids; ++p)
    if (g.npoints[p] < 2 || !(g.dx[p] > 0) || !(g.max[p] > g.min[p])) return false;
  return true;
}

// SetRhoPMin, eos_spiner_common.hpp:59-109.  P is [NR][NT]; rho_at_pmin is [NT].
double SetRhoPMin(const double *P, const Grid1D &gT, const Grid1D &gRho, double *rho_at_pmin,
                  const bool pmin_vapor_dome, const double VAPOR_DPDR_THRESH,
                  const double lRhoOffset) {
  double PMin = std::numeric_limits<double>::max();
  const int NT = gT.n_total();
  const int NR = gRho.n_total();
  for (int i = 0; i < NT; ++i) {
    double PMin_at_T = std::numeric_limits<double>::max();
    int jmax = 0;
    for (int j = 0; j < NR; ++j) {
      const double Pji = P[(size_t)j * NT + i];
      if (Pji < PMin_at_T) {
        PMin_at_T = Pji;
        jmax = j;
      }
      if ((j > 1) && pmin_vapor_dome) {
        const double dP = Pji - P[(size_t)(j - 2) * NT + i];
        const double dr = from_log(gRho.x(j), lRhoOffset) - from_log(gRho.x(j - 2), lRhoOffset);
        const double dpdr = ratio(dP, dr);
        if (dpdr < VAPOR_DPDR_THRESH) jmax = j;
      }
    }
    if ((PMin_at_T > 0) && !pmin_vapor_dome) {
      rho_at_pmin[i] = 0;
      PMin_at_T = 0;
    } else {
      rho_at_pmin[i] = from_log(gRho.x(jmax), lRhoOffset);
    }
    PMin = std::min(PMin_at_T, PMin);
  }
  if (pmin_vapor_dome) {
    for (int i = NT - 2; i >= 0; i--) {
      if (rho_at_pmin[i] < rho_at_pmin[i + 1]) rho_at_pmin[i] = rho_at_pmin[i + 1];
    }
  }
  return PMin;
}

} // namespace

int build_spiner_rhot(const sgx_sp5_tables *raw, int reproducible, int pmin_vapor_dome,
                      sgx_eos_s *e) {
  if (!raw) return fail(SGX_ERR_INVALID, "null sgx_sp5_tables");
  if (!grid_ok(raw->lrho) || !grid_ok(raw->lT))
    return fail(SGX_ERR_INVALID, "malformed lrho/lT grid");
  const double *req[] = {raw->P,      raw->sie,   raw->bMod,    raw->dPdRho,  raw->dPdE,
                         raw->dTdRho, raw->dTdE,  raw->dEdRho,  raw->dEdT,    raw->PCold,
                         raw->sieCold, raw->bModCold, raw->dPdRhoCold};
  for (const double *p : req)
    if (!p) return fail(SGX_ERR_INVALID, "SpinerEOSDependsRhoT needs all (rho,T) and cold tables");

  SpinerRhoT &s = e->rhot;
  TableStore &ts = e->tables;
  s.gRho = make_grid(raw->lrho);
  s.gT = make_grid(raw->lT);
  const int numRho = s.gRho.n_total(), numT = s.gT.n_total();
  s.numRho = numRho;
  s.numT = numT;
  const size_t n2 = (size_t)numRho * numT;
  s.lRhoOffset = std::abs(raw->rho_offset);
  s.lTOffset = std::abs(raw->T_offset);
  double rhoNormal = std::abs(raw->normal_density);
  s.reproducible = reproducible ? 1 : 0;

  ts.add("P", raw->P, n2);
  ts.add("sie", raw->sie, n2);
  ts.add("bMod", raw->bMod, n2);
  ts.add("dPdRho", raw->dPdRho, n2);
  ts.add("dPdE", raw->dPdE, n2);
  ts.add("dTdRho", raw->dTdRho, n2);
  ts.add("dTdE", raw->dTdE, n2);
  ts.add("dEdRho", raw->dEdRho, n2);
  ts.add("dEdT", raw->dEdT, n2);
  ts.add("PCold", raw->PCold, numRho);
  ts.add("sieCold", raw->sieCold, numRho);
  ts.add("bModCold", raw->bModCold, numRho);
  ts.add("dPdRhoCold", raw->dPdRhoCold, numRho);
  for (const char *nm : {"dPdECold", "dTdRhoCold", "dTdECold", "dEdTCold", "lTColdCrit", "PMax",
                         "dEdTMax", "gm1Max", "sielTMax"})
    ts.add(nm, nullptr, numRho);
  ts.add("rho_at_pmin", nullptr, numT);

  // NB: take the pointers only after every add() (the vector may have been reallocated)
  double *P = ts.h("P"), *sie = ts.h("sie"), *bMod = ts.h("bMod"), *dPdRho = ts.h("dPdRho"),
         *dPdE = ts.h("dPdE"), *dTdRho = ts.h("dTdRho"), *dTdE = ts.h("dTdE"),
         *dEdRho = ts.h("dEdRho"), *dEdT = ts.h("dEdT");
  double *sieCold = ts.h("sieCold"), *bModCold = ts.h("bModCold");
  double *dPdECold = ts.h("dPdECold"), *dTdRhoCold = ts.h("dTdRhoCold"),
         *dTdECold = ts.h("dTdECold"), *dEdTCold = ts.h("dEdTCold"),
         *lTColdCrit = ts.h("lTColdCrit");
  double *PMax = ts.h("PMax"), *dEdTMax = ts.h("dEdTMax"), *gm1Max = ts.h("gm1Max"),
         *sielTMax = ts.h("sielTMax");

  // bounds (:607-613)
  s.lRhoMin = s.gRho.gmin();
  s.lRhoMax = s.gRho.gmax();
  s.rhoMax = from_log(s.lRhoMax, s.lRhoOffset);
  s.lTMin = s.gT.gmin();
  s.lTMax = s.gT.gmax();
  s.TMax = from_log(s.lTMax, s.lTOffset);
  auto rho_ = [&](const double lRho) {
    const double r = from_log(lRho, s.lRhoOffset);
    return r < 0 ? 0 : r;
  };
  auto I2 = [&](const double *tab, double lRho, double lT) {
    return interp2d(tab, s.gRho, s.gT, numT, lRho, lT);
  };

  // fixBulkModulus_ (:689-714)
  for (int j = 0; j < numRho; j++) {
    const double lRho = s.gRho.x(j);
    const double rho = rho_(lRho);
    for (int i = 0; i < numT; i++) {
      const double lT = s.gT.x(i);
      const double press = I2(P, lRho, lT);
      const double DPDR_E = I2(dPdRho, lRho, lT);
      const double DPDE_R = I2(dPdE, lRho, lT);
      const double DEDR_T = I2(dEdRho, lRho, lT);
      const double DPDR_T = DPDR_E + DPDE_R * DEDR_T;
      double bm;
      if (DPDE_R > 0.0 && rho > 0.0) {
        bm = rho * DPDR_E + DPDE_R * (press / rho);
      } else if (rho > 0.0) {
        bm = std::max(rho * DPDR_T, 0.0);
      } else {
        bm = 0.0;
      }
      bMod[(size_t)j * numT + i] = std::max(bm, std::abs(kEps));
    }
  }

  // setlTColdCrit_ (:716-775)
  for (int j = 0; j < numRho; j++) {
    const double lRho = s.gRho.x(j);
    const double sieColdj = interp1d(sieCold, s.gRho, lRho);
    std::vector<int> crossings;
    int last_pos_crossing = -1;
    for (int i = 0; i < numT - 1; i++) {
      const double sign_curr = sie[(size_t)j * numT + i] - sieColdj;
      const double sign_next = sie[(size_t)j * numT + i + 1] - sieColdj;
      if (sign_curr < 0 && sign_next > 0) {
        crossings.push_back(i);
        last_pos_crossing = (int)crossings.size() - 1;
      } else if (sign_curr > 0 && sign_next < 0) {
        crossings.push_back(i);
      }
    }
    if (last_pos_crossing <= 0) {
      lTColdCrit[j] = s.lTMin;
    } else {
      auto sieFunc = [&](const double x) { return I2(sie, lRho, x); };
      double lT;
      const int ilast = crossings[last_pos_crossing];
      const double lTlower = s.gT.x(ilast) - 1.0e-14;
      const double lTupper = s.gT.x(ilast + 1) + 1.0e-14;
      const double lTGuess = 0.5 * (lTlower + lTupper);
      const auto status =
          root::regula_falsi(sieFunc, sieColdj, lTGuess, lTlower, lTupper,
                             SpinerRhoT::ROOT_THRESH, SpinerRhoT::ROOT_THRESH, lT);
      if (status != root::SUCCESS) lT = lTGuess;
      lTColdCrit[j] = lT;
    }
    lTColdCrit[j] = std::max(s.lTMin, lTColdCrit[j]);
  }

  // minimum pressure as a function of temperature (:625)
  s.PMin = SetRhoPMin(P, s.gT, s.gRho, ts.h("rho_at_pmin"), pmin_vapor_dome != 0, 1e-8,
                      s.lRhoOffset);

  // cold-curve extrapolation tables (:638-649)
  for (int j = 0; j < numRho; j++) {
    const double lRho = s.gRho.x(j);
    const double lT = lTColdCrit[j];
    bModCold[j] = I2(bMod, lRho, lT);
    dPdECold[j] = I2(dPdE, lRho, lT);
    dTdRhoCold[j] = I2(dTdRho, lRho, lT);
    dTdECold[j] = I2(dTdE, lRho, lT);
    dEdTCold[j] = I2(dEdT, lRho, lT);
  }
  // top-of-table curves (:652-663)
  for (int j = 0; j < numRho; j++) {
    const double lRho = s.gRho.x(j);
    const double rho = rho_(lRho);
    PMax[j] = P[(size_t)j * numT + numT - 1];
    dEdTMax[j] = dEdT[(size_t)j * numT + numT - 1];
    gm1Max[j] = ratio(dPdE[(size_t)j * numT + numT - 1], rho);
    sielTMax[j] = sie[(size_t)j * numT + numT - 1];
  }
  // reference state (:665-684)
  double lRhoNormal = to_log(rhoNormal, s.lRhoOffset);
  if (!(s.lRhoMin < lRhoNormal && lRhoNormal < s.lRhoMax)) {
    lRhoNormal = 0.5 * (s.lRhoMin + s.lRhoMax);
    rhoNormal = rho_(lRhoNormal);
  }
  double TNormal = ROOM_TEMPERATURE;
  double lTNormal = to_log(TNormal, s.lTOffset);
  if (!(s.lTMin < lTNormal && lTNormal < s.lTMax)) {
    lTNormal = 0.5 * (s.lTMin + s.lTMax);
    TNormal = from_log(lTNormal, s.lTOffset);
  }
  auto &sc = ts.scalars;
  sc["rhoNormal"] = rhoNormal;
  sc["TNormal"] = TNormal;
  sc["sieNormal"] = I2(sie, lRhoNormal, lTNormal);
  sc["PNormal"] = I2(P, lRhoNormal, lTNormal);
  sc["CvNormal"] = I2(dEdT, lRhoNormal, lTNormal);
  sc["bModNormal"] = I2(bMod, lRhoNormal, lTNormal);
  sc["dPdENormal"] = I2(dPdE, lRhoNormal, lTNormal);
  const double dPdR = I2(dPdRho, lRhoNormal, lTNormal);
  sc["dVdTNormal"] = sc["dPdENormal"] * sc["CvNormal"] / (rhoNormal * rhoNormal * dPdR);
  sc["lRhoMin"] = s.lRhoMin;
  sc["lRhoMax"] = s.lRhoMax;
  sc["rhoMax"] = s.rhoMax;
  sc["lTMin"] = s.lTMin;
  sc["lTMax"] = s.lTMax;
  sc["TMax"] = s.TMax;
  sc["PMin"] = s.PMin;
  sc["lRhoOffset"] = s.lRhoOffset;
  sc["lTOffset"] = s.lTOffset;
  sc["meanAtomicMass"] = raw->mean_atomic_mass;
  sc["meanAtomicNumber"] = raw->mean_atomic_number;

  // CheckParams (:94-106)
  if (!(numRho > 0 && numT > 0 && s.lRhoMax > s.lRhoMin && s.rhoMax > 0 &&
        s.lTMax > s.lTMin && s.TMax > 0))
    return fail(SGX_ERR_INVALID, "SpinerEOSDependsRhoT::CheckParams failed");

  static const char *k2d[] = {"P", "sie", "bMod", "dPdRho", "dPdE", "dTdRho", "dTdE", "dEdRho",
                              "dEdT"};
  for (const char *nm : k2d) add_packed(ts, nm, numRho, numT);
  // Search hints of the direct T(rho, sie) inversion (sgx_spiner_root.cu; coordinate described at
  // RootRowMeta, sgx_internal.hpp).  A pair (j, j+1) of density rows qualifies when every
  // T-segment of both rows rises strictly (junction nodes of a piecewise grid are stored twice
  // and may repeat): the blend of the two rows is then strictly monotone in lT, the root unique.
  // For bin k the table holds the range [lo, hi] of table cells r with row(r) <= e < row(r+1)
  // for an energy of that bin, over both rows each in its own coordinate, widened by one cell
  // (the blended row sits between the two).  It is a hint: the kernel tests the table itself.
  {
    std::vector<char> junction(numT, 0);
    for (int g = 1; g < s.gT.ngrids; ++g) junction[s.gT.ptot[g] - 1] = 1;
    std::vector<char> rising(numRho, 1);
    const double *sie_h = ts.h("sie");
    for (int j = 0; j < numRho; ++j)
      for (int i = 0; i + 1 < numT; ++i) {
        const double a = sie_h[(size_t)j * numT + i], b = sie_h[(size_t)j * numT + i + 1];
        if (!(b > a || (junction[i] && b == a))) rising[j] = 0;
      }
    const int npair = numRho > 1 ? numRho - 1 : 1;
    std::vector<RootRowMeta> meta(numRho, RootRowMeta{0.0, 1.0, 0.0, 0.0});
    std::vector<unsigned int> hint((size_t)npair * kRootHintBins, 1u); // lo = 1, hi = 0: empty
    const bool fits = numT <= 65535 && numT >= 3 &&
                      (size_t)numRho * (sizeof(RootRowMeta) + 2 * sizeof(double)) <= 96 * 1024;
    for (int j = 0; j < numRho; ++j) {
      meta[j].base = sie_h[(size_t)j * numT];
      meta[j].s = sie_h[(size_t)j * numT + 1] - sie_h[(size_t)j * numT];
    }
    for (int j = 0; j + 1 < numRho && fits; ++j) {
      if (!(rising[j] && rising[j + 1])) continue;
      const double *rows[2] = {sie_h + (size_t)j * numT, sie_h + (size_t)(j + 1) * numT};
      const double span0 = (rows[0][numT - 1] - meta[j].base) + meta[j].s;
      const double span1 = (rows[1][numT - 1] - meta[j + 1].base) + meta[j + 1].s;
      const double umin = fastmath::log10(std::min(meta[j].s, meta[j + 1].s));
      const double umax = fastmath::log10(std::max(span0, span1));
      if (!(meta[j].s > 0.0) || !(meta[j + 1].s > 0.0) || !(umax > umin)) continue;
      const double inv_du = (double)kRootHintBins / (umax - umin);
      std::vector<int> lo(kRootHintBins, numT), hi(kRootHintBins, -1);
      auto cover = [&](int k0, int k1, int r) {
        for (int k = k0; k <= k1; ++k) {
          lo[k] = std::min(lo[k], r);
          hi[k] = std::max(hi[k], r);
        }
      };
      for (int q = 0; q < 2; ++q) {
        const double *row = rows[q];
        const RootRowMeta &m = meta[j + q];
        for (int r = 0; r + 1 < numT; ++r)
          cover(root_hint_bin(m.base, m.s, umin, inv_du, row[r]),
                root_hint_bin(m.base, m.s, umin, inv_du, row[r + 1]), r);
        // energies below the row's first node / above its last clamp to the end cells
        cover(0, root_hint_bin(m.base, m.s, umin, inv_du, row[0]), 0);
        cover(root_hint_bin(m.base, m.s, umin, inv_du, row[numT - 1]), kRootHintBins - 1,
              numT - 2);
      }
      unsigned int *hj = hint.data() + (size_t)j * kRootHintBins;
      for (int k = 0; k < kRootHintBins; ++k) {
        const int l = std::max(lo[k] - 1, 0), h = std::min(hi[k] + 1, numT - 2);
        hj[k] = (unsigned)l | ((unsigned)h << 16);
      }
      meta[j].umin = umin;
      meta[j].inv_du = inv_du;
    }
    static_assert(sizeof(RootRowMeta) == 4 * sizeof(double), "RootRowMeta is 4 doubles");
    static_assert(kRootHintBins % 2 == 0, "hint rows are stored as doubles");
    ts.add("sie#rootmeta", reinterpret_cast<const double *>(meta.data()), (size_t)numRho * 4);
    ts.add("sie#hint", reinterpret_cast<const double *>(hint.data()),
           (size_t)npair * kRootHintBins / 2);
  }
  int rc = upload_tables(e);
  if (rc) return rc;
  if (cudaMalloc(&e->root_stats, 2 * sizeof(uint64_t)) != cudaSuccess ||
      cudaMemset(e->root_stats, 0, 2 * sizeof(uint64_t)) != cudaSuccess)
    return fail(SGX_ERR_CUDA, "root statistics buffer: %s", cudaGetErrorString(cudaGetLastError()));
  s.P = ts.d("P#cells");
  s.sie = ts.d("sie#cells");
  s.bMod = ts.d("bMod#cells");
  s.dPdRho = ts.d("dPdRho#cells");
  s.dPdE = ts.d("dPdE#cells");
  s.dTdRho = ts.d("dTdRho#cells");
  s.dTdE = ts.d("dTdE#cells");
  s.dEdRho = ts.d("dEdRho#cells");
  s.dEdT = ts.d("dEdT#cells");
  s.PMax = ts.d("PMax");
  s.sielTMax = ts.d("sielTMax");
  s.dEdTMax = ts.d("dEdTMax");
  s.gm1Max = ts.d("gm1Max");
  s.lTColdCrit = ts.d("lTColdCrit");
  s.PCold = ts.d("PCold");
  s.sieCold = ts.d("sieCold");
  s.bModCold = ts.d("bModCold");
  s.dPdRhoCold = ts.d("dPdRhoCold");
  s.dPdECold = ts.d("dPdECold");
  s.dTdRhoCold = ts.d("dTdRhoCold");
  s.dTdECold = ts.d("dTdECold");
  s.dEdTCold = ts.d("dEdTCold");
  s.rho_at_pmin = ts.d("rho_at_pmin");
  SGX_CUDA_OK(cudaMalloc(&e->dev_struct, sizeof(SpinerRhoT)));
  SGX_CUDA_OK(cudaMemcpy(e->dev_struct, &s, sizeof(SpinerRhoT), cudaMemcpyHostToDevice));
  return SGX_OK;
}

int build_spiner_rhosie(const sgx_sp5_tables *raw, int reproducible, int pmin_vapor_dome,
                        sgx_eos_s *e) {
  if (!raw) return fail(SGX_ERR_INVALID, "null sgx_sp5_tables");
  if (!grid_ok(raw->lrho) || !grid_ok(raw->lT) || !grid_ok(raw->lE))
    return fail(SGX_ERR_INVALID, "malformed lrho/lT/lE grid");
  const double *req[] = {raw->sie,      raw->P,       raw->bMod,    raw->dPdRho,  raw->dPdE,
                         raw->dTdRho,   raw->dTdE,    raw->dEdRho,  raw->T_re,    raw->P_re,
                         raw->bMod_re,  raw->dPdRho_re, raw->dPdE_re, raw->dTdRho_re,
                         raw->dTdE_re,  raw->dEdRho_re, raw->PCold,  raw->sieCold,
                         raw->bModCold, raw->dPdRhoCold};
  for (const double *p : req)
    if (!p) return fail(SGX_ERR_INVALID, "SpinerEOSDependsRhoSie needs both table sets");

  SpinerRhoSie &s = e->rhosie;
  TableStore &ts = e->tables;
  s.gRho = make_grid(raw->lrho);
  s.gT = make_grid(raw->lT);
  s.gE = make_grid(raw->lE);
  const int numRho = s.gRho.n_total(), numT = s.gT.n_total(), numE = s.gE.n_total();
  s.numRho = numRho;
  s.numT = numT;
  s.numE = numE;
  const size_t nT2 = (size_t)numRho * numT, nE2 = (size_t)numRho * numE;
  s.lRhoOffset = std::abs(raw->rho_offset);
  s.lTOffset = std::abs(raw->T_offset);
  s.lEOffset = std::abs(raw->sie_offset);
  double rhoNormal = std::abs(raw->normal_density);
  s.reproducible = reproducible ? 1 : 0;

  ts.add("sie", raw->sie, nT2);
  ts.add("P", raw->P, nT2);
  ts.add("bMod", raw->bMod, nT2);
  ts.add("dPdRho", raw->dPdRho, nT2);
  ts.add("dPdE", raw->dPdE, nT2);
  ts.add("dTdRho", raw->dTdRho, nT2);
  ts.add("dTdE", raw->dTdE, nT2);
  ts.add("dEdRho", raw->dEdRho, nT2);
  ts.add("T_re", raw->T_re, nE2);
  ts.add("P_re", raw->P_re, nE2);
  ts.add("bMod_re", raw->bMod_re, nE2);
  ts.add("dPdRho_re", raw->dPdRho_re, nE2);
  ts.add("dPdE_re", raw->dPdE_re, nE2);
  ts.add("dTdRho_re", raw->dTdRho_re, nE2);
  ts.add("dTdE_re", raw->dTdE_re, nE2);
  ts.add("dEdRho_re", raw->dEdRho_re, nE2);
  ts.add("PCold", raw->PCold, numRho);
  ts.add("sieCold", raw->sieCold, numRho);
  ts.add("bModCold", raw->bModCold, numRho);
  ts.add("dPdRhoCold", raw->dPdRhoCold, numRho);
  ts.add("rho_at_pmin", nullptr, numT);
  ts.add("PlRhoMax", nullptr, numT);
  ts.add("dPdRhoMax", nullptr, numT);

  // calcBMod_ (:579-601) on both table sets
  auto calcBMod = [&](const char *sfx, int n1) {
    const std::string x(sfx);
    double *P = ts.h("P" + x), *bMod = ts.h("bMod" + x), *dPdRho = ts.h("dPdRho" + x),
           *dPdE = ts.h("dPdE" + x), *dEdRho = ts.h("dEdRho" + x);
    for (int j = 0; j < numRho; j++) {
      const double lRho = s.gRho.x(j);
      const double rho = from_log(lRho, s.lRhoOffset);
      for (int i = 0; i < n1; i++) {
        const size_t k = (size_t)j * n1 + i;
        const double press = P[k];
        const double DPDR_E = dPdRho[k];
        const double DPDE_R = dPdE[k];
        const double DEDR_T = dEdRho[k];
        const double DPDR_T = DPDR_E + DPDE_R * DEDR_T;
        double bm;
        if (DPDE_R > 0.0 && rho > 0.0) {
          bm = rho * DPDR_E + DPDE_R * (press / rho);
        } else if (rho > 0.0) {
          bm = std::max(rho * DPDR_T, 0.0);
        } else {
          bm = 0.0;
        }
        bMod[k] = std::max(bm, kEps);
      }
    }
  };
  calcBMod("", numT);
  calcBMod("_re", numE);

  s.lRhoMin = s.gRho.gmin();
  s.lRhoMax = s.gRho.gmax();
  s.rhoMax = from_log(s.lRhoMax, s.lRhoOffset);
  // slices at the maximum density (:529-530)
  {
    const double *P = ts.h("P"), *dPdRho = ts.h("dPdRho");
    double *a = ts.h("PlRhoMax"), *b = ts.h("dPdRhoMax");
    for (int i = 0; i < numT; ++i) {
      a[i] = P[(size_t)(numRho - 1) * numT + i];
      b[i] = dPdRho[(size_t)(numRho - 1) * numT + i];
    }
  }
  s.PMin = SetRhoPMin(ts.h("P"), s.gT, s.gRho, ts.h("rho_at_pmin"), pmin_vapor_dome != 0, 1e-8,
                      s.lRhoOffset);
  // reference state (:536-560)
  double lRhoNormal = to_log(rhoNormal, s.lRhoOffset);
  if (!(s.lRhoMin < lRhoNormal && lRhoNormal < s.lRhoMax)) {
    lRhoNormal = 0.5 * (s.lRhoMin + s.lRhoMax);
    rhoNormal = from_log(lRhoNormal, s.lRhoOffset);
  }
  double TNormal = ROOM_TEMPERATURE;
  double lTNormal = to_log(TNormal, s.lTOffset);
  const double lTMin = s.gT.gmin(), lTMax = s.gT.gmax();
  if (!(lTMin < lTNormal && lTNormal < lTMax)) {
    lTNormal = 0.5 * (lTMin + lTMax);
    TNormal = from_log(lTNormal, s.lTOffset);
  }
  auto I2 = [&](const char *nm) {
    return interp2d(ts.h(nm), s.gRho, s.gT, numT, lRhoNormal, lTNormal);
  };
  auto &sc = ts.scalars;
  sc["rhoNormal"] = rhoNormal;
  sc["TNormal"] = TNormal;
  sc["sieNormal"] = I2("sie");
  sc["PNormal"] = I2("P");
  sc["CvNormal"] = 1. / I2("dTdE");
  sc["bModNormal"] = I2("bMod");
  sc["dPdENormal"] = I2("dPdE");
  const double dPdR = I2("dPdRho");
  sc["dVdTNormal"] = sc["dPdENormal"] * sc["CvNormal"] / (rhoNormal * rhoNormal * dPdR);
  sc["lRhoMin"] = s.lRhoMin;
  sc["lRhoMax"] = s.lRhoMax;
  sc["rhoMax"] = s.rhoMax;
  sc["PMin"] = s.PMin;
  sc["lRhoOffset"] = s.lRhoOffset;
  sc["lTOffset"] = s.lTOffset;
  sc["lEOffset"] = s.lEOffset;
  sc["meanAtomicMass"] = raw->mean_atomic_mass;
  sc["meanAtomicNumber"] = raw->mean_atomic_number;

  static const char *k2dT[] = {"sie", "P", "bMod", "dPdRho", "dPdE", "dTdRho", "dTdE", "dEdRho"};
  static const char *k2dE[] = {"T_re", "P_re", "bMod_re", "dPdRho_re", "dPdE_re", "dTdRho_re",
                               "dTdE_re", "dEdRho_re"};
  for (const char *nm : k2dT) add_packed(ts, nm, numRho, numT);
  for (const char *nm : k2dE) add_packed(ts, nm, numRho, numE);
  int rc = upload_tables(e);
  if (rc) return rc;
  s.sie = ts.d("sie#cells");
  s.P = ts.d("P#cells");
  s.bMod = ts.d("bMod#cells");
  s.dPdRho = ts.d("dPdRho#cells");
  s.dPdE = ts.d("dPdE#cells");
  s.dTdRho = ts.d("dTdRho#cells");
  s.dTdE = ts.d("dTdE#cells");
  s.dEdRho = ts.d("dEdRho#cells");
  s.T_re = ts.d("T_re#cells");
  s.P_re = ts.d("P_re#cells");
  s.bMod_re = ts.d("bMod_re#cells");
  s.dPdRho_re = ts.d("dPdRho_re#cells");
  s.dPdE_re = ts.d("dPdE_re#cells");
  s.dTdRho_re = ts.d("dTdRho_re#cells");
  s.dTdE_re = ts.d("dTdE_re#cells");
  s.dEdRho_re = ts.d("dEdRho_re#cells");
  s.PCold = ts.d("PCold");
  s.sieCold = ts.d("sieCold");
  s.bModCold = ts.d("bModCold");
  s.dPdRhoCold = ts.d("dPdRhoCold");
  s.rho_at_pmin = ts.d("rho_at_pmin");
  s.PlRhoMax = ts.d("PlRhoMax");
  s.dPdRhoMax = ts.d("dPdRhoMax");
  SGX_CUDA_OK(cudaMalloc(&e->dev_struct, sizeof(SpinerRhoSie)));
  SGX_CUDA_OK(cudaMemcpy(e->dev_struct, &s, sizeof(SpinerRhoSie), cudaMemcpyHostToDevice));
  return SGX_OK;
}

} // namespace sgx
