// TEST INFRASTRUCTURE ONLY.  Builds the in-memory .sp5 tree (oracle/hdf5_shim) that the
// reference's SpinerEOSDependsRhoT / SpinerEOSDependsRhoSie constructors read, from the raw
// datasets of an sgx_sp5_tables.  Shared by oracle/ref_harness.cpp and tests/cpp/test_plugin.cpp.
// Layout: base/sp5/singularity_eos_sp5.hpp:20-72, utils/spiner/spiner/databox.hpp:853-906.
#ifndef SGX_ORACLE_SP5_TREE_HPP_
#define SGX_ORACLE_SP5_TREE_HPP_
#include <fstream>
#include <string>
#include <unistd.h>

#include <hdf5.h>
#include <hdf5_hl.h>
#include <singularity-eos/base/fast-math/logs.hpp>
#include <singularity-eos/base/sp5/singularity_eos_sp5.hpp>

#include "../include/sgx_b200.h"

namespace sgx_oracle {
using namespace singularity;
// ---- .sp5 tree builder ------------------------------------------------------------
inline void put_grid(h5shim::Node *grids, int axis_k /*1-based*/, const sgx_grid1d &g) {
  const std::string name = "grid_[" + std::to_string(axis_k) + "]";
  auto put_regular = [](h5shim::Node *parent, const std::string &nm, double mn, double mx,
                        double dx, int np) {
    const double range[3] = {mn, mx, dx};
    h5shim::set_dataset(parent, nm, range, 3);
    h5shim::set_attr_int(parent->child(nm, false), "npoints", {np});
  };
  if (g.ngrids == 1) {
    put_regular(grids, name, g.min[0], g.max[0], g.dx[0], g.npoints[0]);
  } else {
    h5shim::Node *grp = grids->child(name, true);
    h5shim::set_attr_int(grp, "ngrids", {g.ngrids});
    for (int p = 0; p < g.ngrids; ++p) {
      put_regular(grp, "patch_[" + std::to_string(p + 1) + "]", g.min[p], g.max[p], g.dx[p],
                  g.npoints[p]);
    }
  }
}
inline int grid_total(const sgx_grid1d &g) {
  int n = 0;
  for (int p = 0; p < g.ngrids; ++p)
    n += g.npoints[p];
  return n;
}
// rank-2 databox: dims (fastest first) = {n_fast, n_slow}
inline void put_db2(h5shim::Node *parent, const char *name, const double *data,
             const sgx_grid1d &fast, const sgx_grid1d &slow) {
  if (data == nullptr) return;
  h5shim::Node *db = parent->child(name, true);
  const int nf = grid_total(fast), ns = grid_total(slow);
  h5shim::set_attr_int(db, "rank", {2});
  h5shim::set_attr_int(db, "dims", {nf, ns});
  h5shim::set_attr_int(db, "index_types", {0, 0});
  h5shim::set_dataset(db, "data", data, static_cast<std::size_t>(nf) * ns);
  h5shim::Node *grids = db->child("grids", true);
  put_grid(grids, 1, fast);
  put_grid(grids, 2, slow);
}
inline void put_db1(h5shim::Node *parent, const char *name, const double *data,
             const sgx_grid1d &g) {
  if (data == nullptr) return;
  h5shim::Node *db = parent->child(name, true);
  const int n = grid_total(g);
  h5shim::set_attr_int(db, "rank", {1});
  h5shim::set_attr_int(db, "dims", {n});
  h5shim::set_attr_int(db, "index_types", {0});
  h5shim::set_dataset(db, "data", data, n);
  h5shim::Node *grids = db->child("grids", true);
  put_grid(grids, 1, g);
}

inline int g_file_counter = 0;

inline std::string build_sp5(const sgx_sp5_tables *raw, int matid) {
  // the constructors insist the path exists on disk (eos_spiner_common.hpp:128-134)
  std::string fname = "/tmp/sgx_oracle_" + std::to_string(::getpid()) + "_" +
                      std::to_string(g_file_counter++) + ".sp5";
  { std::ofstream touch(fname); }
  h5shim::drop_file(fname);
  h5shim::Node *root = h5shim::file_root(fname, true);
  h5shim::set_attr_int(root, SP5::logType, {static_cast<int>(FastMath::Settings::log_type)});
  h5shim::Node *mat = root->child(std::to_string(matid), true);
  h5shim::set_attr_double(mat, SP5::Offsets::rho, raw->rho_offset);
  h5shim::set_attr_double(mat, SP5::Offsets::T, raw->T_offset);
  h5shim::set_attr_double(mat, SP5::Offsets::sie, raw->sie_offset);
  h5shim::set_attr_double(mat, SP5::Material::normalDensity, raw->normal_density);
  h5shim::set_attr_double(mat, SP5::Material::meanAtomicMass, raw->mean_atomic_mass);
  h5shim::set_attr_double(mat, SP5::Material::meanAtomicNumber, raw->mean_atomic_number);
  h5shim::set_attr_int(mat, SP5::Material::matid, {matid});

  h5shim::Node *lT = mat->child(SP5::Depends::logRhoLogT, true);
  put_db2(lT, SP5::Fields::P, raw->P, raw->lT, raw->lrho);
  put_db2(lT, SP5::Fields::sie, raw->sie, raw->lT, raw->lrho);
  put_db2(lT, SP5::Fields::bMod, raw->bMod, raw->lT, raw->lrho);
  put_db2(lT, SP5::Fields::dPdRho, raw->dPdRho, raw->lT, raw->lrho);
  put_db2(lT, SP5::Fields::dPdE, raw->dPdE, raw->lT, raw->lrho);
  put_db2(lT, SP5::Fields::dTdRho, raw->dTdRho, raw->lT, raw->lrho);
  put_db2(lT, SP5::Fields::dTdE, raw->dTdE, raw->lT, raw->lrho);
  put_db2(lT, SP5::Fields::dEdRho, raw->dEdRho, raw->lT, raw->lrho);
  put_db2(lT, SP5::Fields::dEdT, raw->dEdT, raw->lT, raw->lrho);

  h5shim::Node *lE = mat->child(SP5::Depends::logRhoLogSie, true);
  put_db2(lE, SP5::Fields::T, raw->T_re, raw->lE, raw->lrho);
  put_db2(lE, SP5::Fields::P, raw->P_re, raw->lE, raw->lrho);
  put_db2(lE, SP5::Fields::bMod, raw->bMod_re, raw->lE, raw->lrho);
  put_db2(lE, SP5::Fields::dPdRho, raw->dPdRho_re, raw->lE, raw->lrho);
  put_db2(lE, SP5::Fields::dPdE, raw->dPdE_re, raw->lE, raw->lrho);
  put_db2(lE, SP5::Fields::dTdRho, raw->dTdRho_re, raw->lE, raw->lrho);
  put_db2(lE, SP5::Fields::dTdE, raw->dTdE_re, raw->lE, raw->lrho);
  put_db2(lE, SP5::Fields::dEdRho, raw->dEdRho_re, raw->lE, raw->lrho);

  h5shim::Node *cold = mat->child(SP5::Depends::coldCurve, true);
  put_db1(cold, SP5::Fields::P, raw->PCold, raw->lrho);
  put_db1(cold, SP5::Fields::sie, raw->sieCold, raw->lrho);
  put_db1(cold, SP5::Fields::bMod, raw->bModCold, raw->lrho);
  put_db1(cold, SP5::Fields::dPdRho, raw->dPdRhoCold, raw->lrho);
  return fname;
}
} // namespace sgx_oracle
#endif // SGX_ORACLE_SP5_TREE_HPP_
