// Factories for the plugin types of sgx_device_eos.hpp: build the reference's host-side EOS and
// the matching libsgx_b200 handle from the same parameters / tables.
#ifndef SGX_PLUGIN_MAKE_HPP_
#define SGX_PLUGIN_MAKE_HPP_

#include <string>

#include <singularity-eos/eos/eos_davis.hpp>
#include <singularity-eos/eos/eos_gruneisen.hpp>
#include <singularity-eos/eos/eos_ideal.hpp>
#ifdef SINGULARITY_USE_SPINER_WITH_HDF5
#include <singularity-eos/eos/eos_spiner_rho_sie.hpp>
#include <singularity-eos/eos/eos_spiner_rho_temp.hpp>
#endif

#include "sgx_device_eos.hpp"

namespace singularity {

using SgxIdealGas = SgxDevice<IdealGas>;
using SgxGruneisen = SgxDevice<Gruneisen>;
using SgxDavisReactants = SgxDevice<DavisReactants>;
using SgxDavisProducts = SgxDevice<DavisProducts>;

inline SgxIdealGas MakeSgxIdealGas(const Real gm1, const Real Cv) {
  sgx_eos_t h = nullptr;
  sgx_plugin::check(sgx_eos_create_ideal(gm1, Cv, &h));
  return SgxIdealGas(IdealGas(gm1, Cv), h);
}
inline SgxGruneisen MakeSgxGruneisen(const Real C0, const Real s1, const Real s2, const Real s3,
                                     const Real G0, const Real b, const Real rho0, const Real T0,
                                     const Real P0, const Real Cv) {
  sgx_eos_t h = nullptr;
  sgx_plugin::check(sgx_eos_create_gruneisen(C0, s1, s2, s3, G0, b, rho0, T0, P0, Cv, -1.0, &h));
  return SgxGruneisen(Gruneisen(C0, s1, s2, s3, G0, b, rho0, T0, P0, Cv), h);
}
inline SgxDavisReactants MakeSgxDavisReactants(const Real rho0, const Real e0, const Real P0,
                                               const Real T0, const Real A, const Real B,
                                               const Real C, const Real G0, const Real Z,
                                               const Real alpha, const Real Cv0) {
  sgx_eos_t h = nullptr;
  sgx_plugin::check(
      sgx_eos_create_davis_reactants(rho0, e0, P0, T0, A, B, C, G0, Z, alpha, Cv0, &h));
  return SgxDavisReactants(DavisReactants(rho0, e0, P0, T0, A, B, C, G0, Z, alpha, Cv0), h);
}
inline SgxDavisProducts MakeSgxDavisProducts(const Real a, const Real b, const Real k,
                                             const Real n, const Real vc, const Real pc,
                                             const Real Cv) {
  sgx_eos_t h = nullptr;
  sgx_plugin::check(sgx_eos_create_davis_products(a, b, k, n, vc, pc, Cv, &h));
  return SgxDavisProducts(DavisProducts(a, b, k, n, vc, pc, Cv), h);
}

#ifdef SINGULARITY_USE_SPINER_WITH_HDF5
using SgxSpinerRhoT = SgxDevice<SpinerEOSDependsRhoT>;
using SgxSpinerRhoSie = SgxDevice<SpinerEOSDependsRhoSie>;
// `raw` holds the .sp5 datasets of (filename, matid) exactly as stored in the file -- what
// SpinerEOSDependsRhoT::loadDataboxes_ reads (eos_spiner_rho_temp.hpp:545-583) -- read by the
// host code with H5LTread_dataset.  Both sides then perform the reference's load-time fix-ups
// themselves (bit-identical derived tables: tests/test_gpu_parity.py::test_spiner_load_time_tables).
inline SgxSpinerRhoT MakeSgxSpinerRhoT(const std::string &filename, const int matid,
                                       const sgx_sp5_tables &raw,
                                       const bool reproducibility_mode = false,
                                       const bool pmin_vapor_dome = false) {
  sgx_eos_t h = nullptr;
  sgx_plugin::check(sgx_eos_create_spiner_rhot(&raw, reproducibility_mode, pmin_vapor_dome, &h));
  return SgxSpinerRhoT(
      SpinerEOSDependsRhoT(filename, matid, reproducibility_mode, pmin_vapor_dome), h);
}
inline SgxSpinerRhoSie MakeSgxSpinerRhoSie(const std::string &filename, const int matid,
                                           const sgx_sp5_tables &raw,
                                           const bool reproducibility_mode = false,
                                           const bool pmin_vapor_dome = false) {
  sgx_eos_t h = nullptr;
  sgx_plugin::check(
      sgx_eos_create_spiner_rhosie(&raw, reproducibility_mode, pmin_vapor_dome, &h));
  return SgxSpinerRhoSie(
      SpinerEOSDependsRhoSie(filename, matid, reproducibility_mode, pmin_vapor_dome), h);
}
#endif // SINGULARITY_USE_SPINER_WITH_HDF5

} // namespace singularity
#endif // SGX_PLUGIN_MAKE_HPP_
